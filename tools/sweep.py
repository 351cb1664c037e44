#!/usr/bin/env python
"""Bit-width / encoding sweep of the decode kernels (BASELINE config 4 shapes at a bounded size): per (type, encoding,
width) the algorithmic GB/s (SURVEY.md 8d bytes / CUDA-event time) and its fraction of the measured HBM peak.
Usage: python tools/sweep.py [--values 33554432] [--out profiles/rXX_sweep.json]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests"), ROOT):
    sys.path.insert(0, p)
import flsgpu  # noqa: E402
from bench import algorithmic_bytes  # noqa: E402

WIDTHS = [0, 1, 2, 3, 4, 6, 8, 10, 12, 16, 20, 24, 32, 40, 48, 56, 63, 64]


def float_column(kind, tname, w, n, rng):
    """ALP shapes: ALP0 = digits of w bits, no exceptions; ALPX = i/1000.0 (the bench's DOUBLE column, ~8 exceptions per
    vector); ALPRD = full-entropy mantissas."""
    fdt = np.float64 if tname == "DOUBLE" else np.float32
    base_n = 1 << 20
    if kind == "ALP0":
        d = rng.integers(0, (1 << w) - 1, base_n, endpoint=True).astype(np.float64)
        x = (d / 100.0).astype(fdt) if tname == "DOUBLE" else d.astype(fdt)
        if tname == "DOUBLE":  # keep only values ALP round-trips exactly at (e=2,f=0): make them so by construction
            x = (np.round(x * 100.0) / 100.0).astype(fdt)
    elif kind == "ALPX":
        i = np.arange(1, base_n + 1, dtype=np.int64) + (1 << w)
        x = (i / 1000.0).astype(fdt) if tname == "DOUBLE" else (i.astype(np.float32) / np.float32(100))
    else:
        x = (rng.random(base_n) * 2 - 1).astype(fdt)
    return np.tile(x, n // base_n)


def column(kind, tname, w, n, rng):
    if kind.startswith("ALP"):
        return float_column(kind, tname, w, n, rng)
    dt = np.dtype({"INT64": np.int64, "INT32": np.int32, "INT16": np.int16, "INT8": np.int8}[tname])
    bits = dt.itemsize * 8
    udt = np.dtype(f"u{dt.itemsize}")
    base_n = 1 << 20
    raw = rng.integers(0, (1 << w) - 1, base_n, dtype=np.uint64, endpoint=True) if w else np.zeros(base_n, dtype=np.uint64)
    raw[::1024] = (1 << w) - 1 if w else 0  # every vector really needs w bits
    if kind == "FFOR":
        x = raw.astype(udt)
    elif kind == "DELTA":
        x = np.cumsum(raw.astype(udt), dtype=udt)
    elif kind == "DICT":  # w = log2(cardinality)
        keys = rng.integers(0, 2 ** (bits - 2), 1 << w, dtype=np.uint64).astype(udt)
        x = keys[raw.astype(np.int64)]
    elif kind == "RLE":  # w = log2(mean run length)
        n_runs = 2 * base_n // (1 << w) + 2
        runs = rng.integers(0, 1000, n_runs, dtype=np.uint64).astype(udt)
        x = np.repeat(runs, rng.integers(1, 2 << w, n_runs))[:base_n]  # mean run length 2^w, unaligned
    else:
        raise ValueError(kind)
    return np.tile(x, n // base_n).view(dt)


def measure(t, img, col, n_iter=10):
    import torch
    w = t.column(col).phys_bits // 8
    out = torch.empty(t.out_elems(t.n_rowgroups) * w, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        t.decode_column(col, out.data_ptr(), out.numel(), stream=stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_iter):
        t.decode_column(col, out.data_ptr(), out.numel(), stream=stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n_iter
    rd, wr = algorithmic_bytes(t, img, col)
    return ms, rd, wr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--values", type=int, default=1 << 25)
    ap.add_argument("--out", default=None)
    ap.add_argument("--kinds", default="FFOR,DELTA,DICT,RLE")
    ap.add_argument("--types", default="INT64,INT32,INT16,INT8")
    ap.add_argument("--widths", default=None, help="comma list overriding the per-kind width list")
    args = ap.parse_args()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    flsgpu.init((0,))
    rng = np.random.default_rng(42)
    rows = []
    for tname in args.types.split(","):
        bits = {"INT64": 64, "INT32": 32, "INT16": 16, "INT8": 8, "DOUBLE": 64, "FLOAT": 32}[tname]
        for kind in args.kinds.split(","):
            ws = [w for w in WIDTHS if w <= bits]
            if kind == "DICT":
                ws = [1, 4, 8, 12, 16]
            if kind == "RLE":
                ws = [1, 3, 6, 10]
            if kind.startswith("ALP"):
                ws = {"ALP0": [4, 10, 20], "ALPX": [10, 24], "ALPRD": [0]}[kind]
            if kind.startswith("ALP") != (tname in ("DOUBLE", "FLOAT")):
                continue
            if args.widths:
                ws = [int(v) for v in args.widths.split(",")]
            for w in ws:
                if kind == "DICT" and w >= bits:
                    continue
                try:
                    x = column(kind, tname, w, args.values, rng)
                    force = {"ALP0": "ALP", "ALPX": "ALP", "ALPRD": "ALP_RD"}.get(kind, kind)
                    img = flsgpu.write_table([("c", tname, force)], [x])
                    del x
                    t = flsgpu.Table(img)
                    t.stage()
                    ms, rd, wr = measure(t, img, 0)
                    t.close()
                    del t, img
                except Exception as e:  # a shape the writer refuses (or that does not fit): report and carry on
                    print(dict(type=tname, enc=kind, w=w, error=str(e)[:200]), flush=True)
                    continue
                gbps = (rd + wr) / ms / 1e6
                rows.append(dict(type=tname, enc=kind, w=w, ms=round(ms, 4), alg_read_MB=round(rd / 1e6, 1), alg_write_MB=round(wr / 1e6, 1),
                                 alg_GBps=round(gbps, 1), frac=round(gbps / peak, 3), decoded_GBps=round(wr / ms / 1e6, 1)))
                print(rows[-1], flush=True)
    if args.out:
        json.dump(dict(peak_GBps=peak, values=args.values, rows=rows), open(args.out, "w"), indent=0)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""String decode throughput on the device (resident chunks): flat FSST (TPC-H l_comment-like text), flat dictionary,
dictionary form, uncompressed.  Prints one JSON line per case: rows/s, decoded string bytes/s, algorithmic GB/s
(compressed bytes + offsets read, string_t + heap written).

  python tools/str_bench.py [--rows 20000000] [--iters 10]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests")]

WORDS = [b"furiously", b"regular", b"deposits", b"sleep", b"quickly", b"above", b"the", b"slyly", b"ironic", b"packages",
         b"carefully", b"final", b"accounts", b"boost", b"blithely", b"pending", b"requests", b"haggle", b"express", b"theodolites",
         b"across", b"foxes", b"instructions", b"dependencies", b"platelets", b"nag", b"cajole", b"even", b"bold", b"unusual"]


def comment_column(n, seed=1):
    """(offsets, bytes): 2..8 random words per row, built without Python string objects."""
    rng = np.random.default_rng(seed)
    wl = np.array([len(w) + 1 for w in WORDS], dtype=np.int64)  # word + trailing space
    mat = np.zeros((len(WORDS), wl.max()), dtype=np.uint8)
    for i, w in enumerate(WORDS):
        mat[i, :len(w) + 1] = np.frombuffer(w + b" ", dtype=np.uint8)
    k = rng.integers(2, 9, n)
    total = int(k.sum())
    widx = rng.integers(0, len(WORDS), total)
    lens = wl[widx]
    data = mat[widx][np.arange(mat.shape[1])[None, :] < lens[:, None]]
    row_len = np.add.reduceat(lens, np.concatenate(([0], np.cumsum(k)[:-1])))
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(row_len, out=off[1:])
    return off, np.ascontiguousarray(data)


def dict_column(n, keys):
    kl = np.array([len(k) for k in keys], dtype=np.int64)
    mat = np.zeros((len(keys), kl.max()), dtype=np.uint8)
    for i, k in enumerate(keys):
        mat[i, :len(k)] = np.frombuffer(k, dtype=np.uint8)
    idx = (np.arange(n) * 7919) % len(keys)
    lens = kl[idx]
    data = mat[idx][np.arange(mat.shape[1])[None, :] < lens[:, None]]
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    return off, np.ascontiguousarray(data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=20_000_000)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--cases", default="fsst,plain,dict_flat,dict_form,fsst_dict_flat")
    args = ap.parse_args()
    import torch
    import flsgpu
    flsgpu.init((0,))
    n = args.rows
    comment = comment_column(n)
    modes = dict_column(n, [b"DELIVER IN PERSON", b"COLLECT COD", b"NONE", b"TAKE BACK RETURN"])
    values = dict_column(n, [b"Value %d" % k for k in range(1000)])
    cases = {
        "fsst": ("VARCHAR", "FSST", comment, "flat"),
        "fsst12": ("VARCHAR", "FSST12", comment, "flat"),
        "plain": ("VARCHAR", "STR_UNCOMPRESSED", comment, "flat"),
        "dict_flat": ("VARCHAR", "DICT_STR", values, "flat"),
        "dict_form": ("VARCHAR", "DICT_STR", values, "dict"),
        "fsst_dict_flat": ("VARCHAR", "FSST_DICT", modes, "flat"),
    }
    stream = torch.cuda.current_stream().cuda_stream
    for name in args.cases.split(","):
        tname, enc, col, form = cases[name]
        w = flsgpu.Writer([("s", tname, enc)], 64)
        step = 64 * 65536
        off, data = col
        for a in range(0, n, step):
            b = min(n, a + step)
            w.append([(off[a:b + 1] - off[a], data[int(off[a]):int(off[b])])])
        img = w.finish()
        t = flsgpu.Table(img)
        t.stage(stream=stream)
        rows = t.out_elems(t.n_rowgroups)
        hb = t.string_heap_bytes(0)
        heap = torch.empty(max(16, hb), dtype=torch.uint8, device="cuda")
        comp = sum(t.chunk(r, 0).size for r in range(t.n_rowgroups))
        if form == "flat":
            out = torch.empty(rows * 16, dtype=torch.uint8, device="cuda")
            run = lambda: t.decode_strings(0, out.data_ptr(), out.numel(), heap.data_ptr(), heap.numel(), heap.data_ptr(), stream=stream)
            written = rows * 16 + hb
        else:
            out = torch.empty(rows * 4, dtype=torch.uint8, device="cuda")
            keys = torch.empty(max(16, t.string_dict_keys(0) * 16), dtype=torch.uint8, device="cuda")

            def run():
                t.decode_string_dict(0, keys.data_ptr(), keys.numel(), heap.data_ptr(), heap.numel(), heap.data_ptr(), stream=stream)
                t.decode_string_codes(0, out.data_ptr(), out.numel(), stream=stream)
            written = rows * 4 + keys.numel() + hb
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        raw = int(off[-1])
        # pushed-down string filter (flsgpu_filter_mask on the column): WHERE s >= / = constant -> selection mask
        const, op = ((b"furiously regular", "GE") if col is comment else (bytes(data[int(off[1]):int(off[2])]), "EQ"))
        n_words = t.n_rowgroups * t.vec_per_rowgroup * 32
        mask = torch.empty(n_words, dtype=torch.int32, device="cuda")
        filt = [(0, op, const)]
        for _ in range(3):
            t.filter_mask(filt, mask.data_ptr(), n_words * 4, stream=stream)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.iters):
            t.filter_mask(filt, mask.data_ptr(), n_words * 4, stream=stream)
        e1.record()
        torch.cuda.synchronize()
        fms = e0.elapsed_time(e1) / args.iters
        selected = int(np.unpackbits(mask.cpu().numpy().view(np.uint8)).sum())
        print(json.dumps({"case": name, "encoding": enc, "form": form, "rows": n, "ms": round(ms, 4), "rows_per_s": round(n / ms * 1e3, 1),
                          "raw_string_MB": round(raw / 1e6, 1), "compressed_MB": round(comp / 1e6, 1),
                          "string_GBps": round(raw / ms / 1e6, 1), "alg_GBps": round((comp + written) / ms / 1e6, 1),
                          "filter": {"op": op, "constant": const.decode(), "ms": round(fms, 4), "rows_per_s": round(n / fms * 1e3, 1),
                                     "selected": selected}}), flush=True)
        t.close()
        w.close()
        del out, heap


if __name__ == "__main__":
    main()

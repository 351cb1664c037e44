#!/usr/bin/env python
"""BASELINE configs 3/5 on one GPU: a TPC-H lineitem-shaped table (tests/lineitem_cases.py: dbgen's distributions restated,
dbgen is not in the image) written with the writer's own encoding selection, then, resident in HBM:
  * every column scanned on its own (numeric: flsgpu_decode_column; strings: dictionary form where the column is dictionary
    coded, flat string_t + heap otherwise) — ms, decoded GB/s, algorithmic GB/s;
  * SELECT SUM(c), COUNT(*) over every integer / decimal column (fused decode + aggregate);
  * TPC-H Q1's scan (l_shipdate <= date -> SUMs of quantity, extendedprice, discount, tax, COUNT) and Q6's scan
    (three range predicates -> SUM(extendedprice), SUM(discount)) through filter masks and masked fused aggregates;
all checked against numpy sums kept while generating; plus the oracle's reference-shaped scan on the host cores.
One JSON document on stdout.   python tools/lineitem_bench.py [--sf 10] [--iters 10] [--out profiles/rXX_lineitem.json]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests"), ROOT]

from filter_cases import bits  # noqa: E402
from lineitem_cases import ROWS_PER_SF, lineitem, specs  # noqa: E402

AGG_COLS = ["l_orderkey", "l_partkey", "l_suppkey", "l_linenumber", "l_quantity", "l_extendedprice", "l_discount", "l_tax"]
Q_COLS = ["l_quantity", "l_extendedprice", "l_discount", "l_tax"]


def timed(run, iters, torch):
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=1.0)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    import flsgpu
    import oracle_lib
    flsgpu.init((0,))
    n_rows = int(ROWS_PER_SF * args.sf)
    threads = os.cpu_count() or 1

    # ---- generate + encode batch by batch; keep the numpy answers of the queries
    t0 = time.perf_counter()
    w, names, want = None, None, {"sum": {}, "q1": {}, "q6": {}, "q1_count": 0, "q6_count": 0}
    batch, done, first_order, raw_bytes = 64 * 65536, 0, 0, 0
    while done < n_rows:
        n = min(batch, n_rows - done)
        cols = lineitem(n, sf=args.sf, seed=1000 + done // batch, first_order=first_order)
        last_key = int(cols[0][3][-1]) - 1  # next batch goes on with the next order
        first_order = (last_key // 32) * 8 + last_key % 8 + 1
        if w is None:
            names = [c[0] for c in cols]
            w = flsgpu.Writer(specs(cols), 64)
        by = {c[0]: c[3] for c in cols}
        q1 = by["l_shipdate"] <= 10471
        q6 = ((by["l_shipdate"] >= 8766) & (by["l_shipdate"] < 9131) & (by["l_discount"] >= 5) & (by["l_discount"] <= 7) &
              (by["l_quantity"] < 2400))
        for c in AGG_COLS:
            want["sum"][c] = want["sum"].get(c, 0) + int(by[c].sum(dtype=np.int64))
        for c in Q_COLS:
            want["q1"][c] = want["q1"].get(c, 0) + int(by[c][q1].sum())
            want["q6"][c] = want["q6"].get(c, 0) + int(by[c][q6].sum())
        want["q1_count"] += int(q1.sum())
        want["q6_count"] += int(q6.sum())
        raw_bytes += sum(int(c[3][0][-1]) + 4 * n if c[1] == "VARCHAR" else np.asarray(c[3]).nbytes for c in cols)
        w.append([c[3] for c in cols], threads)
        done += n
    img = w.finish()
    build_s = time.perf_counter() - t0
    doc = measure(img, want, n_rows, args.iters, args.cpu_seconds)
    doc = dict({"workload": "TPC-H lineitem shape, SF%g (%d rows, %d row groups), writer's own encoding selection" % (args.sf, n_rows, doc["row_groups"]),
                "raw_MB": round(raw_bytes / 1e6, 1), "build_s": round(build_s, 1)}, **doc)
    print(json.dumps(doc, indent=1))
    if args.out:
        with open(args.out, "w") as f:
            json.dump(doc, f, indent=1)
    w.close()


def measure(img, want, n_rows, iters, cpu_seconds):
    """The resident measurements over a lineitem container image (numpy uint8): per-column scans, the whole scan as one
    region, fused SUM/COUNT per numeric column, Q1 / Q6 filter scans - every result checked against `want` - and the
    oracle's scan on the host cores.  Used with the numpy generator (main) and with dbgen's table (tools/lineitem_dbgen.py)."""
    import torch
    import flsgpu
    import oracle_lib

    class _A:
        pass
    args = _A()
    args.iters, args.cpu_seconds = iters, cpu_seconds
    flsgpu.init((0,))
    threads = os.cpu_count() or 1
    t = flsgpu.Table(img)
    names = [t.column(c).name.decode() for c in range(t.n_columns)]
    stream = torch.cuda.current_stream().cuda_stream
    t.stage(stream=stream)
    torch.cuda.synchronize()
    n_rg, rows_out = t.n_rowgroups, t.out_elems(t.n_rowgroups)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]

    # ---- per-column scans
    per_column, scan_ms, scan_bytes = [], 0.0, 0
    runs = []
    for c, name in enumerate(names):
        encs = sorted({flsgpu.ENC_NAME[t.chunk(r, c).encoding] for r in range(n_rg)})
        comp = sum(t.chunk(r, c).size for r in range(n_rg))
        if t.column(c).phys_bits == 128:
            hb = t.string_heap_bytes(c)
            heap = torch.empty(max(16, hb), dtype=torch.uint8, device="cuda")
            if all(e in ("DICT_STR", "FSST_DICT", "FSST12_DICT", "STR_CONSTANT") for e in encs):
                form = "dictionary"
                out = torch.empty(rows_out * 4, dtype=torch.uint8, device="cuda")
                keys = torch.empty(max(16, t.string_dict_keys(c) * 16), dtype=torch.uint8, device="cuda")

                def run(c=c, out=out, keys=keys, heap=heap):
                    t.decode_string_dict(c, keys.data_ptr(), keys.numel(), heap.data_ptr(), heap.numel(), heap.data_ptr(), stream=stream)
                    t.decode_string_codes(c, out.data_ptr(), out.numel(), stream=stream)
                written = t.n_tuples * 4 + t.string_dict_keys(c) * 16 + hb
            else:
                form = "flat"
                out = torch.empty(rows_out * 16, dtype=torch.uint8, device="cuda")

                def run(c=c, out=out, heap=heap):
                    t.decode_strings(c, out.data_ptr(), out.numel(), heap.data_ptr(), heap.numel(), heap.data_ptr(), stream=stream)
                written = t.n_tuples * 16 + hb
        else:
            form = "numeric"
            wd = t.column(c).phys_bits // 8
            out = torch.empty(rows_out * wd, dtype=torch.uint8, device="cuda")

            def run(c=c, out=out):
                t.decode_column(c, out.data_ptr(), out.numel(), stream=stream)
            written = t.n_tuples * wd
        ms = timed(run, args.iters, torch)
        runs.append(run)
        scan_ms += ms
        scan_bytes += written
        per_column.append({"column": name, "encodings": encs, "form": form, "compressed_MB": round(comp / 1e6, 2),
                           "decoded_MB": round(written / 1e6, 2), "ms": round(ms, 4), "decoded_GBps": round(written / ms / 1e6, 1),
                           "container_plus_decoded_GBps": round((comp + written) / ms / 1e6, 1),
                           "frac_of_hbm_peak": round((comp + written) / ms / 1e6 / peak, 3)})
        del out

    # ---- the whole 16-column scan as the reader issues it: every column back to back on one stream, timed as one region
    # (consecutive decode launches are chained by programmatic dependent launch)
    chained_ms = timed(lambda: [r() for r in runs], args.iters, torch)

    # ---- SUM / COUNT scans (fused decode + aggregate), unfiltered
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")

    def agg_of(buf):
        r = flsgpu.AggResult.from_buffer_copy(buf.cpu().numpy().tobytes())
        s = (r.sum_hi << 64) | r.sum_lo
        return (s - (1 << 128) if s >= (1 << 127) else s), r.count
    sums = []
    for name in AGG_COLS:
        c = names.index(name)

        def run(c=c):
            flsgpu.agg_reset(res.data_ptr(), stream)
            t.aggregate_column(c, res.data_ptr(), stream=stream)
        ms = timed(run, args.iters, torch)
        s, cnt = agg_of(res)
        assert s == want["sum"][name] and cnt == n_rows, (name, s, want["sum"][name], cnt)
        sums.append({"column": name, "ms": round(ms, 4), "values_per_s": round(n_rows / ms * 1e3, 1), "sum": str(s)})

    # ---- Q1 / Q6 scans: filter mask + masked aggregates
    sd, di, qu = names.index("l_shipdate"), names.index("l_discount"), names.index("l_quantity")
    f_q1 = [(sd, flsgpu.CMP["LE"], bits(np.int32(10471), np.int32))]
    f_q6 = [(sd, flsgpu.CMP["GE"], bits(np.int32(8766), np.int32)), (sd, flsgpu.CMP["LT"], bits(np.int32(9131), np.int32)),
            (di, flsgpu.CMP["GE"], bits(np.int64(5), np.int64)), (di, flsgpu.CMP["LE"], bits(np.int64(7), np.int64)),
            (qu, flsgpu.CMP["LT"], bits(np.int64(2400), np.int64))]
    n_words = n_rg * t.vec_per_rowgroup * 32
    mask = torch.zeros(n_words, dtype=torch.int32, device="cuda")
    res4 = torch.zeros(64 * 4, dtype=torch.uint8, device="cuda")
    queries = []
    for qname, filt, agg_cols in (("q1", f_q1, Q_COLS), ("q6", f_q6, ["l_extendedprice", "l_discount"])):
        def run(filt=filt, agg_cols=agg_cols):
            t.filter_mask(filt, mask.data_ptr(), n_words * 4, 0, stream=stream)
            for i, name in enumerate(agg_cols):
                flsgpu.agg_reset(res4.data_ptr() + 64 * i, stream)
                t.aggregate_column_masked(names.index(name), mask.data_ptr(), res4.data_ptr() + 64 * i, 0, stream=stream)
        ms = timed(run, args.iters, torch)
        got = {}
        for i, name in enumerate(agg_cols):
            s, cnt = agg_of(res4[64 * i:64 * i + 64])
            assert s == want[qname][name] and cnt == want[qname + "_count"], (qname, name, s, want[qname][name], cnt)
            got[name] = str(s)
        queries.append({"query": qname, "ms": round(ms, 4), "rows_per_s": round(n_rows / ms * 1e3, 1), "count": want[qname + "_count"],
                        "sums": got})

    # ---- host cores: the oracle's reference-shaped scan of all 16 columns (one row group per thread)
    o = oracle_lib.Oracle(img)
    cols_all = list(range(len(names)))
    sample_rg = n_rg
    t0 = time.perf_counter()
    passes = 0
    while True:
        o.scan(cols_all, 0, sample_rg, threads)
        passes += 1
        if time.perf_counter() - t0 >= args.cpu_seconds:
            break
    cpu_s = (time.perf_counter() - t0) / passes
    doc = {"rows": n_rows, "row_groups": n_rg, "container_MB": round(len(img) / 1e6, 1),
           "hbm_peak_GBps": peak, "per_column": per_column,
           "full_scan": {"ms": round(scan_ms, 4), "decoded_GBps": round(scan_bytes / scan_ms / 1e6, 1),
                         "rows_per_s": round(n_rows / scan_ms * 1e3, 1), "note": "sum of the per-column times above"},
           "full_scan_one_region": {"ms": round(chained_ms, 4), "decoded_GBps": round(scan_bytes / chained_ms / 1e6, 1),
                                    "rows_per_s": round(n_rows / chained_ms * 1e3, 1),
                                    "note": "all 16 columns back to back on one stream, one pair of CUDA events"},
           "sum_count_scans": sums, "queries": queries,
           "cpu_oracle_scan": {"threads": threads, "s_per_pass": round(cpu_s, 3), "decoded_GBps": round(scan_bytes / cpu_s / 1e9, 2),
                               "passes": passes}}
    t.close()
    return doc


if __name__ == "__main__":
    main()

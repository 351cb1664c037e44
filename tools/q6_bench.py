#!/usr/bin/env python
"""BASELINE config 3 shapes on one GPU: a lineitem-like table (TPC-H column generators, SF10 row count by default) resident
in HBM; (a) SELECT SUM(l_quantity), SUM(l_extendedprice), SUM(l_discount), SUM(l_tax), SUM(l_orderkey), COUNT(*) with the
fused decode+aggregate kernels, (b) the Q6 predicate pushed down: zone-map pruning + device selection mask + masked
SUM(l_extendedprice), COUNT(*).  Results are checked against numpy; the oracle (CPU, 1 thread, bounded sample) is timed
beside them.  One JSON line.

  python tools/q6_bench.py [--rows 59986052] [--iters 10] [--sorted-dates]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests")]


def lineitem(n, first, rng, sorted_dates):
    """dbgen-shaped columns (duckdb/extension/tpch/dbgen): physical integers (DECIMAL(15,2) unscaled, DATE days)."""
    orderkey = (first + np.arange(n, dtype=np.int64)) // 4 * 8 // 7 + 1   # sparse, non-decreasing
    quantity = rng.integers(1, 51, n).astype(np.int64) * 100
    price = rng.integers(90_000, 10_495_000, n).astype(np.int64)
    discount = rng.integers(0, 11, n).astype(np.int64)
    tax = rng.integers(0, 9, n).astype(np.int64)
    ship = rng.integers(8036, 10562, n).astype(np.int32)                   # 1992-01-02 .. 1998-12-01
    if sorted_dates:
        ship = np.sort(ship)
    return [orderkey, quantity, price, discount, tax, ship]


COLS = [("l_orderkey", "INT64", None), ("l_quantity", "DECIMAL", None, 15, 2), ("l_extendedprice", "DECIMAL", None, 15, 2),
        ("l_discount", "DECIMAL", None, 15, 2), ("l_tax", "DECIMAL", None, 15, 2), ("l_shipdate", "DATE", None)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=59_986_052)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--sorted-dates", action="store_true", help="cluster l_shipdate (zone maps then prune most row groups)")
    args = ap.parse_args()
    import torch
    import flsgpu
    import oracle_lib
    flsgpu.init((0,))
    rng = np.random.default_rng(42)
    w = flsgpu.Writer(COLS, 64)
    sums = np.zeros(5, dtype=object)
    q6_sum, q6_cnt = 0, 0
    lo, hi = 8766, 9131  # 1994-01-01 <= l_shipdate < 1995-01-01
    step = 64 * 65536
    for a in range(0, args.rows, step):
        n = min(step, args.rows - a)
        cols = lineitem(n, a, rng, False)
        if args.sorted_dates:  # globally clustered: dates grow with the row number
            cols[5] = (8036 + ((a + np.arange(n, dtype=np.int64)) * 2526 // args.rows)).astype(np.int32)
        for k in range(5):
            sums[k] += int(cols[k].sum())
        m = (cols[5] >= lo) & (cols[5] < hi) & (cols[3] >= 5) & (cols[3] <= 7) & (cols[1] < 2400)
        q6_sum += int(cols[2][m].sum())
        q6_cnt += int(m.sum())
        w.append(cols)
    img = w.finish()
    t = flsgpu.Table(img)
    stream = torch.cuda.current_stream().cuda_stream
    t.stage(stream=stream)
    torch.cuda.synchronize()
    res = torch.zeros(5 * 64, dtype=torch.uint8, device="cuda")

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.iters

    def agg5():
        for k in range(5):
            flsgpu.agg_reset(res.data_ptr() + 64 * k, stream)
            t.aggregate_column(k, res.data_ptr() + 64 * k, stream=stream)

    ms_agg = timed(agg5)
    host = res.cpu().numpy()
    for k in range(5):
        r = flsgpu.AggResult.from_buffer_copy(host[64 * k:64 * k + 64].tobytes())
        assert ((r.sum_hi << 64) | r.sum_lo) == sums[k] and r.count == args.rows, f"column {k}: wrong SUM/COUNT"

    filt = [(5, flsgpu.CMP["GE"], lo), (5, flsgpu.CMP["LT"], hi), (3, flsgpu.CMP["GE"], 5), (3, flsgpu.CMP["LE"], 7), (1, flsgpu.CMP["LT"], 2400)]
    keep = t.rowgroup_prune(filt)
    kd = torch.from_numpy(keep).cuda()
    n_words = t.n_rowgroups * t.vec_per_rowgroup * 32
    mask = torch.zeros(n_words, dtype=torch.int32, device="cuda")

    def q6():
        t.filter_mask(filt, mask.data_ptr(), n_words * 4, kd.data_ptr(), stream=stream)
        flsgpu.agg_reset(res.data_ptr(), stream)
        t.aggregate_column_masked(2, mask.data_ptr(), res.data_ptr(), kd.data_ptr(), stream=stream)

    ms_q6 = timed(q6)
    r = flsgpu.AggResult.from_buffer_copy(res.cpu().numpy()[:64].tobytes())
    assert r.sum_lo == q6_sum and r.count == q6_cnt, "Q6: wrong SUM/COUNT"

    # CPU oracle, one thread, bounded sample of row groups
    o = oracle_lib.Oracle(img)
    n_rg = min(t.n_rowgroups, 24)
    t0 = time.perf_counter()
    for k in range(5):
        o.aggregate(k, 0, n_rg)
    cpu_agg = time.perf_counter() - t0
    t0 = time.perf_counter()
    m = o.filter_mask(filt, 0, n_rg, t.vec_per_rowgroup)
    o.aggregate_masked(2, 0, n_rg, m)
    cpu_q6 = time.perf_counter() - t0
    rows_s = sum(t.rowgroup(r_).n_tuples for r_ in range(n_rg))
    comp = sum(t.chunk(r_, c).size for r_ in range(t.n_rowgroups) for c in range(6))
    print(json.dumps({
        "workload": "lineitem-like, %d rows, 6 columns, %s l_shipdate" % (args.rows, "clustered" if args.sorted_dates else "random"),
        "compressed_MB": round(comp / 1e6, 1), "row_groups": t.n_rowgroups, "row_groups_kept_by_zone_maps": int(keep.sum()),
        "sum5_count": {"ms": round(ms_agg, 4), "values_per_s": round(5 * args.rows / ms_agg * 1e3, 1),
                       "cpu_oracle_1thread_values_per_s": round(5 * rows_s / cpu_agg, 1)},
        "q6_filter_sum": {"ms": round(ms_q6, 4), "rows_per_s": round(args.rows / ms_q6 * 1e3, 1), "selected_rows": q6_cnt,
                          "cpu_oracle_1thread_rows_per_s": round(rows_s / cpu_q6, 1)},
        "checked": "SUMs, COUNTs and the Q6 result equal numpy on the generated data"}), flush=True)
    t.close()
    w.close()


if __name__ == "__main__":
    main()

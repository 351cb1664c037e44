#!/usr/bin/env python
"""BASELINE config 5: a TPC-H lineitem-shaped table of scale factor SF with its row groups sharded across the ranks of one
box (one process per GPU; launch with torch.distributed.run for N > 1), per-GPU partial aggregates combined with NCCL.
Rank r builds and keeps resident rows [r, r+1) * SF * 6 M / N of the table (the 11 integer / decimal / date columns - the
query reads nothing else); one step = the config's query
    SELECT SUM(l_orderkey), SUM(l_partkey), SUM(l_suppkey), SUM(l_linenumber), SUM(l_quantity), SUM(l_extendedprice),
           SUM(l_discount), SUM(l_tax), COUNT(*) FROM lineitem
as eight fused decode+aggregate launches per rank, the 128-bit partial sums split into 32-bit limbs on the device, ONE
ncclAllReduce of 8 x 5 int64 lanes, one 320-byte D2H copy.  Also timed: the decode-only scan of the same 11 columns (no
exchange) and the Q1 / Q6 filter scans (mask + masked aggregates + the same all-reduce).  Every reduced result is checked
against the all-reduced numpy sums kept while generating.  Timing: barrier + synchronize on both sides, CUDA events, max
over ranks.  Rank 0 prints one JSON document.
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \\
       tools/lineitem_dist.py --sf 100 [--iters 20] [--out profiles/rXX_lineitem_sf100_n8.json]"""
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
Q1_COLS = ["l_quantity", "l_extendedprice", "l_discount", "l_tax"]
Q6_COLS = ["l_extendedprice", "l_discount"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=1.0)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":  # keeps NCCL's version banner off stdout
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import flsgpu
    from flsgpu import sharding
    flsgpu.init((local,))
    n_total = int(ROWS_PER_SF * args.sf)
    lo, hi = rank * n_total // world, (rank + 1) * n_total // world
    n_rows = hi - lo
    threads = max(1, (os.cpu_count() or 1) // world)

    # ---- this rank's shard: generate + encode batch by batch, keep the numpy answers
    t0 = time.perf_counter()
    w, names = None, None
    want = {"sum": dict.fromkeys(AGG_COLS, 0), "q1": dict.fromkeys(Q1_COLS, 0), "q6": dict.fromkeys(Q6_COLS, 0), "q1_count": 0, "q6_count": 0}
    batch, done, raw_bytes = 64 * 65536, 0, 0
    first_order = rank * 2 * n_total  # order keys ascend across ranks like across one dbgen run
    while done < n_rows:
        n = min(batch, n_rows - done)
        cols = lineitem(n, sf=args.sf, seed=1000 + 100000 * rank + done // batch, first_order=first_order, strings=False)
        last_key = int(cols[0][3][-1]) - 1
        first_order = (last_key // 32) * 8 + last_key % 8 + 1
        if w is None:
            names = [c[0] for c in cols]
            w = flsgpu.Writer(specs(cols), 64)
        by = {c[0]: c[3] for c in cols}
        q1 = by["l_shipdate"] <= 10471
        q6 = ((by["l_shipdate"] >= 8766) & (by["l_shipdate"] < 9131) & (by["l_discount"] >= 5) & (by["l_discount"] <= 7) &
              (by["l_quantity"] < 2400))
        for c in AGG_COLS:
            want["sum"][c] += int(by[c].sum(dtype=np.int64))
        for c in Q1_COLS:
            want["q1"][c] += int(by[c][q1].sum())
        for c in Q6_COLS:
            want["q6"][c] += int(by[c][q6].sum())
        want["q1_count"] += int(q1.sum())
        want["q6_count"] += int(q6.sum())
        raw_bytes += sum(np.asarray(c[3]).nbytes for c in cols)
        w.append([c[3] for c in cols], threads)
        done += n
    img = w.finish()
    build_s = time.perf_counter() - t0
    t = flsgpu.Table(img)
    stream = torch.cuda.current_stream().cuda_stream
    t.stage(stream=stream)
    torch.cuda.synchronize()
    n_rg, rows_out = t.n_rowgroups, t.out_elems(t.n_rowgroups)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step):
        """ms per step: CUDA events around args.iters steps between barriers, max over ranks."""
        for _ in range(args.warmup):
            step()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            step()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / args.iters], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def expect(local_pairs):
        """all-reduced numpy answers: python ints travel as objects (test infrastructure, not the data path)"""
        if dist is None:
            return local_pairs
        box = [None] * world
        dist.all_gather_object(box, local_pairs)
        return [(sum(b[i][0] for b in box), sum(b[i][1] for b in box)) for i in range(len(local_pairs))]

    # ---- the config's query: 8 fused aggregates per rank -> limbs on the device -> one all-reduce -> one D2H
    res = torch.zeros(64 * len(AGG_COLS), dtype=torch.uint8, device="cuda")
    idx = [names.index(c) for c in AGG_COLS]
    out = {}

    def q_sum():
        for i, c in enumerate(idx):
            flsgpu.agg_reset(res.data_ptr() + 64 * i, stream)
            t.aggregate_column(c, res.data_ptr() + 64 * i, stream=stream)
        out["sum"] = sharding.all_reduce_results(res, dist)
    sum_ms = timed(q_sum)
    assert out["sum"] == expect([(want["sum"][c], n_rows) for c in AGG_COLS]), "reduced SUM/COUNT differs from numpy"

    def q_sum_local():  # the same launches without the exchange and the D2H: what the collective costs
        for i, c in enumerate(idx):
            flsgpu.agg_reset(res.data_ptr() + 64 * i, stream)
            t.aggregate_column(c, res.data_ptr() + 64 * i, stream=stream)
    sum_local_ms = timed(q_sum_local)

    # ---- Q1 / Q6 scans
    sd, di, qu = names.index("l_shipdate"), names.index("l_discount"), names.index("l_quantity")
    f_q1 = [(sd, flsgpu.CMP["LE"], bits(np.int32(10471), np.int32))]
    f_q6 = [(sd, flsgpu.CMP["GE"], bits(np.int32(8766), np.int32)), (sd, flsgpu.CMP["LT"], bits(np.int32(9131), np.int32)),
            (di, flsgpu.CMP["GE"], bits(np.int64(5), np.int64)), (di, flsgpu.CMP["LE"], bits(np.int64(7), np.int64)),
            (qu, flsgpu.CMP["LT"], bits(np.int64(2400), np.int64))]
    n_words = n_rg * t.vec_per_rowgroup * 32
    mask = torch.zeros(n_words, dtype=torch.int32, device="cuda")
    queries = []
    for qname, filt, agg_cols in (("q1", f_q1, Q1_COLS), ("q6", f_q6, Q6_COLS)):
        resq = torch.zeros(64 * len(agg_cols), dtype=torch.uint8, device="cuda")

        def q(filt=filt, agg_cols=agg_cols, resq=resq, qname=qname):
            t.filter_mask(filt, mask.data_ptr(), n_words * 4, 0, stream=stream)
            for i, name in enumerate(agg_cols):
                flsgpu.agg_reset(resq.data_ptr() + 64 * i, stream)
                t.aggregate_column_masked(names.index(name), mask.data_ptr(), resq.data_ptr() + 64 * i, 0, stream=stream)
            out[qname] = sharding.all_reduce_results(resq, dist)
        ms = timed(q)
        exp = expect([(want[qname][c], want[qname + "_count"]) for c in agg_cols])
        assert out[qname] == exp, qname + ": reduced result differs from numpy"
        queries.append({"query": qname, "ms": round(ms, 4), "rows_per_s": round(n_total / ms * 1e3, 1), "count": exp[0][1],
                        "sums": {c: str(v[0]) for c, v in zip(agg_cols, exp)}})

    # ---- decode-only scan of the 11 columns into this rank's HBM (no exchange)
    widths = [t.column(c).phys_bits // 8 for c in range(len(names))]
    outs = [torch.empty(rows_out * wd, dtype=torch.uint8, device="cuda") for wd in widths]

    def scan():
        for c, o_ in enumerate(outs):
            t.decode_column(c, o_.data_ptr(), o_.numel(), stream=stream)
    scan_ms = timed(scan)
    dec_local = sum(n_rows * wd for wd in widths)
    sizes = torch.tensor([dec_local, len(img), raw_bytes, int(build_s * 1e3)], dtype=torch.int64, device="cuda")
    mx = sizes.clone()
    if dist is not None:
        dist.all_reduce(sizes, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dec_all, cont_all, raw_all, _ = sizes.tolist()
    agg_bytes = dec_all * sum(widths[i] for i in idx) // sum(widths)  # decoded-equivalent bytes of the 8 aggregated columns
    if rank == 0:
        doc = {"workload": "TPC-H lineitem shape SF%g, integer/decimal/date columns (%d rows in total, %d row groups per GPU), "
                           "row groups sharded across %d GPU(s), strong scaling" % (args.sf, n_total, n_rg, world),
               "n_gpus": world, "rows": n_total, "raw_MB": round(raw_all / 1e6, 1), "container_MB": round(cont_all / 1e6, 1),
               "build_s_max": round(mx.tolist()[3] / 1e3, 1),
               "sum_count_query": {"sql": "SELECT SUM(l_orderkey), SUM(l_partkey), SUM(l_suppkey), SUM(l_linenumber), SUM(l_quantity), "
                                          "SUM(l_extendedprice), SUM(l_discount), SUM(l_tax), COUNT(*)",
                                   "ms": round(sum_ms, 4), "ms_without_allreduce_and_d2h": round(sum_local_ms, 4),
                                   "rows_per_s": round(n_total / sum_ms * 1e3, 1),
                                   "values_per_s": round(n_total * len(AGG_COLS) / sum_ms * 1e3, 1),
                                   "decoded_equivalent_GBps": round(agg_bytes / sum_ms / 1e6, 1),
                                   "collective": "1 ncclAllReduce(SUM) of %d int64 lanes + 1 D2H of %d B per query" % (5 * len(AGG_COLS), 40 * len(AGG_COLS)),
                                   "count": out["sum"][0][1], "sums": {c: str(v[0]) for c, v in zip(AGG_COLS, out["sum"])}},
               "queries": queries,
               "decode_scan": {"columns": len(names), "ms": round(scan_ms, 4), "decoded_GBps": round(dec_all / scan_ms / 1e6, 1),
                               "rows_per_s": round(n_total / scan_ms * 1e3, 1)}}
        print(json.dumps(doc, indent=1))
        if args.out:
            with open(args.out, "w") as f:
                json.dump(doc, f, indent=1)
    t.close()
    w.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""World-size-2 (gloo, CPU) coverage of the multi-GPU host logic: row-group sharding + reduction of partial aggregates.
Each rank aggregates ITS row groups (here with the oracle standing in for the device, there is no GPU in this
container); the all-reduced result must equal the single-rank aggregate over the whole table."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from flsgpu import sharding


def test_shard_rowgroups_partition():
    for n in (0, 1, 7, 64, 1526, 9156):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_rowgroups(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1


def test_partial_pack_roundtrip_and_carries():
    for s in (0, 1, -1, 2 ** 64, -(2 ** 100), 2 ** 126, -(2 ** 127)):
        assert sharding.unpack_partial(sharding.pack_partial(s, 5)) == (s, 5)
    # limb-wise SUM of many partials then carry propagation == big-int sum
    vals = [2 ** 95 - 3, -(2 ** 90), 2 ** 64 - 1, -1, 12345678901234567890]
    t = sum(sharding.pack_partial(v, 1) for v in vals)
    assert sharding.unpack_partial(t) == (sum(vals), len(vals))


def _results_buffer(partials):
    """k (sum128, count) pairs as the device would hold them: k flsgpu_agg_result structs in one uint8 tensor."""
    import flsgpu
    raw = b""
    for s, c in partials:
        u = s & ((1 << 128) - 1)
        raw += bytes(flsgpu.AggResult(sum_lo=u & (2 ** 64 - 1), sum_hi=u >> 64, count=c))
    return torch.frombuffer(bytearray(raw), dtype=torch.uint8)


def test_limbs_of_result_structs_match_pack_partial():
    partials = [(0, 0), (1, 1), (-1, 7), (2 ** 64, 3), (-(2 ** 100), 2 ** 40), (2 ** 126 + 12345, 9), (-(2 ** 127), 1)]
    limbs = sharding.limbs_of_results(_results_buffer(partials))
    for row, (s, c) in zip(limbs, partials):
        assert row.tolist() == sharding.pack_partial(s, c).tolist()
    assert sharding.all_reduce_results(_results_buffer(partials)) == partials


def _worker(rank, world, port, img_path, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.join(os.path.dirname(here), "duckdb-fastlanes_b200"), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import flsgpu
    import oracle_lib
    dist.init_process_group("gloo", rank=rank, world_size=world)
    img = np.fromfile(img_path, dtype=np.uint8)
    t = flsgpu.Table(img)
    b, e = sharding.shard_rowgroups(t.n_rowgroups, rank, world)
    o = oracle_lib.Oracle(img)
    out = {}
    for col in range(t.n_columns):
        part = o.aggregate(col, b, e - b) if e > b else dict(sum=0, count=0, min=2 ** 63 - 1, max=-(2 ** 63))
        s, c = sharding.all_reduce_partial(part["sum"], part["count"], dist)
        mn, mx = sharding.all_reduce_minmax(part["min"], part["max"], dist)
        out[col] = (s, c, mn, mx)
    # the same reduction with the partials left in a buffer of result structs: one all-reduce for every column
    parts = []
    for col in range(t.n_columns):
        part = o.aggregate(col, b, e - b) if e > b else dict(sum=0, count=0)
        parts.append((part["sum"], part["count"]))
    batched = sharding.all_reduce_results(_results_buffer(parts), dist)
    assert batched == [(out[col][0], out[col][1]) for col in range(t.n_columns)]
    if rank == 0:
        q.put(out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_aggregate_equals_single_rank(tmp_path):
    import flsgpu
    import oracle_lib
    n = 9 * 2048 + 700
    r = np.random.default_rng(8)
    a = r.integers(-2 ** 62, 2 ** 62, n, dtype=np.int64)  # sums overflow 64 bits
    b = np.cumsum(r.integers(0, 100, n)).astype(np.int64)
    c = r.integers(-100, 100, n).astype(np.int16)
    img = flsgpu.write_table([("a", "INT64", None), ("b", "INT64", None), ("c", "INT16", None)], [a, b, c], vec_per_rowgroup=2)
    path = str(tmp_path / "t.fls")
    img.tofile(path)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(rk, 2, port, path, q)) for rk in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for col, src in enumerate((a, b, c)):
        s, cnt, mn, mx = got[col]
        assert s == int(src.astype(object).sum()) and cnt == n
        assert (mn, mx) == (int(src.min()), int(src.max()))

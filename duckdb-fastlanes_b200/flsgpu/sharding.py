"""Row-group sharding across ranks and the reduction of partial scan aggregates.

Row groups are independent units (the reference hands them to independent threads, src/reader/fls_reader.cpp:337-355),
so rank r decodes its own contiguous range and nothing is exchanged on the data path.  The only collective is the
reduction of the per-rank partial aggregates: a 128-bit SUM travels as four 32-bit limbs in int64 lanes so that a plain
SUM all-reduce over up to 2^31 ranks cannot overflow a lane; carries are propagated after the reduce.
"""
import torch


def shard_rowgroups(n_rowgroups, rank, world):
    """Contiguous, balanced [begin, end) of row groups for `rank` (sizes differ by at most one)."""
    base, extra = divmod(n_rowgroups, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def pack_partial(sum128, count):
    """(signed 128-bit python int, count) -> int64[5]: four 32-bit limbs (two's complement) + count."""
    u = sum128 & ((1 << 128) - 1)
    return torch.tensor([(u >> (32 * k)) & 0xFFFFFFFF for k in range(4)] + [count], dtype=torch.int64)


def unpack_partial(t):
    """Inverse of pack_partial after a SUM reduce: propagate carries, wrap to 128 bits, reinterpret as signed."""
    limbs = [int(x) for x in t[:4].tolist()]
    u = sum(l << (32 * k) for k, l in enumerate(limbs)) & ((1 << 128) - 1)
    if u >= 1 << 127:
        u -= 1 << 128
    return u, int(t[4])


def all_reduce_partial(sum128, count, dist=None, device="cpu"):
    """Combine per-rank (SUM, COUNT) partials: one all_reduce(SUM) of five int64 lanes (NCCL on GPU, gloo on CPU)."""
    t = pack_partial(sum128, count).to(device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return unpack_partial(t.cpu())


def limbs_of_results(res):
    """uint8[k * 64] tensor of k flsgpu_agg_result structs (include/flsgpu.h: sum_lo, sum_hi, count, ...) -> int64[k, 5]
    (four 32-bit limbs of the 128-bit SUM + COUNT), computed where the tensor lives: the partials of a query never visit
    the host before the reduce."""
    v = res.view(torch.int64).view(-1, 8)
    lo, hi, m = v[:, 0], v[:, 1], 0xFFFFFFFF
    return torch.stack([lo & m, (lo >> 32) & m, hi & m, (hi >> 32) & m, v[:, 2]], dim=1).contiguous()


def all_reduce_results(res, dist=None):
    """Combine the (SUM, COUNT) partials of k aggregates held in a device buffer of k flsgpu_agg_result structs: ONE
    all_reduce(SUM) of k x 5 int64 lanes, then one D2H copy.  Returns [(sum128 as signed python int, count)] * k."""
    limbs = limbs_of_results(res)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(limbs, op=dist.ReduceOp.SUM)
    host = limbs.cpu()
    return [unpack_partial(host[i]) for i in range(host.shape[0])]


def all_reduce_minmax(mn, mx, dist=None, device="cpu"):
    lo = torch.tensor([mn], dtype=torch.int64, device=device)
    hi = torch.tensor([mx], dtype=torch.int64, device=device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    return int(lo.item()), int(hi.item())

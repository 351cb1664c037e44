"""GPU: filter pushdown through the C ABI (selection masks, zone-map pruned launches, masked SUM/COUNT) == oracle."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from filter_cases import bits, filter_sets, numpy_mask, table

pytestmark = pytest.mark.gpu


def _agg(buf):
    r = flsgpu.AggResult.from_buffer_copy(buf.cpu().numpy().tobytes())
    s = (r.sum_hi << 64) | r.sum_lo
    return (s - (1 << 128) if s >= (1 << 127) else s), r.count, r.fsum


@pytest.mark.parametrize("n,vpr,force", [(5000, 2, None), (200000, 16, None), (70000, 64, "FFOR"), (70000, 8, "DICT"), (70000, 8, "RLE")])
def test_masks_and_masked_aggregates_match_oracle(n, vpr, force):
    import torch
    cols = table(n)
    specs = []
    for c in cols:
        enc = c[2]
        if force and not (c[1] in ("DOUBLE", "FLOAT")):
            enc = force
        specs.append((c[0], c[1], enc) + tuple(c[4:6]))
    try:
        img = flsgpu.write_table(specs, [c[3] for c in cols], vec_per_rowgroup=vpr)
    except flsgpu.FlsError:
        pytest.skip(f"{force} not applicable to every column of this table")
    names = [c[0] for c in cols]
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    n_words = t.n_rowgroups * t.vec_per_rowgroup * 32
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    for fname, fs in filter_sets().items():
        filt = [(names.index(c), flsgpu.CMP[op], bits(k, k.dtype)) for c, op, k in fs]
        exp = o.filter_mask(filt, 0, t.n_rowgroups, t.vec_per_rowgroup)
        for use_prune in (False, True):
            mask = torch.full((n_words,), -1, dtype=torch.int32, device="cuda")
            keep_ptr = 0
            if use_prune:
                keep = torch.from_numpy(t.rowgroup_prune(filt)).cuda()
                keep_ptr = keep.data_ptr()
            t.filter_mask(filt, mask.data_ptr(), n_words * 4, keep_ptr)
            flsgpu.sync()
            got = mask.cpu().numpy().view(np.uint32)
            assert got.tobytes() == exp.tobytes(), (fname, use_prune)
            for col in ("price", "small", "dbl"):
                ci = names.index(col)
                flsgpu.agg_reset(res.data_ptr())
                t.aggregate_column_masked(ci, mask.data_ptr(), res.data_ptr(), keep_ptr)
                flsgpu.sync()
                s, cnt, fsum = _agg(res)
                es, ecnt, efs = o.aggregate_masked(ci, 0, t.n_rowgroups, exp)
                assert cnt == ecnt, (fname, col)
                if col == "dbl":
                    assert (np.isnan(fsum) and np.isnan(efs)) or abs(fsum - efs) <= 1e-9 * max(1.0, abs(efs)), (fname, col)
                else:
                    assert s == es, (fname, col)
        # value level
        want = numpy_mask(cols, fs)
        assert int(np.unpackbits(got.view(np.uint8)).sum()) == int(want.sum()), fname
    t.close()


def test_subrange_and_q6_value():
    """SELECT SUM(price * 1), COUNT(*) WHERE Q6 predicates, over a sub-range of row groups, pruned launches."""
    import torch
    n = 40 * 2048 + 77
    cols = table(n, seed=9)
    names = [c[0] for c in cols]
    img = flsgpu.write_table([(c[0], c[1], c[2]) + tuple(c[4:6]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    fs = filter_sets()["q6"]
    filt = [(names.index(c), flsgpu.CMP[op], bits(k, k.dtype)) for c, op, k in fs]
    rg0, nrg = 7, 25
    keep = t.rowgroup_prune(filt, rg0, nrg)
    assert 0 < keep.sum() < nrg
    kd = torch.from_numpy(keep).cuda()
    n_words = nrg * t.vec_per_rowgroup * 32
    mask = torch.zeros(n_words, dtype=torch.int32, device="cuda")
    t.filter_mask(filt, mask.data_ptr(), n_words * 4, kd.data_ptr(), rg0, nrg)
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    flsgpu.agg_reset(res.data_ptr())
    t.aggregate_column_masked(names.index("price"), mask.data_ptr(), res.data_ptr(), kd.data_ptr(), rg0, nrg)
    flsgpu.sync()
    s, cnt, _ = _agg(res)
    rows = slice(rg0 * 2048, (rg0 + nrg) * 2048)
    want = numpy_mask(cols, fs)[rows]
    assert cnt == int(want.sum()) and s == int(cols[names.index("price")][3][rows][want].sum())
    assert mask.cpu().numpy().view(np.uint32).tobytes() == o.filter_mask(filt, rg0, nrg, 2).tobytes()
    # string columns are not filtered on the device
    t.close()


@pytest.mark.parametrize("n,vpr,slab", [(5000, 2, None), (300000, 4, "3"), (70000, 64, None)])
def test_aggregate_to_host_matches_numpy(n, vpr, slab, monkeypatch):
    """flsgpu_aggregate_to_host: SELECT SUM / COUNT / MIN / MAX [WHERE ...] from host memory to host memory (slabs of row
    groups on two streams, nothing staged beforehand) == plain numpy over the input."""
    if slab:
        monkeypatch.setenv("FLSGPU_AGG_SLAB_RG", slab)  # many slabs: both streams' result sets get used
    cols = table(n)
    names = [c[0] for c in cols]
    img = flsgpu.write_table([(c[0], c[1], c[2]) + tuple(c[4:6]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=vpr)
    by = {c[0]: c[3] for c in cols}
    agg_names = ["price", "discount", "small", "u16", "runs", "dic", "shipdate", "dbl"]
    ids = [names.index(c) for c in agg_names]

    def check(res, rows, with_minmax):
        for name, r in zip(agg_names, res):
            v = by[name][rows]
            assert r.count == len(v), name
            if name == "dbl":
                es = float(np.sum(v))
                assert (np.isnan(r.fsum) and np.isnan(es)) or abs(r.fsum - es) <= 1e-9 * max(1.0, abs(es)), name
                continue
            s = (r.sum_hi << 64) | r.sum_lo
            s -= (1 << 128) if s >= (1 << 127) else 0
            assert s == int(v.astype(object).sum()) if len(v) else s == 0, name
            if with_minmax and len(v):
                mn, mx = (r.min_val, r.max_val) if v.dtype.kind == "i" else (r.min_val & (2 ** 64 - 1), r.max_val & (2 ** 64 - 1))
                assert (mn, mx) == (int(v.min()), int(v.max())), name

    t = flsgpu.Table(img)  # never staged by hand: the call stages what it needs
    everything = np.ones(n, dtype=bool)
    check(t.aggregate_to_host(ids), everything, True)
    check(t.aggregate_to_host(ids), everything, True)  # resident now: same answer
    for fname, fs in filter_sets().items():
        filt = [(names.index(c), flsgpu.CMP[op], bits(k, k.dtype)) for c, op, k in fs]
        check(t.aggregate_to_host(ids, filt), numpy_mask(cols, fs), False)
    # a sub-range of row groups
    if t.n_rowgroups > 3:
        rows = np.zeros(n, dtype=bool)
        stride = t.vec_per_rowgroup * 1024
        rows[stride:min(n, (t.n_rowgroups - 1) * stride)] = True
        check(t.aggregate_to_host(ids, rg_begin=1, n_rg=t.n_rowgroups - 2), rows, True)
    with pytest.raises(flsgpu.FlsError):
        t.aggregate_to_host([99])
    t.close()

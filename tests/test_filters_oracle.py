"""CPU: the oracle's filter pushdown (constant comparisons -> selection mask, zone-map pruning, masked SUM/COUNT) against
plain numpy on the generated data."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from filter_cases import bits, filter_sets, numpy_mask, table


def build(n, vpr):
    cols = table(n)
    img = flsgpu.write_table([(c[0], c[1], c[2]) + tuple(c[4:6]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=vpr)
    return cols, img


def mask_to_bool(t, mask):
    out = []
    for a, b in t.valid_mask_slices():
        w = mask[a // 32:(b + 31) // 32]
        out.append(np.unpackbits(w.view(np.uint8), bitorder="little")[:b - a].astype(bool))
    return np.concatenate(out)


@pytest.mark.parametrize("n,vpr", [(5000, 2), (70000, 16)])
def test_oracle_mask_matches_numpy(n, vpr):
    cols, img = build(n, vpr)
    names = [c[0] for c in cols]
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    for fname, fs in filter_sets().items():
        filt = [(names.index(c), flsgpu.CMP[op], bits(k, k.dtype)) for c, op, k in fs]
        mask = o.filter_mask(filt, 0, t.n_rowgroups, t.vec_per_rowgroup)
        want = numpy_mask(cols, fs)
        got = mask_to_bool(t, mask)
        assert got.tolist() == want.tolist(), fname
        # padding rows never pass
        assert int(np.unpackbits(mask.view(np.uint8)).sum()) == int(want.sum()), fname
        # masked SUM / COUNT of the price column
        s, cnt, _ = o.aggregate_masked(names.index("price"), 0, t.n_rowgroups, mask)
        assert cnt == int(want.sum()) and s == int(cols[names.index("price")][3][want].sum()), fname
        # zone-map pruning never drops a row group that has a passing row (product and oracle restatements agree)
        keep = t.rowgroup_prune([(c, op, k) for c, op, k in filt])
        for rg, (a, b) in enumerate(t.valid_mask_slices()):
            has = bool(got[sum(t.rowgroup(r).n_tuples for r in range(rg)):][:t.rowgroup(rg).n_tuples].any())
            assert keep[rg] or not has, (fname, rg)
            assert o.rowgroup_may_satisfy(filt, rg) or not has, (fname, rg)
            if not any(op == flsgpu.CMP["EQ"] for _, op, _ in filt):
                assert bool(keep[rg]) == o.rowgroup_may_satisfy(filt, rg), (fname, rg)
    # the clustered date column really prunes
    q6 = [(names.index(c), flsgpu.CMP[op], bits(k, k.dtype)) for c, op, k in filter_sets()["q6"]]
    if t.n_rowgroups > 2:
        assert 0 < int(t.rowgroup_prune(q6).sum()) < t.n_rowgroups
    t.close()

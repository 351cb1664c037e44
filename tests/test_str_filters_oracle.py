"""CPU: the oracle's string filters (constant comparisons on VARCHAR / BLOB columns -> selection mask) against plain Python
bytes comparisons on the generated data; every string encoding the writer produces."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from str_filter_cases import filter_sets, python_mask, table
from test_filters_oracle import mask_to_bool


@pytest.mark.parametrize("n,vpr", [(3000, 2), (20000, 8)])
def test_oracle_string_masks_match_python(n, vpr):
    specs, data = table(n)
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=vpr)
    names = [s[0] for s in specs]
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    for fname, fs in filter_sets(data).items():
        filt = [(names.index(c), flsgpu.CMP[op], k) for c, op, k in fs]
        mask = o.filter_mask(filt, 0, t.n_rowgroups, t.vec_per_rowgroup)
        want = python_mask(specs, data, fs)
        assert mask_to_bool(t, mask).tolist() == want.tolist(), fname
        assert int(np.unpackbits(mask.view(np.uint8)).sum()) == int(want.sum()), fname  # padding rows never pass
        # string chunks carry no zone maps: nothing is pruned on their account
        if all(isinstance(k, bytes) for _, _, k in fs):
            assert all(o.rowgroup_may_satisfy(filt, rg) for rg in range(t.n_rowgroups)), fname
            assert t.rowgroup_prune(filt).all(), fname
    t.close()


def test_string_filter_without_bytes_is_rejected():
    specs, data = table(100)
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    f = flsgpu.make_filters([(0, "EQ", b"AIR")])
    f[0].str_data = None
    keep = np.zeros(t.n_rowgroups, dtype=np.uint8)
    rc = flsgpu.lib.flsgpu_rowgroup_prune(t._h, f, 1, 0, t.n_rowgroups, keep.ctypes.data_as(flsgpu.C.c_void_p))
    assert rc != 0
    t.close()

"""GPU: flsgpu_filter_mask_string_dict_static - the allocation-free variant of the dictionary string filter (scratch in a
static device array of FLSGPU_STATIC_SCRATCH_BYTES instead of the stream-ordered pool) - against the existing entry point
flsgpu_filter_mask, through the C ABI, on calls in a row (the array keeps the last call's contents).  No counterpart in the
reference.  Collected last on purpose."""
import numpy as np
import pytest

import flsgpu
from str_filter_cases import PY_OP, table

pytestmark = pytest.mark.gpu

STATIC_BYTES = 4 << 20  # include/flsgpu.h FLSGPU_STATIC_SCRATCH_BYTES


def _both(t, torch, filt, n_words, keep_ptr=0, rg_begin=0, n_rg=None):
    """(mask of flsgpu_filter_mask, mask of the static variant) for one filter; both buffers start as all ones."""
    ref = torch.full((n_words,), -1, dtype=torch.int32, device="cuda")
    got = torch.full((n_words,), -1, dtype=torch.int32, device="cuda")
    t.filter_mask([filt], ref.data_ptr(), n_words * 4, keep_ptr, rg_begin=rg_begin, n_rg=n_rg)
    flsgpu.sync()
    t.filter_mask_string_dict_static(filt, got.data_ptr(), n_words * 4, keep_ptr, rg_begin=rg_begin, n_rg=n_rg)
    flsgpu.sync()
    return ref.cpu().numpy().view(np.uint32), got.cpu().numpy().view(np.uint32)


def test_static_scratch_filter_matches_filter_mask():
    import torch
    specs, data = table(20000)
    names = [s[0] for s in specs]
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=4)
    t = flsgpu.Table(img)
    t.stage()
    n_words = t.n_rowgroups * t.vec_per_rowgroup * 32
    # calls in a row over three dictionary encodings (DICT_STR, FSST_DICT, FSST12_DICT): keys, dictionary bytes and flags of
    # the previous call are still in the array when the next one starts
    calls = [("shipmode", "EQ", b"REG AIR"), ("shipmode", "NE", b"AIR"), ("instruct", "LT", b"DELIVER IN PERSOO"),
             ("instruct", "EQ", b"DELIVER IN PERSON"), ("instruct12", "EQ", b"NONE"), ("shipmode", "GT", b"REG"),
             ("shipmode", "GE", b""), ("shipmode", "LT", b"")]
    selected = []
    for c, op, k in calls:
        filt = (names.index(c), flsgpu.CMP[op], k)
        ref, got = _both(t, torch, filt, n_words)
        assert got.tobytes() == ref.tobytes(), (c, op, k)
        col = data[names.index(c)]
        want = sum(1 for v in col if PY_OP[op](bytes(v), k))  # DuckDB's string_t order is Python's bytes order
        assert int(np.unpackbits(got.view(np.uint8)).sum()) == want, (c, op, k)
        selected.append(want)
    assert any(selected) and selected[-1] == 0 and selected[-2] == len(data[0])
    # the same call twice gives the same mask
    filt = (names.index("shipmode"), flsgpu.CMP["EQ"], b"RAIL")
    a = _both(t, torch, filt, n_words)
    b = _both(t, torch, filt, n_words)
    assert a[1].tobytes() == a[0].tobytes() == b[1].tobytes()
    # pruned row groups and a sub-range of row groups
    keep_host = (np.arange(t.n_rowgroups) % 2).astype(np.uint8)
    keep = torch.from_numpy(keep_host).cuda()
    ref, got = _both(t, torch, (names.index("shipmode"), flsgpu.CMP["NE"], b"AIR"), n_words, keep.data_ptr())
    assert got.tobytes() == ref.tobytes()
    assert not got.reshape(t.n_rowgroups, -1)[keep_host == 0].any()
    sub_words = (t.n_rowgroups - 2) * t.vec_per_rowgroup * 32
    ref, got = _both(t, torch, (names.index("instruct"), flsgpu.CMP["GE"], b"NONE"), sub_words, 0, rg_begin=1, n_rg=t.n_rowgroups - 2)
    assert got.tobytes() == ref.tobytes() and got.any()
    # what the variant refuses: a column that is not dictionary encoded, a numeric column
    mask = torch.zeros(n_words, dtype=torch.int32, device="cuda")
    with pytest.raises(flsgpu.FlsError, match="not dictionary encoded"):
        t.filter_mask_string_dict_static((names.index("comment"), flsgpu.CMP["EQ"], b"x"), mask.data_ptr(), n_words * 4)
    with pytest.raises(flsgpu.FlsError, match="VARCHAR / BLOB"):
        t.filter_mask_string_dict_static((names.index("v"), flsgpu.CMP["EQ"], 5), mask.data_ptr(), n_words * 4)
    t.close()
    # a table that was never staged has no device state: creating it would allocate
    t = flsgpu.Table(img)
    with pytest.raises(flsgpu.FlsError, match="not staged"):
        t.filter_mask_string_dict_static((names.index("shipmode"), flsgpu.CMP["EQ"], b"AIR"), mask.data_ptr(), n_words * 4)
    t.close()


def test_static_scratch_limit_is_enforced():
    import torch
    n = 150000  # unique 24-byte keys: 16 B string_t + 24 B + 1 flag per key > 4 MiB over the whole table, well under it per row group
    strs = [b"key-%019d" % (i * 2654435761 % 10**19) for i in range(n)]
    img = flsgpu.write_table([("k", "VARCHAR", "DICT_STR")], [strs], vec_per_rowgroup=16)
    t = flsgpu.Table(img)
    t.stage()
    n_words = t.n_rowgroups * t.vec_per_rowgroup * 32
    mask = torch.full((n_words,), -1, dtype=torch.int32, device="cuda")
    with pytest.raises(flsgpu.FlsError, match="static array holds %d" % STATIC_BYTES):
        t.filter_mask_string_dict_static((0, flsgpu.CMP["GE"], strs[7]), mask.data_ptr(), n_words * 4)
    # row group by row group it fits, and the pieces put together equal the pooled entry point's mask
    ref = torch.full((n_words,), -1, dtype=torch.int32, device="cuda")
    t.filter_mask([(0, flsgpu.CMP["GE"], strs[7])], ref.data_ptr(), n_words * 4)
    per_rg = t.vec_per_rowgroup * 32
    for r in range(t.n_rowgroups):
        t.filter_mask_string_dict_static((0, flsgpu.CMP["GE"], strs[7]), mask.data_ptr() + r * per_rg * 4, per_rg * 4, rg_begin=r, n_rg=1)
    flsgpu.sync()
    assert mask.cpu().numpy().tobytes() == ref.cpu().numpy().tobytes()
    want = sum(1 for s in strs if s >= strs[7])
    assert int(np.unpackbits(mask.cpu().numpy().view(np.uint8)).sum()) == want
    t.close()

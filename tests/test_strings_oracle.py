"""CPU: string columns written by the product writer and decoded by the oracle reproduce the input (value-level round
trip, the reference's own test contract for VARCHAR/BLOB: test/all_types_single_threaded.test:300-326)."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
import str_cases


@pytest.mark.parametrize("n,vpr", [(1, 2), (1024, 2), (5000, 2), (70000, 64)])
def test_string_roundtrip_through_oracle(n, vpr):
    cs = str_cases.cases(n)
    names = list(cs)
    img = flsgpu.write_table([(k, cs[k][0], cs[k][1]) for k in names], [cs[k][2] for k in names], vec_per_rowgroup=vpr)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    heap_base = 0x7F0000001000
    for c, k in enumerate(names):
        assert t.column(c).phys_bits == 128
        st, heap = o.decode_strings(c, 0, t.n_rowgroups, t.vec_per_rowgroup, heap_base)
        st = st.view(flsgpu.STRING_T).reshape(-1)
        got = []
        for a, b in t.valid_mask_slices():
            got += flsgpu.strings_from(st[a:b], heap, heap_base)
        assert got == cs[k][2], f"column {k}"
        # string_t rules: inlined strings are zero padded, long ones carry their first 4 bytes as prefix
        for s, v in zip(np.concatenate([st[a:b] for a, b in t.valid_mask_slices()])[:200], cs[k][2][:200]):
            if len(v) <= 12:
                assert s["body"].tobytes() == v + b"\0" * (12 - len(v))
            else:
                assert s["body"][:4].tobytes() == v[:4]
    t.close()


def test_string_encodings_selected():
    cs = str_cases.cases(5000)
    names = list(cs)
    img = flsgpu.write_table([(k, cs[k][0], cs[k][1]) for k in names], [cs[k][2] for k in names], vec_per_rowgroup=64)
    t = flsgpu.Table(img)
    enc = {k: flsgpu.ENC_NAME[t.chunk(0, c).encoding] for c, k in enumerate(names)}
    assert enc["value_i"] == "DICT_STR" and enc["constant"] == "STR_CONSTANT" and enc["empty"] == "STR_CONSTANT"
    assert enc["comment"] == "FSST" and enc["value_unique"] == "FSST"
    # FSST really compresses TPC-H-like text
    ch = t.chunk(0, names.index("comment"))
    assert ch.seg_size[2] - 2304 < 0.6 * ch.const_val
    t.close()


def test_string_dictionary_form():
    cs = str_cases.cases(3000)
    img = flsgpu.write_table([("s", "VARCHAR", "DICT_STR"), ("f", "VARCHAR", "FSST_DICT")], [cs["value_i"][2], cs["instruct"][2]], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    for col, src in ((0, cs["value_i"][2]), (1, cs["instruct"][2])):
        row0 = 0
        for rg in range(t.n_rowgroups):
            ch = t.chunk(rg, col)
            keys, heap = o.decode_string_dict(col, rg, ch.aux_count, ch.const_val)
            keys = flsgpu.strings_from(keys.view(flsgpu.STRING_T).reshape(-1), heap)
            n = t.rowgroup(rg).n_tuples
            codes = np.concatenate([o.decode_codes(rg, col, v) for v in range(ch.n_vec)])[:n]
            assert [keys[c] for c in codes] == src[row0:row0 + n]
            row0 += n
    t.close()


def test_writer_buffers_string_rows_across_appends():
    vals = str_cases.comment_like(5000)
    ids = np.arange(5000, dtype=np.int64)
    w = flsgpu.Writer([("id", "INT64", None), ("s", "VARCHAR", None)], 2)
    for a, b in ((0, 100), (100, 2048), (2048, 2049), (2049, 5000)):
        w.append([ids[a:b], vals[a:b]])
    img = w.finish().copy()
    w.close()
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    st, heap = o.decode_strings(1, 0, t.n_rowgroups, t.vec_per_rowgroup)
    st = st.view(flsgpu.STRING_T).reshape(-1)
    got = []
    for a, b in t.valid_mask_slices():
        got += flsgpu.strings_from(st[a:b], heap)
    assert got == vals
    t.close()


def test_string_statistics_come_from_the_dictionaries():
    """flsgpu_column_string_stats (host only): longest string, ASCII flag and min / max prefix of columns whose distinct strings
    sit uncompressed in the container; unavailable - not wrong - for per-row and FSST-compressed encodings."""
    cs = str_cases.cases(5000)
    names = list(cs)
    img = flsgpu.write_table([(k, cs[k][0], cs[k][1]) for k in names], [cs[k][2] for k in names], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    seen = 0
    for c, k in enumerate(names):
        encs = {flsgpu.ENC_NAME[t.chunk(r, c).encoding] for r in range(t.n_rowgroups)}
        st = t.string_stats(c)
        if encs <= {"DICT_STR", "STR_CONSTANT", "RLE_STR", "CROSS_RLE_STR", "FREQUENCY_STR"}:
            vals = [bytes(v) for v in cs[k][2]]
            assert st == (max(len(v) for v in vals), any(b >= 0x80 for v in vals for b in v), min(vals)[:8], max(vals)[:8]), k
            seen += 1
        else:
            assert st is None, (k, encs)
    assert seen >= 6
    with pytest.raises(flsgpu.FlsError):
        flsgpu.Table(flsgpu.write_table([("v", "INT64", None)], [np.arange(10, dtype=np.int64)])).string_stats(0)
    t.close()

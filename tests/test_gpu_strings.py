"""GPU: string columns (VARCHAR / BLOB) decoded by the CUDA string kernels through the C ABI == oracle, bit for bit:
string_t rows (length, inlined bytes / prefix + pointer), string heap bytes, dictionary keys and selection vectors."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
import str_cases

pytestmark = pytest.mark.gpu

HEAP_BASE = 0x7F0000001000  # pretend destination address of the heap (what a host consumer would pass)


def _table(n, vpr, only=None):
    cs = str_cases.cases(n)
    names = [k for k in cs if only is None or k in only]
    img = flsgpu.write_table([(k, cs[k][0], cs[k][1]) for k in names], [cs[k][2] for k in names], vec_per_rowgroup=vpr)
    return cs, names, img


def gpu_strings(t, col, rg_begin=0, n_rg=None, heap_base=HEAP_BASE):
    import torch
    n_rg = t.n_rowgroups - rg_begin if n_rg is None else n_rg
    n_rows = t.out_elems(n_rg)
    hb = t.string_heap_bytes(col, rg_begin, n_rg)
    rows = torch.full((n_rows * 16,), 0xA5, dtype=torch.uint8, device="cuda")
    heap = torch.zeros((max(16, hb),), dtype=torch.uint8, device="cuda")
    t.decode_strings(col, rows.data_ptr(), rows.numel(), heap.data_ptr(), heap.numel(), heap_base, rg_begin, n_rg)
    flsgpu.sync()
    torch.cuda.synchronize()
    return rows.cpu().numpy().reshape(-1, 16), heap.cpu().numpy(), hb


@pytest.mark.parametrize("n,vpr", [(1, 2), (1024, 2), (5000, 2), (70000, 64), (3 * 65536 + 77, 64)])
def test_flat_strings_match_oracle(n, vpr):
    cs, names, img = _table(n, vpr)
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    for c, k in enumerate(names):
        exp_rows, exp_heap = o.decode_strings(c, 0, t.n_rowgroups, t.vec_per_rowgroup, HEAP_BASE)
        got_rows, got_heap, hb = gpu_strings(t, c)
        assert hb == o.string_heap_bytes(c, 0, t.n_rowgroups)
        # every row the oracle writes (all rows of existing vectors) is identical, bit for bit
        for r in range(t.n_rowgroups):
            a = r * t.vec_per_rowgroup * 1024
            b = a + t.rowgroup(r).n_vec * 1024
            assert got_rows[a:b].tobytes() == exp_rows[a:b].tobytes(), (k, r)
        # heap: identical up to each chunk's decoded size (padding bytes between chunks are unspecified)
        off = 0
        for r in range(t.n_rowgroups):
            used = t.chunk(r, c).const_val
            assert got_heap[off:off + used].tobytes() == exp_heap[off:off + used].tobytes(), (k, r)
            off += (used + 15) // 16 * 16
        # value level: the strings are the input
        st = got_rows.copy().view(flsgpu.STRING_T).reshape(-1)
        got = []
        for a, b in t.valid_mask_slices():
            got += flsgpu.strings_from(st[a:b], got_heap, HEAP_BASE)
        assert got == cs[k][2], k
    t.close()


def test_subrange_heap_starts_at_the_range():
    cs, names, img = _table(9 * 2048 + 5, 2, only=("comment", "long", "value_i", "instruct"))
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    for c, k in enumerate(names):
        exp_rows, exp_heap = o.decode_strings(c, 3, 4, t.vec_per_rowgroup, HEAP_BASE)
        got_rows, got_heap, hb = gpu_strings(t, c, 3, 4)
        assert hb == o.string_heap_bytes(c, 3, 4)
        assert got_rows.tobytes() == exp_rows.tobytes(), k
    t.close()


def test_dictionary_form_matches_oracle():
    import torch
    cs = str_cases.cases(3000)
    img = flsgpu.write_table([("s", "VARCHAR", "DICT_STR"), ("f", "VARCHAR", "FSST_DICT"), ("c", "VARCHAR", "FSST"),
                              ("g", "VARCHAR", "FSST12_DICT")],
                             [cs["value_i"][2], cs["instruct"][2], cs["comment"][2], cs["instruct12"][2]], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    for col, src in ((0, cs["value_i"][2]), (1, cs["instruct"][2]), (3, cs["instruct12"][2])):
        codes = torch.zeros((t.out_elems(t.n_rowgroups),), dtype=torch.int32, device="cuda")
        t.decode_string_codes(col, codes.data_ptr(), codes.numel() * 4)
        flsgpu.sync()
        codes = codes.cpu().numpy().view(np.uint32)
        # all dictionaries of the column in one call: keys back to back, dictionary bytes in the chunks' heap slices
        n_keys, hb = t.string_dict_keys(col), t.string_heap_bytes(col)
        keys = torch.zeros((n_keys * 16,), dtype=torch.uint8, device="cuda")
        heap = torch.zeros((max(16, hb),), dtype=torch.uint8, device="cuda")
        t.decode_string_dict(col, keys.data_ptr(), keys.numel(), heap.data_ptr(), heap.numel(), HEAP_BASE)
        flsgpu.sync()
        all_keys = keys.cpu().numpy().reshape(-1, 16)
        all_heap = heap.cpu().numpy()
        row0 = k0 = h0 = 0
        for rg in range(t.n_rowgroups):
            ch = t.chunk(rg, col)
            exp_keys, exp_heap = o.decode_string_dict(col, rg, ch.aux_count, ch.const_val, HEAP_BASE + h0)
            got_keys = all_keys[k0:k0 + ch.aux_count]
            got_heap = all_heap[h0:h0 + ch.const_val]
            assert got_keys.tobytes() == exp_keys.tobytes()
            assert got_heap.tobytes() == exp_heap[:ch.const_val].tobytes()
            # single row group form: same keys, heap starting at this chunk
            k1 = torch.zeros((ch.aux_count * 16,), dtype=torch.uint8, device="cuda")
            h1 = torch.zeros((max(16, (ch.const_val + 15) // 16 * 16),), dtype=torch.uint8, device="cuda")
            t.decode_string_dict(col, k1.data_ptr(), k1.numel(), h1.data_ptr(), h1.numel(), HEAP_BASE + h0, rg, 1)
            flsgpu.sync()
            assert k1.cpu().numpy().tobytes() == exp_keys.tobytes()
            for v in range(ch.n_vec):
                a = (rg * t.vec_per_rowgroup + v) * 1024
                assert codes[a:a + 1024].tolist() == o.decode_codes(rg, col, v).tolist()
            n = t.rowgroup(rg).n_tuples
            kk = flsgpu.strings_from(got_keys.copy().view(flsgpu.STRING_T).reshape(-1), got_heap, HEAP_BASE + h0)
            a = rg * t.vec_per_rowgroup * 1024
            assert [kk[c] for c in codes[a:a + n]] == src[row0:row0 + n]
            row0 += n
            k0 += ch.aux_count
            h0 += (ch.const_val + 15) // 16 * 16
    # chunks that are not dictionary encoded have no dictionary form
    buf = torch.zeros((t.out_elems(t.n_rowgroups),), dtype=torch.int32, device="cuda")
    with pytest.raises(flsgpu.FlsError):
        t.decode_string_codes(2, buf.data_ptr(), buf.numel() * 4)
    # and a numeric decode of a string column is refused
    with pytest.raises(flsgpu.FlsError):
        t.decode_column(0, buf.data_ptr(), buf.numel() * 4)
    t.close()


def test_scan_strings_to_host_pointers_are_host_addresses():
    """End to end (stage + decode + D2H): the string_t pointers point INTO the caller's heap buffer."""
    import ctypes as C
    n = 5 * 65536 + 1234
    vals = str_cases.comment_like(n, seed=3)
    img = flsgpu.write_table([("s", "VARCHAR", None)], [vals])
    t = flsgpu.Table(img)
    n_rows = t.out_elems(t.n_rowgroups)
    hb = t.string_heap_bytes(0)
    rows = np.zeros((n_rows, 16), dtype=np.uint8)
    heap = np.zeros(max(16, hb), dtype=np.uint8)
    t.scan_strings_to_host(0, rows.ctypes.data, rows.nbytes, heap.ctypes.data, heap.nbytes)
    st = rows.view(flsgpu.STRING_T).reshape(-1)
    got = []
    for a, b in t.valid_mask_slices():
        for s in st[a:b]:
            ln = int(s["length"])
            if ln <= 12:
                got.append(s["body"][:ln].tobytes())
            else:
                ptr = int(s["body"][4:].view("<u8")[0])
                assert heap.ctypes.data <= ptr and ptr + ln <= heap.ctypes.data + heap.nbytes
                got.append(C.string_at(ptr, ln))  # dereference the pointer as DuckDB would
    assert got == vals
    t.close()


def test_long_strings_and_wide_dictionaries():
    r = np.random.default_rng(5)
    n = 4096
    # one very long string per vector (a lane decodes it alone), keys > 65536 distinct -> 32-bit code lanes
    longs = [bytes(r.integers(97, 123, 20000 if i % 1024 == 7 else int(r.integers(0, 20)), dtype=np.uint8)) for i in range(n)]
    n2 = 70000
    wide = [b"key-%06d" % (i * 7919 % 66000) for i in range(n2)]
    for spec, vals in ((("l", "VARCHAR", "FSST"), longs), (("w", "VARCHAR", "DICT_STR"), wide), (("w2", "VARCHAR", "FSST_DICT"), wide)):
        img = flsgpu.write_table([spec], [vals], vec_per_rowgroup=128)
        t = flsgpu.Table(img)
        t.stage()
        o = oracle_lib.Oracle(img)
        exp_rows, exp_heap = o.decode_strings(0, 0, t.n_rowgroups, t.vec_per_rowgroup, HEAP_BASE)
        got_rows, got_heap, _ = gpu_strings(t, 0)
        for rg in range(t.n_rowgroups):
            a = rg * t.vec_per_rowgroup * 1024
            b = a + t.rowgroup(rg).n_vec * 1024
            assert got_rows[a:b].tobytes() == exp_rows[a:b].tobytes(), spec[0]
        st = got_rows.copy().view(flsgpu.STRING_T).reshape(-1)
        got = []
        for a, b in t.valid_mask_slices():
            got += flsgpu.strings_from(st[a:b], got_heap, HEAP_BASE)
        assert got == vals
        t.close()


def test_projection_scan_mixes_numeric_flat_and_dictionary_strings():
    """flsgpu_scan_projection_to_host: one pipelined scan returns an integer column, a flat FSST string column and a
    dictionary-form string column; everything equals the oracle / the input."""
    import ctypes as C
    n = 3 * 65536 + 999
    ids = np.arange(n, dtype=np.int64) * 3
    comments = str_cases.comment_like(n, seed=9)
    modes = [[b"AIR", b"RAIL", b"TRUCK", b"REG AIR", b"SHIP", b"FOB", b"MAIL"][i % 7] for i in range(n)]
    img = flsgpu.write_table([("id", "INT64", None), ("c", "VARCHAR", "FSST"), ("m", "VARCHAR", "DICT_STR")], [ids, comments, modes])
    t = flsgpu.Table(img)
    rows = t.out_elems(t.n_rowgroups)
    out_id = np.zeros(rows, dtype=np.int64)
    out_c = np.zeros((rows, 16), dtype=np.uint8)
    heap_c = np.zeros(max(16, t.string_heap_bytes(1)), dtype=np.uint8)
    out_m = np.zeros(rows, dtype=np.uint32)
    keys_m = np.zeros((t.string_dict_keys(2), 16), dtype=np.uint8)
    heap_m = np.zeros(max(16, t.string_heap_bytes(2)), dtype=np.uint8)
    cols = [
        flsgpu.ScanColumn(col=0, out_host=out_id.ctypes.data, out_capacity_bytes=out_id.nbytes),
        flsgpu.ScanColumn(col=1, string_form=flsgpu.STRING_FLAT, out_host=out_c.ctypes.data, out_capacity_bytes=out_c.nbytes,
                          heap_host=heap_c.ctypes.data, heap_capacity_bytes=heap_c.nbytes),
        flsgpu.ScanColumn(col=2, string_form=flsgpu.STRING_DICTIONARY, out_host=out_m.ctypes.data, out_capacity_bytes=out_m.nbytes,
                          heap_host=heap_m.ctypes.data, heap_capacity_bytes=heap_m.nbytes, keys_host=keys_m.ctypes.data,
                          keys_capacity_bytes=keys_m.nbytes),
    ]
    t.scan_projection_to_host(cols)
    sl = t.valid_mask_slices()
    assert np.concatenate([out_id[a:b] for a, b in sl]).tobytes() == ids.tobytes()
    st = out_c.view(flsgpu.STRING_T).reshape(-1)
    got = []
    for a, b in sl:
        got += flsgpu.strings_from(st[a:b], heap_c, heap_c.ctypes.data)
    assert got == comments
    km = keys_m.view(flsgpu.STRING_T).reshape(-1)
    k0, got = 0, []
    for rg, (a, b) in enumerate(sl):
        nk = t.chunk(rg, 2).aux_count
        kk = flsgpu.strings_from(km[k0:k0 + nk], heap_m, heap_m.ctypes.data)
        got += [kk[c] for c in out_m[a:b]]
        k0 += nk
    assert got == modes
    # numeric-only entry point refuses string columns
    with pytest.raises(flsgpu.FlsError):
        t.scan_columns_to_host([1], [out_c.ctypes.data], [out_c.nbytes])
    t.close()

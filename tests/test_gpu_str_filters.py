"""GPU: filters on VARCHAR / BLOB columns through the C ABI (flsgpu_filter_mask: slabs decoded to string_t rows in scratch
memory, compared on the device) == oracle == plain Python, combined with numeric filters, with pruned row groups, through the
host-to-host filtered scan and through the reader mirror."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from str_filter_cases import COLUMNS, filter_sets, python_mask, table

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,vpr,slab_mb,dict_path", [(3000, 2, None, "1"), (20000, 8, None, "1"), (70000, 4, "1", "0"), (20000, 8, None, "0")])
def test_string_masks_match_oracle(n, vpr, slab_mb, dict_path, monkeypatch):
    import torch
    if slab_mb:
        monkeypatch.setenv("FLSGPU_STR_FILTER_SLAB_MB", slab_mb)  # many slabs per pass
    # dictionary-encoded ranges: "1" = comparison per key + flag lookup per row, "0" = the generic decode-and-compare path
    monkeypatch.setenv("FLSGPU_STR_FILTER_DICT", dict_path)
    specs, data = table(n)
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=vpr)
    names = [s[0] for s in specs]
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    n_words = t.n_rowgroups * t.vec_per_rowgroup * 32
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    vi = names.index("v")
    for fname, fs in filter_sets(data).items():
        filt = [(names.index(c), flsgpu.CMP[op], k) for c, op, k in fs]
        exp = o.filter_mask(filt, 0, t.n_rowgroups, t.vec_per_rowgroup)
        want = python_mask(specs, data, fs)
        assert int(np.unpackbits(exp.view(np.uint8)).sum()) == int(want.sum()), fname
        mask = torch.full((n_words,), -1, dtype=torch.int32, device="cuda")
        t.filter_mask(filt, mask.data_ptr(), n_words * 4)
        flsgpu.sync()
        got = mask.cpu().numpy().view(np.uint32)
        assert got.tobytes() == exp.tobytes(), fname
        # SELECT SUM(v), COUNT(*) WHERE <string filters>
        flsgpu.agg_reset(res.data_ptr())
        t.aggregate_column_masked(vi, mask.data_ptr(), res.data_ptr())
        flsgpu.sync()
        r = flsgpu.AggResult.from_buffer_copy(res.cpu().numpy().tobytes())
        assert r.count == int(want.sum()) and r.sum_lo == int(np.asarray(data[vi])[want].sum()), fname
    # pruned row groups select nothing, whatever the strings say
    keep_host = (np.arange(t.n_rowgroups) % 2).astype(np.uint8)
    keep = torch.from_numpy(keep_host).cuda()
    filt = [(names.index("shipmode"), flsgpu.CMP["NE"], b"AIR")]
    exp = o.filter_mask(filt, 0, t.n_rowgroups, t.vec_per_rowgroup).reshape(t.n_rowgroups, -1).copy()
    exp[keep_host == 0] = 0
    mask = torch.full((n_words,), -1, dtype=torch.int32, device="cuda")
    t.filter_mask(filt, mask.data_ptr(), n_words * 4, keep.data_ptr())
    flsgpu.sync()
    assert mask.cpu().numpy().view(np.uint32).tobytes() == exp.tobytes()
    # a sub-range of row groups
    if t.n_rowgroups > 3:
        filt = [(names.index("comment"), flsgpu.CMP["GE"], b"furiously"), (names.index("blob"), flsgpu.CMP["LT"], b"\x80")]
        exp = o.filter_mask(filt, 1, t.n_rowgroups - 2, t.vec_per_rowgroup)
        mask = torch.full((exp.size,), -1, dtype=torch.int32, device="cuda")
        t.filter_mask(filt, mask.data_ptr(), exp.size * 4, 0, rg_begin=1, n_rg=t.n_rowgroups - 2)
        flsgpu.sync()
        assert mask.cpu().numpy().view(np.uint32).tobytes() == exp.tobytes()
    t.close()


@pytest.mark.parametrize("threads", [1, 3])
def test_string_filters_through_reader(tmp_path, threads):
    """WHERE l_shipmode = 'RAIL' AND ... pushed into the scan (read_fls.cpp:39-42 filter_pushdown = true; the reference
    evaluates them in FilterExecutor::Apply after the decode): chunks come back sliced; result == plain Python."""
    import reader_lib
    n = 9 * 4096 + 55
    specs, data = table(n)
    names = [s[0] for s in specs]
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=4)
    p = str(tmp_path / "sf.fls")
    img.tofile(p)
    r = reader_lib.Reader(p, rowgroups_per_batch=2)
    proj = [names.index("v"), names.index("shipmode"), names.index("comment"), names.index("blob")]
    sets = filter_sets(data)
    for fname in ("dict_eq", "dict_range", "fsst_ge", "mixed", "two_strings", "blob_lt_empty", "blob_gt_ff"):
        fs = sets[fname]
        # filters address the projection; project the filter columns that are not in it
        p2 = list(proj)
        for c, _, _ in fs:
            if names.index(c) not in p2:
                p2.append(names.index(c))
        filt = [(p2.index(names.index(c)), flsgpu.CMP[op], k) for c, op, k in fs]
        got, row_ids, st = r.scan(p2, filt, n_threads=threads, stride=128)
        want = python_mask(specs, data, fs)
        idx = np.nonzero(want)[0]
        assert st["rows"] == len(idx), fname
        assert np.array_equal(row_ids, idx), fname
        assert got[0].tobytes() == np.asarray(data[names.index("v")])[want].tobytes(), fname
        for k, col in ((1, "shipmode"), (2, "comment"), (3, "blob")):
            src = data[names.index(col)]
            assert got[k] == [src[i] for i in idx], (fname, col)
    r.close()


@pytest.mark.parametrize("n,vpr,slab_mb", [(3000, 2, None), (9 * 4096 + 55, 4, "1"), (200000, 16, None)])
def test_selected_rows_scan_matches_python(n, vpr, slab_mb, monkeypatch):
    """flsgpu_scan_selected_to_host: the filter's selection applied on the device (fls_select_plan / fls_select_compact kernels),
    only the rows that pass come back, packed in row order: == plain Python over the input, for numeric columns, flat strings of
    every encoding (dictionary encoded ones included) and the row index."""
    if slab_mb:
        monkeypatch.setenv("FLSGPU_SLAB_MB", slab_mb)  # several slabs: both streams, offsets carried from slab to slab
    specs, data = table(n)
    names = [s[0] for s in specs]
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=vpr)
    t = flsgpu.Table(img)  # not staged by hand
    rows = t.out_elems(t.n_rowgroups)
    proj = ["v", "shipmode", "comment", "blob", "flag_runs", "freq_none"]
    sets = filter_sets(data)
    for fname in ("mixed", "dict_eq", "fsst_ge", "blob_lt_empty", "blob_ge_empty", "two_strings", "rle_str_eq_long", "frequency_str_ne"):
        fs = sets[fname]
        filt = [(names.index(c), flsgpu.CMP[op], k) for c, op, k in fs]
        bufs, cols = [], []
        for name in proj:
            c = names.index(name)
            if t.column(c).phys_bits == 128:
                out = np.zeros(rows * 16, dtype=np.uint8)
                heap = np.zeros(max(16, t.string_heap_bytes(c)), dtype=np.uint8)
                cols.append(flsgpu.ScanColumn(col=c, string_form=flsgpu.STRING_FLAT, out_host=out.ctypes.data, out_capacity_bytes=out.nbytes,
                                              heap_host=heap.ctypes.data, heap_capacity_bytes=heap.nbytes))
                bufs.append((out, heap))
            else:
                out = np.zeros(rows, dtype=np.int64)
                cols.append(flsgpu.ScanColumn(col=c, out_host=out.ctypes.data, out_capacity_bytes=out.nbytes))
                bufs.append((out, None))
        # dictionary form of dictionary-encoded columns: packed entries into ONE dictionary over all row groups of the range
        dict_bufs = []
        for name in ("shipmode", "instruct"):
            c = names.index(name)
            codes = np.zeros(rows, dtype=np.uint32)
            keys = np.zeros(max(1, t.string_dict_keys(c)) * 16, dtype=np.uint8)
            heap = np.zeros(max(16, t.string_heap_bytes(c)), dtype=np.uint8)
            cols.append(flsgpu.ScanColumn(col=c, string_form=flsgpu.STRING_DICTIONARY, out_host=codes.ctypes.data, out_capacity_bytes=codes.nbytes,
                                          heap_host=heap.ctypes.data, heap_capacity_bytes=heap.nbytes, keys_host=keys.ctypes.data,
                                          keys_capacity_bytes=keys.nbytes))
            dict_bufs.append((name, codes, keys, heap))
        row_index = np.full(rows, 0xFFFFFFFF, dtype=np.uint32)
        n_sel = t.scan_selected_to_host(cols, filt, row_index.ctypes.data, row_index.nbytes)
        want = python_mask(specs, data, fs)
        idx = np.nonzero(want)[0]
        assert n_sel == len(idx), fname
        # row index: position inside the padded range layout -> row number of the table
        stride = t.vec_per_rowgroup * 1024
        got_rows = (row_index[:n_sel] // stride).astype(np.int64) * stride + row_index[:n_sel] % stride
        assert np.array_equal(got_rows, idx), fname
        for name, (out, heap) in zip(proj, bufs):
            src = data[names.index(name)]
            if heap is None:
                assert out[:n_sel].tobytes() == np.asarray(src)[want].tobytes(), (fname, name)
            else:
                st = out[:n_sel * 16].view(flsgpu.STRING_T).reshape(-1)
                assert flsgpu.strings_from(st, heap, heap.ctypes.data) == [bytes(src[i]) for i in idx], (fname, name)
        for name, codes, keys, heap in dict_bufs:
            src = data[names.index(name)]
            all_keys = flsgpu.strings_from(keys.view(flsgpu.STRING_T).reshape(-1), heap, heap.ctypes.data)
            assert [all_keys[k] for k in codes[:n_sel]] == [bytes(src[i]) for i in idx], (fname, name, "dictionary form")
    with pytest.raises(flsgpu.FlsError):
        t.scan_selected_to_host(cols, [])
    with pytest.raises(flsgpu.FlsError):  # dictionary form of a column that is not dictionary encoded
        c = names.index("comment")
        t.scan_selected_to_host([flsgpu.ScanColumn(col=c, string_form=flsgpu.STRING_DICTIONARY, out_host=codes.ctypes.data,
                                                   out_capacity_bytes=codes.nbytes, heap_host=heap.ctypes.data, heap_capacity_bytes=1 << 40,
                                                   keys_host=keys.ctypes.data, keys_capacity_bytes=keys.nbytes)], filt)
    t.close()

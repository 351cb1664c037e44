"""GPU: EXP_EQUAL chunks (a column stored as a reference to another column of the row group) through the C ABI and the reader
mirror: decode / aggregate / filter the referencing column alone (only its own chunks staged) == oracle == input."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from equal_cases import table

pytestmark = pytest.mark.gpu


def test_equal_columns_decode_aggregate_filter():
    import torch
    vpr, n = 4, 6 * 4096 + 77
    specs, data = table(n, vpr * 1024)
    names = [s[0] for s in specs]
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=vpr)
    o = oracle_lib.Oracle(img)
    # host-to-host scan of the referencing column alone: only its own chunks (= the shared bytes) are asked for
    t = flsgpu.Table(img)
    bi = names.index("b")
    host = np.zeros(t.out_elems(t.n_rowgroups), dtype=np.int64)
    t.scan_column_to_host(bi, host.ctypes.data, host.nbytes)
    assert np.concatenate([host[a:b] for a, b in t.valid_mask_slices()]).tobytes() == data[bi].tobytes()
    t.close()
    for only in (["h", "e"], names):
        t = flsgpu.Table(img)
        cols = [names.index(c) for c in only]
        t.stage()
        for c in cols:
            src = data[c]
            nbytes = t.out_elems(t.n_rowgroups) * src.dtype.itemsize
            out = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
            t.decode_column(c, out.data_ptr(), nbytes)
            flsgpu.sync()
            got = out.cpu().numpy().view(src.dtype)
            exp = o.decode_column(c, 0, t.n_rowgroups, 0, src.dtype, vpr)
            for a, b in t.valid_mask_slices():
                assert got[a:b].tobytes() == exp[a:b].tobytes(), names[c]
            flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
            assert flat.tobytes() == src.tobytes(), names[c]
        t.close()
    # SUM(b) WHERE b >= k  (b's chunks are references to a's)
    t = flsgpu.Table(img)
    t.stage()
    k = int(data[bi][n // 3])
    filt = [(bi, flsgpu.CMP["GE"], k)]
    n_words = t.n_rowgroups * vpr * 32
    mask = torch.zeros(n_words, dtype=torch.int32, device="cuda")
    t.filter_mask(filt, mask.data_ptr(), n_words * 4)
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    flsgpu.agg_reset(res.data_ptr())
    t.aggregate_column_masked(bi, mask.data_ptr(), res.data_ptr())
    flsgpu.sync()
    r = flsgpu.AggResult.from_buffer_copy(res.cpu().numpy().tobytes())
    sel = data[bi] >= k
    assert r.count == int(sel.sum()) and r.sum_lo == int(data[bi][sel].sum())
    assert mask.cpu().numpy().view(np.uint32).tobytes() == o.filter_mask(filt, 0, t.n_rowgroups, vpr).tobytes()
    t.close()


def test_equal_columns_through_reader(tmp_path):
    import reader_lib
    vpr, n = 4, 5 * 4096 + 9
    specs, data = table(n, vpr * 1024)
    names = [s[0] for s in specs]
    p = str(tmp_path / "eq.fls")
    flsgpu.write_table(specs, data, vec_per_rowgroup=vpr).tofile(p)
    r = reader_lib.Reader(p, rowgroups_per_batch=2)
    got, _, st = r.scan([names.index("h"), names.index("c"), names.index("b")], n_threads=2)
    assert st["rows"] == n
    for g, c in zip(got, ("h", "c", "b")):
        assert g.tobytes() == data[names.index(c)].tobytes(), c
    r.close()

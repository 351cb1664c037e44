"""GPU, BASELINE configs 3/5 shape: every column of a TPC-H lineitem-shaped table decoded by the CUDA path == oracle, bit
for bit, and == the generated input; Q1 / Q6 style SUM/COUNT scans (filter mask -> masked fused aggregates) == numpy."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from filter_cases import bits
from lineitem_cases import lineitem, specs, string_list
from test_gpu_parity import gpu_decode
from test_gpu_strings import HEAP_BASE, gpu_strings
from test_lineitem_oracle import q1_q6

pytestmark = pytest.mark.gpu


def test_lineitem_columns_and_queries():
    import torch
    n = 5 * 65536 + 999
    cols = lineitem(n, sf=0.1, seed=21)
    names = [c[0] for c in cols]
    img = flsgpu.write_table(specs(cols), [c[3] for c in cols], vec_per_rowgroup=64)
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    for i, c in enumerate(cols):
        if c[1] == "VARCHAR":
            exp_rows, exp_heap = o.decode_strings(i, 0, t.n_rowgroups, 64, HEAP_BASE)
            got_rows, got_heap, hb = gpu_strings(t, i)
            for a, b in t.valid_mask_slices():
                assert got_rows[a:b].tobytes() == exp_rows[a:b].tobytes(), c[0]
            st = got_rows.copy().view(flsgpu.STRING_T).reshape(-1)
            got = []
            for a, b in t.valid_mask_slices():
                got += flsgpu.strings_from(st[a:b], got_heap, HEAP_BASE)
            assert got == string_list(c[3]), c[0]
        else:
            dt = np.dtype(t.native_dtype(i))
            exp = o.decode_column(i, 0, t.n_rowgroups, 0, dt, 64)
            got = gpu_decode(t, i, 0, dt)
            for a, b in t.valid_mask_slices():
                assert got[a:b].tobytes() == exp[a:b].tobytes(), c[0]
            flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
            assert flat.tobytes() == np.ascontiguousarray(c[3]).astype(dt).tobytes(), c[0]
    by, q1, q6 = q1_q6(cols)
    sd, di, qu = names.index("l_shipdate"), names.index("l_discount"), names.index("l_quantity")
    f_q1 = [(sd, flsgpu.CMP["LE"], bits(np.int32(10471), np.int32))]
    f_q6 = [(sd, flsgpu.CMP["GE"], bits(np.int32(8766), np.int32)), (sd, flsgpu.CMP["LT"], bits(np.int32(9131), np.int32)),
            (di, flsgpu.CMP["GE"], bits(np.int64(5), np.int64)), (di, flsgpu.CMP["LE"], bits(np.int64(7), np.int64)),
            (qu, flsgpu.CMP["LT"], bits(np.int64(2400), np.int64))]
    n_words = t.n_rowgroups * 64 * 32
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    for filt, want in ((f_q1, q1), (f_q6, q6)):
        mask = torch.zeros(n_words, dtype=torch.int32, device="cuda")
        t.filter_mask(filt, mask.data_ptr(), n_words * 4, 0)
        flsgpu.sync()
        assert mask.cpu().numpy().view(np.uint32).tobytes() == o.filter_mask(filt, 0, t.n_rowgroups, 64).tobytes()
        for col in ("l_quantity", "l_extendedprice", "l_discount", "l_tax"):
            flsgpu.agg_reset(res.data_ptr())
            t.aggregate_column_masked(names.index(col), mask.data_ptr(), res.data_ptr(), 0)
            flsgpu.sync()
            r = flsgpu.AggResult.from_buffer_copy(res.cpu().numpy().tobytes())
            assert r.count == int(want.sum()) and ((r.sum_hi << 64) | r.sum_lo) == int(by[col][want].sum()), col
    t.close()

"""BASELINE configs 3/5 shape on CPU: a TPC-H lineitem-shaped table (tests/lineitem_cases.py) goes writer -> container ->
oracle and comes back value for value; the SUM/COUNT scans of Q1 / Q6 over the oracle's filter masks equal numpy."""
import numpy as np

import flsgpu
import oracle_lib
from filter_cases import bits
from lineitem_cases import lineitem, specs, string_list

HEAP_BASE = 0x7F0000001000


def q1_q6(cols):
    by = {c[0]: c[3] for c in cols}
    q1 = by["l_shipdate"] <= 10471  # 1998-12-01 - 90 days
    q6 = ((by["l_shipdate"] >= 8766) & (by["l_shipdate"] < 9131) & (by["l_discount"] >= 5) & (by["l_discount"] <= 7) &
          (by["l_quantity"] < 2400))
    return by, q1, q6


def test_lineitem_roundtrip_and_q1_q6_on_the_oracle():
    n = 2 * 65536 + 4321
    cols = lineitem(n, sf=0.05)
    names = [c[0] for c in cols]
    img = flsgpu.write_table(specs(cols), [c[3] for c in cols], vec_per_rowgroup=64)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    assert t.n_tuples == n and t.n_rowgroups == 3
    for i, c in enumerate(cols):
        if c[1] == "VARCHAR":
            rows, heap = o.decode_strings(i, 0, t.n_rowgroups, 64, HEAP_BASE)
            st = rows.copy().view(flsgpu.STRING_T).reshape(-1)
            got = []
            for a, b in t.valid_mask_slices():
                got += flsgpu.strings_from(st[a:b], heap, HEAP_BASE)
            assert got == string_list(c[3]), c[0]
        else:
            dt = np.dtype(t.native_dtype(i))
            out = o.decode_column(i, 0, t.n_rowgroups, 0, dt, 64)
            flat = np.concatenate([out[a:b] for a, b in t.valid_mask_slices()])
            assert flat.tobytes() == np.ascontiguousarray(c[3]).astype(dt).tobytes(), c[0]
    by, q1, q6 = q1_q6(cols)
    f_q1 = [(names.index("l_shipdate"), flsgpu.CMP["LE"], bits(np.int32(10471), np.int32))]
    f_q6 = [(names.index("l_shipdate"), flsgpu.CMP["GE"], bits(np.int32(8766), np.int32)),
            (names.index("l_shipdate"), flsgpu.CMP["LT"], bits(np.int32(9131), np.int32)),
            (names.index("l_discount"), flsgpu.CMP["GE"], bits(np.int64(5), np.int64)),
            (names.index("l_discount"), flsgpu.CMP["LE"], bits(np.int64(7), np.int64)),
            (names.index("l_quantity"), flsgpu.CMP["LT"], bits(np.int64(2400), np.int64))]
    for filt, want in ((f_q1, q1), (f_q6, q6)):
        mask = o.filter_mask(filt, 0, t.n_rowgroups, 64)
        for col in ("l_quantity", "l_extendedprice", "l_discount", "l_tax"):
            s, cnt, _ = o.aggregate_masked(names.index(col), 0, t.n_rowgroups, mask)
            assert cnt == int(want.sum()) and s == int(by[col][want].sum()), col
    t.close()

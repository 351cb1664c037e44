"""CPU: product encoder -> container -> oracle decode == input, bit for bit (SURVEY.md 8c "oracle definition").

This is the value-level round-trip contract of the reference's own tests (test/all_types_single_threaded.test:280-352)
applied per type AND per encoding.
"""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from fls_cases import column_spec, standard_cases, width_sweep_cases, INT_TYPES


def roundtrip(cases, vec_per_rg=2):
    cols = [column_spec(c) for c in cases]
    arrs = [c[3] for c in cases]
    img = flsgpu.write_table(cols, arrs, vec_per_rowgroup=vec_per_rg)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    for i, c in enumerate(cases):
        dt = t.native_dtype(i)
        out = o.decode_column(i, 0, t.n_rowgroups, 0, dt, t.vec_per_rowgroup)
        got = np.concatenate([out[a:b] for a, b in t.valid_mask_slices()])
        exp = np.ascontiguousarray(c[3]).view(got.dtype) if got.dtype.itemsize == c[3].dtype.itemsize else c[3].astype(dt)
        assert got.tobytes() == exp.tobytes(), c[0]
        if c[2] is not None:
            assert flsgpu.ENC_NAME[t.chunk(0, i).encoding] == c[2], c[0]
    t.close()


def test_standard_cases_roundtrip():
    roundtrip(standard_cases())


@pytest.mark.parametrize("tname", list(INT_TYPES))
def test_width_sweep_roundtrip(tname):
    roundtrip(width_sweep_cases(tname), vec_per_rg=64)


def test_exact_vector_multiple_and_single_row():
    for n in (1, 1023, 1024, 1025, 2048):
        roundtrip(standard_cases(n=n, seed=n), vec_per_rg=1)


def test_widening_matches_numpy_static_cast():
    # materializer CastNumeric (reference fls_type_resolver.hpp:22-33): static_cast from the column's own signedness
    n = 2048 + 77
    r = np.random.default_rng(5)
    src = r.integers(-128, 127, n, dtype=np.int8, endpoint=True)
    usrc = r.integers(0, 255, n, dtype=np.uint8, endpoint=True)
    img = flsgpu.write_table([("s", "INT8", None), ("u", "UINT8", None)], [src, usrc], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    for out_name, dt in (("I16", np.int16), ("I32", np.int32), ("I64", np.int64), ("U64", np.uint64)):
        for col, a in ((0, src), (1, usrc)):
            out = o.decode_column(col, 0, t.n_rowgroups, flsgpu.OUT[out_name], dt, t.vec_per_rowgroup)
            got = np.concatenate([out[x:y] for x, y in t.valid_mask_slices()])
            assert np.array_equal(got, a.astype(dt)), (out_name, col)

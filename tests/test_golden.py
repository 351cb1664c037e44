"""CPU: pins the oracle.

1. ALP / ALP-RD arithmetic against golden vectors generated from the reference tree's own implementation (DuckDB's
   vendored alp.hpp / alprd.hpp, compiled by oracle/Makefile into oracle/_ref; generator: oracle/make_golden.py).
2. The reference's own test tables (test/all_types_single_threaded.test) round-trip through writer + oracle, at the
   reference's 1024 rows and at BASELINE config 1's 1 000 000 rows.
3. FSST: the oracle's decoder against code streams produced by the reference tree's own FSST compressor
   (duckdb/third_party/fsst, oracle/_ref/libfsstref.so; generator: oracle/make_golden_fsst.py), and the product encoder's
   streams against what the reference's fsst_decompress makes of them.
Byte-level parity with the reference's .fls files stays UNPINNED (third-party FastLanes library absent).
"""
import ctypes as C
import os

import numpy as np
import pytest

import flsgpu
import oracle_lib
from ref_fixtures import reference_string_tables, reference_test_tables

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "alp_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def test_alp_decode_f64_matches_reference_vectors(golden):
    got = np.array([oracle_lib.lib.fls_oracle_alp_decode_f64(int(d), int(e), int(f))
                    for d, e, f in zip(golden["f64_digits"], golden["f64_e"], golden["f64_f"])], dtype=np.float64)
    assert got.view(np.uint64).tolist() == golden["f64_bits"].tolist()


def test_alp_decode_f32_matches_reference_vectors(golden):
    got = np.array([oracle_lib.lib.fls_oracle_alp_decode_f32(int(d), int(e), int(f))
                    for d, e, f in zip(golden["f32_digits"], golden["f32_e"], golden["f32_f"])], dtype=np.float32)
    assert got.view(np.uint32).tolist() == golden["f32_bits"].tolist()


@pytest.mark.parametrize("bits", [64, 32])
def test_alprd_container_matches_reference_glue(golden, bits):
    img = golden[f"rd{bits}_image"]
    udt = np.uint64 if bits == 64 else np.uint32
    o = oracle_lib.Oracle(img)
    out = o.decode_column(0, 0, 1, 0, udt, 2)
    assert out.tobytes() == golden[f"rd{bits}_expected"].tobytes()


def test_writer_exceptions_agree_with_reference_classification(golden):
    """Values the reference calls impossible to encode (NaN, inf, -0.0, out of range; alp.hpp:100-103) must come back
    bit-exact through an ALP chunk, i.e. be carried as exceptions."""
    vals = golden["enc64_values"]
    assert golden["enc64_impossible"][:7].tolist() == [0, 1, 1, 1, 1, 1, 1]
    img = flsgpu.write_table([("v", "DOUBLE", "ALP")], [vals], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    out = oracle_lib.Oracle(img).decode_column(0, 0, t.n_rowgroups, 0, np.float64, 2)
    assert out[: len(vals)].tobytes() == vals.tobytes()
    n_exc = sum(t.chunk(r, 0).aux_count for r in range(t.n_rowgroups))
    assert n_exc >= int(golden["enc64_impossible"].sum())
    t.close()


def test_golden_file_is_not_stale():
    ref = os.path.join(ROOT, "oracle", "_ref", "libalpref.so")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref not built here (needs /root/reference)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "oracle", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    fresh = mg.generate()
    committed = np.load(GOLDEN)
    assert sorted(fresh) == sorted(committed.files)
    for k in fresh:
        assert fresh[k].tobytes() == committed[k].tobytes(), k


FSST_GOLDEN = os.path.join(ROOT, "tests", "golden", "fsst_golden.npz")


@pytest.mark.parametrize("corpus", ["comment", "blob", "value"])
def test_fsst_oracle_decodes_reference_streams(corpus):
    g = np.load(FSST_GOLDEN)
    off, data = g[f"ref_{corpus}_offsets"], g[f"ref_{corpus}_bytes"]
    dec, ends = oracle_lib.fsst_decode_stream(g[f"ref_{corpus}_symbol"], g[f"ref_{corpus}_len"], g[f"ref_{corpus}_codes"],
                                              g[f"ref_{corpus}_comp_end"], int(off[-1]))
    assert dec.tobytes() == data.tobytes()
    assert ends.tolist() == off[1:].tolist()


@pytest.mark.parametrize("corpus", ["comment", "blob", "value"])
def test_fsst_product_streams_decode_like_the_reference(corpus):
    """The reference decoder's output over the product encoder's table + streams (stored in the golden file) equals what
    the oracle makes of the same streams."""
    g = np.load(FSST_GOLDEN)
    want, want_ends = g[f"prod_{corpus}_ref_decoded"], g[f"prod_{corpus}_ref_ends"]
    dec, ends = oracle_lib.fsst_decode_stream(g[f"prod_{corpus}_symbol"], g[f"prod_{corpus}_len"], g[f"prod_{corpus}_codes"],
                                              g[f"prod_{corpus}_comp_end"], int(want_ends[-1]))
    assert dec.tobytes() == want.tobytes() and ends.tolist() == want_ends.tolist()


def test_fsst_golden_file_is_not_stale():
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libfsstref.so")):
        pytest.skip("oracle/_ref not built here (needs /root/reference)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_fsst", os.path.join(ROOT, "oracle", "make_golden_fsst.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    fresh = mg.generate()
    committed = np.load(FSST_GOLDEN)
    assert sorted(fresh) == sorted(committed.files)
    for k in fresh:
        assert fresh[k].tobytes() == committed[k].tobytes(), k


@pytest.mark.parametrize("n", [1024, 1_000_000])
def test_reference_test_tables_roundtrip(n):
    """compare_roundtrip of the reference test (:7-34, :280-352) must return zero rows: ids and values identical."""
    tables = reference_test_tables(n)
    ids = np.arange(1, n + 1, dtype=np.int64)
    for name, spec in tables.items():
        tname, col = spec[0], spec[1]
        img = flsgpu.write_table([("id", "INT64", None), ("col", tname, None)], [ids, col])
        t = flsgpu.Table(img)
        assert t.n_tuples == n and t.n_rowgroups == (n + 65535) // 65536
        o = oracle_lib.Oracle(img)
        for c, src in ((0, ids), (1, col)):
            dt = t.native_dtype(c)
            out = o.decode_column(c, 0, t.n_rowgroups, 0, dt, t.vec_per_rowgroup, n_threads=4)
            got = np.concatenate([out[a:b] for a, b in t.valid_mask_slices()])
            assert got.tobytes() == np.ascontiguousarray(src).tobytes(), (name, c)
        t.close()


@pytest.mark.parametrize("n", [1024, 100_000])
def test_reference_string_tables_roundtrip(n):
    """VARCHAR, BLOB, HUGEINT and UHUGEINT (written as decimal strings) tables of the reference test (:78-90,113-118,155-160,
    :300-326,350-352): writer + oracle give back every value."""
    ids = np.arange(1, n + 1, dtype=np.int64)
    for name, (tname, vals) in reference_string_tables(n).items():
        img = flsgpu.write_table([("id", "INT64", None), ("col", tname, None)], [ids, vals])
        t = flsgpu.Table(img)
        o = oracle_lib.Oracle(img)
        st, heap = o.decode_strings(1, 0, t.n_rowgroups, t.vec_per_rowgroup)
        st = st.view(flsgpu.STRING_T).reshape(-1)
        got = []
        for a, b in t.valid_mask_slices():
            got += flsgpu.strings_from(st[a:b], heap)
        assert got == vals, name
        t.close()


def test_committed_container_still_decodes():
    """tests/golden/container_v1.npz (oracle/make_golden_container.py): a container written at format version 1 - every numeric
    encoding on INT32 / DOUBLE columns, one column per string encoding - with the rows it was written from.  Today's validator
    must accept the committed bytes and today's oracle must decode them to those rows: the on-disk format stays readable."""
    import flsgpu
    import oracle_lib
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "container_v1.npz"))
    img = np.ascontiguousarray(g["image"])
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    assert t.n_columns == int(g["n_columns"]) == 35 and t.vec_per_rowgroup == 1 and t.n_tuples == 1124
    seen = set()
    for c in range(t.n_columns):
        seen |= {flsgpu.ENC_NAME[t.chunk(r, c).encoding] for r in range(t.n_rowgroups)}
        if str(g["types"][c]) in ("VARCHAR", "BLOB"):
            off, data = g[f"off{c}"], g[f"col{c}"].tobytes()
            want = [data[int(off[i]):int(off[i + 1])] for i in range(len(off) - 1)]
            st, heap = o.decode_strings(c, 0, t.n_rowgroups, t.vec_per_rowgroup, 0x7F0000001000)
            st = st.view(flsgpu.STRING_T).reshape(-1)
            got = []
            for a, b in t.valid_mask_slices():
                got += flsgpu.strings_from(st[a:b], heap, 0x7F0000001000)
            assert got == want, str(g["names"][c])
        else:
            want = g[f"col{c}"]
            out = o.decode_column(c, 0, t.n_rowgroups, 0, want.dtype, t.vec_per_rowgroup)
            got = np.concatenate([out[a:b] for a, b in t.valid_mask_slices()])
            assert got.tobytes() == want.tobytes(), str(g["names"][c])
    t.close()
    assert {"UNCOMPRESSED", "CONSTANT", "FFOR", "DELTA", "DICT", "RLE", "ALP", "ALP_RD", "CROSS_RLE", "FREQUENCY", "SLPATCH", "DICT_STR",
            "STR_UNCOMPRESSED", "STR_CONSTANT", "FSST", "FSST_DICT", "FSST12", "FSST12_DICT", "RLE_STR", "CROSS_RLE_STR",
            "FREQUENCY_STR"} <= seen, seen

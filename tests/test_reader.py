"""The C++ reader mirror (csrc/reader: FastLanesReader::Initialize / TryInitializeScan / Scan ...) driven the way
DuckDB's MultiFileScan drives the reference's reader.  CPU tests cover bind-time behaviour; GPU tests are the
reference's round-trip SQL tests (test/all_types_{single,multi}_threaded.test) restated over the reader."""
import numpy as np
import pytest

import flsgpu
from ref_fixtures import reference_string_tables, reference_test_tables


def write(tmp_path, name, cols, arrs, vec_per_rg=64):
    w = flsgpu.Writer(cols, vec_per_rg)
    w.append(arrs)
    w.finish()
    p = str(tmp_path / name)
    w.save(p)
    w.close()
    return p


def test_bind_schema_and_type_translation(tmp_path):
    import reader_lib
    n = 3000
    i = np.arange(n)
    cols = [("a", "INT8", None), ("b", "UINT8", None), ("c", "UINT32", None), ("d", "DECIMAL", None, 10, 2), ("e", "DATE", None),
            ("f", "TIMESTAMP", None), ("g", "BOOLEAN", None), ("h", "FLOAT", None), ("k", "DECIMAL", None, 4, 1)]
    arrs = [i.astype(np.int8), i.astype(np.uint8), i.astype(np.uint32), i.astype(np.int64), i.astype(np.int32), i.astype(np.int64),
            (i % 2).astype(np.uint8), i.astype(np.float32), i.astype(np.int16)]
    r = reader_lib.Reader(write(tmp_path, "t.fls", cols, arrs, 2), file_row_number=True)
    # reference src/reader/translation_utils.cpp:19-24: every unsigned width reads back as UBIGINT
    assert [c["type"] for c in r.columns] == ["TINYINT", "UBIGINT", "UBIGINT", "DECIMAL", "DATE", "TIMESTAMP", "BOOLEAN", "FLOAT",
                                              "DECIMAL", "BIGINT"]
    assert r.columns[3]["width"] == 10 and r.columns[3]["scale"] == 2 and r.columns[3]["bytes"] == 8
    assert r.columns[8]["bytes"] == 2 and r.columns[-1]["name"] == "file_row_number"
    assert r.cardinality == n and r.total_tuples == n
    ps = r.partition_stats()
    assert ps[0] == (0, 2048) and ps[-1] == (2048, n - 2048) and sum(c for _, c in ps) == n
    st = r.statistics("a")
    assert st["has_stats"] and (st["min_i"], st["max_i"]) == (-128, 127)
    st = r.statistics("h")
    assert (st["min_d"], st["max_d"]) == (0.0, float(n - 1))
    assert not r.statistics("nope")["has_stats"]
    r.close()
    assert reader_lib.Reader(write(tmp_path, "u.fls", cols, arrs, 2), explicit_cardinality=77).cardinality == 77


def test_open_errors(tmp_path):
    import reader_lib
    with pytest.raises(reader_lib.ReaderError) as e:
        reader_lib.Reader(str(tmp_path / "missing.fls"))
    assert e.value.status == -2  # reference: std::runtime_error on I/O failure
    (tmp_path / "junk.fls").write_bytes(b"not a container" * 100)
    with pytest.raises(reader_lib.ReaderError) as e:
        reader_lib.Reader(str(tmp_path / "junk.fls"))
    assert e.value.status == -1


def test_scan_without_device_fails_loudly(tmp_path):
    import torch
    import reader_lib
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    p = write(tmp_path, "t.fls", [("a", "INT32", None)], [np.arange(5000, dtype=np.int32)], 2)
    r = reader_lib.Reader(p)
    with pytest.raises(reader_lib.ReaderError) as e:
        r.scan_all()
    assert e.value.status == -3


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [1, 2, 8])
@pytest.mark.parametrize("n", [1024, 200_000])
def test_reference_roundtrip_tables_through_reader(tmp_path, n, threads):
    """test/all_types_single_threaded.test and its multi-threaded twin: SELECT * FROM read_fls(...) equals the table."""
    import reader_lib
    ids = np.arange(1, n + 1, dtype=np.int64)
    for name, spec in reference_test_tables(n).items():
        p = write(tmp_path, f"{name}_{n}_{threads}.fls", [("id", "INT64", None), ("col", spec[0], None)], [ids, spec[1]], 8)
        r = reader_lib.Reader(p, rowgroups_per_batch=3)
        (got_id, got_col), st = r.scan_all(n_threads=threads)
        assert st["rows"] == n and st["max_chunk_rows"] <= 2048
        assert np.array_equal(got_id, ids), name
        exp = spec[1].astype(np.uint64) if spec[0].startswith("UINT") else spec[1]
        assert got_col.tobytes() == np.ascontiguousarray(exp).astype(got_col.dtype).tobytes(), name
        assert abs(r.progress - 100.0) < 1e-9
        r.close()


@pytest.mark.gpu
def test_projection_virtual_column_and_chunking(tmp_path):
    import reader_lib
    n = 5 * 8192 + 1500  # row groups of 8 vectors; the last one has a partial last vector
    r0 = np.random.default_rng(4)
    a = r0.integers(0, 200, n).astype(np.uint8)
    b = np.cumsum(r0.integers(0, 9, n)).astype(np.int64)
    c = np.round(r0.random(n) * 50, 1)
    p = write(tmp_path, "p.fls", [("a", "UINT8", None), ("b", "INT64", None), ("c", "DOUBLE", None)], [a, b, c], 8)
    r = reader_lib.Reader(p, file_row_number=True, rowgroups_per_batch=2)
    (frn, gc, ga, ga2), st = r.scan_all([3, 2, 0, 0], n_threads=3)  # virtual column + a duplicated projection
    assert np.array_equal(frn, np.arange(n)) and gc.tobytes() == c.tobytes()
    assert np.array_equal(ga, a.astype(np.uint64)) and np.array_equal(ga2, ga)
    # chunks never cross a row group: 5 full row groups x 4 chunks + ceil(1500/2048) = 21
    assert st["chunks"] == 5 * 4 + 1 and st["rows"] == n
    with pytest.raises(reader_lib.ReaderError):
        r.scan_all([9])
    r.close()


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [1, 4])
@pytest.mark.parametrize("n", [1024, 50_000])
def test_reference_string_tables_through_reader(tmp_path, n, threads):
    """varchar_table / blob_table / hugeint_table / uhugeint_table of the reference test: SELECT * FROM read_fls(...)."""
    import reader_lib
    ids = np.arange(1, n + 1, dtype=np.int64)
    for name, (tname, vals) in reference_string_tables(n).items():
        p = write(tmp_path, f"{name}_{n}_{threads}.fls", [("id", "INT64", None), ("col", tname, None)], [ids, vals], 8)
        r = reader_lib.Reader(p, rowgroups_per_batch=3)
        assert r.columns[1]["type"] == ("BLOB" if tname == "BLOB" else "VARCHAR")
        (got_id, got_col), _, st = r.scan(n_threads=threads)
        assert st["rows"] == n and np.array_equal(got_id, ids), name
        assert got_col == vals, name
        r.close()


@pytest.mark.gpu
def test_dictionary_vectors_and_mixed_encodings(tmp_path):
    """Dictionary-encoded string chunks reach the consumer as DICTIONARY_VECTORs (dictionary_kernel.hpp:80-101); a batch
    that mixes dictionary and FSST chunks is materialised flat; both give the same strings."""
    import reader_lib
    n = 6 * 8192
    low = [[b"AIR", b"RAIL", b"TRUCK", b"REG AIR", b"SHIP", b"FOB", b"MAIL"][i % 7] for i in range(n)]
    rng = np.random.default_rng(8)
    mixed = [b"k%d" % (i % 13) if (i // 8192) % 2 == 0 else b"unique-%d-%d" % (i, int(rng.integers(0, 1 << 30))) for i in range(n)]
    p = write(tmp_path, "d.fls", [("low", "VARCHAR", None), ("mixed", "VARCHAR", None)], [low, mixed], 8)
    t = flsgpu.Table(path=p)
    encs = {flsgpu.ENC_NAME[t.chunk(r, 1).encoding] for r in range(t.n_rowgroups)}
    assert "DICT_STR" in encs and len(encs) > 1
    t.close()
    r = reader_lib.Reader(p, rowgroups_per_batch=2)
    (got_low,), _, st = r.scan([0], n_threads=2)
    assert got_low == low and st["dictionary_vectors"] == st["chunks"] > 0
    (got_mixed,), _, st = r.scan([1], n_threads=2)
    assert got_mixed == mixed and st["dictionary_vectors"] == 0
    r2 = reader_lib.Reader(p, rowgroups_per_batch=1)  # one row group per batch: the dictionary row groups keep their form
    (got_mixed,), _, st = r2.scan([1], n_threads=3)
    assert got_mixed == mixed and 0 < st["dictionary_vectors"] < st["chunks"]
    r.close()
    r2.close()


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [1, 4])
def test_filter_pushdown_through_reader(tmp_path, threads):
    """WHERE clauses pushed into the scan (read_fls.cpp:39-42): zone maps prune row groups, the device evaluates the
    comparisons, chunks come back sliced by a selection vector; result == numpy."""
    import reader_lib
    from filter_cases import bits, filter_sets, numpy_mask, table
    n = 20 * 8192 + 333
    cols = table(n, seed=21)
    names = [c[0] for c in cols]
    comments = [b"row %d" % i for i in range(n)]
    p = write(tmp_path, "f.fls", [(c[0], c[1], c[2]) + tuple(c[4:6]) for c in cols] + [("note", "VARCHAR", None)],
              [c[3] for c in cols] + [comments], 8)
    r = reader_lib.Reader(p, rowgroups_per_batch=4)
    proj = list(range(len(cols) + 1))
    for fname in ("q6", "dbl_nan_gt", "small_neg", "runs_band", "none_pass", "all_pass", "dic_eq"):
        fs = filter_sets()[fname]
        filt = [(names.index(c), flsgpu.CMP[op], bits(k, k.dtype)) for c, op, k in fs]
        got, row_ids, st = r.scan(proj, filt, n_threads=threads)
        want = numpy_mask(cols, fs)
        assert st["rows"] == int(want.sum()), fname
        assert np.array_equal(row_ids, np.nonzero(want)[0]), fname
        for ci in (names.index("price"), names.index("shipdate"), names.index("dbl")):
            assert got[ci].tobytes() == cols[ci][3][want].tobytes(), (fname, names[ci])
        assert got[-1] == [comments[i] for i in np.nonzero(want)[0]], fname
        if fname == "q6":
            assert st["pruned_rowgroups"] > 0
        if fname == "none_pass":
            assert st["pruned_rowgroups"] == 21 and st["chunks"] == 0
    r.close()


def test_empty_table_scans_to_zero_rows_without_touching_the_device(tmp_path):
    """0-row COPY ... (FORMAT fls): the file has a schema and no row groups; read_fls returns nothing (and, having nothing to
    decode, needs no device)."""
    import reader_lib
    p = write(tmp_path, "empty.fls", [("a", "INT32", None), ("s", "VARCHAR", None)], [np.zeros(0, dtype=np.int32), []], 2)
    r = reader_lib.Reader(p, file_row_number=True)
    assert [c["type"] for c in r.columns] == ["INTEGER", "VARCHAR", "BIGINT"] and r.total_tuples == 0 and r.cardinality == 0
    assert r.partition_stats() == [] and r.max_threads == 1
    cols, _, st = r.scan(n_threads=2)
    assert st["rows"] == 0 and st["chunks"] == 0 and len(cols[0]) == 0 and cols[1] == []
    cols, ids, st = r.scan([0], [(0, flsgpu.CMP["GT"], 5)])
    assert st["rows"] == 0 and len(ids) == 0
    r.close()
    m = reader_lib.MultiReader([p, p], filename=True)
    assert m.cardinality == 84  # fls_multi_file_info.cpp:162-163: max(initial, 42) * file_count
    cols, _, _, st = m.scan()
    assert st["rows"] == 0
    m.close()

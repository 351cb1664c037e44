"""read_fls over several files: the C++ mirror of FastLanesMultiFileInfo + DuckDB's MultiFileReader semantics
(reference src/reader/fls_multi_file_info.cpp, src/read_fls.cpp:32-46): globs and lists, named parameters, bind with and
without union_by_name, cardinality / MaxThreads, `filename` and `file_row_number`, filters across files."""
import numpy as np
import pytest

import flsgpu


def write(tmp_path, name, cols, arrs, vec_per_rg=2):
    w = flsgpu.Writer(cols, vec_per_rg)
    w.append(arrs)
    w.finish()
    p = str(tmp_path / name)
    w.save(p)
    w.close()
    return p


def three_files(tmp_path):
    """part_0/1: (id BIGINT, v INTEGER, s VARCHAR); part_2: (id BIGINT, s VARCHAR, extra DOUBLE): a different column set."""
    data = {}
    row = 0
    for k, n in enumerate((5000, 3000)):
        ids = np.arange(row, row + n, dtype=np.int64)
        v = (ids % 97).astype(np.int32)
        s = [b"file%d-row%d" % (k, i) for i in range(n)]
        data[k] = dict(id=ids, v=v, s=s)
        write(tmp_path, f"part_{k}.fls", [("id", "INT64", None), ("v", "INT32", None), ("s", "VARCHAR", None)], [ids, v, s])
        row += n
    n = 4100
    ids = np.arange(row, row + n, dtype=np.int64)
    s = [b"tail-%d" % (i % 5) for i in range(n)]
    extra = ids / 8.0
    data[2] = dict(id=ids, s=s, extra=extra)
    write(tmp_path, "part_2.fls", [("id", "INT64", None), ("s", "VARCHAR", None), ("extra", "DOUBLE", None)], [ids, s, extra])
    return data


def test_bind_options_and_cardinality(tmp_path):
    import reader_lib
    data = three_files(tmp_path)
    g = str(tmp_path / "part_[01].fls")
    r = reader_lib.MultiReader(g, file_row_number=True, filename=True)
    assert [f.rsplit("/", 1)[1] for f, _ in r.files] == ["part_0.fls", "part_1.fls"] and [n for _, n in r.files] == [5000, 3000]
    assert [(c["name"], c["type"]) for c in r.columns] == [("id", "BIGINT"), ("v", "INTEGER"), ("s", "VARCHAR"), ("filename", "VARCHAR"),
                                                             ("file_row_number", "BIGINT")]
    # fls_multi_file_info.cpp:99-110 / :152-164: several files -> no thread limit, cardinality = max(first, 42) * files
    assert r.max_threads is None and r.cardinality == 5000 * 2
    r.close()
    one = reader_lib.MultiReader([str(tmp_path / "part_1.fls")])
    assert one.cardinality == 3000 and one.max_threads == 1  # 2 row groups fit one default batch
    one.close()
    assert reader_lib.MultiReader(g, explicit_cardinality=12345).cardinality == 12345
    # a list of patterns keeps the list order, each glob sorted
    r = reader_lib.MultiReader([str(tmp_path / "part_2.fls"), g], union_by_name=True)
    assert [f.rsplit("/", 1)[1] for f, _ in r.files] == ["part_2.fls", "part_0.fls", "part_1.fls"]
    assert [c["name"] for c in r.columns] == ["id", "s", "extra", "v"]  # union by name, first appearance order
    r.close()
    # errors: NULL argument (BinderException), unknown parameter, no match, schema mismatch without union_by_name
    with pytest.raises(reader_lib.ReaderError) as e:
        reader_lib.MultiReader(g, file_row_number=None)
    assert "Cannot use NULL as argument to file_row_number" in str(e.value)
    with pytest.raises(reader_lib.ReaderError):
        reader_lib.MultiReader(g, no_such_option=1)
    with pytest.raises(reader_lib.ReaderError) as e:
        reader_lib.MultiReader(str(tmp_path / "nothing_*.fls"))
    assert e.value.status == -2 and "No files found" in str(e.value)
    with pytest.raises(reader_lib.ReaderError) as e:
        reader_lib.MultiReader(str(tmp_path / "part_*.fls"))
    assert "does not contain column" in str(e.value)
    assert data[0]["id"][0] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [1, 4])
def test_glob_scan_with_virtual_columns(tmp_path, threads):
    import reader_lib
    data = three_files(tmp_path)
    r = reader_lib.MultiReader(str(tmp_path / "part_[01].fls"), file_row_number=True, filename=True)
    (ids, v, s, fn, frn), file_ids, row_ids, st = r.scan(n_threads=threads)
    assert st["rows"] == 8000
    assert np.array_equal(ids, np.concatenate([data[0]["id"], data[1]["id"]]))
    assert np.array_equal(v, np.concatenate([data[0]["v"], data[1]["v"]]))
    assert s == data[0]["s"] + data[1]["s"]
    assert np.array_equal(frn, np.concatenate([np.arange(5000), np.arange(3000)])) and np.array_equal(frn, row_ids)
    assert [x.decode().rsplit("/", 1)[1] for x in fn[:1] + fn[-1:]] == ["part_0.fls", "part_1.fls"]
    assert all(f == fn[0] for f in fn[:5000]) and np.array_equal(file_ids, np.repeat([0, 1], [5000, 3000]))
    r.close()


@pytest.mark.gpu
def test_union_by_name_and_filters_across_files(tmp_path):
    import reader_lib
    data = three_files(tmp_path)
    r = reader_lib.MultiReader(str(tmp_path / "part_*.fls"), union_by_name=True)
    assert [c["name"] for c in r.columns] == ["id", "v", "s", "extra"]
    (ids, v, s, extra), file_ids, _, st = r.scan(n_threads=3)
    assert st["rows"] == 12100 and np.array_equal(ids.data, np.arange(12100))
    # part_2 has no `v`, part_0/1 have no `extra`: NULL there
    assert v.mask[:8000].sum() == 0 and v.mask[8000:].all() and np.array_equal(v.data[:8000], np.concatenate([data[0]["v"], data[1]["v"]]))
    assert extra.mask[:8000].all() and extra.mask[8000:].sum() == 0 and np.array_equal(extra.data[8000:], data[2]["extra"])
    assert s == data[0]["s"] + data[1]["s"] + data[2]["s"]
    r.close()
    # WHERE id >= 4000 AND id < 9000 AND v < 50: a comparison with the missing (NULL) column drops part_2 entirely
    r = reader_lib.MultiReader(str(tmp_path / "part_*.fls"), union_by_name=True)
    filt = [(0, flsgpu.CMP["GE"], 4000), (0, flsgpu.CMP["LT"], 9000), (1, flsgpu.CMP["LT"], 50)]
    (ids, v), file_ids, _, st = r.scan([0, 1], filt, n_threads=2)
    all_ids = np.arange(8000)
    all_v = np.concatenate([data[0]["v"], data[1]["v"]])
    want = (all_ids >= 4000) & (all_ids < 9000) & (all_v < 50)
    assert np.array_equal(ids.data, all_ids[want]) and np.array_equal(v.data, all_v[want]) and set(file_ids) == {0, 1}
    assert st["pruned_rowgroups"] >= 1  # zone maps of `id` rule out the first row group of part_0
    r.close()

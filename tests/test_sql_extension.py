"""The real DuckDB `fastlanes` extension (duckdb-fastlanes_b200/extension): read_fls / COPY ... (FORMAT fls) linked into
DuckDB v1.4.0 built from the reference tree's vendored sources.

GPU tests shell out to DuckDB's own SQLLogicTest runner (extension/bin/unittest) on
  * the REFERENCE's test files, unmodified (staged by `make -C oracle ref` into oracle/_ref/test: the reference's only
    parity contract, test/all_types_{single,multi}_threaded.test), and their 1 M-row variants (BASELINE config 1);
  * the extension's own tests (extension/test/sql/*.test).
A run only counts when nothing was skipped: with the extension missing, `require fastlanes` turns a file into a skip that
exits green.
CPU tests cover what needs no device: the writer, bind/metadata, COUNT(*) from partition stats, and that a scan without a
device fails loudly instead of falling back to the CPU.
"""
import os
import re
import shutil
import subprocess
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXT = os.path.join(ROOT, "duckdb-fastlanes_b200", "extension")
BIN = os.path.join(EXT, "bin")
REF_TESTS = os.path.join(ROOT, "oracle", "_ref")
ENV = dict(os.environ, LD_LIBRARY_PATH=BIN + ":" + os.path.join(ROOT, "duckdb-fastlanes_b200") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))

needs_binaries = pytest.mark.skipif(not os.path.exists(os.path.join(BIN, "unittest")),
                                    reason="extension binaries not built (make -C duckdb-fastlanes_b200/extension)")


def run_sql(sql, threads=None):
    pre = f"SET threads={threads};\n" if threads else ""
    p = subprocess.run([os.path.join(BIN, "duckdb"), "-csv", "-noheader"], input=pre + sql, capture_output=True, text=True, env=ENV,
                       timeout=600)
    return p.returncode, p.stdout, p.stderr


def run_unittest(test_dir, pattern, log_name):
    tmp = tempfile.mkdtemp(prefix="fls_sqltest_")
    shutil.rmtree(tmp)  # the runner insists on creating it
    p = subprocess.run([os.path.join(BIN, "unittest"), "--test-dir", test_dir, "--test-temp-dir", tmp, pattern], capture_output=True,
                       text=True, env=ENV, timeout=900)
    out = p.stdout + p.stderr
    log_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(log_dir, exist_ok=True)
    with open(os.path.join(log_dir, log_name), "w") as f:
        f.write(f"$ unittest --test-dir {os.path.relpath(test_dir, ROOT)} \"{pattern}\"\nexit code {p.returncode}\n{out}")
    shutil.rmtree(tmp, ignore_errors=True)
    return p.returncode, out


def assert_all_passed(rc, out, n_files):
    assert rc == 0, out[-4000:]
    assert "skipped" not in out.lower() and "require fastlanes" not in out, "tests were SKIPPED, not passed:\n" + out[-2000:]
    m = re.search(r"All tests passed \((\d+) assertions? in (\d+) test cases?\)", out)
    assert m, out[-2000:]
    assert int(m.group(2)) == n_files, f"expected {n_files} test files to run:\n" + out[-2000:]
    return int(m.group(1))


@needs_binaries
def test_copy_and_bind_work_without_a_device_and_scans_fail_loudly(tmp_path):
    path = str(tmp_path / "t.fls")
    rc, out, err = run_sql(f"""
        CREATE TABLE t AS SELECT i AS id, CAST(i * 10 AS BIGINT) AS col, 'Value ' || (i % 100) AS s, i / 1000.0 AS d,
                                 CAST(i % 200 AS UTINYINT) AS u, CAST(i AS DECIMAL(9,2)) AS dec FROM range(1, 100001) tbl(i);
        COPY t TO '{path}' (FORMAT fls);
        SELECT column_name || ':' || column_type FROM (DESCRIBE SELECT * FROM read_fls('{path}'));
        SELECT count(*) FROM read_fls('{path}');
    """)
    assert rc == 0, err
    lines = out.replace('"', '').split()
    assert lines[-7:-1] == ["id:BIGINT", "col:BIGINT", "s:VARCHAR", "d:DOUBLE", "u:UBIGINT", "dec:DECIMAL(9,2)"], out
    assert lines[-1] == "100000"  # answered from the footer's partition statistics: no decode
    assert os.path.getsize(path) > 1024
    import torch
    if not torch.cuda.is_available():
        rc, out, err = run_sql(f"SELECT sum(col) FROM read_fls('{path}');")
        assert rc != 0 and "no CUDA device" in err and "no CPU fallback" in err, (out, err)


@needs_binaries
def test_the_files_copy_writes_are_the_engines_container(tmp_path):
    """COPY ... (FORMAT fls) goes through flsgpu_writer_*: the oracle (and the ctypes binding) read the file."""
    import numpy as np

    import flsgpu
    import oracle_lib

    path = str(tmp_path / "t.fls")
    rc, out, err = run_sql(f"""
        COPY (SELECT CAST(i AS BIGINT) * 7 AS a, CAST(i % 977 AS INTEGER) AS b, i / 8.0 AS c FROM range(0, 150000) t(i))
        TO '{path}' (FORMAT fls, row_group_size 32768);
    """)
    assert rc == 0, err
    img = np.fromfile(path, dtype=np.uint8)
    t = flsgpu.Table(img)
    assert (t.n_rowgroups, t.vec_per_rowgroup, t.n_tuples) == (5, 32, 150000)
    oracle = oracle_lib.Oracle(img)
    i = np.arange(150000)
    for c, exp in enumerate(((i * 7).astype(np.int64), (i % 977).astype(np.int32), i / 8.0)):
        got = oracle.decode_column(c, 0, t.n_rowgroups, 0, exp.dtype, t.vec_per_rowgroup)
        flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
        assert flat.tobytes() == exp.tobytes()
    t.close()


@needs_binaries
def test_aggregate_pushdown_plans_without_a_device(tmp_path):
    """The optimizer rewrite is host logic: EXPLAIN shows FLS_AGGREGATE_SCAN for the supported shape and the ordinary scan for
    everything else (extension/test/sql/fls_aggregate_pushdown.test holds the executed twin of this, on the GPU)."""
    path = str(tmp_path / "p.fls")
    rc, out, err = run_sql(f"""
        COPY (SELECT i AS id, CAST(i % 100 AS INTEGER) AS k, i / 4.0 AS x, 'm' || (i % 4) AS mode FROM range(0, 50000) t(i))
        TO '{path}' (FORMAT fls, row_group_size 8192);
    """)
    assert rc == 0, err

    def plan(sql, pre=""):
        rc, out, err = run_sql(f"{pre}EXPLAIN {sql};")
        assert rc == 0, err
        return out

    src = f"read_fls('{path}')"
    assert "FLS_AGGREGATE_SCAN" in plan(f"SELECT sum(id), count(*), min(k), max(k) FROM {src}")
    assert "FLS_AGGREGATE_SCAN" in plan(f"SELECT sum(id), count(*) FROM {src} WHERE k < 10 AND mode = 'm1'")
    for sql in (f"SELECT mode, sum(id) FROM {src} GROUP BY mode", f"SELECT sum(id * 2) FROM {src}", f"SELECT sum(x) FROM {src}",
                f"SELECT count(DISTINCT k) FROM {src}", f"SELECT min(id) FROM {src} WHERE k > 3", f"SELECT avg(id) FROM {src}",
                f"SELECT sum(id) FILTER (WHERE k > 3) FROM {src}", f"SELECT sum(id) FROM (SELECT id FROM {src} LIMIT 10)",
                f"SELECT sum(id) FROM read_fls(['{path}', '{path}'])", f"SELECT sum(id) FROM {src} WHERE k IN (1, 5, 7) OR id = 3",
                f"SELECT sum(a.id), count(*) FROM {src} a, {src} b WHERE a.id = b.id"):
        assert "FLS_AGGREGATE_SCAN" not in plan(sql), sql
    assert "FLS_AGGREGATE_SCAN" not in plan(f"SELECT sum(id) FROM {src}", "SET fastlanes_aggregate_pushdown = false;\n")


@needs_binaries
def test_copy_modes_and_options_write_the_source_rows(tmp_path):
    """Every COPY path of extension/test/sql/fls_copy_options.test (row_group_size, sidecar footer, file rotation, parallel and
    ordered batch copy, overwriting an existing file, an empty result) checked on the CPU: the streamed files open and the
    oracle decodes them to the source table."""
    import glob

    import numpy as np

    import flsgpu
    import oracle_lib

    d = str(tmp_path)
    rc, out, err = run_sql(f"""
        CREATE TABLE src AS SELECT i AS id, CAST(i * 3 AS INTEGER) AS v, 'row ' || i AS s FROM range(0, 100000) t(i);
        COPY src TO '{d}/rg2048.fls' (FORMAT fls, row_group_size 2048);
        COPY src TO '{d}/sidecar.fls' (FORMAT fls, inline_footer false);
        COPY src TO '{d}/rotated' (FORMAT fls, row_group_size 8192, row_groups_per_file 2);
        SET preserve_insertion_order = false;
        COPY src TO '{d}/parallel.fls' (FORMAT fls, row_group_size 4096);
        SET preserve_insertion_order = true;
        COPY src TO '{d}/batch.fls' (FORMAT fls, row_group_size 4096);
        COPY src TO '{d}/batch.fls' (FORMAT fls, row_group_size 4096);
        COPY (SELECT i AS id, i * 2 AS v FROM range(0, 10) t(i) WHERE i > 100) TO '{d}/empty.fls' (FORMAT fls);
    """, threads=4)
    assert rc == 0, err

    def rows_of(paths):
        ids, vs, strs = [], [], []
        for p in paths:
            img = np.fromfile(p, dtype=np.uint8)
            if os.path.exists(p + ".footer"):
                img = np.concatenate([img, np.fromfile(p + ".footer", dtype=np.uint8)])
            t = flsgpu.Table(path=p)
            o = oracle_lib.Oracle(img)
            a = o.decode_column(0, 0, t.n_rowgroups, 0, np.int64, t.vec_per_rowgroup)
            b = o.decode_column(1, 0, t.n_rowgroups, 0, np.int32, t.vec_per_rowgroup)
            st, heap = o.decode_strings(2, 0, t.n_rowgroups, t.vec_per_rowgroup, 0x7F0000001000)
            st = st.view(flsgpu.STRING_T).reshape(-1)
            for x, y in t.valid_mask_slices():
                ids.append(a[x:y])
                vs.append(b[x:y])
                strs += flsgpu.strings_from(st[x:y], heap, 0x7F0000001000)
            t.close()
        return np.concatenate(ids), np.concatenate(vs), strs

    want = np.arange(100000)
    for name, ordered in (("rg2048", True), ("sidecar", True), ("batch", True), ("parallel", False)):
        ids, vs, strs = rows_of([f"{d}/{name}.fls"])
        order = np.arange(len(ids)) if ordered else np.argsort(ids, kind="stable")
        assert np.array_equal(ids[order], want) and np.array_equal(vs[order], want * 3), name
        assert [strs[k] for k in order[:3000]] == [b"row %d" % k for k in range(3000)], name
    files = sorted(glob.glob(f"{d}/rotated/*.fls"))
    assert len(files) == 7  # 13 row groups of 8192 rows, two per file
    ids, vs, _ = rows_of(files)
    assert np.array_equal(np.sort(ids), want) and np.array_equal(vs, ids * 3)
    t = flsgpu.Table(path=f"{d}/empty.fls")
    assert (t.n_rowgroups, t.n_columns, t.n_tuples) == (0, 2, 0)
    t.close()


@needs_binaries
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["all_types_single_threaded", "all_types_multi_threaded"])
def test_reference_sqllogictests_unmodified(name):
    path = os.path.join(REF_TESTS, "test", name + ".test")
    assert os.path.exists(path), "reference tests not staged: make -C oracle ref (needs /root/reference)"
    rc, out = run_unittest(REF_TESTS, f"test/{name}.test", f"sqltest_reference_{name}.log")
    assert assert_all_passed(rc, out, 1) >= 13


@needs_binaries
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["all_types_single_threaded_1m", "all_types_multi_threaded_1m"])
def test_reference_sqllogictests_at_one_million_rows(name):
    rc, out = run_unittest(REF_TESTS, f"test/{name}.test", f"sqltest_reference_{name}.log")
    assert assert_all_passed(rc, out, 1) >= 13


@needs_binaries
@pytest.mark.gpu
def test_extension_sqllogictests():
    n_files = len([f for f in os.listdir(os.path.join(EXT, "test", "sql")) if f.endswith(".test")])
    rc, out = run_unittest(EXT, "test/sql/*", "sqltest_extension.log")
    assert_all_passed(rc, out, n_files)

"""Writes tests/golden/container_v1.npz: ONE container written by the product writer at format version 1 (every numeric
encoding on INT32 / DOUBLE columns and one column per string encoding, 1 124 rows in row groups of one vector) next to
the rows it was written from.  tests/test_golden.py::test_committed_container_still_decodes opens it with today's validator
and decodes it with today's oracle: a change of the on-disk format that forgets older files fails there.
Run from the repo root: python oracle/make_golden_container.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import flsgpu  # noqa: E402
import fls_cases  # noqa: E402
import str_cases  # noqa: E402


def tables(n=1124):
    cs = [c for c in fls_cases.standard_cases(n=n, seed=5) if c[1] in ("INT32", "DOUBLE")]
    sc = str_cases.cases(n)
    keys = ["shipmode", "comment", "constant", "instruct", "blob_plain", "comment12", "instruct12", "flag_runs", "status_cross", "freq_none"]
    cols = [fls_cases.column_spec(c) for c in cs] + [(k, sc[k][0], sc[k][1]) for k in keys]
    arrs = [c[3] for c in cs] + [sc[k][2] for k in keys]
    return cols, arrs


def main():
    cols, arrs = tables()
    img = flsgpu.write_table(cols, arrs, vec_per_rowgroup=1, n_threads=1)
    out = {"image": img, "n_columns": np.array(len(cols)), "names": np.array([c[0] for c in cols]), "types": np.array([c[1] for c in cols])}
    for i, a in enumerate(arrs):
        if isinstance(a, list):
            out[f"off{i}"] = np.concatenate([[0], np.cumsum([len(s) for s in a])]).astype(np.uint64)
            out[f"col{i}"] = np.frombuffer(b"".join(a), dtype=np.uint8)
        else:
            out[f"col{i}"] = np.ascontiguousarray(a)
    path = os.path.join(ROOT, "tests", "golden", "container_v1.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes; image", img.size, "bytes,", len(cols), "columns")


if __name__ == "__main__":
    main()

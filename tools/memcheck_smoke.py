"""A small host-to-host scan of every numeric encoding (INT16, INT64, DOUBLE) and three string encodings, no torch: short
enough to run under `compute-sanitizer --tool memcheck --error-exitcode 9 python tools/memcheck_smoke.py` where that tool is
available (it is closed on the pool this repo was measured on; the device side is covered there by the bit-exact parity
tests, the random-table soak and the kernels' own clamps instead)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import flsgpu  # noqa: E402
import fls_cases  # noqa: E402
import str_cases  # noqa: E402

n = 2 * 1024 + 77
cs = [c for c in fls_cases.standard_cases(n=n) if c[1] in ("INT64", "DOUBLE", "INT16")]
img = flsgpu.write_table([fls_cases.column_spec(c) for c in cs], [c[3] for c in cs], vec_per_rowgroup=2)
t = flsgpu.Table(img)
rows = t.out_elems(t.n_rowgroups)
for i, c in enumerate(cs):
    out = np.zeros(rows, dtype=c[3].dtype)
    t.scan_column_to_host(i, out.ctypes.data, out.nbytes)
    got = np.concatenate([out[a:b] for a, b in t.valid_mask_slices()])
    assert got.tobytes() == c[3].tobytes(), c[0]
res = t.aggregate_to_host([0, 1])
t.close()
sc = str_cases.cases(n)
names = ["comment", "shipmode", "flag_runs"]
img = flsgpu.write_table([(k, sc[k][0], sc[k][1]) for k in names], [sc[k][2] for k in names], vec_per_rowgroup=2)
t = flsgpu.Table(img)
for i, k in enumerate(names):
    out = np.zeros((t.out_elems(t.n_rowgroups), 16), dtype=np.uint8)
    heap = np.zeros(max(16, t.string_heap_bytes(i)), dtype=np.uint8)
    t.scan_strings_to_host(i, out.ctypes.data, out.nbytes, heap.ctypes.data, heap.nbytes)
    st = out.view(flsgpu.STRING_T).reshape(-1)
    got = []
    for a, b in t.valid_mask_slices():
        got += flsgpu.strings_from(st[a:b], heap, heap.ctypes.data)
    assert got == sc[k][2], k
t.close()
print("memcheck smoke ok:", len(cs), "numeric columns,", len(names), "string columns")

#!/usr/bin/env python
"""A/B of two builds of libflsgpu.so on one box: the chained resident scan of a bench-shaped table (no reader, no oracle),
ms per step.  FLSGPU_LIB picks the library.  Usage: FLSGPU_LIB=... python tools/ab_scan.py [rows] [steps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests"), ROOT):
    sys.path.insert(0, p)
import torch  # noqa: E402
import flsgpu  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
GEN = [("tiny", "INT8", lambda i: ((i % 256) - 128).astype(np.int8)), ("small", "INT16", lambda i: ((i % 65536) - 32768).astype(np.int16)),
       ("int", "INT32", lambda i: i.astype(np.int32)), ("big", "INT64", lambda i: i * 10), ("ubig", "UINT64", lambda i: (i * 10).astype(np.uint64)),
       ("flt", "FLOAT", lambda i: i.astype(np.float32) / np.float32(100)), ("dbl", "DOUBLE", lambda i: i / 1000.0),
       ("dec", "INT64", lambda i: (i % 10_000_000) * 100), ("date", "INT32", lambda i: (8116 + (i % 10000)).astype(np.int32)),
       ("ts", "INT64", lambda i: 1_735_689_600_000_000 + i * 1_000_000), ("rnd", "INT64", None)]
flsgpu.init((0,))
w = flsgpu.Writer([(n, t, None) for n, t, _ in GEN], 64)
rng = np.random.default_rng(1)
done = 0
while done < rows:
    n = min(64 * 65536, rows - done)
    i = np.arange(done + 1, done + n + 1, dtype=np.int64)
    w.append([g(i) if g else rng.integers(0, 1 << 20, n) for _, _, g in GEN])  # rnd: FFOR W = 20
    done += n
img = w.finish()
t = flsgpu.Table(img)
t.stage()
ncol = t.n_columns
outs = [torch.empty(t.out_elems(t.n_rowgroups) * (t.column(c).phys_bits // 8), dtype=torch.uint8, device="cuda") for c in range(ncol)]
st = torch.cuda.current_stream().cuda_stream


def step():
    for c in range(ncol):
        t.decode_column(c, outs[c].data_ptr(), outs[c].numel(), stream=st)


for _ in range(5):
    step()
torch.cuda.synchronize()
res = []
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    res.append(round(e0.elapsed_time(e1) / steps, 4))
print(os.path.basename(flsgpu.LIB_PATH), "ms per step:", res)
os._exit(0)

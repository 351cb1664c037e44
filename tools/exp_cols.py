#!/usr/bin/env python
"""Experiment: N identical INT64 DELTA columns; per-column decode time in round-robin order and back to back alone."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests"), ROOT):
    sys.path.insert(0, p)
import torch
import flsgpu

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
kinds = (sys.argv[2] if len(sys.argv) > 2 else "big,big,big,big").split(",")
ncol = len(kinds)
GEN = {"big": ("INT64", lambda i: i * 10), "dbl": ("DOUBLE", lambda i: i / 1000.0), "date": ("INT32", lambda i: (8116 + (i % 10000)).astype(np.int32)),
       "tiny": ("INT8", lambda i: ((i % 256) - 128).astype(np.int8)), "small": ("INT16", lambda i: ((i % 65536) - 32768).astype(np.int16)),
       "flt": ("FLOAT", lambda i: i.astype(np.float32) / np.float32(100)), "dec": ("INT64", lambda i: (i % 10_000_000) * 100),
       "ts": ("INT64", lambda i: 1_735_689_600_000_000 + i * 1_000_000),
       "int": ("INT32", lambda i: i.astype(np.int32)), "code": ("UINT16", lambda i: (i % 1000).astype(np.uint16))}
flsgpu.init((0,))
w = flsgpu.Writer([(f"c{k}", GEN[kd][0], "DICT" if kd == "code" else None) for k, kd in enumerate(kinds)], 64)
done = 0
while done < rows:
    n = min(64 * 65536, rows - done)
    i = np.arange(done + 1, done + n + 1, dtype=np.int64)
    w.append([GEN[kd][1](i) for kd in kinds])
    done += n
img = w.finish()
t = flsgpu.Table(img)
t.stage()
outs = [torch.empty(t.out_elems(t.n_rowgroups) * (t.column(c).phys_bits // 8), dtype=torch.uint8, device="cuda") for c in range(ncol)]
st = torch.cuda.current_stream().cuda_stream
print("out ptrs", [hex(o.data_ptr()) for o in outs])

def run(order, steps=20):
    for c in order * 3:
        t.decode_column(c, outs[c].data_ptr(), outs[c].numel(), stream=st)
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in order] for _ in range(steps)]
    for s in range(steps):
        for j, c in enumerate(order):
            ev[s][j][0].record()
            t.decode_column(c, outs[c].data_ptr(), outs[c].numel(), stream=st)
            ev[s][j][1].record()
    torch.cuda.synchronize()
    return [round(float(np.mean([ev[s][j][0].elapsed_time(ev[s][j][1]) for s in range(steps)])), 4) for j in range(len(order))]

print(kinds)
print("round robin fwd:", run(list(range(ncol))))
print("round robin rev:", run(list(range(ncol))[::-1])[::-1])
import os as _os
_os._exit(0)

#!/usr/bin/env python
"""Flat decode (heap + string_t rows) of every string column of a .fls file, resident: ms, text GB/s, algorithmic GB/s.
Usage: python tools/fsst_file_bench.py file.fls [iters]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests")]
import torch  # noqa: E402
import flsgpu  # noqa: E402

img = np.fromfile(sys.argv[1], dtype=np.uint8)
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
flsgpu.init((0,))
t = flsgpu.Table(img)
t.stage()
st = torch.cuda.current_stream().cuda_stream
rows = t.out_elems(t.n_rowgroups)
for c in range(t.n_columns):
    if t.column(c).phys_bits != 128:
        continue
    hb = t.string_heap_bytes(c)
    out = torch.empty(rows * 16, dtype=torch.uint8, device="cuda")
    heap = torch.empty(max(16, hb), dtype=torch.uint8, device="cuda")

    def run():
        t.decode_strings(c, out.data_ptr(), out.numel(), heap.data_ptr(), heap.numel(), heap.data_ptr(), stream=st)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    comp = sum(t.chunk(r, c).size for r in range(t.n_rowgroups))
    encs = sorted({flsgpu.ENC_NAME[t.chunk(r, c).encoding] for r in range(t.n_rowgroups)})
    print(json.dumps({"column": t.column(c).name.decode(), "encodings": encs, "balance": os.environ.get("FLSGPU_STR_BALANCE", "1"), "ms": round(ms, 4),
                      "text_GBps": round(hb / ms / 1e6, 1), "alg_GBps": round((comp + hb + t.n_tuples * 16) / ms / 1e6, 1)}))
os._exit(0)

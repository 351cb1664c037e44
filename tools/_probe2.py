import sys, os, time
sys.path.insert(0, "duckdb-fastlanes_b200"); sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch, flsgpu, bench
flsgpu.init((0,))
w, img = bench.build_table_image(30_000_000, 16)
np.asarray(img).tofile("/dev/shm/p.fls")
for name, t in (("memory(writer image)", flsgpu.Table(img)), ("file", flsgpu.Table(path="/dev/shm/p.fls"))):
    for i in range(3):
        t.evict(); torch.cuda.synchronize()
        t0 = time.perf_counter(); t.stage(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print(name, i, round(t.resident_bytes() / dt / 1e9, 1), "GB/s H2D", round(dt * 1e3, 2), "ms")
    t.close()
os.remove("/dev/shm/p.fls")

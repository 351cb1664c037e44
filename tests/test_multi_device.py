"""Single-process multi-GPU: flsgpu_init keeps every listed device, one table handle per device over the same host image
(flsgpu_table_share + flsgpu_table_bind_device), the reader sends batch k of row groups to device k mod G.  On a box with
one GPU the list is (0, 0): two handles, two arenas, two sets of streams on the same device — the same code path.
Runs in a child process: the device list is process-wide state."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import sys, os
sys.path[:0] = [os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests"), ROOT]
import numpy as np, torch
import flsgpu, oracle_lib, reader_lib
from test_vector_forms import columns

n_dev = torch.cuda.device_count()
devs = tuple(range(n_dev)) if n_dev > 1 else (0, 0)
flsgpu.init(devs)
assert flsgpu.n_devices() == len(devs)
n = 700000
cols = columns(n)
img = flsgpu.write_table([(c[0], c[1], c[2]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=8)
path = os.path.join(TMP, "multi.fls")
img.tofile(path)

# 1. C ABI: a handle per device; every handle decodes ITS share of the row groups on ITS device; together = the oracle
t0 = flsgpu.Table(img)
handles = [t0] + [flsgpu.Table(share=t0) for _ in devs[1:]]
for g, h in enumerate(handles):
    h.bind_device(g)
o = oracle_lib.Oracle(img)
c = [x[0] for x in cols].index("rand")
per = (t0.n_rowgroups + len(devs) - 1) // len(devs)
for g, h in enumerate(handles):
    rg0, nrg = g * per, min(per, t0.n_rowgroups - g * per)
    with torch.cuda.device(devs[g]):
        out = torch.zeros(h.out_elems(nrg) * 4, dtype=torch.uint8, device=f"cuda:{devs[g]}")
        h.stage(rg0, nrg)
        h.decode_column(c, out.data_ptr(), out.numel(), rg0, nrg)
        torch.cuda.synchronize()
        got = out.cpu().numpy().view(np.int32)
    exp = o.decode_column(c, rg0, nrg, 0, np.int32, t0.vec_per_rowgroup)
    for a, b in h.valid_mask_slices(rg0, nrg):
        assert got[a:b].tobytes() == exp[a:b].tobytes(), ("device", g)
    assert h.resident_bytes() > 0
try:
    handles[0].bind_device(1)  # too late: the handle has device state
    raise SystemExit("bind_device after staging must fail")
except flsgpu.FlsError:
    pass
for h in handles[1:] + [t0]:
    h.close()

# 2. the reader: batches alternate between the devices, the rows a consumer reads are the input
for threads in (1, 4):
    r = reader_lib.Reader(path, rowgroups_per_batch=3)
    got, _, stats = r.scan(n_threads=threads)
    for ci, col in enumerate(cols):
        exp = col[3].astype(np.uint64) if col[3].dtype.kind == "u" else col[3]
        assert got[ci].view(exp.dtype).tobytes() == exp.tobytes(), (col[0], threads)
    assert stats["rows"] == n
    r.close()
print("multi-device ok:", devs)
'''


@pytest.mark.gpu
def test_one_process_scans_a_table_across_all_devices(tmp_path):
    code = f"ROOT = {ROOT!r}\nTMP = {str(tmp_path)!r}\n" + CHILD
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "multi-device ok" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]

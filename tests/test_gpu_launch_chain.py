"""GPU: decode launches of one stream are chained by programmatic dependent launch (a launch releases its successor at once
and waits for its predecessor before it exits).  Whatever follows in the stream must still see EVERY earlier decode complete:
a long decode followed by a very short one followed by a consumer of the LONG one's output, no synchronisation in between."""
import numpy as np
import pytest

import flsgpu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("pdl", ["1", "0"])
def test_consumer_after_short_launch_sees_the_long_one(pdl):
    import subprocess
    import sys
    import os
    # FLSGPU_PDL is read when the library is loaded: run the body in a fresh interpreter
    code = r'''
import os, sys
import numpy as np, torch
sys.path[:0] = [os.path.join(os.getcwd(), "duckdb-fastlanes_b200"), os.path.join(os.getcwd(), "tests")]
import flsgpu
flsgpu.init((0,))
rng = np.random.default_rng(3)
big = np.cumsum(rng.integers(0, 1 << 20, 6_000_000)).astype(np.int64)
small = (np.arange(2048) % 100).astype(np.int8)
tb = flsgpu.Table(flsgpu.write_table([("b", "INT64", "DELTA")], [big]))
ts = flsgpu.Table(flsgpu.write_table([("s", "INT8", "FFOR")], [small], vec_per_rowgroup=2))
stream = torch.cuda.current_stream().cuda_stream
tb.stage(stream=stream); ts.stage(stream=stream)
nb = tb.out_elems(tb.n_rowgroups) * 8
out_b = torch.zeros(nb, dtype=torch.uint8, device="cuda")
out_s = torch.zeros(ts.out_elems(ts.n_rowgroups), dtype=torch.uint8, device="cuda")
copies = []
for it in range(12):
    out_b.zero_()
    tb.decode_column(0, out_b.data_ptr(), nb, stream=stream)
    for _ in range(1 + it % 3):
        ts.decode_column(0, out_s.data_ptr(), out_s.numel(), stream=stream)
    copies.append(out_b.clone())  # ordinary kernel right behind the short launch(es)
torch.cuda.synchronize()
exp = big.tobytes()
for c in copies:
    got = c.cpu().numpy().view(np.int64)
    flat = np.concatenate([got[a:b] for a, b in tb.valid_mask_slices()])
    assert flat.tobytes() == exp
assert out_s.cpu().numpy().view(np.int8)[:2048].tobytes() == small.tobytes()
print("chain ok")
'''
    env = dict(os.environ, FLSGPU_PDL=pdl)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "chain ok" in r.stdout, r.stderr[-2000:]

"""GPU: random tables (tests/random_tables.py) go writer -> container -> CUDA decode through the C ABI and come back bit for
bit (and equal the oracle's decode of the same bytes on every row of every existing vector)."""
import os

import numpy as np
import pytest

import flsgpu
import oracle_lib
import random_tables

pytestmark = pytest.mark.gpu
HEAP_BASE = 0x7F0000001000


def test_random_tables_roundtrip_through_cuda():
    import torch
    first, last = (int(v) for v in os.environ.get("FLS_RANDOM_SEEDS", "1000:1150").split(":"))  # a longer soak by hand
    for seed in range(first, last):
        cols, arrs, vpr = random_tables.table(seed)
        img = flsgpu.write_table(cols, arrs, vec_per_rowgroup=vpr, n_threads=1)
        t = flsgpu.Table(img)
        t.stage()
        o = oracle_lib.Oracle(img)
        n_rows = t.out_elems(t.n_rowgroups)
        for c, (col, a) in enumerate(zip(cols, arrs)):
            if col[1] in ("VARCHAR", "BLOB"):
                hb = t.string_heap_bytes(c)
                rows = torch.full((n_rows * 16,), 0xA5, dtype=torch.uint8, device="cuda")
                heap = torch.zeros((max(16, hb),), dtype=torch.uint8, device="cuda")
                t.decode_strings(c, rows.data_ptr(), rows.numel(), heap.data_ptr(), heap.numel(), HEAP_BASE)
                flsgpu.sync()
                torch.cuda.synchronize()
                st = rows.cpu().numpy().view(flsgpu.STRING_T).reshape(-1)
                heap = heap.cpu().numpy()
                got = []
                for x, y in t.valid_mask_slices():
                    got += flsgpu.strings_from(st[x:y], heap, HEAP_BASE)
                assert got == a, (seed, col)
                exp_rows, _ = o.decode_strings(c, 0, t.n_rowgroups, t.vec_per_rowgroup, HEAP_BASE)
                for r in range(t.n_rowgroups):
                    x = r * t.vec_per_rowgroup * 1024
                    y = x + t.rowgroup(r).n_vec * 1024
                    assert rows.cpu().numpy().reshape(-1, 16)[x:y].tobytes() == exp_rows[x:y].tobytes(), (seed, col, r)
            else:
                nbytes = n_rows * a.dtype.itemsize
                buf = torch.full((nbytes,), 0xA5, dtype=torch.uint8, device="cuda")
                t.decode_column(c, buf.data_ptr(), nbytes)
                flsgpu.sync()
                torch.cuda.synchronize()
                out = buf.cpu().numpy().view(a.dtype)
                got = np.concatenate([out[x:y] for x, y in t.valid_mask_slices()])
                assert got.tobytes() == a.tobytes(), (seed, col)
                exp = o.decode_column(c, 0, t.n_rowgroups, 0, a.dtype, t.vec_per_rowgroup)
                for x, y in t.valid_mask_slices():
                    assert out[x:y].tobytes() == exp[x:y].tobytes(), (seed, col)
        t.close()

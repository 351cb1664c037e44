"""CPU: random tables (tests/random_tables.py) go writer -> container -> oracle and come back bit for bit."""
import numpy as np

import flsgpu
import oracle_lib
import random_tables

HEAP_BASE = 0x7F0000001000


def test_random_tables_roundtrip_through_the_oracle():
    for seed in range(400):
        cols, arrs, vpr = random_tables.table(seed)
        img = flsgpu.write_table(cols, arrs, vec_per_rowgroup=vpr, n_threads=1)
        t = flsgpu.Table(img)
        o = oracle_lib.Oracle(img)
        for c, (col, a) in enumerate(zip(cols, arrs)):
            if col[1] in ("VARCHAR", "BLOB"):
                st, heap = o.decode_strings(c, 0, t.n_rowgroups, t.vec_per_rowgroup, HEAP_BASE)
                st = st.view(flsgpu.STRING_T).reshape(-1)
                got = []
                for x, y in t.valid_mask_slices():
                    got += flsgpu.strings_from(st[x:y], heap, HEAP_BASE)
                assert got == a, (seed, col)
            else:
                out = o.decode_column(c, 0, t.n_rowgroups, 0, a.dtype, t.vec_per_rowgroup)
                got = np.concatenate([out[x:y] for x, y in t.valid_mask_slices()])
                assert got.tobytes() == a.tobytes(), (seed, col)
        t.close()

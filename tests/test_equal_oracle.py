"""CPU: EXP_EQUAL chunks written by the writer decode (oracle) to the input; the bytes exist once."""
import numpy as np

import flsgpu
import oracle_lib
from equal_cases import table


def equal_map(t):
    return {(r, c): (t.chunk(r, c).is_equal, t.chunk(r, c).equal_col) for r in range(t.n_rowgroups) for c in range(t.n_columns)}


def test_equal_chunks_share_bytes_and_decode(monkeypatch):
    vpr, n = 4, 4 * 4096 + 100
    specs, data = table(n, vpr * 1024)
    img = flsgpu.write_table(specs, data, vec_per_rowgroup=vpr)
    monkeypatch.setenv("FLSGPU_WRITER_EQUAL", "0")
    plain = flsgpu.write_table(specs, data, vec_per_rowgroup=vpr)
    monkeypatch.delenv("FLSGPU_WRITER_EQUAL")
    assert img.size < plain.size * 0.75
    t = flsgpu.Table(img)
    names = [s[0] for s in specs]
    eq = equal_map(t)
    for r in range(t.n_rowgroups):
        assert eq[(r, names.index("a"))] == (0, 0)
        assert eq[(r, names.index("b"))] == (1, names.index("a"))
        assert eq[(r, names.index("c"))] == ((0, 0) if r == 1 else (1, names.index("a")))
        assert eq[(r, names.index("e"))] == (1, names.index("d"))
        assert eq[(r, names.index("h"))] == (1, names.index("g"))
        assert eq[(r, names.index("f"))][0] == 0 and eq[(r, names.index("forced"))][0] == 0
        assert t.chunk(r, names.index("b")).offset == t.chunk(r, names.index("a")).offset
    tp = flsgpu.Table(plain)
    assert not any(v[0] for v in equal_map(tp).values())
    o = oracle_lib.Oracle(img)
    for c, src in enumerate(data):
        got = o.decode_column(c, 0, t.n_rowgroups, 0, src.dtype, vpr)
        flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
        assert flat.tobytes() == src.tobytes(), names[c]
    t.close()
    tp.close()

"""Random tables for the round-trip property tests (CPU: writer -> oracle, GPU: writer -> CUDA): per seed a table of 1..5
columns of random types, value shapes (full-range bits, narrow ranges around a random base at every width, sorted with random
gaps, dictionaries, runs, constants, sparse outliers, extremes alternating; decimals, random bit patterns, NaN / inf / -0.0 /
denormals for floats; binary, low-cardinality, long runs, unique, empty and escape-heavy strings) and forced or self-chosen
encodings, at row counts around the vector size and row groups of 1 / 2 / 3 / 64 vectors."""
import numpy as np

INT = {"INT8": np.int8, "INT16": np.int16, "INT32": np.int32, "INT64": np.int64, "UINT8": np.uint8, "UINT16": np.uint16, "UINT32": np.uint32, "UINT64": np.uint64}
INT_ENC = ["UNCOMPRESSED", "FFOR", "DELTA", "DICT", "RLE", "CROSS_RLE", "FREQUENCY", "SLPATCH", None, None]
FLT_ENC = ["ALP", "ALP_RD", "UNCOMPRESSED", "DICT", "RLE", None, None, "CROSS_RLE", "FREQUENCY"]
STR_ENC = ["DICT_STR", "STR_UNCOMPRESSED", "FSST", "FSST_DICT", "FSST12", "FSST12_DICT", "RLE_STR", "CROSS_RLE_STR", "FREQUENCY_STR", None, None]

def gen_int(r, dt, n):
    info = np.iinfo(dt); bits = np.dtype(dt).itemsize * 8
    kind = r.integers(0, 9)
    if kind == 0:
        return r.integers(0, 256, n * np.dtype(dt).itemsize, dtype=np.uint8).view(dt).copy()
    if kind == 1:
        w = int(r.integers(0, bits + 1)); base = int(r.integers(0, 256, 8, dtype=np.uint8).view(np.uint64)[0])
        raw = r.integers(0, (1 << w) - 1, n, dtype=np.uint64, endpoint=True) if w else np.zeros(n, dtype=np.uint64)
        return (raw + np.uint64(base & ((1 << 64) - 1))).astype(np.dtype(f"u{np.dtype(dt).itemsize}")).view(dt)
    if kind == 2:
        w = int(r.integers(0, bits)); step = r.integers(0, (1 << min(w, 62)) - 1, n, dtype=np.uint64, endpoint=True) if w else np.zeros(n, dtype=np.uint64)
        return np.cumsum(step, dtype=np.uint64).astype(np.dtype(f"u{np.dtype(dt).itemsize}")).view(dt)
    if kind == 3:
        k = int(r.integers(1, 300)); vals = r.integers(0, 256, k * np.dtype(dt).itemsize, dtype=np.uint8).view(dt)
        return vals[r.integers(0, k, n)]
    if kind == 4:
        k = int(r.integers(1, 50)); vals = r.integers(0, 256, k * np.dtype(dt).itemsize, dtype=np.uint8).view(dt)
        lens = r.integers(1, int(r.integers(2, 3000)), n + 1)
        return np.repeat(vals[r.integers(0, k, n + 1)], lens)[:n].copy()
    if kind == 5:
        return np.full(n, r.integers(0, 256, np.dtype(dt).itemsize, dtype=np.uint8).view(dt)[0], dtype=dt)
    if kind == 6:
        x = np.full(n, 3, dtype=dt); m = r.random(n) < r.random() * 0.2
        x[m] = r.integers(0, 256, int(m.sum()) * np.dtype(dt).itemsize, dtype=np.uint8).view(dt); return x
    if kind == 7:
        x = np.zeros(n, dtype=dt); x[::2] = info.max; x[1::2] = info.min; return x
    return (np.arange(n) % int(r.integers(1, 5000))).astype(dt)

def gen_flt(r, dt, n):
    kind = r.integers(0, 8)
    if kind == 0: x = r.integers(-10**6, 10**6, n) / 10.0 ** int(r.integers(0, 8))
    elif kind == 1: x = r.standard_normal(n) * 10.0 ** int(r.integers(-30, 30))
    elif kind == 2: x = r.integers(0, 256, n * 8, dtype=np.uint8).view(np.float64).copy()
    elif kind == 3: x = np.round(r.random(n) * 1000, int(r.integers(0, 4)))
    elif kind == 4: x = np.full(n, r.random())
    elif kind == 5:
        x = r.integers(0, 1000, n) / 100.0; m = r.random(n) < 0.05
        x[m] = r.choice([np.nan, np.inf, -np.inf, -0.0, 5e-324, 1.7976931348623157e308], int(m.sum()))
    elif kind == 6: x = np.repeat(r.random(n // 10 + 1), 10)[:n]
    else: x = (np.arange(n) * 0.01)
    if dt == np.float32:
        if kind == 2: return r.integers(0, 256, n * 4, dtype=np.uint8).view(np.float32).copy()
        with np.errstate(over='ignore', invalid='ignore'): return x.astype(np.float32)
    return x.astype(np.float64)

def gen_str(r, n):
    kind = r.integers(0, 7)
    if kind == 0: return [bytes(r.integers(0, 256, int(r.integers(0, 40)), dtype=np.uint8)) for _ in range(n)]
    if kind == 1:
        k = int(r.integers(1, 60)); vals = [bytes(r.integers(32, 127, int(r.integers(0, 30)), dtype=np.uint8)) for _ in range(k)]
        return [vals[int(i)] for i in r.integers(0, k, n)]
    if kind == 2:
        k = int(r.integers(1, 10)); vals = [bytes(r.integers(97, 123, int(r.integers(0, 300)), dtype=np.uint8)) for _ in range(k)]
        out = []
        while len(out) < n: out += [vals[int(r.integers(0, k))]] * int(r.integers(1, 2000))
        return out[:n]
    if kind == 3: return [b"Value %d" % i for i in range(n)]
    if kind == 4: return [b""] * n
    if kind == 5:
        words = [b"the", b"quick", b"brown", b"fox", b"jumps", b"\xff\xff", b"\x00", b"lazy dog " * 5]
        return [b" ".join(words[int(j)] for j in r.integers(0, len(words), int(r.integers(0, 9)))) for _ in range(n)]
    v = bytes(r.integers(0, 256, int(r.integers(0, 70)), dtype=np.uint8)); return [v] * n



def table(seed):
    """(column specs, arrays, vec_per_rowgroup) of one random table."""
    r = np.random.default_rng(seed)
    n = int(r.choice([1, 2, 1023, 1024, 1025, int(r.integers(1, 7000))]))
    vpr = int(r.choice([1, 2, 3, 64]))
    cols, arrs = [], []
    for k in range(int(r.integers(1, 6))):
        t = int(r.integers(0, 4))
        if t <= 1:
            name = list(INT)[int(r.integers(0, 8))]
            enc = INT_ENC[int(r.integers(0, len(INT_ENC)))]
            cols.append((f"c{k}", name, enc))
            arrs.append(gen_int(r, INT[name], n))
        elif t == 2:
            name = ["FLOAT", "DOUBLE"][int(r.integers(0, 2))]
            enc = FLT_ENC[int(r.integers(0, len(FLT_ENC)))]
            cols.append((f"c{k}", name, enc))
            arrs.append(gen_flt(r, np.float32 if name == "FLOAT" else np.float64, n))
        else:
            name = ["VARCHAR", "BLOB"][int(r.integers(0, 2))]
            enc = STR_ENC[int(r.integers(0, len(STR_ENC)))]
            cols.append((f"c{k}", name, enc))
            arrs.append(gen_str(r, n))
    return cols, arrs, vpr

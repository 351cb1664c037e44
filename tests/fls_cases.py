"""Seeded input generators shared by the CPU (oracle) and GPU (parity) tests.

Every case is (name, type_name, force_encoding|None, numpy array[, precision, scale]).  The cases cover what the
reference's own tests write (test/all_types_single_threaded.test:36-160: ids 1..N, i%256-128, i/1000.0, ...), plus
the edges SURVEY.md section 7 lists: W = 0, W = T, partial last vector, INT_MIN/INT_MAX, NaN / -0.0 / inf for ALP,
exceptions, single-run and 1024-run RLE vectors.
"""
import numpy as np

INT_TYPES = {"INT8": np.int8, "INT16": np.int16, "INT32": np.int32, "INT64": np.int64,
             "UINT8": np.uint8, "UINT16": np.uint16, "UINT32": np.uint32, "UINT64": np.uint64}
INT_ENCODINGS = ["UNCOMPRESSED", "FFOR", "DELTA", "DICT", "RLE", "CROSS_RLE", "FREQUENCY", "SLPATCH"]


def _rng(seed):
    return np.random.default_rng(seed)


def int_column(type_name, kind, n, seed=0):
    """A column of `n` values of an integer type with a given statistical shape."""
    dt = np.dtype(INT_TYPES[type_name])
    info = np.iinfo(dt)
    bits = dt.itemsize * 8
    r = _rng(seed + bits)
    if kind == "sorted":  # ids 1..N / timestamps
        step = r.integers(0, 4, n, dtype=np.int64)
        x = np.cumsum(step)
        return (x % (int(info.max) + 1)).astype(dt) if bits < 64 else x.astype(dt)
    if kind == "narrow":  # FFOR with a small width around a big base
        base = int(info.max) // 2
        return (base + r.integers(0, 1 << min(bits - 1, 5), n, dtype=np.int64)).astype(dt)
    if kind == "full":  # every bit used: W = T
        raw = r.integers(0, 256, n * dt.itemsize, dtype=np.uint8)
        x = raw.view(dt).copy()
        x[0] = info.min
        x[-1] = info.max
        return x
    if kind == "runs":  # low cardinality, long runs
        vals = r.integers(max(info.min, -1000), min(info.max, 1000), 64, dtype=np.int64, endpoint=True)
        lens = r.integers(1, 200, n // 50 + 2)
        x = np.repeat(vals[r.integers(0, 64, len(lens))], lens)[:n]
        if len(x) < n:
            x = np.concatenate([x, np.full(n - len(x), x[-1])])
        return x.astype(dt)
    if kind == "lowcard":  # dictionary, no run structure
        vals = r.integers(info.min, info.max, 37, dtype=dt, endpoint=True)
        return vals[r.integers(0, 37, n)]
    if kind == "outliers":  # narrow values plus a few huge ones (SLPATCH / FREQUENCY)
        x = r.integers(0, 16, n, dtype=np.int64).astype(dt)
        x[r.random(n) < 0.7] = 3
        pos = r.integers(0, n, max(1, n // 97))
        x[pos] = (np.full(len(pos), info.max, dtype=dt) - r.integers(0, 5, len(pos)).astype(dt))
        return x
    if kind == "constant":
        return np.full(n, info.min + 7, dtype=dt)
    if kind == "alternating":  # 1024 runs per vector
        return (np.arange(n) & 1).astype(dt)
    raise ValueError(kind)


def float_column(type_name, kind, n, seed=0):
    dt = np.float32 if type_name == "FLOAT" else np.float64
    r = _rng(seed + 99)
    if kind == "decimal":  # i / 1000.0, the reference test's DOUBLE generator (:98-104)
        x = (np.arange(1, n + 1, dtype=np.float64) / (100.0 if dt == np.float32 else 1000.0))
        return x.astype(dt)
    if kind == "prices":
        return np.round(r.random(n) * 1000.0, 2).astype(dt)
    if kind == "real":  # full-entropy mantissas: ALP-RD territory
        return (r.random(n) * 2.0 - 1.0).astype(dt)
    if kind == "specials":
        x = np.round(r.random(n) * 100.0, 1).astype(dt)
        sp = np.array([np.nan, np.inf, -np.inf, -0.0, 0.0, np.finfo(dt).max, np.finfo(dt).tiny, 1e-300 if dt == np.float64 else 1e-38],
                      dtype=dt)
        pos = r.integers(0, n, max(8, n // 61))
        x[pos] = sp[r.integers(0, len(sp), len(pos))]
        return x
    if kind == "constant":
        return np.full(n, 2.5, dtype=dt)
    raise ValueError(kind)


def standard_cases(n=3 * 1024 + 517, seed=0):
    """One case per (type, encoding) the encoder supports, on data that suits the encoding."""
    cases = []
    for tname in INT_TYPES:
        for enc, kind in [("UNCOMPRESSED", "full"), ("FFOR", "narrow"), ("FFOR", "full"), ("DELTA", "sorted"),
                          ("DELTA", "full"), ("DICT", "lowcard"), ("RLE", "runs"), ("RLE", "alternating"),
                          ("CROSS_RLE", "runs"), ("FREQUENCY", "outliers"), ("SLPATCH", "outliers"),
                          ("CONSTANT", "constant"), (None, "sorted"), (None, "runs")]:
            cases.append((f"{tname}-{enc}-{kind}", tname, enc, int_column(tname, kind, n, seed)))
    for tname in ("FLOAT", "DOUBLE"):
        for enc, kind in [("ALP", "decimal"), ("ALP", "prices"), ("ALP", "specials"), ("ALP_RD", "real"),
                          ("ALP_RD", "specials"), ("UNCOMPRESSED", "real"), ("DICT", "prices"), ("RLE", "constant"),
                          ("CONSTANT", "constant"), (None, "decimal"), (None, "real")]:
            cases.append((f"{tname}-{enc}-{kind}", tname, enc, float_column(tname, kind, n, seed)))
    cases.append(("BOOLEAN-None", "BOOLEAN", None, (np.arange(n) % 3 == 0).astype(np.uint8)))
    cases.append(("DATE-None", "DATE", None, (8000 + np.arange(n) % 10000).astype(np.int32)))
    cases.append(("TIMESTAMP-None", "TIMESTAMP", None, (1_735_689_600_000_000 + np.arange(n, dtype=np.int64) * 1_000_000)))
    cases.append(("DECIMAL10-None", "DECIMAL", None, (np.arange(n) % 10_000_000 * 100).astype(np.int64), 10, 2))
    cases.append(("DECIMAL4-None", "DECIMAL", None, (np.arange(n) % 9999 - 5000).astype(np.int16), 4, 1))
    cases.append(("DECIMAL9-None", "DECIMAL", None, (np.arange(n) * 7919 % 999_999_999).astype(np.int32), 9, 3))
    return cases


def column_spec(case):
    name, tname, enc, arr = case[:4]
    extra = tuple(case[4:6]) if len(case) > 4 else ()
    return (name[:60], tname, enc) + extra


def width_sweep_cases(type_name, seed=0):
    """FFOR and DELTA at every bit width 0..T (BASELINE config 4's sweep) on one vector each + a partial tail."""
    dt = np.dtype(INT_TYPES[type_name])
    bits = dt.itemsize * 8
    r = _rng(seed)
    udt = np.dtype(f"u{dt.itemsize}")
    cases = []
    n = 1024 + 300
    for w in range(bits + 1):
        hi = (1 << w) - 1
        raw = r.integers(0, hi, n, dtype=np.uint64, endpoint=True) if w else np.zeros(n, dtype=np.uint64)
        raw[0] = hi
        raw[1] = 0
        x = (raw + np.uint64(5 if w < bits else 0)).astype(udt).view(udt)
        cases.append((f"{type_name}-FFOR-w{w}", type_name, "FFOR", x.view(dt).copy()))
        gaps = raw.copy()
        cs = np.cumsum(gaps.astype(udt), dtype=udt)  # wraps: DELTA uses modular arithmetic
        cases.append((f"{type_name}-DELTA-w{w}", type_name, "DELTA", cs.view(dt).copy()))
    return cases

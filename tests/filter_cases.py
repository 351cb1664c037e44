"""Filter pushdown cases shared by the CPU and GPU tests: a lineitem-shaped table (TPC-H Q6 columns) + edge-case columns."""
import numpy as np


def bits(x, dtype):
    """Physical constant of a filter: integers sign/zero-extended to 64 bits, floats as raw IEEE bits."""
    dt = np.dtype(dtype)
    if dt.kind == "f":
        return int(np.array([x], dtype=dt).view(np.uint32 if dt.itemsize == 4 else np.uint64)[0])
    return int(np.array([x], dtype=dt).astype(np.int64 if dt.kind == "i" else np.uint64).view(np.uint64)[0])


def table(n, seed=5):
    r = np.random.default_rng(seed)
    shipdate = np.sort(r.integers(8036, 10592, n)).astype(np.int32)  # clustered dates: zone maps prune
    rnd_date = r.integers(8036, 10592, n).astype(np.int32)
    discount = r.integers(0, 11, n).astype(np.int64)                  # DECIMAL(15,2) unscaled
    quantity = (r.integers(1, 51, n) * 100).astype(np.int64)
    price = r.integers(90000, 10500000, n).astype(np.int64)
    small = r.integers(-128, 128, n).astype(np.int8)
    u16 = r.integers(0, 65536, n).astype(np.uint16)
    dbl = np.round(r.random(n) * 1000, 2)
    dbl[::97] = np.nan
    dbl[5::131] = -0.0
    flt = (r.integers(-500, 500, n) / 4).astype(np.float32)
    runs = np.repeat(r.integers(0, 1000, n // 300 + 1), 300)[:n].astype(np.int32)  # RLE / CROSS_RLE friendly
    dic = r.integers(0, 7, n).astype(np.int64) * 1_000_003                         # DICT
    cols = [("shipdate", "DATE", None, shipdate), ("rnd_date", "DATE", None, rnd_date), ("discount", "DECIMAL", None, discount, 15, 2),
            ("quantity", "DECIMAL", None, quantity, 15, 2), ("price", "DECIMAL", None, price, 15, 2), ("small", "INT8", None, small),
            ("u16", "UINT16", None, u16), ("dbl", "DOUBLE", None, dbl), ("flt", "FLOAT", None, flt), ("runs", "INT32", None, runs),
            ("dic", "INT64", None, dic)]
    return cols


def filter_sets():
    """name -> [(column name, op, numpy-typed constant)]"""
    return {
        "q6": [("shipdate", "GE", np.int32(8766)), ("shipdate", "LT", np.int32(9131)), ("discount", "GE", np.int64(5)),
               ("discount", "LE", np.int64(7)), ("quantity", "LT", np.int64(2400))],
        "rnd_date_eq": [("rnd_date", "EQ", np.int32(9000))],
        "small_neg": [("small", "LT", np.int8(-100))],
        "small_ne": [("small", "NE", np.int8(0))],
        "u16_hi": [("u16", "GT", np.uint16(65000))],
        "dbl_nan_gt": [("dbl", "GT", np.float64(999.0))],     # NaN rows pass (NaN is the greatest value)
        "dbl_nan_eq": [("dbl", "EQ", np.float64("nan"))],
        "dbl_zero": [("dbl", "EQ", np.float64(0.0))],         # -0.0 == 0.0
        "dbl_le": [("dbl", "LE", np.float64(10.5))],
        "flt_ge": [("flt", "GE", np.float32(100.25))],
        "runs_band": [("runs", "GE", np.int32(200)), ("runs", "LE", np.int32(400))],
        "dic_eq": [("dic", "EQ", np.int64(3 * 1_000_003))],
        "none_pass": [("price", "LT", np.int64(0))],
        "all_pass": [("price", "GE", np.int64(0))],
    }


def numpy_mask(cols, fs):
    by_name = {c[0]: c[3] for c in cols}
    n = len(cols[0][3])
    m = np.ones(n, dtype=bool)
    for name, op, k in fs:
        v = by_name[name]
        if v.dtype.kind == "f":
            vn, kn = np.isnan(v), bool(np.isnan(k))
            eq = (vn & kn) | (~vn & (not kn) & (v == k))
            lt = ~vn & (np.ones(n, dtype=bool) if kn else (v < k))
        else:
            eq, lt = v == k, v < k
        m &= {"EQ": eq, "NE": ~eq, "LT": lt, "LE": lt | eq, "GT": ~lt & ~eq, "GE": ~lt}[op]
    return m

"""EXP_EQUAL (multi-column compression: a column that repeats another column of the same row group is stored as a reference;
the reference checks for it in FastLanesReader::HasMccEncoding, src/reader/fls_reader.cpp:43-82) — shared test table."""
import numpy as np


def table(n, rg_rows):
    rng = np.random.default_rng(5)
    a = np.cumsum(rng.integers(0, 50, n)).astype(np.int64)
    c = a.copy()
    if n > rg_rows:  # equal in every row group but the second
        c[rg_rows + 3] += 1
    d = (rng.integers(-1000, 1000, n)).astype(np.int32)
    g = np.round(rng.random(n) * 1000, 2)
    specs = [("a", "INT64", None), ("b", "INT64", None), ("c", "INT64", None), ("d", "INT32", None), ("e", "INT32", None),
             ("f", "UINT64", None), ("g", "DOUBLE", None), ("h", "DOUBLE", None), ("forced", "INT64", "FFOR")]
    data = [a, a.copy(), c, d, d.copy(), a.astype(np.uint64), g, g.copy(), a.copy()]
    # expected source column per (column): b -> a everywhere, c -> a except row group 1, e -> d, h -> g;
    # f differs in signedness and `forced` carries an explicit encoding: both stay ordinary chunks
    return specs, data

"""TPC-H lineitem-shaped table (BASELINE configs 3 and 5): the 16 columns with dbgen's value distributions restated in numpy
(dbgen itself is not in the image), generated in order-key order like dbgen does.  Strings are (offsets u64[n+1], bytes u8)
pairs, the vectorised form the writer takes.  Used by tests/test_lineitem*.py at a few row groups and by
tools/lineitem_bench.py at SF1..SF10."""
import numpy as np

ROWS_PER_SF = 6_000_000  # dbgen: 6 001 215 at SF1
DATE_LO, DATE_HI = 8035, 10440  # 1992-01-01 .. 1998-08-02 (order dates), days since 1970-01-01

INSTRUCT = [b"DELIVER IN PERSON", b"COLLECT COD", b"NONE", b"TAKE BACK RETURN"]
MODE = [b"REG AIR", b"AIR", b"RAIL", b"SHIP", b"TRUCK", b"MAIL", b"FOB"]
_WORDS = (b"furiously carefully quickly slyly blithely ironic final regular express special pending bold even silent unusual "
          b"deposits requests packages accounts instructions theodolites foxes pinto beans dependencies platelets ideas asymptotes "
          b"sleep wake nag haggle cajole boost detect integrate use about above across after against along among the of are").split()


def strings_from_choices(choices, idx):
    """(offsets, bytes) of choices[idx[i]] for every i, without Python string objects."""
    lens_c = np.array([len(c) for c in choices], dtype=np.int64)
    width = int(lens_c.max())
    mat = np.zeros((len(choices), width), dtype=np.uint8)
    for k, c in enumerate(choices):
        mat[k, :len(c)] = np.frombuffer(c, dtype=np.uint8)
    lens = lens_c[idx]
    off = np.zeros(len(idx) + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    data = mat[idx][np.arange(width)[None, :] < lens[:, None]]
    return off, np.ascontiguousarray(data)


_POOL = {}


def _comment_pool(n_pool=1 << 18):
    """dbgen draws l_comment from a fixed text pool; here: n_pool comments of 10..43 characters built from a small grammar
    vocabulary (words separated by blanks, cut to a random length), as a [n_pool, 44] byte matrix + lengths."""
    if n_pool not in _POOL:
        r = np.random.default_rng(1234)
        n_words = 9  # enough words to cover 43 characters
        widx = r.integers(0, len(_WORDS), (n_pool, n_words))
        wl = np.array([len(w) + 1 for w in _WORDS], dtype=np.int64)  # with the trailing space
        width = int(wl.max())
        wmat = np.full((len(_WORDS), width), ord(" "), dtype=np.uint8)
        for k, w in enumerate(_WORDS):
            wmat[k, :len(w)] = np.frombuffer(w, dtype=np.uint8)
        out = np.zeros((n_pool, 44 + width), dtype=np.uint8)
        pos = np.zeros(n_pool, dtype=np.int64)
        rows = np.arange(n_pool)
        for j in range(n_words):
            l = wl[widx[:, j]]
            for b in range(width):
                m = (b < l) & (pos + b < 44)
                out[rows[m], (pos + b)[m]] = wmat[widx[m, j], b]
            pos += l
        _POOL[n_pool] = (np.ascontiguousarray(out[:, :44]), r.integers(10, 44, n_pool))
    return _POOL[n_pool]


def comments(n, r):
    """l_comment of n rows: (offsets, bytes), sampled from the text pool."""
    mat, pool_len = _comment_pool()
    idx = r.integers(0, len(pool_len), n)
    lens = pool_len[idx]
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    data = mat[idx][np.arange(44)[None, :] < lens[:, None]]
    return off, np.ascontiguousarray(data)


def lineitem(n_rows, sf=1.0, seed=7, first_order=0, strings=True):
    """[(name, fls type, forced encoding (None = the writer's selection), data[, width, scale])] of n_rows lineitem rows.
    strings=False: the 11 integer / decimal / date columns only (what BASELINE config 5's SUM/COUNT scans read)."""
    r = np.random.default_rng(seed)
    # orders of 1..7 lines; order keys are sparse like dbgen's (8 used out of every 32), ascending
    n_orders = n_rows // 3 + 64  # 4 lines per order on average: comfortably more orders than the rows need
    lines = r.integers(1, 8, n_orders)
    order_of = np.repeat(np.arange(n_orders, dtype=np.int64), lines)[:n_rows]
    first = np.concatenate(([0], np.cumsum(lines)))[:-1]
    linenumber = (np.arange(n_rows, dtype=np.int64) - first[order_of] + 1).astype(np.int32)
    o = order_of + first_order
    orderkey = (o // 8) * 32 + o % 8 + 1
    n_part, n_supp = max(1, int(200_000 * sf)), max(1, int(10_000 * sf))
    partkey = r.integers(1, n_part + 1, n_rows).astype(np.int64)
    suppkey = ((partkey + r.integers(0, 4, n_rows) * (n_supp // 4 + (partkey - 1) // n_supp)) % n_supp + 1).astype(np.int64)
    quantity = r.integers(1, 51, n_rows).astype(np.int64)
    retail = 90000 + (partkey // 10) % 20001 + 100 * (partkey % 1000)  # cents (dbgen's part retail price)
    extendedprice = quantity * retail
    discount = r.integers(0, 11, n_rows).astype(np.int64)
    tax = r.integers(0, 9, n_rows).astype(np.int64)
    orderdate = r.integers(DATE_LO, DATE_HI + 1, n_orders)[order_of]
    shipdate = (orderdate + r.integers(1, 122, n_rows)).astype(np.int32)
    commitdate = (orderdate + r.integers(30, 91, n_rows)).astype(np.int32)
    receiptdate = (shipdate + r.integers(1, 31, n_rows)).astype(np.int32)
    today = 9298  # 1995-06-17: dbgen's CURRENTDATE
    returnflag = np.where(receiptdate <= today, r.integers(0, 2, n_rows), 2)  # R / A, else N
    linestatus = (shipdate > today).astype(np.int64)  # F / O
    if not strings:
        return [
            ("l_orderkey", "INT64", None, orderkey),
            ("l_partkey", "INT64", None, partkey),
            ("l_suppkey", "INT64", None, suppkey),
            ("l_linenumber", "INT32", None, linenumber),
            ("l_quantity", "DECIMAL", None, quantity * 100, 15, 2),
            ("l_extendedprice", "DECIMAL", None, extendedprice, 15, 2),
            ("l_discount", "DECIMAL", None, discount, 15, 2),
            ("l_tax", "DECIMAL", None, tax, 15, 2),
            ("l_shipdate", "DATE", None, shipdate),
            ("l_commitdate", "DATE", None, commitdate),
            ("l_receiptdate", "DATE", None, receiptdate),
        ]
    return [
        ("l_orderkey", "INT64", None, orderkey),
        ("l_partkey", "INT64", None, partkey),
        ("l_suppkey", "INT64", None, suppkey),
        ("l_linenumber", "INT32", None, linenumber),
        ("l_quantity", "DECIMAL", None, quantity * 100, 15, 2),
        ("l_extendedprice", "DECIMAL", None, extendedprice, 15, 2),
        ("l_discount", "DECIMAL", None, discount, 15, 2),
        ("l_tax", "DECIMAL", None, tax, 15, 2),
        ("l_returnflag", "VARCHAR", None, strings_from_choices([b"R", b"A", b"N"], returnflag)),
        ("l_linestatus", "VARCHAR", None, strings_from_choices([b"F", b"O"], linestatus)),
        ("l_shipdate", "DATE", None, shipdate),
        ("l_commitdate", "DATE", None, commitdate),
        ("l_receiptdate", "DATE", None, receiptdate),
        ("l_shipinstruct", "VARCHAR", None, strings_from_choices(INSTRUCT, r.integers(0, 4, n_rows))),
        ("l_shipmode", "VARCHAR", None, strings_from_choices(MODE, r.integers(0, 7, n_rows))),
        ("l_comment", "VARCHAR", None, comments(n_rows, r)),
    ]


def specs(cols):
    return [(c[0], c[1], c[2]) + tuple(c[4:6]) for c in cols]


def string_list(col_data):
    off, data = col_data
    b = data.tobytes()
    return [b[int(off[i]):int(off[i + 1])] for i in range(len(off) - 1)]

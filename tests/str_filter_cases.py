"""String-filter test inputs shared by the CPU (oracle vs plain Python) and GPU (C ABI vs oracle) tests: constant
comparisons on VARCHAR / BLOB columns as FilterExecutor::Apply evaluates them (reference src/reader/filter_executor.cpp:19-60
-> ColumnSegment::FilterSelection; DuckDB string_t order: bytewise, the shorter string first on a tie)."""
import operator

import numpy as np

from str_cases import cases

PY_OP = dict(EQ=operator.eq, NE=operator.ne, LT=operator.lt, LE=operator.le, GT=operator.gt, GE=operator.ge)
COLUMNS = ["shipmode", "instruct", "comment", "blob", "blob_plain", "len_edges", "long", "constant", "empty", "comment12", "instruct12",
           "flag_runs", "status_cross", "freq_none"]


def table(n):
    """(specs, data): the string columns of str_cases.cases plus an INT64 column for mixed filter sets."""
    cs = cases(n)
    specs = [(name, cs[name][0], cs[name][1]) for name in COLUMNS] + [("v", "INT64", None)]
    data = [cs[name][2] for name in COLUMNS] + [(np.arange(n, dtype=np.int64) * 7) % 1000]
    return specs, data


def filter_sets(data):
    """name -> [(column name, op, constant)]; constants are bytes for string columns, ints for `v`."""
    blob = data[COLUMNS.index("blob")]
    comment = data[COLUMNS.index("comment")]
    long_col = data[COLUMNS.index("long")]
    some_blob = next(b for b in blob if len(b) > 13)
    sets = {
        "dict_eq": [("shipmode", "EQ", b"REG AIR")],
        "dict_ne": [("shipmode", "NE", b"AIR")],
        "dict_range": [("shipmode", "GE", b"MAIL"), ("shipmode", "LT", b"SHIP")],
        "dict_prefix_of_key": [("shipmode", "GT", b"REG")],            # "REG" < "REG AIR": decided by length
        "fsst_dict_eq_long": [("instruct", "EQ", b"DELIVER IN PERSON")],  # > 12 bytes: walks the heap
        "fsst_dict_lt": [("instruct", "LT", b"DELIVER IN PERSOO")],
        "fsst_eq_row": [("comment", "EQ", comment[len(comment) // 2])],
        "fsst_ge": [("comment", "GE", b"furiously regular")],
        "fsst_lt_prefix": [("comment", "LT", b"slyl")],                 # exactly the 4 prefix bytes
        "blob_eq": [("blob", "EQ", some_blob)],
        "blob_gt_ff": [("blob", "GT", b"\xff")],
        "blob_le_nul": [("blob_plain", "LE", b"\x00\x00")],
        "blob_ge_empty": [("blob_plain", "GE", b"")],                   # everything passes
        "blob_lt_empty": [("blob_plain", "LT", b"")],                   # nothing passes
        "blob_eq_empty": [("blob_plain", "EQ", b"")],
        "len_edges_12": [("len_edges", "GT", b"x" * 12)],
        "len_edges_13": [("len_edges", "LE", b"x" * 13)],
        "len_edges_eq": [("len_edges", "EQ", b"x" * 29)],
        "long_eq": [("long", "EQ", long_col[min(len(long_col) - 1, 1234)])],
        "long_lt": [("long", "LT", b"00100:abcdefghijabcdefghij")],
        "constant_eq": [("constant", "EQ", b"DELIVER IN PERSON")],
        "constant_ne": [("constant", "NE", b"DELIVER IN PERSON")],
        "empty_eq": [("empty", "EQ", b"")],
        "empty_gt": [("empty", "GT", b"")],
        "fsst12_ge": [("comment12", "GE", b"quickly")],
        "fsst12_dict_eq": [("instruct12", "EQ", b"NONE")],
        "rle_str_eq_long": [("flag_runs", "EQ", b"RETURNED TO SENDER")],
        "rle_str_lt": [("flag_runs", "LT", b"N")],
        "cross_rle_str_ge": [("status_cross", "GE", b"P")],
        "frequency_str_ne": [("freq_none", "NE", b"NONE")],
        "mixed": [("v", "GE", 100), ("shipmode", "EQ", b"RAIL"), ("v", "LT", 900), ("comment", "GE", b"c")],
        "two_strings": [("shipmode", "NE", b"FOB"), ("instruct", "NE", b"NONE")],
    }
    return sets


def python_mask(specs, data, fs):
    names = [s[0] for s in specs]
    n = len(data[0])
    keep = np.ones(n, dtype=bool)
    for col, op, k in fs:
        values = data[names.index(col)]
        f = PY_OP[op]
        if isinstance(k, (bytes, bytearray)):
            keep &= np.fromiter((f(v, k) for v in values), dtype=bool, count=n)
        else:
            keep &= f(np.asarray(values), k)
    return keep

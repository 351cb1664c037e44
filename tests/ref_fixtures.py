"""The numeric tables of the reference's own round-trip test (reference test/all_types_single_threaded.test:36-160),
restated as numpy generators: every table has `id BIGINT = 1..N` plus one `col`.  N = 1024 in the reference; BASELINE
config 1 scales the same pattern to 1 000 000 rows."""
import numpy as np


def reference_test_tables(n=1024):
    i = np.arange(1, n + 1, dtype=np.int64)
    return {
        # name: (fls type, col values[, precision, scale])
        "bool_table": ("BOOLEAN", (i % 2 == 0).astype(np.uint8)),                      # :36-41
        "tinyint_table": ("INT8", ((i % 256) - 128).astype(np.int8)),                  # :43-48
        "smallint_table": ("INT16", ((i % 65536) - 32768).astype(np.int16)),           # :50-55
        "int_table": ("INT32", i.astype(np.int32)),                                    # :57-62
        "bigint_table": ("INT64", i * 10),                                             # :64-69
        "ubigint_table": ("UINT64", (i * 10).astype(np.uint64)),                       # :71-76
        "float_table": ("FLOAT", i.astype(np.float32) / np.float32(100.0)),            # :92-97
        "double_table": ("DOUBLE", i.astype(np.float64) / 1000.0),                     # :99-104
        "decimal_table": ("DOUBLE", (i.astype(np.float64) / 10.0)),                    # :106-111 (really DOUBLE)
        "date_table": ("TIMESTAMP", (8116 + i) * 86_400_000_000),                      # :120-125 (DATE + INTERVAL -> TIMESTAMP)
        "timestamp_table": ("TIMESTAMP", 1_735_689_600_000_000 + i * 1_000_000),       # :127-132
    }


def reference_string_tables(n=1024):
    """The string-typed tables of the same test: VARCHAR (:113-118), BLOB (:155-160), and HUGEINT / UHUGEINT (:78-90), which
    the reference writes as decimal strings (src/writer/fls_view_writer.cpp:20-46, Int128ViewWriterFactory)."""
    import hashlib
    return {
        "varchar_table": ("VARCHAR", [b"Value %d" % i for i in range(1, n + 1)]),
        "blob_table": ("BLOB", [hashlib.md5(str(i).encode()).hexdigest().encode() for i in range(1, n + 1)]),
        "hugeint_table": ("VARCHAR", [str(i * -469231731687303715884105728).encode() for i in range(1, n + 1)]),
        "uhugeint_table": ("VARCHAR", [str(i * 469231731687303715884105728).encode() for i in range(1, n + 1)]),
    }

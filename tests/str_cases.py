"""String column test inputs shared by the CPU (oracle) and GPU parity tests.  The VARCHAR / BLOB generators restate the
reference's own test tables (test/all_types_single_threaded.test: 'Value ' || i style VARCHAR, BLOB of encoded text)."""
import numpy as np

WORDS = [b"furiously", b"regular", b"deposits", b"sleep", b"quickly", b"above", b"the", b"slyly", b"ironic", b"packages",
         b"carefully", b"final", b"accounts", b"boost", b"blithely", b"pending", b"requests", b"haggle", b"express", b"theodolites"]


def comment_like(n, seed=7):
    """TPC-H l_comment-shaped text: 2..8 words, high cardinality -> FSST."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        k = int(rng.integers(2, 9))
        out.append(b" ".join(WORDS[int(j)] for j in rng.integers(0, len(WORDS), k)))
    return out


def runs_of(n, values, max_run, seed):
    """n strings in runs of 1..max_run rows drawn from `values` (neighbouring runs differ)."""
    rng = np.random.default_rng(seed)
    out, prev = [], -1
    while len(out) < n:
        k = int(rng.integers(0, len(values)))
        if k == prev:
            k = (k + 1) % len(values)
        prev = k
        out += [values[k]] * int(rng.integers(1, max_run + 1))
    return out[:n]


def mostly(n, frequent, others, every, seed):
    """`frequent` everywhere except about one row in `every`, which draws from `others`."""
    rng = np.random.default_rng(seed)
    hit = rng.integers(0, every, n) == 0
    pick = rng.integers(0, len(others), n)
    return [others[int(pick[i])] if hit[i] else frequent for i in range(n)]


def cases(n):
    rng = np.random.default_rng(11)
    binary = [bytes(rng.integers(0, 256, int(rng.integers(0, 40)), dtype=np.uint8)) for _ in range(n)]  # includes 0xFF bytes
    return {
        "value_i": ("VARCHAR", None, [b"Value %d" % (i % 1000) for i in range(n)]),           # reference test generator, dict
        "value_unique": ("VARCHAR", None, [b"Value %d" % i for i in range(n)]),                 # all distinct, short (inlined)
        "comment": ("VARCHAR", None, comment_like(n)),                                           # FSST
        "constant": ("VARCHAR", None, [b"DELIVER IN PERSON"] * n),                               # STR_CONSTANT (> 12 bytes)
        "empty": ("VARCHAR", None, [b""] * n),
        "shipmode": ("VARCHAR", "DICT_STR", [[b"AIR", b"RAIL", b"TRUCK", b"REG AIR", b"SHIP", b"FOB", b"MAIL"][i % 7] for i in range(n)]),
        "instruct": ("VARCHAR", "FSST_DICT", [[b"DELIVER IN PERSON", b"COLLECT COD", b"NONE", b"TAKE BACK RETURN"][(i * 7) % 4] for i in range(n)]),
        "blob": ("BLOB", "FSST", binary),                                                         # escapes, 0xFF literals
        "blob_plain": ("BLOB", "STR_UNCOMPRESSED", binary),
        "comment12": ("VARCHAR", "FSST12", comment_like(n, seed=13)),                              # 12-bit codes, odd / even code counts
        "blob12": ("BLOB", "FSST12", binary),                                                     # mostly literal codes (< 256)
        "instruct12": ("VARCHAR", "FSST12_DICT", [[b"DELIVER IN PERSON", b"COLLECT COD", b"NONE", b"TAKE BACK RETURN"][(i * 5) % 4] for i in range(n)]),
        "len_edges": ("VARCHAR", "FSST", [b"x" * (i % 30) for i in range(n)]),                   # lengths around the 12-byte inline limit
        # the three VALUE encodings (rle_map_kernel.hpp:26-61, cross_rle_kernel.hpp:55-97, frequency_kernel.hpp:23-69)
        "flag_runs": ("VARCHAR", "RLE_STR", runs_of(n, [b"A", b"N", b"R", b"RETURNED TO SENDER", b""], 200, 21)),
        "rle_unique": ("VARCHAR", "RLE_STR", [b"Value %d" % i for i in range(n)]),               # 1024 runs per vector
        "status_cross": ("VARCHAR", "CROSS_RLE_STR", runs_of(n, [b"F", b"O", b"P", b"PENDING REVIEW BY CLERK"], 5000, 22)),
        "cross_short": ("BLOB", "CROSS_RLE_STR", runs_of(n, [b"\x00", b"\xff\xfe", b"abcdefghijklm", b""], 3, 23)),
        "freq_none": ("VARCHAR", "FREQUENCY_STR", mostly(n, b"NONE", [b"DELIVER IN PERSON", b"COLLECT COD", b"", b"x"], 30, 24)),
        "freq_const": ("VARCHAR", "FREQUENCY_STR", [b"TAKE BACK RETURN"] * n),                    # no exceptions at all
        "long": ("VARCHAR", "STR_UNCOMPRESSED", [b"%05d:" % i + b"abcdefghij" * (1 + i % 50) for i in range(n)]),
    }

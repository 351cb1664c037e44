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
        "long": ("VARCHAR", "STR_UNCOMPRESSED", [b"%05d:" % i + b"abcdefghij" * (1 + i % 50) for i in range(n)]),
    }

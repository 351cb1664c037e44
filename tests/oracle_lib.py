"""ctypes wrapper of oracle/liboracle.so — the CPU restatement used as the CHECKER in tests (never by the product)."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")


def build():
    srcs = ["fls_oracle.c", "fls_oracle_scan.c", "fls_oracle_str.c", "fls_oracle_tmpl.h", "fls_oracle.h"]
    stale = not os.path.exists(LIB) or any(
        os.path.getmtime(os.path.join(ORACLE_DIR, s)) > os.path.getmtime(LIB) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)


def build_native():
    """The oracle compiled for THIS machine's cores (-march=native) for the timed CPU baselines of bench.py: the library
    that travels with the repo is built -march=x86-64-v3 so that it runs wherever the snapshot lands.  Returns the path, or
    None when no compiler is at hand (the portable build is then timed, and the bench line says so)."""
    out = os.path.join(ORACLE_DIR, "_ref", "liboracle_native.so")
    try:
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["gcc", "-O3", "-march=native", "-ffp-contract=off", "-fPIC", "-std=gnu11", "-I" + os.path.join(ROOT, "include"),
                               "-shared", "-o", out] + [os.path.join(ORACLE_DIR, f) for f in ("fls_oracle.c", "fls_oracle_scan.c", "fls_oracle_str.c")]
                              + ["-lpthread"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        return out
    except (OSError, subprocess.CalledProcessError):
        return None


build()
NATIVE = build_native() if os.environ.get("FLS_ORACLE_NATIVE") == "1" else None
# FLS_ORACLE_LIB: another build of the same sources (tools/sanitize.sh: AddressSanitizer / UBSan)
lib = C.CDLL(os.environ.get("FLS_ORACLE_LIB") or NATIVE or LIB)
MARCH = "native" if NATIVE else "x86-64-v3"


class OracleTable(C.Structure):
    _fields_ = [("bytes", C.c_void_p), ("size", C.c_uint64), ("header", C.c_uint8 * 64), ("columns", C.c_void_p),
                ("rowgroups", C.c_void_p), ("chunks", C.c_void_p)]


lib.fls_oracle_alp_decode_f64.restype = C.c_double
lib.fls_oracle_alp_decode_f64.argtypes = [C.c_int64, C.c_int, C.c_int]
lib.fls_oracle_alp_decode_f32.restype = C.c_float
lib.fls_oracle_alp_decode_f32.argtypes = [C.c_int32, C.c_int, C.c_int]


class OracleFilter(C.Structure):
    _fields_ = [("col", C.c_uint32), ("op", C.c_int32), ("constant", C.c_uint64), ("str_data", C.c_char_p), ("str_len", C.c_uint64)]


def _oracle_filters(filters):
    """[(col, op, constant)]: an int (physical bits) for numeric columns, bytes for VARCHAR / BLOB columns."""
    out = []
    for c, op, k in filters:
        if isinstance(k, (bytes, bytearray)):
            out.append(OracleFilter(c, op, 0, bytes(k), len(k)))
        else:
            out.append(OracleFilter(c, op, int(k) & 0xFFFFFFFFFFFFFFFF, None, 0))
    return (OracleFilter * len(out))(*out)


class Oracle:
    def __init__(self, image):
        self.image = np.ascontiguousarray(image, dtype=np.uint8)
        self.t = OracleTable()
        rc = lib.fls_oracle_open(C.c_void_p(self.image.ctypes.data), C.c_uint64(self.image.size), C.byref(self.t))
        if rc:
            raise RuntimeError(f"fls_oracle_open failed: {rc}")

    def decode_column(self, col, rg_begin, n_rg, out_type, out_dtype, vec_per_rg, n_threads=1):
        out = np.zeros(n_rg * vec_per_rg * 1024, dtype=out_dtype)
        rc = lib.fls_oracle_decode_column(C.byref(self.t), col, rg_begin, n_rg, out_type, C.c_void_p(out.ctypes.data),
                                          n_threads)
        if rc:
            raise RuntimeError(f"fls_oracle_decode_column failed: {rc}")
        return out

    def scan(self, cols, rg_begin, n_rg, n_threads):
        ca = (C.c_uint32 * len(cols))(*cols)
        nbytes, cks = C.c_uint64(), C.c_uint64()
        rc = lib.fls_oracle_scan(C.byref(self.t), ca, len(cols), rg_begin, n_rg, n_threads, C.byref(nbytes), C.byref(cks))
        if rc:
            raise RuntimeError(f"fls_oracle_scan failed: {rc}")
        return nbytes.value, cks.value

    def decode_codes(self, rg, col, vec):
        codes = np.zeros(1024, dtype=np.uint32)
        if lib.fls_oracle_decode_codes(C.byref(self.t), rg, col, vec, C.c_void_p(codes.ctypes.data)):
            raise RuntimeError("fls_oracle_decode_codes failed")
        return codes

    def string_heap_bytes(self, col, rg_begin, n_rg):
        n = C.c_uint64()
        if lib.fls_oracle_string_heap_bytes(C.byref(self.t), col, rg_begin, n_rg, C.byref(n)):
            raise RuntimeError("fls_oracle_string_heap_bytes failed")
        return n.value

    def decode_strings(self, col, rg_begin, n_rg, vec_per_rg, heap_base=0):
        """(string_t array as a (n, 16) uint8 view source, heap bytes): flat materialisation of a string column."""
        out = np.zeros((n_rg * vec_per_rg * 1024, 16), dtype=np.uint8)
        heap = np.zeros(max(16, self.string_heap_bytes(col, rg_begin, n_rg)), dtype=np.uint8)
        rc = lib.fls_oracle_decode_strings(C.byref(self.t), col, rg_begin, n_rg, C.c_void_p(out.ctypes.data),
                                           C.c_void_p(heap.ctypes.data), C.c_uint64(heap_base))
        if rc:
            raise RuntimeError(f"fls_oracle_decode_strings failed: {rc}")
        return out, heap

    def decode_string_dict(self, col, rg, n_keys, heap_bytes, heap_base=0):
        out = np.zeros((n_keys, 16), dtype=np.uint8)
        heap = np.zeros(max(16, heap_bytes), dtype=np.uint8)
        rc = lib.fls_oracle_decode_string_dict(C.byref(self.t), col, rg, C.c_void_p(out.ctypes.data),
                                               C.c_void_p(heap.ctypes.data), C.c_uint64(heap_base))
        if rc:
            raise RuntimeError(f"fls_oracle_decode_string_dict failed: {rc}")
        return out, heap

    def filter_mask(self, filters, rg_begin, n_rg, vec_per_rg):
        """filters: [(col, op, constant_bits)] -> uint32 mask [n_rg * vec_per_rg * 32]."""
        arr = _oracle_filters(filters)
        mask = np.zeros(n_rg * vec_per_rg * 32, dtype=np.uint32)
        rc = lib.fls_oracle_filter_mask(C.byref(self.t), arr, len(filters), rg_begin, n_rg, C.c_void_p(mask.ctypes.data))
        if rc:
            raise RuntimeError(f"fls_oracle_filter_mask failed: {rc}")
        return mask

    def rowgroup_may_satisfy(self, filters, rg):
        arr = _oracle_filters(filters)
        return bool(lib.fls_oracle_rowgroup_may_satisfy(C.byref(self.t), arr, len(filters), rg))

    def aggregate_masked(self, col, rg_begin, n_rg, mask):
        lo, hi, cnt, fs = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_double()
        mask = np.ascontiguousarray(mask, dtype=np.uint32)
        rc = lib.fls_oracle_aggregate_masked(C.byref(self.t), col, rg_begin, n_rg, C.c_void_p(mask.ctypes.data), C.byref(lo),
                                             C.byref(hi), C.byref(cnt), C.byref(fs))
        if rc:
            raise RuntimeError(f"fls_oracle_aggregate_masked failed: {rc}")
        s = (hi.value << 64) | lo.value
        return (s - (1 << 128) if s >= (1 << 127) else s), cnt.value, fs.value

    def aggregate(self, col, rg_begin, n_rg):
        lo, hi, cnt = C.c_uint64(), C.c_uint64(), C.c_uint64()
        mn, mx, fs = C.c_int64(), C.c_int64(), C.c_double()
        rc = lib.fls_oracle_aggregate(C.byref(self.t), col, rg_begin, n_rg, C.byref(lo), C.byref(hi), C.byref(cnt),
                                      C.byref(mn), C.byref(mx), C.byref(fs))
        if rc:
            raise RuntimeError(f"fls_oracle_aggregate failed: {rc}")
        s = (hi.value << 64) | lo.value
        if s >= 1 << 127:
            s -= 1 << 128
        return dict(sum=s, count=cnt.value, min=mn.value, max=mx.value, fsum=fs.value)


def fl_index(row, lane):
    return lib.fls_oracle_index(row, lane)


def fl_transpose(pos):
    return lib.fls_oracle_transpose(pos)


def unpack(packed, T, W):
    packed = np.ascontiguousarray(packed)
    out = np.zeros(1024, dtype=np.dtype(f"u{T // 8}"))
    lib.fls_oracle_unpack(C.c_void_p(packed.ctypes.data), T, W, C.c_void_p(out.ctypes.data))
    return out


lib.fls_oracle_fsst_decompress.restype = C.c_uint64
lib.fls_oracle_fsst_decompress.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]


def fsst_decode_stream(symbol, length, codes, comp_end, dec_cap):
    """Oracle FSST decoder (fls_oracle_str.c) over back-to-back code streams: (decoded bytes, decoded end offsets)."""
    table = np.zeros(256 * 8 + 256, dtype=np.uint8)  # fls_fsst_table: symbol[256] u64 | len[256] u8
    table[:2048] = np.ascontiguousarray(symbol, dtype="<u8").view(np.uint8)
    table[2048:] = length
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    out = np.zeros(dec_cap + 16, dtype=np.uint8)
    ends = np.zeros(len(comp_end), dtype=np.uint64)
    pos = prev = 0
    for i, e in enumerate(comp_end):
        e = int(e)
        pos += lib.fls_oracle_fsst_decompress(C.c_void_p(table.ctypes.data), C.c_void_p(codes.ctypes.data + prev), e - prev,
                                              C.c_void_p(out.ctypes.data + pos))
        ends[i] = pos
        prev = e
    return out[:pos].copy(), ends

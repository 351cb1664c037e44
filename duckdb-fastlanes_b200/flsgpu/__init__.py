"""ctypes binding of libflsgpu.so (the C ABI in include/flsgpu.h).

This is the Python test/bench driver over the same entry points the DuckDB extension's C++ reader calls; it adds
no decode logic of its own.  There is no CPU fallback: if the shared library is missing the import fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FLSGPU_LIB", os.path.join(os.path.dirname(_HERE), "libflsgpu.so"))
if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python __graft_entry__.py build` "
        "(flsgpu has no CPU fallback and will not run without its CUDA library)"
    )
lib = C.CDLL(LIB_PATH)

# ---- enums (include/fls_format.h, include/flsgpu.h) -------------------------------------------------------------
TYPE = dict(INT8=1, INT16=2, INT32=3, INT64=4, UINT8=5, UINT16=6, UINT32=7, UINT64=8, FLOAT=9, DOUBLE=10,
            BOOLEAN=11, DATE=12, TIMESTAMP=13, DECIMAL=14, VARCHAR=15, BLOB=16)
ENC = dict(UNCOMPRESSED=0, CONSTANT=1, FFOR=2, DELTA=3, DICT=4, RLE=5, ALP=6, ALP_RD=7, CROSS_RLE=8, FREQUENCY=9,
           SLPATCH=10, DICT_STR=11, STR_UNCOMPRESSED=12, STR_CONSTANT=13, FSST=14, FSST_DICT=15, FSST12=16, FSST12_DICT=17,
           RLE_STR=18, CROSS_RLE_STR=19, FREQUENCY_STR=20)
STRING_TYPES = (15, 16)
# DuckDB string_t (string_type.hpp:236-247): u32 length + 12 bytes (inlined | 4 prefix bytes + pointer)
STRING_T = np.dtype([("length", "<u4"), ("body", "u1", 12)])
ENC_NAME = {v: k for k, v in ENC.items()}
OUT = dict(NATIVE=0, I8=1, I16=2, I32=3, I64=4, U8=5, U16=6, U32=7, U64=8, F32=9, F64=10)
OUT_NP = {1: np.int8, 2: np.int16, 3: np.int32, 4: np.int64, 5: np.uint8, 6: np.uint16, 7: np.uint32, 8: np.uint64,
          9: np.float32, 10: np.float64}
_NP_OF_TYPE = {1: np.int8, 2: np.int16, 3: np.int32, 4: np.int64, 5: np.uint8, 6: np.uint16, 7: np.uint32,
               8: np.uint64, 9: np.float32, 10: np.float64, 11: np.uint8, 12: np.int32, 13: np.int64}
VEC = 1024


class ColumnSpec(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data_type", C.c_int32), ("decimal_precision", C.c_int32),
                ("decimal_scale", C.c_int32), ("force_encoding", C.c_int32)]


class ScanColumn(C.Structure):  # flsgpu_scan_column
    _fields_ = [("col", C.c_uint32), ("out_type", C.c_int32), ("string_form", C.c_int32), ("compact", C.c_uint32),
                ("out_host", C.c_void_p), ("out_capacity_bytes", C.c_uint64), ("heap_host", C.c_void_p),
                ("heap_capacity_bytes", C.c_uint64), ("keys_host", C.c_void_p), ("keys_capacity_bytes", C.c_uint64),
                ("forms", C.c_void_p)]


STRING_FLAT, STRING_DICTIONARY = 0, 1
FORM_FLAT, FORM_CONSTANT, FORM_SEQUENCE = 0, 1, 2
VECTOR_FORM = np.dtype([("start", "<i8"), ("increment", "<i8"), ("form", "<u4"), ("flat_slot", "<u4")])  # flsgpu_vector_form
CMP = dict(EQ=0, NE=1, LT=2, LE=3, GT=4, GE=5)


class Filter(C.Structure):  # flsgpu_filter
    _fields_ = [("col", C.c_uint32), ("op", C.c_int32), ("constant", C.c_uint64), ("str_data", C.c_char_p), ("str_len", C.c_uint64)]


def make_filters(filters):
    """[(col, 'GE', constant)] -> ctypes array; constant = physical bits (int) for numeric columns, bytes for VARCHAR / BLOB."""
    out = []
    for c, op, k in filters:
        op = CMP[op] if isinstance(op, str) else op
        if isinstance(k, (bytes, bytearray)):
            out.append(Filter(c, op, 0, bytes(k), len(k)))
        else:
            out.append(Filter(c, op, int(k) & 0xFFFFFFFFFFFFFFFF, None, 0))
    return (Filter * len(out))(*out)


class StringColumn(C.Structure):
    _fields_ = [("offsets", C.c_void_p), ("bytes", C.c_void_p)]


def pack_strings(values):
    """list of str/bytes (or an (offsets, bytes) pair) -> (uint64 offsets[n+1], uint8 bytes)."""
    if isinstance(values, tuple):
        off, data = values
        return np.ascontiguousarray(off, dtype=np.uint64), np.ascontiguousarray(data, dtype=np.uint8)
    enc = [v.encode() if isinstance(v, str) else bytes(v) for v in values]
    off = np.zeros(len(enc) + 1, dtype=np.uint64)
    np.cumsum([len(e) for e in enc], out=off[1:])
    data = np.frombuffer(b"".join(enc), dtype=np.uint8)
    return off, (data if data.size else np.zeros(1, dtype=np.uint8))


def strings_from(string_t, heap, heap_base=0):
    """Python bytes of every element of a STRING_T array whose pointers are heap_base + offset into heap."""
    out = []
    hb = heap.tobytes() if isinstance(heap, np.ndarray) else bytes(heap)
    for s in string_t:
        n = int(s["length"])
        if n <= 12:
            out.append(s["body"][:n].tobytes())
        else:
            off = int(s["body"][4:].view("<u8")[0]) - heap_base
            out.append(hb[off:off + n])
    return out


class ColumnDesc(C.Structure):
    _fields_ = [("name", C.c_char * 64), ("data_type", C.c_uint8), ("phys_bits", C.c_uint8), ("is_signed", C.c_uint8),
                ("is_float", C.c_uint8), ("decimal_precision", C.c_uint8), ("decimal_scale", C.c_uint8),
                ("reserved", C.c_uint8 * 10)]


class RowgroupDesc(C.Structure):
    _fields_ = [("first_tuple", C.c_uint64), ("n_tuples", C.c_uint32), ("n_vec", C.c_uint32), ("reserved", C.c_uint64)]


class ChunkDesc(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("size", C.c_uint64), ("seg_off", C.c_uint64 * 4), ("seg_size", C.c_uint64 * 4),
                ("const_val", C.c_uint64), ("min_val", C.c_uint64), ("max_val", C.c_uint64), ("n_vec", C.c_uint32),
                ("n_tuples", C.c_uint32), ("aux_count", C.c_uint32), ("encoding", C.c_uint8), ("phys_bits", C.c_uint8),
                ("idx_bits", C.c_uint8), ("max_width", C.c_uint8), ("has_stats", C.c_uint8), ("is_equal", C.c_uint8), ("equal_col_bytes", C.c_uint8 * 2),
                ("rd_dict", C.c_uint16 * 8), ("max_blob_bytes", C.c_uint32)]

    @property
    def equal_col(self):
        """EXP_EQUAL chunks (is_equal): the column of the same row group whose values (and bytes) this chunk shares."""
        return self.equal_col_bytes[0] | (self.equal_col_bytes[1] << 8)


class TableInfo(C.Structure):
    _fields_ = [("n_columns", C.c_uint32), ("n_rowgroups", C.c_uint32), ("vec_per_rowgroup", C.c_uint32),
                ("reserved", C.c_uint32), ("n_tuples", C.c_uint64), ("n_vectors", C.c_uint64), ("file_bytes", C.c_uint64)]


class StringStats(C.Structure):  # flsgpu_string_stats
    _fields_ = [("available", C.c_uint32), ("max_length", C.c_uint32), ("has_non_ascii", C.c_uint32), ("min_length", C.c_uint32),
                ("max_value_length", C.c_uint32), ("min_value", C.c_uint8 * 8), ("max_value", C.c_uint8 * 8), ("reserved", C.c_uint32)]


class AggResult(C.Structure):
    _fields_ = [("sum_lo", C.c_uint64), ("sum_hi", C.c_uint64), ("count", C.c_uint64), ("min_val", C.c_int64),
                ("max_val", C.c_int64), ("fsum", C.c_double), ("reserved", C.c_uint64 * 2)]


assert C.sizeof(ColumnDesc) == 80 and C.sizeof(RowgroupDesc) == 24 and C.sizeof(ChunkDesc) == 144
assert C.sizeof(AggResult) == 64

lib.flsgpu_last_error.restype = C.c_char_p
lib.flsgpu_version.restype = C.c_char_p
lib.flsgpu_launch_count.restype = C.c_uint64
for _n in ("flsgpu_table_close", "flsgpu_writer_destroy"):
    getattr(lib, _n).restype = None

# every symbol include/flsgpu.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "flsgpu_last_error", "flsgpu_version", "flsgpu_init", "flsgpu_device_count", "flsgpu_writer_create",
    "flsgpu_writer_append", "flsgpu_writer_finish", "flsgpu_writer_image", "flsgpu_writer_save", "flsgpu_writer_stream_to",
    "flsgpu_writer_destroy", "flsgpu_table_open", "flsgpu_table_open_memory", "flsgpu_table_close",
    "flsgpu_table_info_get", "flsgpu_table_column", "flsgpu_table_rowgroup", "flsgpu_table_chunk",
    "flsgpu_table_stage", "flsgpu_table_evict", "flsgpu_table_unstage", "flsgpu_table_resident_bytes", "flsgpu_decode_column",
    "flsgpu_rowgroup_decode", "flsgpu_aggregate_column", "flsgpu_agg_reset", "flsgpu_scan_column_to_host", "flsgpu_scan_columns_to_host",
    "flsgpu_sync", "flsgpu_host_alloc", "flsgpu_host_free", "flsgpu_launch_count",
    "flsgpu_string_heap_bytes", "flsgpu_decode_strings", "flsgpu_decode_string_codes", "flsgpu_decode_string_dict",
    "flsgpu_string_dict_keys", "flsgpu_scan_strings_to_host", "flsgpu_scan_projection_to_host",
    "flsgpu_rowgroup_prune", "flsgpu_filter_mask", "flsgpu_aggregate_column_masked", "flsgpu_scan_filtered_to_host",
    "flsgpu_writer_encode_rowgroup", "flsgpu_writer_flush_rowgroup", "flsgpu_encoded_rowgroup_free",
    "flsgpu_writer_set_inline_footer", "flsgpu_writer_rowgroups", "flsgpu_column_vector_forms", "flsgpu_transfer_bytes",
    "flsgpu_n_devices", "flsgpu_table_share", "flsgpu_table_bind_device", "flsgpu_filter_mask_string_dict_static",
    "flsgpu_aggregate_to_host", "flsgpu_scan_selected_to_host", "flsgpu_column_string_stats",
]


class FlsError(RuntimeError):
    def __init__(self, status, where):
        msg = lib.flsgpu_last_error().decode(errors="replace")
        super().__init__(f"{where}: status {status}: {msg}")
        self.status = status


def _check(rc, where):
    if rc != 0:
        raise FlsError(rc, where)


def _vp(x):
    return C.c_void_p(int(x) if x else 0)


class Writer:
    """COPY ... (FORMAT fls): flsgpu_writer_* (reference src/writer/fls_writer.cpp)."""

    def __init__(self, columns, vec_per_rowgroup=64):
        # columns: list of (name, type_name, force_encoding_name|None[, precision, scale])
        self._names = [c[0].encode() for c in columns]
        specs = (ColumnSpec * len(columns))()
        self.np_types = []
        for i, c in enumerate(columns):
            prec, scale = (c[3], c[4]) if len(c) > 3 else (0, 0)
            specs[i] = ColumnSpec(self._names[i], TYPE[c[1]], prec, scale, -1 if c[2] is None else ENC[c[2]])
            if c[1] == "DECIMAL":
                self.np_types.append(np.int16 if prec <= 4 else np.int32 if prec <= 9 else np.int64)
            elif TYPE[c[1]] in STRING_TYPES:
                self.np_types.append(None)
            else:
                self.np_types.append(_NP_OF_TYPE[TYPE[c[1]]])
        self._h = C.c_void_p()
        _check(lib.flsgpu_writer_create(len(columns), specs, vec_per_rowgroup, C.byref(self._h)), "writer_create")

    def append(self, arrays, n_threads=None):
        keep, addrs, n = [], [], None
        for a, t in zip(arrays, self.np_types):
            if t is None:  # string column: flsgpu_string_column {offsets, bytes}
                off, data = pack_strings(a)
                sc = StringColumn(off.ctypes.data, data.ctypes.data)
                keep += [off, data, sc]
                addrs.append(C.addressof(sc))
                rows = len(off) - 1
            else:
                arr = np.ascontiguousarray(a, dtype=t)
                keep.append(arr)
                addrs.append(arr.ctypes.data)
                rows = len(arr)
            assert n is None or rows == n, "columns of different lengths"
            n = rows
        ptrs = (C.c_void_p * len(addrs))(*addrs)
        _check(lib.flsgpu_writer_append(self._h, ptrs, C.c_uint64(n), n_threads or (os.cpu_count() or 1)), "writer_append")

    def finish(self):
        _check(lib.flsgpu_writer_finish(self._h), "writer_finish")
        p, n = C.c_void_p(), C.c_uint64()
        _check(lib.flsgpu_writer_image(self._h, C.byref(p), C.byref(n)), "writer_image")
        # zero-copy view; valid until close()
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,))

    def save(self, path):
        _check(lib.flsgpu_writer_save(self._h, os.fsencode(path)), "writer_save")

    def stream_to(self, path):
        """Row groups go to `path` as they fill; finish_stream() completes the file (no in-memory image)."""
        _check(lib.flsgpu_writer_stream_to(self._h, os.fsencode(path)), "writer_stream_to")

    def finish_stream(self):
        _check(lib.flsgpu_writer_finish(self._h), "writer_finish")

    def close(self):
        if self._h:
            lib.flsgpu_writer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        self.close()


def transfer_bytes():
    """(h2d, d2h) bytes handed to the copy engines by this process so far."""
    a, b = C.c_uint64(), C.c_uint64()
    lib.flsgpu_transfer_bytes(C.byref(a), C.byref(b))
    return a.value, b.value


def write_table(columns, arrays, vec_per_rowgroup=64, n_threads=None):
    """Encode arrays into a container image (numpy uint8 copy)."""
    w = Writer(columns, vec_per_rowgroup)
    w.append(arrays, n_threads)
    img = w.finish().copy()
    w.close()
    return img


class Table:
    """TableMetadata + row-group readers: flsgpu_table_* (reference src/reader/table_metadata.cpp, fls_reader.cpp)."""

    def __init__(self, image=None, path=None, copy=False, share=None):
        self._h = C.c_void_p()
        self._image = None
        if share is not None:  # another handle over the same host image (one per device of a single-process scan)
            self._image = share._image
            self._parent = share
            _check(lib.flsgpu_table_share(share._h, C.byref(self._h)), "table_share")
        elif path is not None:
            _check(lib.flsgpu_table_open(os.fsencode(path), C.byref(self._h)), "table_open")
        else:
            self._image = np.ascontiguousarray(image, dtype=np.uint8)
            _check(lib.flsgpu_table_open_memory(_vp(self._image.ctypes.data), C.c_uint64(self._image.size), int(copy),
                                                C.byref(self._h)), "table_open_memory")
        info = TableInfo()
        _check(lib.flsgpu_table_info_get(self._h, C.byref(info)), "table_info")
        self.info = info
        self.n_columns, self.n_rowgroups = info.n_columns, info.n_rowgroups
        self.vec_per_rowgroup, self.n_tuples = info.vec_per_rowgroup, info.n_tuples

    def bind_device(self, index):
        """Bind the handle to the index-th device of flsgpu.init's list (before its first staging / decode call)."""
        _check(lib.flsgpu_table_bind_device(self._h, int(index)), "table_bind_device")

    def column(self, col):
        d = ColumnDesc()
        _check(lib.flsgpu_table_column(self._h, col, C.byref(d)), "table_column")
        return d

    def rowgroup(self, rg):
        d = RowgroupDesc()
        _check(lib.flsgpu_table_rowgroup(self._h, rg, C.byref(d)), "table_rowgroup")
        return d

    def chunk(self, rg, col):
        d = ChunkDesc()
        _check(lib.flsgpu_table_chunk(self._h, rg, col, C.byref(d)), "table_chunk")
        return d

    def native_dtype(self, col):
        d = self.column(col)
        if d.is_float:
            return np.float32 if d.phys_bits == 32 else np.float64
        return np.dtype(f"{'i' if d.is_signed else 'u'}{d.phys_bits // 8}").type

    def out_elems(self, n_rg):
        return n_rg * self.vec_per_rowgroup * VEC

    def stage(self, rg_begin=0, n_rg=None, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_table_stage(self._h, rg_begin, n_rg, _vp(stream)), "table_stage")

    def evict(self):
        _check(lib.flsgpu_table_evict(self._h), "table_evict")

    def unstage(self):
        _check(lib.flsgpu_table_unstage(self._h), "table_unstage")

    def resident_bytes(self):
        n = C.c_uint64()
        _check(lib.flsgpu_table_resident_bytes(self._h, C.byref(n)), "resident_bytes")
        return n.value

    def decode_column(self, col, out_ptr, out_bytes, rg_begin=0, n_rg=None, out_type=0, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_decode_column(self._h, col, rg_begin, n_rg, out_type, _vp(out_ptr), C.c_uint64(out_bytes),
                                        _vp(stream)), "decode_column")

    def rowgroup_decode(self, rg, cols, out_ptrs, out_types=None, stream=0):
        n = len(cols)
        ca = (C.c_uint32 * n)(*cols)
        pa = (C.c_void_p * n)(*[int(p) for p in out_ptrs])
        ta = (C.c_int32 * n)(*(out_types or [0] * n))
        _check(lib.flsgpu_rowgroup_decode(self._h, rg, ca, n, pa, ta, _vp(stream)), "rowgroup_decode")

    def aggregate_column(self, col, result_ptr, rg_begin=0, n_rg=None, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_aggregate_column(self._h, col, rg_begin, n_rg, _vp(result_ptr), _vp(stream)), "aggregate_column")

    def scan_column_to_host(self, col, out_host_ptr, out_bytes, rg_begin=0, n_rg=None, out_type=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_scan_column_to_host(self._h, col, rg_begin, n_rg, out_type, _vp(out_host_ptr),
                                              C.c_uint64(out_bytes)), "scan_column_to_host")

    def scan_columns_to_host(self, cols, out_host_ptrs, out_bytes, rg_begin=0, n_rg=None, out_types=None):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        n = len(cols)
        ca = (C.c_uint32 * n)(*cols)
        ta = (C.c_int32 * n)(*(out_types or [0] * n))
        pa = (C.c_void_p * n)(*[int(p) for p in out_host_ptrs])
        ba = (C.c_uint64 * n)(*out_bytes)
        _check(lib.flsgpu_scan_columns_to_host(self._h, ca, n, rg_begin, n_rg, ta, pa, ba), "scan_columns_to_host")

    # ---- string columns (flsgpu.h "string columns") ----
    def string_heap_bytes(self, col, rg_begin=0, n_rg=None):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        n = C.c_uint64()
        _check(lib.flsgpu_string_heap_bytes(self._h, col, rg_begin, n_rg, C.byref(n)), "string_heap_bytes")
        return n.value

    def decode_strings(self, col, out_ptr, out_bytes, heap_ptr, heap_bytes, heap_base_addr, rg_begin=0, n_rg=None, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_decode_strings(self._h, col, rg_begin, n_rg, _vp(out_ptr), C.c_uint64(out_bytes), _vp(heap_ptr),
                                         C.c_uint64(heap_bytes), C.c_uint64(heap_base_addr), _vp(stream)), "decode_strings")

    def decode_string_codes(self, col, out_ptr, out_bytes, rg_begin=0, n_rg=None, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_decode_string_codes(self._h, col, rg_begin, n_rg, _vp(out_ptr), C.c_uint64(out_bytes), _vp(stream)),
               "decode_string_codes")

    def string_dict_keys(self, col, rg_begin=0, n_rg=None):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        n = C.c_uint64()
        _check(lib.flsgpu_string_dict_keys(self._h, col, rg_begin, n_rg, C.byref(n)), "string_dict_keys")
        return n.value

    def decode_string_dict(self, col, keys_ptr, keys_bytes, heap_ptr, heap_bytes, heap_base_addr, rg_begin=0, n_rg=None, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_decode_string_dict(self._h, col, rg_begin, n_rg, _vp(keys_ptr), C.c_uint64(keys_bytes), _vp(heap_ptr),
                                             C.c_uint64(heap_bytes), C.c_uint64(heap_base_addr), _vp(stream)), "decode_string_dict")

    def scan_strings_to_host(self, col, out_host_ptr, out_bytes, heap_host_ptr, heap_bytes, rg_begin=0, n_rg=None):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_scan_strings_to_host(self._h, col, rg_begin, n_rg, _vp(out_host_ptr), C.c_uint64(out_bytes),
                                               _vp(heap_host_ptr), C.c_uint64(heap_bytes)), "scan_strings_to_host")

    # ---- filter pushdown ----
    def rowgroup_prune(self, filters, rg_begin=0, n_rg=None):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        keep = np.zeros(n_rg, dtype=np.uint8)
        _check(lib.flsgpu_rowgroup_prune(self._h, make_filters(filters), len(filters), rg_begin, n_rg, _vp(keep.ctypes.data)), "rowgroup_prune")
        return keep

    def filter_mask(self, filters, mask_ptr, mask_bytes, keep_ptr=0, rg_begin=0, n_rg=None, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_filter_mask(self._h, make_filters(filters), len(filters), rg_begin, n_rg, _vp(keep_ptr), _vp(mask_ptr),
                                      C.c_uint64(mask_bytes), _vp(stream)), "filter_mask")

    def filter_mask_string_dict_static(self, filt, mask_ptr, mask_bytes, keep_ptr=0, rg_begin=0, n_rg=None, stream=0):
        """One string filter over a dictionary-encoded range, scratch in the library's static device array (no allocation)."""
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_filter_mask_string_dict_static(self._h, make_filters([filt]), rg_begin, n_rg, _vp(keep_ptr), _vp(mask_ptr),
                                                         C.c_uint64(mask_bytes), _vp(stream)), "filter_mask_string_dict_static")

    def aggregate_column_masked(self, col, mask_ptr, result_ptr, keep_ptr=0, rg_begin=0, n_rg=None, stream=0):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        _check(lib.flsgpu_aggregate_column_masked(self._h, col, rg_begin, n_rg, _vp(mask_ptr), _vp(keep_ptr), _vp(result_ptr), _vp(stream)),
               "aggregate_column_masked")

    def aggregate_to_host(self, cols, filters=(), rg_begin=0, n_rg=None):
        """Fused SUM / COUNT / MIN / MAX of the listed numeric columns [WHERE filters], host to host -> list of AggResult."""
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        ids = (C.c_uint32 * len(cols))(*cols)
        out = (AggResult * len(cols))()
        _check(lib.flsgpu_aggregate_to_host(self._h, ids, len(cols), make_filters(filters) if filters else None, len(filters), rg_begin, n_rg, out),
               "aggregate_to_host")
        return list(out)

    def scan_filtered_to_host(self, scan_columns, filters, mask_host_ptr, mask_bytes, rg_begin=0, n_rg=None):
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        arr = (ScanColumn * len(scan_columns))(*scan_columns)
        _check(lib.flsgpu_scan_filtered_to_host(self._h, arr, len(scan_columns), make_filters(filters), len(filters), rg_begin, n_rg,
                                                _vp(mask_host_ptr), C.c_uint64(mask_bytes)), "scan_filtered_to_host")

    def scan_selected_to_host(self, scan_columns, filters, row_index_ptr=0, row_index_bytes=0, rg_begin=0, n_rg=None):
        """Filtered scan, selection applied on the device: returns the number of selected rows (packed in every out_host)."""
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        arr = (ScanColumn * len(scan_columns))(*scan_columns)
        n_sel = C.c_uint64()
        _check(lib.flsgpu_scan_selected_to_host(self._h, arr, len(scan_columns), make_filters(filters), len(filters), rg_begin, n_rg,
                                                _vp(row_index_ptr), C.c_uint64(row_index_bytes), C.byref(n_sel)), "scan_selected_to_host")
        return n_sel.value

    def string_stats(self, col):
        """None, or (max_length, has_non_ascii, min prefix, max prefix) of a string column whose distinct strings sit uncompressed."""
        st = StringStats()
        _check(lib.flsgpu_column_string_stats(self._h, col, C.byref(st)), "column_string_stats")
        if not st.available:
            return None
        return st.max_length, bool(st.has_non_ascii), bytes(st.min_value[:st.min_length]), bytes(st.max_value[:st.max_value_length])

    def vector_forms(self, col, rg_begin=0, n_rg=None):
        """(forms: structured array of VECTOR_FORM per vector of the range, n_flat)."""
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        out = np.zeros(n_rg * self.vec_per_rowgroup, dtype=VECTOR_FORM)
        n_flat = C.c_uint64()
        _check(lib.flsgpu_column_vector_forms(self._h, col, rg_begin, n_rg, _vp(out.ctypes.data), C.byref(n_flat)), "column_vector_forms")
        return out, n_flat.value

    def scan_projection_to_host(self, scan_columns, rg_begin=0, n_rg=None):
        """scan_columns: list of ScanColumn."""
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        arr = (ScanColumn * len(scan_columns))(*scan_columns)
        _check(lib.flsgpu_scan_projection_to_host(self._h, arr, len(scan_columns), rg_begin, n_rg), "scan_projection_to_host")

    def valid_mask_slices(self, rg_begin=0, n_rg=None):
        """(start, stop) element ranges of real tuples in the decode_column output layout."""
        n_rg = self.n_rowgroups - rg_begin if n_rg is None else n_rg
        stride = self.vec_per_rowgroup * VEC
        return [(r * stride, r * stride + self.rowgroup(rg_begin + r).n_tuples) for r in range(n_rg)]

    def close(self):
        if self._h:
            lib.flsgpu_table_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        self.close()


def init(device_ids=(0,)):
    arr = (C.c_int * len(device_ids))(*device_ids)
    _check(lib.flsgpu_init(len(device_ids), arr), "init")


def n_devices():
    n = C.c_int()
    _check(lib.flsgpu_n_devices(C.byref(n)), "n_devices")
    return n.value


def agg_reset(ptr, stream=0):
    _check(lib.flsgpu_agg_reset(_vp(ptr), _vp(stream)), "agg_reset")


def sync(stream=0):
    _check(lib.flsgpu_sync(_vp(stream)), "sync")


def host_alloc(nbytes):
    p = C.c_void_p()
    _check(lib.flsgpu_host_alloc(C.c_uint64(nbytes), C.byref(p)), "host_alloc")
    return p.value


def host_free(ptr):
    _check(lib.flsgpu_host_free(_vp(ptr)), "host_free")


def launch_count():
    return int(lib.flsgpu_launch_count())

"""ctypes view of libflsreader.so: the C++ reader mirror (csrc/reader) + its executor stand-in."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "duckdb-fastlanes_b200", "libflsreader.so")
lib = C.CDLL(os.environ.get("FLSREADER_LIB", LIB))  # FLSREADER_LIB: tools/sanitize.sh
lib.flsreader_last_error.restype = C.c_char_p
lib.flsreader_close.restype = None
lib.flsreader_cardinality.restype = C.c_uint64
lib.flsreader_max_threads.restype = C.c_uint64
lib.flsreader_total_tuples.restype = C.c_uint64
lib.flsreader_progress.restype = C.c_double

TYPE_NAMES = ["INVALID", "BOOLEAN", "TINYINT", "SMALLINT", "INTEGER", "BIGINT", "UBIGINT", "FLOAT", "DOUBLE", "DATE",
              "TIMESTAMP", "DECIMAL", "VARCHAR", "BLOB"]


class StringSink(C.Structure):
    _fields_ = [("bytes", C.c_void_p), ("lengths", C.c_void_p), ("stride", C.c_uint32)]


class ReaderFilter(C.Structure):
    _fields_ = [("column_index", C.c_uint64), ("op", C.c_int32), ("constant", C.c_uint64), ("str_data", C.c_char_p),
                ("str_len", C.c_uint64)]


def reader_filters(filters):
    """[(index into column_ids, op, constant)]: physical bits (int) or the bytes of a VARCHAR / BLOB constant."""
    out = []
    for ci, op, k in filters:
        if isinstance(k, (bytes, bytearray)):
            out.append(ReaderFilter(ci, op, 0, bytes(k), len(k)))
        else:
            out.append(ReaderFilter(ci, op, int(k) & 0xFFFFFFFFFFFFFFFF, None, 0))
    return (ReaderFilter * max(1, len(out)))(*out)


class ScanStats(C.Structure):
    _fields_ = [("rows", C.c_uint64), ("chunks", C.c_uint64), ("max_chunk_rows", C.c_uint64), ("dictionary_vectors", C.c_uint64),
                ("pruned_rowgroups", C.c_uint64), ("constant_vectors", C.c_uint64), ("sequence_vectors", C.c_uint64),
                ("assembled_vectors", C.c_uint64)]
NP_OF = {"BOOLEAN": np.uint8, "TINYINT": np.int8, "SMALLINT": np.int16, "INTEGER": np.int32, "BIGINT": np.int64,
         "UBIGINT": np.uint64, "FLOAT": np.float32, "DOUBLE": np.float64, "DATE": np.int32, "TIMESTAMP": np.int64}


class ReaderError(RuntimeError):
    def __init__(self, status):
        super().__init__(f"status {status}: {lib.flsreader_last_error().decode(errors='replace')}")
        self.status = status


class Reader:
    def __init__(self, path, file_row_number=False, explicit_cardinality=0, rowgroups_per_batch=0):
        self._h = C.c_void_p()
        rc = lib.flsreader_open(os.fsencode(path), int(file_row_number), C.c_uint64(explicit_cardinality),
                                C.c_uint64(rowgroups_per_batch), C.byref(self._h))
        if rc:
            raise ReaderError(rc)
        self.columns = []
        for i in range(lib.flsreader_n_columns(self._h)):
            name = C.create_string_buffer(64)
            tid, pb, w, s = C.c_int(), C.c_int(), C.c_int(), C.c_int()
            lib.flsreader_column(self._h, i, name, C.byref(tid), C.byref(pb), C.byref(w), C.byref(s))
            self.columns.append(dict(name=name.value.decode(), type=TYPE_NAMES[tid.value], bytes=pb.value, width=w.value,
                                     scale=s.value))

    def np_dtype(self, i):
        c = self.columns[i]
        if c["type"] == "DECIMAL":
            return {2: np.int16, 4: np.int32, 8: np.int64}[c["bytes"]]
        return NP_OF[c["type"]]

    @property
    def cardinality(self):
        return lib.flsreader_cardinality(self._h)

    @property
    def max_threads(self):
        return lib.flsreader_max_threads(self._h)

    @property
    def total_tuples(self):
        return lib.flsreader_total_tuples(self._h)

    @property
    def progress(self):
        return lib.flsreader_progress(self._h)

    def partition_stats(self):
        n = C.c_uint64()
        lib.flsreader_partition_stats(self._h, None, None, C.c_uint64(0), C.byref(n))
        a = (C.c_uint64 * n.value)()
        b = (C.c_uint64 * n.value)()
        lib.flsreader_partition_stats(self._h, a, b, n, C.byref(n))
        return list(zip(a, b))

    def statistics(self, name):
        has, mi, ma, md, xd = C.c_int(), C.c_int64(), C.c_int64(), C.c_double(), C.c_double()
        rc = lib.flsreader_statistics(self._h, name.encode(), C.byref(has), C.byref(mi), C.byref(ma), C.byref(md), C.byref(xd))
        if rc:
            raise ReaderError(rc)
        return dict(has_stats=bool(has.value), min_i=mi.value, max_i=ma.value, min_d=md.value, max_d=xd.value)

    def scan_all(self, column_ids=None, n_threads=1):
        """SELECT cols FROM read_fls(path) -> (list of numpy columns in file row order, stats dict)."""
        ids = list(range(len(self.columns))) if column_ids is None else list(column_ids)
        n = self.total_tuples
        # ids the reader does not know are passed through so that the C++ side reports them
        outs = [np.zeros(n, dtype=self.np_dtype(i) if i < len(self.columns) else np.uint8) for i in ids]
        ia = (C.c_uint64 * len(ids))(*ids)
        pa = (C.c_void_p * len(ids))(*[o.ctypes.data for o in outs])
        rows, chunks, mx = C.c_uint64(), C.c_uint64(), C.c_uint64()
        rc = lib.flsreader_scan_all(self._h, ia, len(ids), n_threads, pa, C.byref(rows), C.byref(chunks), C.byref(mx))
        if rc:
            raise ReaderError(rc)
        return outs, dict(rows=rows.value, chunks=chunks.value, max_chunk_rows=mx.value)

    def scan_raw(self, column_ids=None, n_threads=1, stride=16):
        """Unfiltered scan into numpy buffers, strings left in their fixed-stride sinks: [array | (bytes[n, stride], lengths)]."""
        ids = list(range(len(self.columns))) if column_ids is None else list(column_ids)
        n = self.total_tuples
        outs, keep, ptrs = [], [], []
        for i in ids:
            if self.columns[i]["type"] in ("VARCHAR", "BLOB"):
                data, lens = np.zeros((n, stride), dtype=np.uint8), np.zeros(n, dtype=np.uint32)
                sink = StringSink(data.ctypes.data, lens.ctypes.data, stride)
                keep.append(sink)
                outs.append((data, lens))
                ptrs.append(C.addressof(sink))
            else:
                a = np.zeros(n, dtype=self.np_dtype(i))
                outs.append(a)
                ptrs.append(a.ctypes.data)
        st = ScanStats()
        rc = lib.flsreader_scan(self._h, (C.c_uint64 * len(ids))(*ids), len(ids), reader_filters(()), 0, n_threads,
                                (C.c_void_p * len(ids))(*ptrs), None, C.byref(st))
        if rc:
            raise ReaderError(rc)
        assert st.rows == n
        return outs

    def scan(self, column_ids=None, filters=(), n_threads=1, stride=64):
        """SELECT cols FROM read_fls(path) WHERE filters.  filters: [(index into column_ids, op, physical constant)].
        Returns (columns, row ids or None, stats): string columns come back as lists of bytes; with filters the rows are
        sorted by file row number."""
        ids = list(range(len(self.columns))) if column_ids is None else list(column_ids)
        n = self.total_tuples
        outs, keep, ptrs = [], [], []
        for i in ids:
            if self.columns[i]["type"] in ("VARCHAR", "BLOB"):
                data = np.zeros((n, stride), dtype=np.uint8)
                lens = np.zeros(n, dtype=np.uint32)
                sink = StringSink(data.ctypes.data, lens.ctypes.data, stride)
                keep.append(sink)
                outs.append((data, lens))
                ptrs.append(C.addressof(sink))
            else:
                a = np.zeros(n, dtype=self.np_dtype(i))
                outs.append(a)
                ptrs.append(a.ctypes.data)
        ia = (C.c_uint64 * len(ids))(*ids)
        pa = (C.c_void_p * len(ids))(*ptrs)
        fa = reader_filters(filters)
        row_ids = np.zeros(n, dtype=np.int64)
        st = ScanStats()
        rc = lib.flsreader_scan(self._h, ia, len(ids), fa, len(filters), n_threads, pa, C.c_void_p(row_ids.ctypes.data), C.byref(st))
        if rc:
            raise ReaderError(rc)
        m = st.rows
        order = np.argsort(row_ids[:m], kind="stable") if filters else np.arange(m)
        cols = []
        for o in outs:
            if isinstance(o, tuple):
                data, lens = o
                cols.append([data[j, :lens[j]].tobytes() for j in order])
            else:
                cols.append(o[:m][order] if filters else o)
        stats = dict(rows=st.rows, chunks=st.chunks, max_chunk_rows=st.max_chunk_rows, dictionary_vectors=st.dictionary_vectors,
                     pruned_rowgroups=st.pruned_rowgroups, constant_vectors=st.constant_vectors, sequence_vectors=st.sequence_vectors,
                     assembled_vectors=st.assembled_vectors)
        return cols, (row_ids[:m][order] if filters else None), stats

    def set_option(self, key, value):
        rc = lib.flsreader_set_option(self._h, key.encode(), C.c_uint64(int(value)))
        if rc:
            raise ReaderError(rc)

    def evict(self):
        rc = lib.flsreader_evict(self._h)
        if rc:
            raise ReaderError(rc)

    def consume(self, column_ids=None, n_threads=1):
        """The scan with a consumer that takes every vector as it is (nothing copied): what the bench times end to end."""
        ids = list(range(len(self.columns))) if column_ids is None else list(column_ids)
        ia = (C.c_uint64 * len(ids))(*ids)
        st = ScanStats()
        rc = lib.flsreader_scan(self._h, ia, len(ids), reader_filters(()), 0, n_threads, None, None, C.byref(st))
        if rc:
            raise ReaderError(rc)
        return {k: getattr(st, k) for k, _ in ScanStats._fields_}

    def close(self):
        if self._h:
            lib.flsreader_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        self.close()


lib.flsmulti_close.restype = None
lib.flsmulti_cardinality.restype = C.c_uint64
lib.flsmulti_max_threads.restype = C.c_uint64


class MultiReader:
    """read_fls(patterns, **named parameters): FastLanesMultiFileInfo + MultiFileReader semantics of the C++ reader mirror.
    A handle binds one scan (open a new one for another projection)."""

    def __init__(self, patterns, **options):
        pats = [patterns] if isinstance(patterns, (str, bytes)) else list(patterns)
        pa = (C.c_char_p * len(pats))(*[os.fsencode(p) for p in pats])
        keys = list(options)
        ka = (C.c_char_p * max(1, len(keys)))(*[k.encode() for k in keys])
        va = (C.c_char_p * max(1, len(keys)))(*[None if options[k] is None else str(options[k]).lower().encode() if isinstance(options[k], bool)
                                               else str(options[k]).encode() for k in keys])
        self._h = C.c_void_p()
        rc = lib.flsmulti_open(pa, len(pats), ka, va, len(keys), C.byref(self._h))
        if rc:
            raise ReaderError(rc)
        self.files = []
        for i in range(lib.flsmulti_n_files(self._h)):
            name = C.create_string_buffer(256)
            n = C.c_uint64()
            lib.flsmulti_file(self._h, i, name, C.byref(n))
            self.files.append((name.value.decode(), n.value))
        self.columns = []
        for i in range(lib.flsmulti_n_columns(self._h)):
            name = C.create_string_buffer(64)
            tid, pb, w, s = C.c_int(), C.c_int(), C.c_int(), C.c_int()
            lib.flsmulti_column(self._h, i, name, C.byref(tid), C.byref(pb), C.byref(w), C.byref(s))
            self.columns.append(dict(name=name.value.decode(), type=TYPE_NAMES[tid.value], bytes=pb.value, width=w.value, scale=s.value))

    @property
    def cardinality(self):
        return lib.flsmulti_cardinality(self._h)

    @property
    def max_threads(self):
        v = lib.flsmulti_max_threads(self._h)
        return None if v == 0xFFFFFFFFFFFFFFFF else v

    def scan(self, column_ids=None, filters=(), n_threads=1, stride=256):
        """-> (columns ordered by (file, file_row_number); None where a union column is missing), file ids, row ids, stats"""
        ids = list(range(len(self.columns))) if column_ids is None else list(column_ids)
        n = sum(t for _, t in self.files)
        outs, keep, ptrs, valids = [], [], [], []
        for i in ids:
            c = self.columns[i]
            if c["type"] in ("VARCHAR", "BLOB"):
                data, lens = np.zeros((n, stride), dtype=np.uint8), np.zeros(n, dtype=np.uint32)
                sink = StringSink(data.ctypes.data, lens.ctypes.data, stride)
                keep.append(sink)
                outs.append((data, lens))
                ptrs.append(C.addressof(sink))
            else:
                dt = {2: np.int16, 4: np.int32, 8: np.int64}[c["bytes"]] if c["type"] == "DECIMAL" else NP_OF[c["type"]]
                a = np.zeros(n, dtype=dt)
                outs.append(a)
                ptrs.append(a.ctypes.data)
            valids.append(np.ones(n, dtype=np.uint8))
        ia = (C.c_uint64 * len(ids))(*ids)
        pa = (C.c_void_p * len(ids))(*ptrs)
        va = (C.c_void_p * len(ids))(*[v.ctypes.data for v in valids])
        fa = reader_filters(filters)
        file_ids, row_ids = np.zeros(n, dtype=np.uint32), np.zeros(n, dtype=np.int64)
        st = ScanStats()
        rc = lib.flsmulti_scan(self._h, ia, len(ids), fa, len(filters), n_threads, pa, va, C.c_void_p(file_ids.ctypes.data),
                               C.c_void_p(row_ids.ctypes.data), C.byref(st))
        if rc:
            raise ReaderError(rc)
        m = st.rows
        order = np.lexsort((row_ids[:m], file_ids[:m]))
        cols = []
        for o, v in zip(outs, valids):
            vv = v[:m][order].astype(bool)
            if isinstance(o, tuple):
                data, lens = o
                cols.append([data[j, :lens[j]].tobytes() if ok else None for j, ok in zip(order, vv)])
            else:
                cols.append(np.ma.masked_array(o[:m][order], mask=~vv))
        stats = dict(rows=st.rows, chunks=st.chunks, max_chunk_rows=st.max_chunk_rows, dictionary_vectors=st.dictionary_vectors,
                     pruned_rowgroups=st.pruned_rowgroups)
        return cols, file_ids[:m][order], row_ids[:m][order], stats

    def close(self):
        if self._h:
            lib.flsmulti_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        self.close()

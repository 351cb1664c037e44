/*
 * flsgpu.h — C ABI of the B200 FastLanes decode engine (libflsgpu.so).
 *
 * This library sits UNDER the read_fls table function of the DuckDB extension: it replaces the absent
 * third-party FastLanes library (reference CMakeLists.txt:56-67) plus the extension's materializer
 * (reference src/include/reader/materializer/ **), i.e. layers L0+L1 of SURVEY.md section 1.  Every entry
 * point cites the reference interface it stands in for.  POD only, no exceptions cross this boundary, every
 * call returns 0 on success or a negative flsgpu_status; flsgpu_last_error() gives the message (thread local).
 * There is no CPU fallback: decode entry points fail with FLSGPU_ERR_CUDA when no device is usable.
 */
#ifndef FLSGPU_H
#define FLSGPU_H

#include <stddef.h>
#include <stdint.h>
#include "fls_format.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum flsgpu_status {
	FLSGPU_OK = 0,
	FLSGPU_ERR_INVALID = -1, /* bad argument / malformed container  (reference: InternalException, BinderException) */
	FLSGPU_ERR_IO = -2,      /* file could not be read/written      (reference: std::runtime_error) */
	FLSGPU_ERR_CUDA = -3,    /* CUDA runtime error, or no device */
	FLSGPU_ERR_UNSUPPORTED = -4, /* encoding/type combination not implemented (reference fls_kernel.hpp:24-34) */
	FLSGPU_ERR_CONVERSION = -5   /* destination type cannot hold the column   (reference fls_type_resolver.hpp:75-82) */
} flsgpu_status;

typedef struct flsgpu_table flsgpu_table;   /* replaces TableMetadata + fastlanes::TableReader (table_metadata.cpp:13-35) */
typedef struct flsgpu_writer flsgpu_writer; /* replaces fastlanes::FileWriter/RowGroupWriter (fls_writer.cpp:271-284) */

const char* flsgpu_last_error(void);
const char* flsgpu_version(void);

/* ---- device ------------------------------------------------------------------------------------------ */
/* The CUDA devices this process uses, by ordinal (an ordinal may repeat).  One process per GPU (torchrun) passes one.  A
 * single-process host with N worker threads - DuckDB (reference fls_multi_file_info.cpp:99-110: one worker per row group
 * inside one process) - passes all of them, opens one handle per device over the same host image (flsgpu_table_share +
 * flsgpu_table_bind_device) and sends each batch of row groups to one handle: row groups are independent, so nothing is
 * exchanged between devices.  Not calling it = {0}.  No reference counterpart. */
int flsgpu_init(int n_devices, const int* device_ids);
int flsgpu_device_count(int* out); /* CUDA devices visible to the process */
int flsgpu_n_devices(int* out);    /* length of flsgpu_init's list */

/* ---- writer: COPY ... (FORMAT fls), reference src/writer/fls_writer.cpp -------------------------------- */
typedef struct flsgpu_column_spec {
	const char* name;
	int32_t data_type;         /* fls_data_type */
	int32_t decimal_precision; /* DECIMAL only; physical width INT16/32/64 by precision (writer/translation_utils.cpp:46-58) */
	int32_t decimal_scale;
	int32_t force_encoding;    /* fls_encoding, or -1: choose the smallest (encoding selection lives in the absent lib) */
} flsgpu_column_spec;

/* Rows of a VARCHAR / BLOB column as handed to flsgpu_writer_append (fls_view_writer.cpp:61-94: offsets + bytes):
 * row i = bytes[offsets[i], offsets[i+1]); offsets has n_tuples + 1 entries. */
typedef struct flsgpu_string_column {
	const uint64_t* offsets;
	const uint8_t* bytes;
} flsgpu_string_column;

/* WithSchema/WithRowGroupSize (fls_writer.cpp:271-284); vec_per_rowgroup 0 -> 64 (fls_writer.hpp:14). */
int flsgpu_writer_create(uint32_t n_columns, const flsgpu_column_spec* cols, uint32_t vec_per_rowgroup,
                         flsgpu_writer** out);
/* Sink (fls_writer.cpp:46-102): columns[i] points at n_tuples values of column i's PHYSICAL type
 * (string columns: columns[i] points at a flsgpu_string_column).  Rows are
 * buffered until a full row group exists; full row groups are encoded on n_threads host threads. */
int flsgpu_writer_append(flsgpu_writer* w, const void* const* columns, uint64_t n_tuples, int n_threads);
/* Finalize (fls_writer.cpp:332-347): flush the tail row group and build the footer. */
int flsgpu_writer_finish(flsgpu_writer* w);
/* Row groups prepared out of order and appended in order: PrepareBatch / FlushBatch and the per-thread row-group writers of
 * Sink / Combine (fls_writer.cpp:46-187,306-330: CreateRowGroupWriter + PrepareRowGroup, then rg_writer->Flush()).
 * encode is thread safe (it only reads the writer's schema) and takes 1..row_group_size tuples, columns as for
 * flsgpu_writer_append; flush appends the row group to the file in call order, takes ownership, and may be called from any
 * thread.  Do not mix with flsgpu_writer_append while rows are pending there. */
typedef struct flsgpu_encoded_rowgroup flsgpu_encoded_rowgroup;
int flsgpu_writer_encode_rowgroup(const flsgpu_writer* w, const void* const* columns, uint64_t n_tuples, flsgpu_encoded_rowgroup** out);
int flsgpu_writer_flush_rowgroup(flsgpu_writer* w, flsgpu_encoded_rowgroup* rg);
void flsgpu_encoded_rowgroup_free(flsgpu_encoded_rowgroup* rg); /* only for row groups that were never flushed */
/* WithInlinedFooter (fls_writer.cpp:229-230,278): 0 -> flsgpu_writer_save writes the footer to "<path>.footer". Default 1. */
int flsgpu_writer_set_inline_footer(flsgpu_writer* w, int inline_footer);
/* Row groups flushed so far / tuples accepted so far (num_row_groups of FastLanesWriteGlobalState: file rotation). */
int flsgpu_writer_rowgroups(const flsgpu_writer* w, uint64_t* n_rowgroups, uint64_t* n_tuples);
/* Streaming: call once, before the first row group.  Every row group then goes to `path` as it is flushed (or, with
 * flsgpu_writer_append, as it fills) and only its descriptors stay in memory; flsgpu_writer_finish writes the footer and the
 * header and closes the file, which is byte for byte what flsgpu_writer_save would have written from an in-memory image.
 * The reference's FileWriter writes row groups as they are flushed too (rg_writer->Flush(), fls_writer.cpp:70,94,130,161,325);
 * without this call the container is assembled in memory (flsgpu_writer_image) - what tests and benchmarks want.
 * After it flsgpu_writer_image fails and flsgpu_writer_save accepts the same path only (a no-op). */
int flsgpu_writer_stream_to(flsgpu_writer* w, const char* path);
int flsgpu_writer_image(const flsgpu_writer* w, const void** bytes, uint64_t* size); /* valid until destroy */
int flsgpu_writer_save(const flsgpu_writer* w, const char* path);
void flsgpu_writer_destroy(flsgpu_writer* w);

/* ---- table metadata: TableMetadata / SchemaBuilder / RowGroupStatistics ------------------------------- */
int flsgpu_table_open(const char* path, flsgpu_table** out);                         /* table_metadata.cpp:13-16 */
int flsgpu_table_open_memory(const void* bytes, uint64_t size, int copy, flsgpu_table** out);
void flsgpu_table_close(flsgpu_table* t);
/* A second handle over the same host image (not copied, not validated again); every handle has its own device state. */
int flsgpu_table_share(const flsgpu_table* src, flsgpu_table** out);
/* Bind the handle to flsgpu_init's device_ids[device_index]: its HBM arena, staging copies and kernels live there.
 * Before the first staging / decode call of the handle; default index 0. */
int flsgpu_table_bind_device(flsgpu_table* t, int device_index);

typedef struct flsgpu_table_info {
	uint32_t n_columns;
	uint32_t n_rowgroups;
	uint32_t vec_per_rowgroup;
	uint32_t reserved;
	uint64_t n_tuples;
	uint64_t n_vectors;
	uint64_t file_bytes;
} flsgpu_table_info;
int flsgpu_table_info_get(const flsgpu_table* t, flsgpu_table_info* out);
int flsgpu_table_column(const flsgpu_table* t, uint32_t col, fls_column_desc* out);            /* schema_builder.cpp:179-243 */
int flsgpu_table_rowgroup(const flsgpu_table* t, uint32_t rg, fls_rowgroup_desc* out);         /* fls_reader.cpp:129-138 */
int flsgpu_table_chunk(const flsgpu_table* t, uint32_t rg, uint32_t col, fls_chunk_desc* out); /* row_group_statistics.cpp:60-127 */

/* String statistics for the host database's optimizer (DuckDB StringStats: the reference reports none for VARCHAR columns,
 * row_group_statistics.cpp:60-127 covers numeric types): longest string, whether any byte is >= 0x80, and the first 8 bytes of
 * the smallest / greatest string (bytewise order, the shorter first on a tie).  available = 0 unless every chunk of the column
 * keeps its distinct strings uncompressed (DICT_STR, STR_CONSTANT, RLE_STR, CROSS_RLE_STR, FREQUENCY_STR): the summary then
 * costs a walk over those few strings, never a decode.  Host only. */
typedef struct flsgpu_string_stats {
	uint32_t available;
	uint32_t max_length;       /* bytes of the longest string */
	uint32_t has_non_ascii;
	uint32_t min_length;       /* bytes of min_value that are valid (<= 8) */
	uint32_t max_value_length; /* bytes of max_value that are valid (<= 8) */
	uint8_t min_value[8];
	uint8_t max_value[8];
	uint32_t reserved;
} flsgpu_string_stats;
int flsgpu_column_string_stats(const flsgpu_table* t, uint32_t col, flsgpu_string_stats* out);

/* ---- vector forms: what can be handed to DuckDB without decoding ---------------------------------------------------------- */
/* The reference hands a dec_constant_opr chunk to DuckDB as a CONSTANT_VECTOR instead of 1024 copies (constant_kernel.hpp:
 * 19-27).  This container records two such facts per vector (fls_format.h FLS_VEC_FORM_*): CONSTANT and arithmetic SEQUENCE
 * (DuckDB SEQUENCE_VECTOR: start + i * increment).  Such vectors need no decode, no HBM traffic and - what matters for a
 * host-fed scan - no PCIe transfer of decoded bytes.  Host only: reads entry points of the container image. */
typedef struct flsgpu_vector_form { /* 24 bytes */
	int64_t start;      /* CONSTANT: the value, SEQUENCE: value(0).  Integers sign- / zero-extended by the column's signedness
	                       (like flsgpu_filter.constant), FLOAT bits in the low word, DOUBLE bits */
	int64_t increment;  /* SEQUENCE: value(i) = start + i * increment for i = 0..1023, exact in the column's type: a vector whose
	                       progression would wrap around inside the type, or leave int64, is reported FLAT */
	uint32_t form;      /* FLS_VEC_FORM_FLAT / _CONSTANT / _SEQUENCE */
	uint32_t flat_slot; /* FLAT: rank of the vector among the FLAT vectors of the range = its 1024-value slot in the output of
	                       a compact scan (flsgpu_scan_column.compact); 0xffffffff otherwise */
} flsgpu_vector_form;
/* out[(r - rg_begin) * vec_per_rowgroup + v] for every vector of the range (vectors a short row group does not have:
 * CONSTANT 0); *n_flat (may be NULL) = number of FLAT vectors.  String columns: every vector FLAT. */
int flsgpu_column_vector_forms(const flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, flsgpu_vector_form* out,
                               uint64_t* n_flat);

/* ---- staging: get_rowgroup_reader (fls_reader.cpp:615-618) -------------------------------------------- */
/* Copy the compressed image of row groups [rg_begin, rg_begin+n_rg) to HBM (pinned host -> device, async on
 * `stream`) and build the device descriptor tables.  No-op for row groups already resident. */
int flsgpu_table_stage(flsgpu_table* t, uint32_t rg_begin, uint32_t n_rg, void* cuda_stream);
/* Forget residency but keep the HBM arena (the next stage copies again): used to time host-fed scans. */
int flsgpu_table_evict(flsgpu_table* t);
/* Free every device resource of the table. */
int flsgpu_table_unstage(flsgpu_table* t);
int flsgpu_table_resident_bytes(const flsgpu_table* t, uint64_t* out);

/* ---- decode: DecodePhysicalColumns (fls_reader.cpp:436-472) + KernelTraits<Op>::Decode ------------------ */
/* Destination physical type of a decoded column = DuckDB PhysicalType the materializer writes
 * (fls_type_resolver.hpp:42-83).  Same-width copy or widening static_cast (CastNumeric, :22-33). */
typedef enum flsgpu_out_type {
	FLSGPU_OUT_NATIVE = 0, /* the column's own physical width */
	FLSGPU_OUT_I8 = 1, FLSGPU_OUT_I16 = 2, FLSGPU_OUT_I32 = 3, FLSGPU_OUT_I64 = 4,
	FLSGPU_OUT_U8 = 5, FLSGPU_OUT_U16 = 6, FLSGPU_OUT_U32 = 7, FLSGPU_OUT_U64 = 8,
	FLSGPU_OUT_F32 = 9, FLSGPU_OUT_F64 = 10
} flsgpu_out_type;

/* Decode column `col` of row groups [rg_begin, rg_begin+n_rg) in ONE launch into out_dev (device memory,
 * 16-byte aligned).  Row group r's vector v lands at element ((r-rg_begin)*vec_per_rowgroup + v)*1024: whole
 * vectors are written, the caller exposes only n_tuples of each row group (fls_reader.cpp:429-431).
 * out_capacity_bytes must cover n_rg*vec_per_rowgroup*1024 elements. Asynchronous on `stream`.
 * Consecutive decode launches of one stream are chained by programmatic dependent launch (they may overlap each other:
 * do not aim two back-to-back decodes at overlapping output ranges); anything else enqueued on the stream afterwards —
 * copies, the caller's kernels, events — sees every earlier decode complete, as usual.  FLSGPU_PDL=0 turns the chaining off. */
int flsgpu_decode_column(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, int32_t out_type,
                         void* out_dev, uint64_t out_capacity_bytes, void* cuda_stream);

/* Per-row-group form used by the reader's TryInitializeScan (one row group per worker, fls_reader.cpp:337-355).
 * out_ptrs[i] receives column col_ids[i]; pointers may be device memory or pinned host memory mapped for the
 * device. */
int flsgpu_rowgroup_decode(flsgpu_table* t, uint32_t rg, const uint32_t* col_ids, uint32_t n_cols,
                           void* const* out_ptrs, const int32_t* out_types, void* cuda_stream);

/* ---- string columns (VARCHAR / BLOB): the materializer's string kernels ------------------------------------------ */
/* The decoded element is DuckDB's 16-byte string_t (duckdb/common/types/string_type.hpp:22-60,236-247): u32 length, then
 * 12 inlined bytes (zero padded) when length <= 12, else 4 prefix bytes + a pointer.  The bytes of non-inlined strings go
 * to ONE string heap per call (the reference allocates per string: uncompressed_kernel.hpp:26-38, fsst_kernel.hpp:42-56);
 * pointers are heap_base_addr + offset, so the caller decides where the heap is going to live (its own host buffer after
 * a D2H copy, DuckDB's VectorStringBuffer, or the device heap itself: heap_base_addr = (uint64_t)heap_dev).
 * Row layout as for flsgpu_decode_column (whole vectors; padding rows are empty strings or repeat the last code). */
/* Bytes of heap a flat decode of [rg_begin, rg_begin + n_rg) needs: every chunk's decoded bytes, padded to 16. Host only. */
int flsgpu_string_heap_bytes(const flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* out);
/* Flat materialisation: dec_fls_str_uncompressed_opr / dec_constant_str_opr / dec_fsst_opr / dec_fsst_dict_opr /
 * dec_dict_opr<fls_string_t> -> string_t per row.  Two launches (heap fill incl. parallel FSST decode, then rows). */
int flsgpu_decode_strings(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_strings_dev,
                          uint64_t out_capacity_bytes, void* heap_dev, uint64_t heap_capacity_bytes, uint64_t heap_base_addr,
                          void* cuda_stream);
/* Dictionary form, as the reference hands dictionary-encoded chunks to DuckDB (DICTIONARY_VECTOR: dictionary_kernel.hpp:
 * 60-101, fsst_dict_kernel.hpp:54-80): the u32 selection vector of every row ... */
int flsgpu_decode_string_codes(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint32_t* out_codes_dev,
                               uint64_t out_capacity_bytes, void* cuda_stream);
/* ... and the string_t of the keys of every chunk of the range: chunk r's keys follow those of the range's earlier
 * chunks (fls_chunk_desc.aux_count keys each; flsgpu_string_dict_keys gives the total), its dictionary bytes sit in its
 * slice of the heap (same layout as the flat form).  FLSGPU_ERR_UNSUPPORTED when a chunk of the range is not
 * DICT_STR / FSST_DICT (the reader then takes the flat form for that row group). */
int flsgpu_string_dict_keys(const flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* out);
int flsgpu_decode_string_dict(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_keys_dev,
                              uint64_t keys_capacity_bytes, void* heap_dev, uint64_t heap_capacity_bytes, uint64_t heap_base_addr,
                              void* cuda_stream);
/* Host to host: stage + decode + D2H of string_t rows and heap, pipelined in slabs; pointers refer to heap_host. */
int flsgpu_scan_strings_to_host(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_strings_host,
                                uint64_t out_capacity_bytes, void* heap_host, uint64_t heap_capacity_bytes);

/* ---- fused decode + aggregate (SUM/COUNT/MIN/MAX over a scan) ---------------------------- */
typedef struct flsgpu_agg_result { /* written to device memory, 64 bytes */
	uint64_t sum_lo;  /* 128-bit two's complement sum (HUGEINT / DECIMAL(38) semantics, sum.cpp:161-210) */
	uint64_t sum_hi;
	uint64_t count;
	int64_t min_val;  /* integer columns: the column's physical value, signed or unsigned 64 bits; not produced for FLOAT / DOUBLE */
	int64_t max_val;
	double fsum;      /* FLOAT/DOUBLE columns: sum in double (not order-stable, reported only) */
	uint64_t reserved[2];
} flsgpu_agg_result;
/* Accumulates into *result_dev (caller zero-initialises with flsgpu_agg_reset). Asynchronous.  SUM / COUNT are folded by the
 * kernels; min_val / max_val are merged from the zone maps of the chunks in the range (fls_chunk_desc.has_stats, min_val,
 * max_val: exact, written from the data; integer columns): they cover the chunks that carry statistics - a caller that needs
 * MIN / MAX of the whole range checks has_stats of its chunks first (the aggregate pushdown of the extension does). */
int flsgpu_aggregate_column(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg,
                            flsgpu_agg_result* result_dev, void* cuda_stream);
int flsgpu_agg_reset(flsgpu_agg_result* result_dev, void* cuda_stream);

/* ---- filter pushdown: RowGroupFilter (row_group_filter.cpp:96-199) + FilterExecutor (filter_executor.cpp:19-60) ------ */
/* Constant comparisons (DuckDB ConstantFilter), combined with AND like a TableFilterSet.  Numeric columns: `constant` is
 * the physical value of the column's type: integers sign- / zero-extended to 64 bits (DATE days, TIMESTAMP micros,
 * DECIMAL unscaled), FLOAT bits in the low word, DOUBLE bits.  Floating point follows DuckDB: NaN equals NaN and is
 * greater than every other value.  VARCHAR / BLOB columns: the constant is the str_len bytes at str_data (host memory,
 * read during the call; `constant` is ignored), compared like DuckDB's string_t: bytewise over the common length, the
 * shorter string first on a tie (string_type.hpp:155-217).  String chunks carry no zone maps: they never prune. */
typedef enum flsgpu_cmp {
	FLSGPU_CMP_EQ = 0, FLSGPU_CMP_NE = 1, FLSGPU_CMP_LT = 2, FLSGPU_CMP_LE = 3, FLSGPU_CMP_GT = 4, FLSGPU_CMP_GE = 5
} flsgpu_cmp;
typedef struct flsgpu_filter {
	uint32_t col;
	int32_t op; /* flsgpu_cmp */
	uint64_t constant;
	const uint8_t* str_data; /* string columns only (may be NULL when str_len == 0) */
	uint64_t str_len;
} flsgpu_filter;
/* RowGroupMaySatisfyFilters: keep_out[r] = 0 when the zone maps (fls_chunk_desc.min_val / max_val) of row group
 * rg_begin + r prove that no row passes.  Host only. */
int flsgpu_rowgroup_prune(const flsgpu_table* t, const flsgpu_filter* filters, uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg,
                          uint8_t* keep_out);
/* FilterExecutor::Apply on the device: decode each filter column (numeric: fused, nothing is written but the mask; string:
 * slabs of row groups are decoded to string_t rows in stream-ordered scratch memory and compared there) and set bit i
 * of word w of vector v (mask_dev[v * 32 + w], vectors laid out as in flsgpu_decode_column) when row v*1024 + w*32 + i
 * passes every filter.  Padding rows never pass.  rg_keep_dev (device copy of flsgpu_rowgroup_prune's output, or NULL):
 * pruned row groups are not decoded and select nothing. */
int flsgpu_filter_mask(flsgpu_table* t, const flsgpu_filter* filters, uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg,
                       const uint8_t* rg_keep_dev, uint32_t* mask_dev, uint64_t mask_capacity_bytes, void* cuda_stream);
/* SUM / COUNT of the rows whose mask bit is set (SELECT SUM(col), COUNT(*) ... WHERE <filters>): accumulates into
 * *result_dev like flsgpu_aggregate_column; min_val / max_val are not produced. */
int flsgpu_aggregate_column_masked(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, const uint32_t* mask_dev,
                                   const uint8_t* rg_keep_dev, flsgpu_agg_result* result_dev, void* cuda_stream);
/* The host-to-host form of the fused aggregates - what a rewrite of SELECT SUM(c), COUNT(*), MIN(c), MAX(c) FROM read_fls(...)
 * [WHERE <constant comparisons>] runs (the reference decodes every value into DuckDB vectors and lets DuckDB's
 * aggregate operators consume them: fls_reader.cpp:436-472 -> ungrouped aggregate).  Stages the compressed chunks of the
 * columns the query touches (slabs of row groups on two streams, H2D overlapping the kernels), evaluates the filters into a
 * device mask, folds every listed numeric column and writes one flsgpu_agg_result per column to results_host[i].  With
 * filters only sum_lo / sum_hi / count / fsum are produced (min_val / max_val stay 0); without, MIN / MAX come from the
 * chunks' zone maps (integer columns whose chunks carry statistics).  Synchronous. */
int flsgpu_aggregate_to_host(flsgpu_table* t, const uint32_t* col_ids, uint32_t n_cols, const flsgpu_filter* filters, uint32_t n_filters,
                             uint32_t rg_begin, uint32_t n_rg, flsgpu_agg_result* results_host);
/* Allocation-free variant of flsgpu_filter_mask for ONE constant comparison on a VARCHAR / BLOB column whose chunks in the
 * range are all dictionary encoded (DICT_STR / FSST_DICT / FSST12_DICT).  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The scratch flsgpu_filter_mask takes from a stream-ordered pool (string_t + bytes of
 * the range's dictionary keys, one flag byte per key, the constant, the row groups' tuple counts) lives in a static device
 * array of FLSGPU_STATIC_SCRATCH_BYTES that every device loading the library holds until exit; the call issues no
 * cudaMalloc / cudaFree / pool allocation.  Requirements: the table is staged already (staging allocates the arena; a
 * table without device state is refused), n_rg <= 1024, and the range's scratch need fits the array - otherwise
 * FLSGPU_ERR_UNSUPPORTED with the sizes in the message (split the range, or use flsgpu_filter_mask).  Same mask layout and
 * rg_keep_dev meaning as flsgpu_filter_mask; the mask is overwritten, not ANDed.  Asynchronous on cuda_stream.  All calls of
 * a device share the one array: calls that overlap on the GPU (different streams, or threads) race - serialise them. */
#define FLSGPU_STATIC_SCRATCH_BYTES (4u << 20)
int flsgpu_filter_mask_string_dict_static(flsgpu_table* t, const flsgpu_filter* filter, uint32_t rg_begin, uint32_t n_rg,
                                          const uint8_t* rg_keep_dev, uint32_t* mask_dev, uint64_t mask_capacity_bytes,
                                          void* cuda_stream);

/* ---- end to end from host buffers: what FastLanesReader::Scan hands to DuckDB -------------------------- */
/* Stage (H2D) + decode + copy back (D2H) into pinned/pageable host memory `out_host`, pipelined in slabs of
 * row groups over two streams.  Synchronous: returns when out_host is complete. */
int flsgpu_scan_column_to_host(flsgpu_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, int32_t out_type,
                               void* out_host, uint64_t out_capacity_bytes);

/* Projection form (FastLanesReaderProjectionPlan, fls_multi_file_info.hpp:48-55): all projected columns of a slab are
 * staged with merged copies, decoded and copied back together. out_hosts[i] receives column col_ids[i]. */
int flsgpu_scan_columns_to_host(flsgpu_table* t, const uint32_t* col_ids, uint32_t n_cols, uint32_t rg_begin, uint32_t n_rg,
                                const int32_t* out_types, void* const* out_hosts, const uint64_t* out_capacity_bytes);

/* General projection (numeric columns and string columns in either form) through the same slab pipeline. */
typedef enum flsgpu_string_form {
	FLSGPU_STRING_FLAT = 0,      /* out_host: string_t per row; heap_host: string bytes */
	FLSGPU_STRING_DICTIONARY = 1 /* out_host: u32 selection entry per row; keys_host: string_t per key; heap_host: dictionary bytes */
} flsgpu_string_form;
typedef struct flsgpu_scan_column {
	uint32_t col;
	int32_t out_type;           /* numeric columns: flsgpu_out_type */
	int32_t string_form;        /* string columns: flsgpu_string_form */
	uint32_t compact;           /* numeric columns, 1: only the FLAT vectors of flsgpu_column_vector_forms are decoded and copied;
	                               vector k of them lands at element flat_slot * 1024 of out_host (CONSTANT / SEQUENCE vectors
	                               cost nothing: the caller builds them from their form).  out_capacity_bytes >= n_flat * 1024 * width */
	void* out_host;             /* rows of the column, layout of flsgpu_decode_column */
	uint64_t out_capacity_bytes;
	void* heap_host;            /* string columns: >= flsgpu_string_heap_bytes; string_t pointers point into it */
	uint64_t heap_capacity_bytes;
	void* keys_host;            /* dictionary form: >= 16 * flsgpu_string_dict_keys */
	uint64_t keys_capacity_bytes;
	const flsgpu_vector_form* forms; /* compact: what flsgpu_column_vector_forms returned for this very range, when the caller has
	                               it already (it needs it to read the result); NULL: the call works the forms out itself */
} flsgpu_scan_column;
int flsgpu_scan_projection_to_host(flsgpu_table* t, const flsgpu_scan_column* cols, uint32_t n_cols, uint32_t rg_begin, uint32_t n_rg);
/* The same scan with pushed-down filters (FastLanesReader::Scan -> FilterExecutor::Apply, fls_reader.cpp:516-547): every
 * projected column is returned in full, plus the selection mask of the filter set (flsgpu_filter_mask layout) so that the
 * reader slices each 2048-row chunk with a selection vector.  Filter columns need not be projected. */
int flsgpu_scan_filtered_to_host(flsgpu_table* t, const flsgpu_scan_column* cols, uint32_t n_cols, const flsgpu_filter* filters,
                                 uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg, uint32_t* mask_host, uint64_t mask_capacity_bytes);

/* The same scan with the selection applied on the device (chunk.Slice(sel, count) of filter_executor.cpp:57-59, done before the
 * copy instead of after it): only the rows that pass every filter are copied to the host, packed back to back over the whole
 * range in row order - out_host of column i holds *n_selected_out rows when the call returns (its capacity must be that of
 * the unfiltered column: how many rows pass is not known beforehand).  String columns: flat (packed string_t rows, the heap
 * copied whole), or dictionary form when every chunk of the range is dictionary encoded - then keys_host / heap_host receive
 * the keys and dictionary bytes of ALL row groups of the range back to back (as flsgpu_decode_string_dict lays them out) and
 * the packed u32 entries index that one range-wide dictionary, so that rows of different row groups can share a DuckDB
 * DICTIONARY_VECTOR.  `compact` / `forms` are ignored.
 * row_index_host (optional): for every selected row its position (rg - rg_begin) * vec_per_rowgroup * 1024 + row inside the
 * range, u32 - what a reader needs for file_row_number.  n_filters must be >= 1.  Synchronous. */
int flsgpu_scan_selected_to_host(flsgpu_table* t, const flsgpu_scan_column* cols, uint32_t n_cols, const flsgpu_filter* filters,
                                 uint32_t n_filters, uint32_t rg_begin, uint32_t n_rg, uint32_t* row_index_host,
                                 uint64_t row_index_capacity_bytes, uint64_t* n_selected_out);

int flsgpu_sync(void* cuda_stream);
/* Pinned host allocations for staging / results (cudaHostAlloc). */
int flsgpu_host_alloc(uint64_t bytes, void** out);
int flsgpu_host_free(void* p);

/* Bytes this process has handed to the copy engines so far: staging of compressed chunks + vector maps (host to device) and
 * the results of the host-to-host scans (device to host).  bench.py reports the per-step difference as e2e h2d / d2h bytes. */
void flsgpu_transfer_bytes(uint64_t* h2d, uint64_t* d2h);
/* Launch counter: number of decode/aggregate kernels this process has launched (bench.py gpu_launches). */
uint64_t flsgpu_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* FLSGPU_H */

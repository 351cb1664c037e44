/* fls_reader_capi.h — C view of the reader mirror (csrc/reader/fls_reader.hpp) for the Python tests and the bench:
 * an executor stand-in that drives TryInitializeScan/Scan on N worker threads the way DuckDB's MultiFileScan does
 * (duckdb/src/include/duckdb/common/multi_file/multi_file_function.hpp:397-472,591-616).  Not part of the drop-in
 * boundary (that is include/flsgpu.h); it exists so that the reader logic is testable without a DuckDB build. */
#ifndef FLS_READER_CAPI_H
#define FLS_READER_CAPI_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct flsreader_handle flsreader_handle;
const char* flsreader_last_error(void);
int flsreader_open(const char* path, int file_row_number, uint64_t explicit_cardinality, uint64_t rowgroups_per_batch,
                   flsreader_handle** out);
void flsreader_close(flsreader_handle* h);
int flsreader_n_columns(const flsreader_handle* h);
/* type_id: flsreader::LogicalTypeId */
int flsreader_column(const flsreader_handle* h, int i, char* name64, int* type_id, int* physical_bytes, int* width, int* scale);
uint64_t flsreader_cardinality(const flsreader_handle* h);
uint64_t flsreader_max_threads(const flsreader_handle* h);
uint64_t flsreader_total_tuples(const flsreader_handle* h);
double flsreader_progress(const flsreader_handle* h);
int flsreader_partition_stats(const flsreader_handle* h, uint64_t* row_start, uint64_t* count, uint64_t capacity, uint64_t* n);
int flsreader_statistics(const flsreader_handle* h, const char* name, int* has_stats, int64_t* min_i, int64_t* max_i, double* min_d,
                         double* max_d);
/* SELECT column_ids FROM read_fls(path): n_threads workers; chunk rows land in out_cols[c] at their file row number.
 * column id == n_columns (with file_row_number enabled) selects the virtual column. */
int flsreader_scan_all(flsreader_handle* h, const uint64_t* column_ids, uint32_t n_cols, int n_threads, void* const* out_cols,
                       uint64_t* rows_out, uint64_t* chunks_out, uint64_t* max_chunk_rows);
/* out_cols (or one of its entries) may be NULL: that column's chunks are produced and counted but not copied anywhere - the
 * consumer takes each vector as it is, like a DuckDB operator (the bench's end-to-end scan).
 * General form: string columns (VARCHAR / BLOB) are written through a fixed-stride sink (out_cols[c] points at a
 * flsreader_string_sink), optional pushed-down filters (column_index addresses column_ids, constants are physical values
 * as in flsgpu_filter).  With filters the passing rows are appended in arrival order and row_ids_out receives their file
 * row numbers; without, rows land at their file row number and row_ids_out is ignored. */
typedef struct flsreader_string_sink {
	uint8_t* bytes;    /* row r occupies bytes[r*stride, r*stride + lengths[r]) */
	uint32_t* lengths;
	uint32_t stride;
} flsreader_string_sink;
typedef struct flsreader_filter {
	uint64_t column_index;
	int32_t op; /* flsgpu_cmp */
	uint64_t constant;
	const uint8_t* str_data; /* VARCHAR / BLOB columns: the constant's bytes (copied by the call), else NULL */
	uint64_t str_len;
} flsreader_filter;
typedef struct flsreader_scan_stats {
	uint64_t rows, chunks, max_chunk_rows;
	uint64_t dictionary_vectors; /* vectors handed over as DICTIONARY_VECTOR (dictionary_kernel.hpp:80-101) */
	uint64_t pruned_rowgroups;   /* row groups the zone maps ruled out (row_group_filter.cpp:80-94) */
	uint64_t constant_vectors;   /* vectors handed over as CONSTANT_VECTOR (constant_kernel.hpp:19-27) */
	uint64_t sequence_vectors;   /* ... as SEQUENCE_VECTOR (start + i * increment) */
	uint64_t assembled_vectors;  /* chunks put together on the host because their two halves had different forms */
} flsreader_scan_stats;
/* Reader options without a named parameter of read_fls: "compressed_vectors" (default 1). */
int flsreader_set_option(flsreader_handle* h, const char* key, uint64_t value);
/* Forget which row groups are resident in HBM (the next scan stages them again): host-fed timing. */
int flsreader_evict(flsreader_handle* h);
int flsreader_scan(flsreader_handle* h, const uint64_t* column_ids, uint32_t n_cols, const flsreader_filter* filters, uint32_t n_filters,
                   int n_threads, void* const* out_cols, int64_t* row_ids_out, flsreader_scan_stats* stats);
/* ---- read_fls over several files: FastLanesMultiFileInfo + MultiFileReader semantics (csrc/reader/fls_multi_file.hpp) ---- */
typedef struct flsmulti_handle flsmulti_handle;
/* patterns: paths or globs (read_fls(VARCHAR | LIST(VARCHAR))); option_keys / option_values: named parameters as text
 * (explicit_cardinality, file_row_number, union_by_name, filename); a NULL value pointer is SQL NULL. */
int flsmulti_open(const char* const* patterns, uint32_t n_patterns, const char* const* option_keys, const char* const* option_values,
                  uint32_t n_options, flsmulti_handle** out);
void flsmulti_close(flsmulti_handle* h);
int flsmulti_n_files(const flsmulti_handle* h);
int flsmulti_file(const flsmulti_handle* h, int i, char* name256, uint64_t* n_tuples);
int flsmulti_n_columns(const flsmulti_handle* h);
int flsmulti_column(const flsmulti_handle* h, int i, char* name64, int* type_id, int* physical_bytes, int* width, int* scale);
uint64_t flsmulti_cardinality(const flsmulti_handle* h);
uint64_t flsmulti_max_threads(const flsmulti_handle* h); /* ~0: no limit */
/* Rows are appended in arrival order; file_ids_out / row_ids_out identify them (file index, file_row_number).
 * valid_out[c] (one byte per row, may be NULL) is 0 where a union_by_name column is missing from the row's file. */
int flsmulti_scan(flsmulti_handle* h, const uint64_t* column_ids, uint32_t n_cols, const flsreader_filter* filters, uint32_t n_filters,
                  int n_threads, void* const* out_cols, uint8_t* const* valid_out, uint32_t* file_ids_out, int64_t* row_ids_out,
                  flsreader_scan_stats* stats);
#ifdef __cplusplus
}
#endif
#endif

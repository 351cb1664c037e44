/*
 * fls_oracle.h — CPU restatement of the reference's read_fls decode path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library; the product (libflsgpu.so) never links or calls it.
 *
 * PARITY STATUS: the reference's decode arithmetic lives in an un-vendored, unpinned third-party library
 * (github.com/sebastiaan-dev/fastlanes-patched @ dev; reference CMakeLists.txt:56-62), so byte-level parity
 * with the reference binary is UNPINNED.  What is pinned:
 *   - ALP / ALP-RD value arithmetic against DuckDB's vendored implementation, which is in the reference tree
 *     (duckdb/src/include/duckdb/storage/compression/alp/...), via oracle/_ref and tests/golden/alp_*.bin;
 *   - the value-level round trip contract of the reference's own tests
 *     (test/all_types_single_threaded.test:280-352): decode(encode(x)) == x for every type;
 *   - materializer semantics that ARE in the reference tree (cited per function in fls_oracle.c).
 */
#ifndef FLS_ORACLE_H
#define FLS_ORACLE_H
#include <stdint.h>
#include "fls_format.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fls_oracle_table {
	const uint8_t* bytes;
	uint64_t size;
	fls_file_header header;
	const fls_column_desc* columns;
	const fls_rowgroup_desc* rowgroups;
	const fls_chunk_desc* chunks;
} fls_oracle_table;

/* out types: same numbering as flsgpu_out_type (0 = native width). */
int fls_oracle_open(const uint8_t* bytes, uint64_t size, fls_oracle_table* t);

/* "Library" step: decode vector `vec` of chunk (rg, col) into scratch (1024 x T bits, physical words).  For DELTA
 * the scratch holds the TRANSPOSED data, exactly like dec_transpose_opr::transposed_data; *is_transposed is set. */
int fls_oracle_decode_vector(const fls_oracle_table* t, uint32_t rg, uint32_t col, uint32_t vec, void* scratch,
                             int* is_transposed);

/* Materializer step (KernelTraits<Op>::Decode): copy / widen / untranspose scratch into out (1024 elements). */
int fls_oracle_materialize(const fls_oracle_table* t, uint32_t col, const void* scratch, int is_transposed, int out_type,
                           void* out);

/* Whole column range, same output layout as flsgpu_decode_column; n_threads host threads, one row group per
 * thread at a time (reference fls_multi_file_info.cpp:99-110). */
int fls_oracle_decode_column(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, int out_type,
                             void* out, int n_threads);

/* Reference-shaped scan used as the timed CPU baseline: every thread owns a 2048-row chunk buffer per column and
 * decodes two vectors per Scan call (fls_reader.cpp:422-472), folding each chunk into a checksum so the work is
 * observable.  Returns the number of decoded BYTES (as materialised) through *decoded_bytes. */
int fls_oracle_scan(const fls_oracle_table* t, const uint32_t* cols, uint32_t n_cols, uint32_t rg_begin, uint32_t n_rg,
                    int n_threads, uint64_t* decoded_bytes, uint64_t* checksum);

/* Layout primitives exposed for known-answer tests. */
int fls_oracle_index(int row, int lane);
int fls_oracle_transpose(int pos);
void fls_oracle_unpack(const void* packed, unsigned T, unsigned W, void* out /* 1024 x T bits by position */);
void fls_oracle_untranspose(const void* transposed, unsigned T, void* out);
double fls_oracle_alp_decode_f64(int64_t digit, int e, int f);
float fls_oracle_alp_decode_f32(int32_t digit, int e, int f);

/* 128-bit SUM / COUNT / MIN / MAX over a decoded column range: checker for the fused aggregate kernels. */
int fls_oracle_aggregate(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* sum_lo,
                         uint64_t* sum_hi, uint64_t* count, int64_t* min_val, int64_t* max_val, double* fsum);

/* ---- filter pushdown (fls_oracle_scan.c): constant comparisons, op numbering of flsgpu_cmp (EQ NE LT LE GT GE) ---------- */
typedef struct fls_oracle_filter {
	uint32_t col;
	int32_t op;
	uint64_t constant; /* physical bits, integers extended to 64 */
	const uint8_t* str_data; /* VARCHAR / BLOB columns: the constant's bytes (compared bytewise, then by length) */
	uint64_t str_len;
} fls_oracle_filter;
int fls_oracle_filter_mask(const fls_oracle_table* t, const fls_oracle_filter* filters, uint32_t n_filters, uint32_t rg_begin,
                           uint32_t n_rg, uint32_t* mask /* n_rg * vec_per_rowgroup * 32 words */);
int fls_oracle_rowgroup_may_satisfy(const fls_oracle_table* t, const fls_oracle_filter* filters, uint32_t n_filters, uint32_t rg);
int fls_oracle_aggregate_masked(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, const uint32_t* mask,
                                uint64_t* sum_lo, uint64_t* sum_hi, uint64_t* count, double* fsum);

/* ---- string columns (fls_oracle_str.c) ---------------------------------------------------------------------------- */
void fls_oracle_rle_index(const uint8_t* blob, unsigned width, uint16_t dmin, uint16_t* idx /* 1024 */);
int fls_oracle_decode_codes(const fls_oracle_table* t, uint32_t rg, uint32_t col, uint32_t vec, uint32_t* codes /* 1024 */);
uint64_t fls_oracle_fsst_decompress(const fls_fsst_table* tab, const uint8_t* in, uint64_t len_in, uint8_t* out);
uint64_t fls_oracle_fsst12_decompress(const fls_fsst12_table* tab, const uint8_t* in, uint64_t len_in, uint8_t* out);
/* Bytes of string heap a flat decode of [rg_begin, rg_begin + n_rg) needs (sum of the chunks' 16-byte padded decoded sizes). */
int fls_oracle_string_heap_bytes(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* out);
/* Flat materialisation: 16-byte string_t per row (same row layout as fls_oracle_decode_column) + heap; pointers of
 * non-inlined strings are heap_base_addr + offset into heap. */
int fls_oracle_decode_strings(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_strings,
                              uint8_t* heap, uint64_t heap_base_addr);
/* Dictionary form: string_t of every key of chunk (rg, col) + the dictionary bytes (const_val of them). */
int fls_oracle_decode_string_dict(const fls_oracle_table* t, uint32_t col, uint32_t rg, void* out_keys, uint8_t* heap,
                                  uint64_t heap_base_addr);

/* One vector of a per-row string chunk (STR_CONSTANT / STR_UNCOMPRESSED / FSST) as the reference materializer decodes it
 * (string by string); -3 for dictionary encodings (those go through decode_string_dict + decode_codes). */
int fls_oracle_decode_string_vector(const fls_oracle_table* t, uint32_t rg, uint32_t col, uint32_t vec, void* out_strings,
                                    uint8_t* heap, uint64_t heap_base_addr, uint64_t* heap_used);

#ifdef __cplusplus
}
#endif
#endif

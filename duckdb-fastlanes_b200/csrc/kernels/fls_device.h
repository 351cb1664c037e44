// fls_device.h — structs shared by the host launcher and the kernels.
#pragma once
#include <stdint.h>
#include "fls_format.h"

// Device-side view of one column chunk (64 bytes, built at stage time from fls_chunk_desc).
struct alignas(16) FlsChunkDev {
	const uint8_t* data;      // DATA segment
	const fls_vec_meta* meta; // META segment
	const uint8_t* aux0;      // dictionary keys | run values | exception values | CROSS_RLE values
	const uint8_t* aux1;      // exception positions (u16) | CROSS_RLE run STARTS (u32 exclusive prefix, n_runs+1)
	uint64_t const_val;
	uint32_t aux_count;
	uint16_t n_vec;
	uint8_t encoding;
	uint8_t idx_bits;
	uint16_t rd_dict[8];
};
static_assert(sizeof(FlsChunkDev) == 64, "FlsChunkDev must be 64 bytes");

// Device-side accumulator of the fused aggregates; same 64 bytes as flsgpu_agg_result (include/flsgpu.h).  min/max are
// kept as ORDER KEYS (value ^ sign bit for signed columns) in key_min/key_max so that a zeroed-then-reset accumulator
// works for either signedness; flsgpu_aggregate_column's finalize step writes min_val/max_val from them.
struct flsgpu_agg_acc {
	unsigned long long sum_lo, sum_hi, count;
	long long min_val, max_val;
	double fsum;
	unsigned long long key_min, key_max;
};
static_assert(sizeof(flsgpu_agg_acc) == 64, "flsgpu_agg_acc must mirror flsgpu_agg_result");

// Launch parameters of one decode launch: column `col` of row groups [rg_begin, rg_begin + n_rg).
struct FlsLaunch {
	const FlsChunkDev* chunks; // &chunk_dev[rg_begin * n_cols + col]
	uint32_t chunk_stride;     // n_cols
	uint32_t vec_per_rg;
	uint32_t total_vecs;       // n_rg * vec_per_rg (vectors beyond a chunk's n_vec are skipped)
	uint32_t slot_bytes;       // shared-memory bytes per pipeline slot (multiple of 128)
	uint32_t stage_bytes;      // per-warp staging area for transposition (0 if no chunk needs it)
	uint32_t n_slots;          // pipeline depth per warp
	uint32_t conv;             // 0: zero-extend / same width, 1: sign-extend, 2: f32 -> f64
	uint32_t family_mask;      // bit e set: this kernel handles encoding e
	uint32_t run_shift;        // log2 of the run of consecutive vectors a warp takes (0..5)
	uint32_t idx_bits;         // DICT launches: only chunks whose codes sit in idx_bits-wide lanes (0: any)
	uint32_t prefetch;         // 1: request a batch's 32 blobs into L2 at batch start (materialising launches)
	uint32_t pdl;              // 1: materialising launch issued with programmatic stream serialization: let the next decode
	                           //    launch of the stream start as this one's CTAs retire (griddepcontrol), see launch_family
	void* out;
	// fused aggregate launches (flsgpu_aggregate_column): out is unused, partials are folded into *agg
	flsgpu_agg_acc* agg;
	uint32_t agg_signed; // integer columns: 1 = two's complement values
	uint32_t agg_float;  // 1 = FLOAT/DOUBLE column (fsum + count only)
	// filter pushdown (aggregate-family launches only; FilterExecutor::Apply, reference filter_executor.cpp:19-60)
	uint32_t mask_mode;  // 0: none; 1: mask = cmp; 2: mask &= cmp; 3: aggregate only rows whose mask bit is set
	uint32_t cmp_op;     // flsgpu_cmp
	uint64_t cmp_const;  // physical bits of the constant (integers sign/zero-extended to 64, float/double raw bits)
	uint32_t cmp_op2;    // second comparison on the same column, ANDed in the same pass (0xff: none): ranges cost one decode
	uint64_t cmp_const2;
	uint32_t* mask;      // [total_vecs * 32] words: bit i of word w of vector v = row v*1024 + w*32 + i
	const uint8_t* rg_keep; // per row group of the range: 0 = pruned by its zone map (skipped), or nullptr
	// compact scans (flsgpu_scan_column.compact): the launch walks total_vecs LISTED vectors, vec_map[k] = vector index inside
	// the range of the k-th one, whose values land at slot k of the output; nullptr = every vector of the range, in place
	const uint32_t* vec_map;
	// 1: the listed vectors are decoded IN PLACE (output position = the vector's own index inside the range, not its slot in the
	// list): launches of a column whose chunks fall into several kernel families walk only their own family's vectors
	uint32_t map_in_place;
};

// Launch parameters of the string kernels (flsgpu_kernels_str.cu): column `col` of row groups [rg_begin, rg_begin + n_rg).
struct FlsStrLaunch {
	const FlsChunkDev* chunks;   // &chunk_dev[rg_begin * n_cols + col]
	const uint64_t* heap_prefix; // &heap_prefix[rg_begin * n_cols + col]: 16-byte padded decoded bytes of the column's chunks before this one
	uint64_t heap_prefix0;       // heap_prefix of row group rg_begin (the heap handed in starts there)
	const uint64_t* key_prefix;  // &key_prefix[rg_begin * n_cols + col]: dictionary keys of the column's chunks before this one
	uint64_t key_prefix0;
	uint8_t* heap;               // device string heap, 16-byte aligned
	uint64_t heap_base_addr;     // address the heap will have where the strings are consumed (string_t pointers)
	void* out;                   // string_t[n_rg * vec_per_rg * 1024] | u32 codes | string_t[n_keys]
	uint32_t chunk_stride;       // n_cols
	uint32_t vec_per_rg;
	uint32_t n_rg;
	uint32_t groups_per_rg;      // heap kernel: CTAs per chunk (set by the launcher)
	uint32_t codes_only;         // rows kernel: emit the u32 selection vector instead of string_t
	uint32_t dict_only;          // heap kernel: fill only the slices of dictionary-encoded chunks
	uint32_t lane_strings;       // FSST decode: strings per lane (a warp tile = 32 * lane_strings strings)
	uint32_t cbuf_bytes;         // per-warp shared-memory staging of a tile's codes / decoded bytes (multiples of 16;
	uint32_t dbuf_bytes;         //  tiles that do not fit are decoded straight to global memory)
	uint32_t fuse_rows;          // FSST chunks: the heap kernel also builds the string_t rows (the rows kernel skips them)
	uint32_t balance;            // FSST decode: lanes split a tile by code bytes (cut at string boundaries), not by string count
	uint32_t fuse_keys;          // dictionary decode of plain DICT_STR chunks: the heap kernel also builds the keys' string_t (from the
	                             // source bytes, which the heap slice copies verbatim), the keys launch is not issued
	// rows kernel, filter mode (dictionary-encoded chunks only): the comparison was evaluated once per dictionary key
	// (key_flags, in key_prefix order), the rows only look their code's flag up -> selection mask words, nothing else written
	const uint8_t* key_flags;    // nullptr: not a filter launch
	const uint32_t* n_tuples;    // tuples of every row group of the range
	const uint8_t* rg_keep;      // per row group: 0 = pruned (selects nothing), or nullptr
	uint32_t* filter_mask;       // [n_rg * vec_per_rg * 32] words
	uint32_t filter_and;         // 0: mask = flags; 1: mask &= flags
	uint32_t pdl;                // 1: launch chaining (programmatic dependent launch) is on, see flsgpu_kernels_str.cu
};

// Launch parameters of the string filter kernel (flsgpu_kernels_str.cu): a constant comparison over the decoded string_t
// rows of a slab of row groups (FilterExecutor::Apply on a VARCHAR / BLOB column, reference filter_executor.cpp:19-60).
struct FlsStrFilterLaunch {
	const void* rows;          // string_t[n_rg * vec_per_rg * 1024]; pointers of non-inlined strings are device addresses
	const uint32_t* n_tuples;  // tuples of every row group of the slab (rows beyond them never pass)
	const uint8_t* rg_keep;    // per row group of the slab: 0 = pruned (selects nothing), or nullptr
	const uint8_t* const_data; // the constant's bytes on the device
	uint32_t const_len;
	uint32_t const_prefix;     // its first 4 bytes, zero padded (string_t prefix layout)
	uint32_t op;               // flsgpu_cmp
	uint32_t and_into;         // 0: mask = cmp; 1: mask &= cmp
	uint32_t* mask;            // [n_rg * vec_per_rg * 32] words of the slab
	uint32_t n_rg;
	uint32_t vec_per_rg;
	// dictionary form: rows = the string_t of n_keys dictionary keys, one flag byte per key goes to key_flags_out
	uint8_t* key_flags_out;
	uint32_t n_keys;
};

// Per-warp shared memory holding the 32 vector descriptors of a batch (64 B each), in front of the pipeline slots.
#define FLS_DESC_BYTES_PER_WARP 2048u

// Encoding families (which kernel template handles which fls_encoding)
#define FLS_FAMILY_MAIN                                                                                                \
	((1u << FLS_ENC_UNCOMPRESSED) | (1u << FLS_ENC_CONSTANT) | (1u << FLS_ENC_FFOR) | (1u << FLS_ENC_DELTA) |          \
	 (1u << FLS_ENC_ALP) | (1u << FLS_ENC_ALP_RD) | (1u << FLS_ENC_CROSS_RLE) | (1u << FLS_ENC_FREQUENCY) |            \
	 (1u << FLS_ENC_SLPATCH))
#define FLS_FAMILY_DICT (1u << FLS_ENC_DICT)
#define FLS_FAMILY_RLE (1u << FLS_ENC_RLE)

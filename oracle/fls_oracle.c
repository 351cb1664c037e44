/*
 * fls_oracle.c — CPU restatement of the reference's read_fls decode path.  TEST INFRASTRUCTURE ONLY.
 * See fls_oracle.h for the parity status ("unpinned" at byte level; ALP pinned against DuckDB's vendored ALP).
 *
 * Structure mirrors the reference's two passes over every vector:
 *   1. the FastLanes library decodes one 1024-value vector into operator scratch  -> fls_oracle_decode_vector
 *   2. the extension's materializer copies/casts/untransposes it into the DuckDB chunk -> fls_oracle_materialize
 *      (reference src/include/reader/materializer/fls_type_resolver.hpp:120-165, kernels/ *.hpp)
 */
#include "fls_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

static const int kOrder[8] = {0, 4, 2, 6, 1, 5, 3, 7};

/* SURVEY.md A-2 (unified transposed layout of the FastLanes paper) */
static inline int oracle_index(int row, int lane) {
	return kOrder[row >> 3] * 16 + (row & 7) * 128 + lane;
}
/* transposed position -> logical position */
static inline int oracle_transpose(int pos) {
	const int lane = pos & 15, order = (pos >> 4) & 7, row = pos >> 7;
	return lane * 64 + kOrder[order] * 8 + row;
}

/* ALP tables: duckdb/src/include/duckdb/storage/compression/alp/alp_constants.hpp:43-131 */
static const int64_t kFact[19] = {1LL,
                                  10LL,
                                  100LL,
                                  1000LL,
                                  10000LL,
                                  100000LL,
                                  1000000LL,
                                  10000000LL,
                                  100000000LL,
                                  1000000000LL,
                                  10000000000LL,
                                  100000000000LL,
                                  1000000000000LL,
                                  10000000000000LL,
                                  100000000000000LL,
                                  1000000000000000LL,
                                  10000000000000000LL,
                                  100000000000000000LL,
                                  1000000000000000000LL};
static const double kFracD[21] = {1.0,
                                  0.1,
                                  0.01,
                                  0.001,
                                  0.0001,
                                  0.00001,
                                  0.000001,
                                  0.0000001,
                                  0.00000001,
                                  0.000000001,
                                  0.0000000001,
                                  0.00000000001,
                                  0.000000000001,
                                  0.0000000000001,
                                  0.00000000000001,
                                  0.000000000000001,
                                  0.0000000000000001,
                                  0.00000000000000001,
                                  0.000000000000000001,
                                  0.0000000000000000001,
                                  0.00000000000000000001};
static const float kFracF[11] = {1.0F,      0.1F,       0.01F,       0.001F,       0.0001F,      0.00001F,
                                 0.000001F, 0.0000001F, 0.00000001F, 0.000000001F, 0.0000000001F};

/* alp.hpp:129-134 DecodeValue: static_cast<T>(encoded) * static_cast<T>(FACT_ARR[f]) * FRAC_ARR[e] */
double fls_oracle_alp_decode_f64(int64_t digit, int e, int f) {
	return (double)digit * (double)kFact[f] * kFracD[e];
}
float fls_oracle_alp_decode_f32(int32_t digit, int e, int f) {
	return (float)digit * (float)kFact[f] * kFracF[e];
}
static inline uint64_t alp_decode_64(int64_t d, int e, int f) {
	const double x = fls_oracle_alp_decode_f64(d, e, f);
	uint64_t u;
	memcpy(&u, &x, 8);
	return u;
}
static inline uint32_t alp_decode_32(int32_t d, int e, int f) {
	const float x = fls_oracle_alp_decode_f32(d, e, f);
	uint32_t u;
	memcpy(&u, &x, 4);
	return u;
}

static void unffor_8(const uint8_t* restrict, unsigned, uint8_t, uint8_t* restrict);
static void unffor_16(const uint16_t* restrict, unsigned, uint16_t, uint16_t* restrict);
static void unffor_32(const uint32_t* restrict, unsigned, uint32_t, uint32_t* restrict);
static void unpack_16(const uint16_t* restrict, unsigned, uint16_t* restrict);
static void rsum_16(const uint16_t* restrict, const uint16_t* restrict, uint16_t, uint16_t* restrict);
static void untranspose_16(const uint16_t* restrict, uint16_t* restrict);

#define U uint8_t
#define S int8_t
#define TB 8
#define SUF 8
#include "fls_oracle_tmpl.h"
#undef U
#undef S
#undef TB
#undef SUF
#define U uint16_t
#define S int16_t
#define TB 16
#define SUF 16
#include "fls_oracle_tmpl.h"
#undef U
#undef S
#undef TB
#undef SUF
#define U uint32_t
#define S int32_t
#define TB 32
#define SUF 32
#include "fls_oracle_tmpl.h"
#undef U
#undef S
#undef TB
#undef SUF
#define U uint64_t
#define S int64_t
#define TB 64
#define SUF 64
#include "fls_oracle_tmpl.h"
#undef U
#undef S
#undef TB
#undef SUF

int fls_oracle_index(int row, int lane) {
	return oracle_index(row, lane);
}
int fls_oracle_transpose(int pos) {
	return oracle_transpose(pos);
}
void fls_oracle_unpack(const void* packed, unsigned T, unsigned W, void* out) {
	switch (T) {
	case 8: unpack_8((const uint8_t*)packed, W, (uint8_t*)out); break;
	case 16: unpack_16((const uint16_t*)packed, W, (uint16_t*)out); break;
	case 32: unpack_32((const uint32_t*)packed, W, (uint32_t*)out); break;
	default: unpack_64((const uint64_t*)packed, W, (uint64_t*)out); break;
	}
}
void fls_oracle_untranspose(const void* in, unsigned T, void* out) {
	switch (T) {
	case 8: untranspose_8((const uint8_t*)in, (uint8_t*)out); break;
	case 16: untranspose_16((const uint16_t*)in, (uint16_t*)out); break;
	case 32: untranspose_32((const uint32_t*)in, (uint32_t*)out); break;
	default: untranspose_64((const uint64_t*)in, (uint64_t*)out); break;
	}
}

int fls_oracle_open(const uint8_t* bytes, uint64_t size, fls_oracle_table* t) {
	if (!bytes || size < sizeof(fls_file_header)) {
		return -1;
	}
	memset(t, 0, sizeof(*t));
	t->bytes = bytes;
	t->size = size;
	memcpy(&t->header, bytes, sizeof(fls_file_header));
	if (memcmp(t->header.magic, FLS_MAGIC, 8) != 0 || t->header.footer_offset + t->header.footer_size > size) {
		return -1;
	}
	const uint8_t* f = bytes + t->header.footer_offset;
	t->columns = (const fls_column_desc*)f;
	f += (size_t)t->header.n_columns * sizeof(fls_column_desc);
	t->rowgroups = (const fls_rowgroup_desc*)f;
	f += (size_t)t->header.n_rowgroups * sizeof(fls_rowgroup_desc);
	t->chunks = (const fls_chunk_desc*)f;
	return 0;
}

int fls_oracle_decode_vector(const fls_oracle_table* t, uint32_t rg, uint32_t col, uint32_t vec, void* scratch,
                             int* is_transposed) {
	if (rg >= t->header.n_rowgroups || col >= t->header.n_columns) {
		return -1;
	}
	const fls_chunk_desc* ch = &t->chunks[(size_t)rg * t->header.n_columns + col];
	if (vec >= ch->n_vec) {
		return -1;
	}
	const uint8_t* base = t->bytes + ch->offset;
	switch (ch->phys_bits) {
	case 8: return decode_vector_8(base, ch, vec, (uint8_t*)scratch, is_transposed);
	case 16: return decode_vector_16(base, ch, vec, (uint16_t*)scratch, is_transposed);
	case 32: return decode_vector_32(base, ch, vec, (uint32_t*)scratch, is_transposed);
	case 64: return decode_vector_64(base, ch, vec, (uint64_t*)scratch, is_transposed);
	default: return -2;
	}
}

/* Run index of one RLE vector (numeric RLE and RLE_STR share the blob): [128 lane bases][128*W deltas] in 16-bit lanes ->
 * unffor + rsum (transposed) -> untranspose_i (rle_map_kernel.hpp:45-58). */
void fls_oracle_rle_index(const uint8_t* blob, unsigned width, uint16_t dmin, uint16_t* idx) {
	uint16_t d16[1024], t16[1024];
	unpack_16((const uint16_t*)(blob + 128), width, d16);
	rsum_16((const uint16_t*)blob, d16, dmin, t16);
	untranspose_16(t16, idx);
}

/* Codes of one vector of a dictionary-encoded chunk (numeric DICT, DICT_STR, FSST_DICT): dec_dict_opr<KEY, IDX>::Index()
 * (reference dictionary_kernel.hpp:14-20,89-99): FFOR-packed in idx_bits lanes, base = smallest code of the vector. */
int fls_oracle_decode_codes(const fls_oracle_table* t, uint32_t rg, uint32_t col, uint32_t vec, uint32_t* codes) {
	if (rg >= t->header.n_rowgroups || col >= t->header.n_columns) {
		return -1;
	}
	const fls_chunk_desc* ch = &t->chunks[(size_t)rg * t->header.n_columns + col];
	if (vec >= ch->n_vec || (ch->encoding != FLS_ENC_DICT && !FLS_ENC_IS_STR_DICT(ch->encoding))) {
		return -1;
	}
	const uint8_t* base = t->bytes + ch->offset;
	const fls_vec_meta* m = (const fls_vec_meta*)(base + ch->seg_off[FLS_SEG_META]) + vec;
	const uint8_t* blob = base + ch->seg_off[FLS_SEG_DATA] + (size_t)m->data_off16 * 16;
	if (ch->idx_bits == 8) {
		uint8_t c8[1024];
		unffor_8((const uint8_t*)blob, m->width, 0, c8);
		for (unsigned i = 0; i < 1024; ++i) codes[i] = c8[i];
	} else if (ch->idx_bits == 16) {
		uint16_t c16[1024];
		unffor_16((const uint16_t*)blob, m->width, 0, c16);
		for (unsigned i = 0; i < 1024; ++i) codes[i] = c16[i];
	} else {
		unffor_32((const uint32_t*)blob, m->width, 0, codes);
	}
	for (unsigned i = 0; i < 1024; ++i) {
		codes[i] += (uint32_t)m->base;
	}
	return 0;
}

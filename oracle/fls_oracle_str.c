/*
 * fls_oracle_str.c — CPU restatement of the string side of the read_fls decode path.  TEST INFRASTRUCTURE ONLY
 * (see fls_oracle.h: only tests/, smoke() and bench.py's CPU legs may load this).
 *
 * Follows the reference materializer's string kernels (src/include/reader/materializer/kernels/):
 *   dec_fls_str_uncompressed_opr  uncompressed_kernel.hpp:26-38   AddString(data + offset, length[i]), offset += length
 *   dec_constant_str_opr          constant_kernel.hpp:40-49       one string for the whole vector
 *   dec_dict_opr<fls_string_t>    dictionary_kernel.hpp:52-102    Prepare: dictionary of string_t from end offsets + bytes;
 *                                                                 Decode: selection vector = the codes
 *   dec_fsst_opr                  fsst_kernel.hpp:19-57           per string: fsst_decompress(decoder, len, in, 8*len, buf)
 *   dec_fsst_dict_opr             fsst_dict_kernel.hpp:20-52,54-80 FSST-decoded dictionary + codes
 *   dec_rle_map_opr<fls_string_t> rle_map_kernel.hpp:26-61        untranspose the run index, row i = run string idxs[i]
 *   dec_cross_rle_opr<fls_string_t> cross_rle_kernel.hpp:55-97,121-144  row-group wide runs, vector v = rows [1024 v, 1024 v + 1024)
 *   dec_frequency_str_opr         frequency_kernel.hpp:23-69      the frequent string, except at the exception positions
 * FSST decoding = duckdb/third_party/fsst/fsst.h:162-230 (in the reference tree): code < 255 -> symbol[code] (len[code]
 * bytes), 255 -> next byte literal.  The decoded element is DuckDB's string_t
 * (duckdb/src/include/duckdb/common/types/string_type.hpp:22-60,236-247): length <= 12 -> inlined, zero padded;
 * otherwise 4 prefix bytes + pointer.  Pointers are heap_base_addr + offset into the heap this function fills, so the
 * result is position independent and comparable bit for bit with the CUDA path.
 *
 * PARITY: the FSST decoder is pinned against the in-tree libfsst (oracle/_ref/libfsstref.so, tests/golden/fsst_golden.npz);
 * the segment layout is this engine's own container (the real .fls bytes live in the absent library): unpinned.
 */
#include <string.h>
#include "fls_oracle.h"

typedef struct {
	uint32_t length;
	uint8_t body[12]; /* inlined[12] | prefix[4] + ptr(8) */
} ostr_t;

static void make_string(ostr_t* s, const uint8_t* data, uint32_t len, uint64_t addr) {
	s->length = len;
	memset(s->body, 0, 12);
	if (len <= 12) {
		if (len) {
			memcpy(s->body, data, len);
		}
	} else {
		memcpy(s->body, data, 4);
		memcpy(s->body + 4, &addr, 8);
	}
}

static uint64_t pad16(uint64_t x) {
	return (x + 15u) & ~(uint64_t)15u;
}

/* fsst.h:162-230, the scalar tail loop (:219-230) applied to the whole string */
uint64_t fls_oracle_fsst_decompress(const fls_fsst_table* tab, const uint8_t* in, uint64_t len_in, uint8_t* out) {
	uint64_t pos_out = 0;
	for (uint64_t p = 0; p < len_in;) {
		const unsigned code = in[p++];
		if (code < FLS_FSST_ESC) {
			memcpy(out + pos_out, &tab->symbol[code], tab->len[code]);
			pos_out += tab->len[code];
		} else {
			out[pos_out++] = in[p++];
		}
	}
	return pos_out;
}

/* fsst12_decompress of the absent library (called at reference fsst12_kernel.hpp:53, fsst12_dict_kernel.hpp:39), restated
 * from cwida/fsst's fsst12.h as published: 12-bit codes, two per 3 bytes (little endian), a trailing single code in 2 bytes;
 * every code indexes symbol[] / len[] (codes < 256 are the bytes themselves).  UNPINNED: no copy of fsst12.h in the tree. */
uint64_t fls_oracle_fsst12_decompress(const fls_fsst12_table* tab, const uint8_t* in, uint64_t len_in, uint8_t* out) {
	uint64_t pos_out = 0, p = 0;
	for (; p + 3 <= len_in; p += 3) {
		const unsigned c0 = in[p] | ((in[p + 1] & 15u) << 8), c1 = (in[p + 1] >> 4) | ((unsigned)in[p + 2] << 4);
		memcpy(out + pos_out, &tab->symbol[c0], tab->len[c0]);
		pos_out += tab->len[c0];
		memcpy(out + pos_out, &tab->symbol[c1], tab->len[c1]);
		pos_out += tab->len[c1];
	}
	if (len_in - p == 2) {
		const unsigned c0 = in[p] | ((in[p + 1] & 15u) << 8);
		memcpy(out + pos_out, &tab->symbol[c0], tab->len[c0]);
		pos_out += tab->len[c0];
	}
	return pos_out;
}

/* one compressed string of a FSST / FSST12 chunk */
static uint64_t fsst_any_decompress(unsigned enc, const uint8_t* aux0, const uint8_t* in, uint64_t len_in, uint8_t* out) {
	return FLS_ENC_IS_FSST12(enc) ? fls_oracle_fsst12_decompress((const fls_fsst12_table*)aux0, in, len_in, out)
	                              : fls_oracle_fsst_decompress((const fls_fsst_table*)aux0, in, len_in, out);
}

/* VALUE encodings (fls_format.h FLS_ENC_IS_STR_VALUES): which of the chunk's value strings every row of vector `vec` is.
 *   RLE_STR        rle_map_kernel.hpp:45-58: unffor + rsum of the u16 run index (left transposed), untranspose_i, then
 *                  value = first run of the vector + idxs[i]
 *   CROSS_RLE_STR  cross_rle_kernel.hpp:55-97: find the run that covers row 1024 * vec, then walk the run lengths
 *   FREQUENCY_STR  frequency_kernel.hpp:44-66: walk the rows with a cursor into the ascending exception positions */
static int value_index_of_vector(const fls_oracle_table* t, const fls_chunk_desc* ch, uint32_t vec, uint32_t* keys) {
	const uint8_t* base = t->bytes + ch->offset;
	const fls_vec_meta* m = (const fls_vec_meta*)(base + ch->seg_off[FLS_SEG_META]) + vec;
	const uint8_t* blob = base + ch->seg_off[FLS_SEG_DATA] + (size_t)m->data_off16 * 16;
	const uint32_t* aux1 = (const uint32_t*)(base + ch->seg_off[FLS_SEG_AUX1]);
	switch (ch->encoding) {
	case FLS_ENC_RLE_STR: {
		uint16_t idx[FLS_VEC_SZ];
		fls_oracle_rle_index(blob, m->width, (uint16_t)m->base, idx);
		for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
			if (idx[i] >= m->aux_cnt) {
				return -1;
			}
			keys[i] = m->aux_off + idx[i];
		}
		return 0;
	}
	case FLS_ENC_CROSS_RLE_STR: {
		const uint32_t* lens = aux1 + ch->aux_count;
		uint64_t target = (uint64_t)vec * FLS_VEC_SZ, pos = 0;
		uint32_t run = 0, produced = 0;
		while (run < ch->aux_count && pos + lens[run] <= target) {
			pos += lens[run];
			++run;
		}
		uint64_t offset = target - pos;
		while (produced < FLS_VEC_SZ && run < ch->aux_count) {
			uint64_t avail = lens[run] - offset;
			uint32_t take = (uint32_t)(avail < FLS_VEC_SZ - produced ? avail : FLS_VEC_SZ - produced);
			for (uint32_t k = 0; k < take; ++k) {
				keys[produced++] = run;
			}
			offset = 0;
			++run;
		}
		return produced == FLS_VEC_SZ ? 0 : -1;
	}
	case FLS_ENC_FREQUENCY_STR: {
		const uint16_t* pos = (const uint16_t*)blob;
		uint32_t e = 0;
		for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
			if (e < m->aux_cnt && pos[e] == i) {
				keys[i] = m->aux_off + 1 + e;
				++e;
			} else {
				keys[i] = m->aux_off;
			}
		}
		return 0;
	}
	default: return -3;
	}
}

int fls_oracle_string_heap_bytes(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, uint64_t* out) {
	uint64_t total = 0;
	if (col >= t->header.n_columns || t->columns[col].phys_bits != 128) {
		return -1;
	}
	for (uint32_t r = rg_begin; r < rg_begin + n_rg; ++r) {
		total += pad16(t->chunks[(uint64_t)r * t->header.n_columns + col].const_val);
	}
	*out = total;
	return 0;
}

/* Decoded bytes of one chunk into its heap region; returns per-string (start, end) lookup through ends[]. */
static void fill_heap(const fls_oracle_table* t, const fls_chunk_desc* ch, uint8_t* heap) {
	const uint8_t* aux0 = t->bytes + ch->offset + ch->seg_off[FLS_SEG_AUX0];
	const uint32_t* aux1 = (const uint32_t*)(t->bytes + ch->offset + ch->seg_off[FLS_SEG_AUX1]);
	if (FLS_ENC_IS_FSST_ANY(ch->encoding)) {
		const uint8_t* codes = aux0 + FLS_FSST_TABLE_BYTES(ch->encoding);
		const uint32_t n = ch->aux_count;
		uint32_t cprev = 0, dprev = 0;
		for (uint32_t i = 0; i < n; ++i) {
			fsst_any_decompress(ch->encoding, aux0, codes + cprev, aux1[i] - cprev, heap + dprev);
			cprev = aux1[i];
			dprev = aux1[n + i];
		}
	} else {
		memcpy(heap, aux0, ch->const_val);
	}
}

int fls_oracle_decode_strings(const fls_oracle_table* t, uint32_t col, uint32_t rg_begin, uint32_t n_rg, void* out_strings,
                              uint8_t* heap, uint64_t heap_base_addr) {
	if (col >= t->header.n_columns || t->columns[col].phys_bits != 128) {
		return -1;
	}
	const uint32_t vpr = t->header.vec_per_rowgroup;
	uint64_t heap_off = 0;
	uint32_t codes[FLS_VEC_SZ];
	for (uint32_t r = 0; r < n_rg; ++r) {
		const fls_chunk_desc* ch = &t->chunks[(uint64_t)(rg_begin + r) * t->header.n_columns + col];
		ostr_t* out = (ostr_t*)out_strings + (uint64_t)r * vpr * FLS_VEC_SZ;
		uint8_t* hp = heap + heap_off;
		const uint64_t haddr = heap_base_addr + heap_off;
		const uint32_t* aux1 = (const uint32_t*)(t->bytes + ch->offset + ch->seg_off[FLS_SEG_AUX1]);
		const int fsst = FLS_ENC_IS_FSST_ANY(ch->encoding);
		const uint32_t* ends = fsst ? aux1 + ch->aux_count : aux1; /* decoded end offsets */
		fill_heap(t, ch, hp);
		const uint32_t n_rows = ch->n_vec * FLS_VEC_SZ;
		switch (ch->encoding) {
		case FLS_ENC_STR_CONSTANT:
			for (uint32_t i = 0; i < n_rows; ++i) {
				make_string(&out[i], hp, ch->aux_count, haddr);
			}
			break;
		case FLS_ENC_STR_UNCOMPRESSED:
		case FLS_ENC_FSST:
		case FLS_ENC_FSST12:
			for (uint32_t i = 0; i < n_rows; ++i) {
				const uint32_t start = i ? ends[i - 1] : 0;
				make_string(&out[i], hp + start, ends[i] - start, haddr + start);
			}
			break;
		case FLS_ENC_DICT_STR:
		case FLS_ENC_FSST_DICT:
		case FLS_ENC_FSST12_DICT:
			for (uint32_t v = 0; v < ch->n_vec; ++v) {
				if (fls_oracle_decode_codes(t, rg_begin + r, col, v, codes)) {
					return -2;
				}
				for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
					const uint32_t k = codes[i] < ch->aux_count ? codes[i] : ch->aux_count - 1;
					const uint32_t start = k ? ends[k - 1] : 0;
					make_string(&out[(uint64_t)v * FLS_VEC_SZ + i], hp + start, ends[k] - start, haddr + start);
				}
			}
			break;
		case FLS_ENC_RLE_STR:
		case FLS_ENC_CROSS_RLE_STR:
		case FLS_ENC_FREQUENCY_STR:
			for (uint32_t v = 0; v < ch->n_vec; ++v) {
				if (value_index_of_vector(t, ch, v, codes)) {
					return -2;
				}
				for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
					const uint32_t k = codes[i];
					const uint32_t start = k ? ends[k - 1] : 0;
					make_string(&out[(uint64_t)v * FLS_VEC_SZ + i], hp + start, ends[k] - start, haddr + start);
				}
			}
			break;
		default: return -3;
		}
		heap_off += pad16(ch->const_val);
	}
	return 0;
}

/* Dictionary form (what the reference hands to DuckDB as a DICTIONARY_VECTOR, dictionary_kernel.hpp:60-78): the string_t
 * of every key of chunk (rg, col) + the dictionary bytes. */
int fls_oracle_decode_string_dict(const fls_oracle_table* t, uint32_t col, uint32_t rg, void* out_keys, uint8_t* heap,
                                  uint64_t heap_base_addr) {
	if (col >= t->header.n_columns || rg >= t->header.n_rowgroups) {
		return -1;
	}
	const fls_chunk_desc* ch = &t->chunks[(uint64_t)rg * t->header.n_columns + col];
	if (!FLS_ENC_IS_STR_DICT(ch->encoding)) {
		return -3;
	}
	const uint32_t* aux1 = (const uint32_t*)(t->bytes + ch->offset + ch->seg_off[FLS_SEG_AUX1]);
	const uint32_t* ends = FLS_ENC_IS_FSST_ANY(ch->encoding) ? aux1 + ch->aux_count : aux1;
	fill_heap(t, ch, heap);
	ostr_t* out = (ostr_t*)out_keys;
	for (uint32_t k = 0; k < ch->aux_count; ++k) {
		const uint32_t start = k ? ends[k - 1] : 0;
		make_string(&out[k], heap + start, ends[k] - start, heap_base_addr + start);
	}
	return 0;
}

/* One vector of a per-row (non-dictionary) string chunk, the way the reference materializer does it: for every row of the
 * vector, decode / copy the string into the vector's heap and build its string_t (uncompressed_kernel.hpp:26-38,
 * constant_kernel.hpp:40-49, fsst_kernel.hpp:42-56: one fsst_decompress call per string).  heap must hold the vector's
 * decoded bytes (at most the chunk's const_val, or 1024 x the constant's length); *heap_used returns them. */
int fls_oracle_decode_string_vector(const fls_oracle_table* t, uint32_t rg, uint32_t col, uint32_t vec, void* out_strings,
                                    uint8_t* heap, uint64_t heap_base_addr, uint64_t* heap_used) {
	if (col >= t->header.n_columns || rg >= t->header.n_rowgroups) {
		return -1;
	}
	const fls_chunk_desc* ch = &t->chunks[(uint64_t)rg * t->header.n_columns + col];
	if (vec >= ch->n_vec) {
		return -1;
	}
	const uint8_t* aux0 = t->bytes + ch->offset + ch->seg_off[FLS_SEG_AUX0];
	const uint32_t* aux1 = (const uint32_t*)(t->bytes + ch->offset + ch->seg_off[FLS_SEG_AUX1]);
	ostr_t* out = (ostr_t*)out_strings;
	uint64_t pos = 0;
	const uint32_t i0 = vec * FLS_VEC_SZ;
	switch (ch->encoding) {
	case FLS_ENC_STR_CONSTANT:
		memcpy(heap, aux0, ch->aux_count);
		for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
			make_string(&out[i], heap, ch->aux_count, heap_base_addr);
		}
		pos = ch->aux_count;
		break;
	case FLS_ENC_STR_UNCOMPRESSED:
		for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
			const uint32_t start = (i0 + i) ? aux1[i0 + i - 1] : 0, len = aux1[i0 + i] - start;
			memcpy(heap + pos, aux0 + start, len);
			make_string(&out[i], heap + pos, len, heap_base_addr + pos);
			pos += len;
		}
		break;
	case FLS_ENC_FSST:
	case FLS_ENC_FSST12: {
		const uint8_t* codes = aux0 + FLS_FSST_TABLE_BYTES(ch->encoding);
		for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
			const uint32_t cstart = (i0 + i) ? aux1[i0 + i - 1] : 0;
			const uint64_t len = fsst_any_decompress(ch->encoding, aux0, codes + cstart, aux1[i0 + i] - cstart, heap + pos);
			make_string(&out[i], heap + pos, (uint32_t)len, heap_base_addr + pos);
			pos += len;
		}
		break;
	}
	case FLS_ENC_RLE_STR:
	case FLS_ENC_CROSS_RLE_STR:
	case FLS_ENC_FREQUENCY_STR: { /* AddString per row, as the reference does: the vector's heap repeats the strings */
		uint32_t keys[FLS_VEC_SZ];
		if (value_index_of_vector(t, ch, vec, keys)) {
			return -2;
		}
		for (uint32_t i = 0; i < FLS_VEC_SZ; ++i) {
			const uint32_t start = keys[i] ? aux1[keys[i] - 1] : 0, len = aux1[keys[i]] - start;
			memcpy(heap + pos, aux0 + start, len);
			make_string(&out[i], heap + pos, len, heap_base_addr + pos);
			pos += len;
		}
		break;
	}
	default: return -3;
	}
	if (heap_used) {
		*heap_used = pos;
	}
	return 0;
}

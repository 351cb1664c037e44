/*
 * fls_oracle_tmpl.h — per-width body of the CPU oracle (TEST INFRASTRUCTURE ONLY; see fls_oracle.h).
 * Included four times by fls_oracle.c with U (unsigned lane type), S (signed twin), TB (bits) and SUF defined.
 *
 * Layout statements follow SURVEY.md Appendix A-1..A-5 (FastLanes paper): value (row,lane) of a vector sits at
 * logical position index(row,lane); packed word k of all lanes is contiguous, so the row-outer / lane-inner loops
 * below are the auto-vectorisable form the FastLanes CPU kernels are built around.
 */
#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUF)
#define LANES (1024 / TB)

/* A-3: bit-unpack W in [0,TB] into out[] by position. */
static void FN(unpack_)(const U* restrict packed, unsigned W, U* restrict out) {
	if (W == 0) {
		memset(out, 0, 1024 * sizeof(U));
		return;
	}
	const U mask = W == TB ? (U)~(U)0 : (U)(((U)1 << W) - 1);
	for (unsigned row = 0; row < TB; ++row) {
		const unsigned bit = row * W, word = bit / TB, sh = bit % TB;
		U* restrict o = out + oracle_index((int)row, 0);
		const U* restrict p0 = packed + (size_t)word * LANES;
		if (sh + W <= TB) {
			for (unsigned lane = 0; lane < LANES; ++lane) {
				o[lane] = (U)((p0[lane] >> sh) & mask);
			}
		} else {
			const U* restrict p1 = p0 + LANES;
			for (unsigned lane = 0; lane < LANES; ++lane) {
				o[lane] = (U)(((p0[lane] >> sh) | (U)(p1[lane] << (TB - sh))) & mask);
			}
		}
	}
}

/* A-4: FFOR = unpack + base (wrap-around in U). dec_unffor_opr. */
static void FN(unffor_)(const U* restrict packed, unsigned W, U base, U* restrict out) {
	FN(unpack_)(packed, W, out);
	for (unsigned i = 0; i < 1024; ++i) {
		out[i] = (U)(out[i] + base);
	}
}

/* A-5: rsum — per-lane running sum over rows; result stays in TRANSPOSED order (dec_transpose_opr input). */
static void FN(rsum_)(const U* restrict bases, const U* restrict deltas /* by position */, U dmin, U* restrict out) {
	U acc[LANES];
	for (unsigned lane = 0; lane < LANES; ++lane) {
		acc[lane] = bases[lane];
	}
	for (unsigned row = 0; row < TB; ++row) {
		const unsigned p = (unsigned)oracle_index((int)row, 0);
		for (unsigned lane = 0; lane < LANES; ++lane) {
			acc[lane] = (U)(acc[lane] + deltas[p + lane] + dmin);
			out[p + lane] = acc[lane];
		}
	}
}

/* untranspose_i (reference fls_type_resolver.hpp:155): out[transpose(pos)] = in[pos]. */
static void FN(untranspose_)(const U* restrict in, U* restrict out) {
	for (int pos = 0; pos < 1024; ++pos) {
		out[oracle_transpose(pos)] = in[pos];
	}
}

/* exceptions: out[pos[k]] = val[k] (alp.hpp:401-404) */
/* out[pos[k]] = value[k] (alp.hpp:401-404).  The exception tail sits behind the packed part of the vector's blob:
 * [n values, padded to 16 B][n positions u16, padded to 16 B] (fls_format.h). */
static void FN(patch_)(U* out, const uint8_t* tail, unsigned n) {
	const U* vals = (const U*)tail;
	const uint16_t* pos = (const uint16_t*)(tail + FLS_PAD16(n * (unsigned)sizeof(U)));
	for (unsigned k = 0; k < n; ++k) {
		out[pos[k]] = vals[k];
	}
}

/* cross_rle_kernel.hpp:8-53 restated with a running start instead of the O(#runs) rescan. */
static void FN(cross_rle_)(const U* vals, const uint32_t* lens, uint32_t n_runs, uint32_t vec, U* out) {
	uint64_t target = (uint64_t)vec * 1024, pos = 0;
	uint32_t i = 0;
	while (i < n_runs && pos + lens[i] <= target) {
		pos += lens[i];
		++i;
	}
	uint64_t offset = target - pos;
	unsigned produced = 0;
	while (produced < 1024 && i < n_runs) {
		uint64_t avail = lens[i] - offset;
		unsigned take = (unsigned)(avail < 1024 - produced ? avail : 1024 - produced);
		for (unsigned k = 0; k < take; ++k) {
			out[produced++] = vals[i];
		}
		offset = 0;
		++i;
	}
}

static int FN(decode_vector_)(const uint8_t* chunk_base, const fls_chunk_desc* ch, uint32_t vec, U* restrict out,
                              int* is_transposed) {
	const fls_vec_meta* m = (const fls_vec_meta*)(chunk_base + ch->seg_off[FLS_SEG_META]) + vec;
	const uint8_t* blob = chunk_base + ch->seg_off[FLS_SEG_DATA] + (size_t)m->data_off16 * 16;
	const U* aux0 = (const U*)(chunk_base + ch->seg_off[FLS_SEG_AUX0]);
	*is_transposed = 0;
	switch (ch->encoding) {
	case FLS_ENC_UNCOMPRESSED: memcpy(out, blob, 1024 * sizeof(U)); return 0;
	case FLS_ENC_CONSTANT:
		for (unsigned i = 0; i < 1024; ++i) {
			out[i] = (U)ch->const_val;
		}
		return 0;
	case FLS_ENC_FFOR:
	case FLS_ENC_SLPATCH:
	case FLS_ENC_FREQUENCY:
		FN(unffor_)((const U*)blob, m->width, (U)m->base, out);
		FN(patch_)(out, blob + 128u * m->width, m->aux_cnt);
		return 0;
	case FLS_ENC_DELTA: {
		U deltas[1024];
		FN(unpack_)((const U*)(blob + 128), m->width, deltas);
		FN(rsum_)((const U*)blob, deltas, (U)m->base, out);
		*is_transposed = 1;
		return 0;
	}
	case FLS_ENC_DICT: { /* dictionary_kernel.hpp:11-22: out[i] = keys[index[i]] */
		uint32_t codes[1024];
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
			const uint32_t code = codes[i] + (uint32_t)m->base;
			if (code >= ch->aux_count) {
				return -1;
			}
			out[i] = aux0[code];
		}
		return 0;
	}
	case FLS_ENC_RLE: { /* rle_map_kernel.hpp:45-58: untranspose the run index, then out[i] = rle_vals[idxs[i]] */
		uint16_t d16[1024], t16[1024], idx[1024];
		unpack_16((const uint16_t*)(blob + 128), m->width, d16);
		rsum_16((const uint16_t*)blob, d16, (uint16_t)m->base, t16);
		untranspose_16(t16, idx);
		const U* rv = aux0 + m->aux_off;
		for (unsigned i = 0; i < 1024; ++i) {
			if (idx[i] >= m->aux_cnt) {
				return -1;
			}
			out[i] = rv[idx[i]];
		}
		return 0;
	}
	case FLS_ENC_CROSS_RLE:
		FN(cross_rle_)(aux0, (const uint32_t*)(chunk_base + ch->seg_off[FLS_SEG_AUX1]), ch->aux_count, vec, out);
		return 0;
#if TB >= 32
	case FLS_ENC_ALP: { /* alp.hpp:377-406: unFOR, decode, patch */
		FN(unffor_)((const U*)blob, m->width, (U)m->base, out);
		for (unsigned i = 0; i < 1024; ++i) {
			out[i] = FN(alp_decode_)((S)out[i], m->e, m->f);
		}
		FN(patch_)(out, blob + 128u * m->width, m->aux_cnt);
		return 0;
	}
	case FLS_ENC_ALP_RD: { /* alprd.hpp:203-236 */
		U left[1024];
		FN(unpack_)((const U*)blob, m->width, out);
		FN(unpack_)((const U*)(blob + 128u * m->width), m->width2, left);
		for (unsigned i = 0; i < 1024; ++i) {
			out[i] = (U)(((U)ch->rd_dict[left[i] & 7] << m->width) | out[i]);
		}
		FN(patch_)(out, blob + 128u * (m->width + m->width2), m->aux_cnt);
		return 0;
	}
#endif
	default: return -2;
	}
}

#undef LANES
#undef FN
#undef CAT
#undef CAT_

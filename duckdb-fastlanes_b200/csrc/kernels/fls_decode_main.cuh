// fls_decode_main.cuh — per-vector decode bodies of the main kernel family (T-lane encodings) and the kernel itself.
#pragma once
#include "fls_decode.cuh"
#include "fls_decode_gather.cuh"

namespace FLS_NS {

// ---- T in {8,16,32} ------------------------------------------------------------------------------------------------------
// FAM: which encodings this instantiation carries (the launcher splits a column's chunks into a "light" launch, an
// ALP launch and an ALP-RD launch so that the register budget of the heavy float paths does not cap the occupancy of
// the integer paths).
constexpr uint32_t kFamLight = FLS_FAMILY_MAIN & ~((1u << FLS_ENC_ALP) | (1u << FLS_ENC_ALP_RD));
constexpr uint32_t kFamAlp = 1u << FLS_ENC_ALP;
constexpr uint32_t kFamAlpRd = 1u << FLS_ENC_ALP_RD;
// T = 64 only: DELTA gets its own instantiation.  With DELTA and the FFOR family in one kernel the compiler folds the four
// switch(W) dispatches (two phases each) into one and needs > 200 registers; apart they fit 128 with no spills.
constexpr uint32_t kFamDelta = 1u << FLS_ENC_DELTA;
constexpr uint32_t kFamLight64 = kFamLight & ~kFamDelta;
// T = 8 written as 8 bits: FFOR and DELTA chunks get a SWAR instantiation of their own (four lanes per register from the
// unpack to the store).  It needs 48 registers and no staging area, so five CTAs are resident per SM; 1-byte values
// leave little work per vector to hide the descriptor / bulk-copy latencies behind, hence the extra warps.
constexpr uint32_t kFamSwar8 = (1u << FLS_ENC_FFOR) | (1u << FLS_ENC_DELTA);

// AGG + part != nullptr: register fast path of the fused aggregates — the vector is complete (n_valid == 1024) and has
// no exceptions, so its values are folded into the partials straight from registers instead of being written anywhere.
template <int T, int OB, uint32_t FAM, bool AGG>
__device__ __forceinline__ void process_small(const VecDesc& d, const uint8_t* slot, uint8_t* stage, uint8_t* out_vec, int lane,
                                              uint32_t conv, AggPartial& part, bool agg_fast, uint32_t agg_signed, uint32_t agg_float) {
	const bool fast = AGG && agg_fast;
	const uint32_t enc = d.misc >> 24, W = (d.misc >> 16) & 0xffu;
	constexpr uint32_t TMASK = T == 32 ? 0xffffffffu : ((1u << (T & 31)) - 1u);
	if constexpr ((FAM & (1u << FLS_ENC_UNCOMPRESSED)) != 0) {
		switch (enc) {
		case FLS_ENC_UNCOMPRESSED: copy_out_linear<T, OB, false>(slot, out_vec, lane, conv); return;
		case FLS_ENC_CONSTANT: fill_constant<T, OB>(out_vec, lane, (uint32_t)d.base, 0, conv); return;
		case FLS_ENC_CROSS_RLE: decode_cross_rle<T, OB>(out_vec, d, lane, conv); return;
		default: break;
		}
	}
	if constexpr (FAM == kFamSwar8) {
		static_assert(T == 8 && OB == 1 && !AGG, "SWAR family: 8-bit lanes written as 8 bits");
		// A thread's four lanes of a row stay packed in one register from the unpack to the store (the per-value form
		// below is instruction bound at ~0.35 of the HBM roofline).
		if (enc == FLS_ENC_FFOR) {
			uint32_t x[8];
			unpack8_swar_dispatch((int)W, reinterpret_cast<const uint32_t*>(slot), lane, x);
			const uint32_t base4 = ((uint32_t)d.base & 0xffu) * 0x01010101u;
			if (W < 8) { // every byte of x is < 128: the low 7 bits cannot carry out of their byte, bit 7 is an XOR
				const uint32_t bl = base4 & 0x7f7f7f7fu, bh = base4 & 0x80808080u;
#pragma unroll
				for (int row = 0; row < 8; ++row) { // value_pos<8>(lane, 4 * row + b) = row * 128 + 4 * lane + b
					st_cs_u32(out_vec + row * 128 + 4 * lane, (x[row] + bl) ^ bh);
				}
			} else {
#pragma unroll
				for (int row = 0; row < 8; ++row) {
					st_cs_u32(out_vec + row * 128 + 4 * lane, add8x4(x[row], base4));
				}
			}
			return;
		}
		if (enc == FLS_ENC_DELTA) {
			const uint32_t* s32 = reinterpret_cast<const uint32_t*>(slot);
			uint32_t x[8];
			unpack8_swar_dispatch((int)W, s32 + 32, lane, x);
			// running sums of the even and the odd lanes in 16-bit fields (8 x (255 + 255) + 255 never reaches the neighbour),
			// the low byte of a field is the value mod 256
			const uint32_t base2 = ((uint32_t)d.base & 0xffu) * 0x00010001u;
			const uint32_t acc = s32[lane]; // the four lane bases
			uint32_t ae = acc & 0x00ff00ffu, ao = __byte_perm(acc, 0u, 0x4341);
#pragma unroll
			for (int row = 0; row < 8; ++row) {
				ae += (x[row] & 0x00ff00ffu) + base2;
				ao += __byte_perm(x[row], 0u, 0x4341) + base2;
				x[row] = __byte_perm(ae, ao, 0x6240); // byte b = value (chain of lane 4*lane + b, position l0_b + row)
			}
			// 4x4 byte transposes: chain b wants its rows 0..3 / 4..7 side by side
			const uint32_t t0 = __byte_perm(x[0], x[1], 0x5140), t1 = __byte_perm(x[0], x[1], 0x7362);
			const uint32_t t2 = __byte_perm(x[2], x[3], 0x5140), t3 = __byte_perm(x[2], x[3], 0x7362);
			const uint32_t u0 = __byte_perm(x[4], x[5], 0x5140), u1 = __byte_perm(x[4], x[5], 0x7362);
			const uint32_t u2 = __byte_perm(x[6], x[7], 0x5140), u3 = __byte_perm(x[6], x[7], 0x7362);
			const uint32_t lo4[4] = {__byte_perm(t0, t2, 0x5410), __byte_perm(t0, t2, 0x7632), __byte_perm(t1, t3, 0x5410),
			                         __byte_perm(t1, t3, 0x7632)};
			const uint32_t hi4[4] = {__byte_perm(u0, u2, 0x5410), __byte_perm(u0, u2, 0x7632), __byte_perm(u1, u3, 0x5410),
			                         __byte_perm(u1, u3, 0x7632)};
			// a chain's 8 values are 8 contiguous output bytes; for a fixed b the 32 lanes of the warp cover four whole
			// 64-byte blocks (8 full sectors per store instruction), so the vector leaves straight from registers
#pragma unroll
			for (int b = 0; b < 4; ++b) {
				st_cs_v2(out_vec + chain_l0(4 * lane + b), lo4[b], hi4[b]); // byte offset == element offset
			}
		}
		return;
	}
	uint32_t v[32];
	const uint32_t base = (uint32_t)d.base;
	if ((FAM & kFamLight) != 0 && enc == FLS_ENC_DELTA) {
		// [128 B lane bases][packed deltas]: per-lane running sum (no cross-lane dependency), then transposition
		const uint32_t* s32 = reinterpret_cast<const uint32_t*>(slot);
		unpack_small_dispatch<T>((int)W, s32 + 32, lane, v);
		const uint32_t bw = s32[lane];
		uint32_t* st32 = reinterpret_cast<uint32_t*>(stage);
		if constexpr (T == 32) {
			uint32_t acc = bw;
#pragma unroll
			for (int j = 0; j < 32; ++j) {
				acc += v[j] + base;
				v[j] = acc;
			}
			const int g0 = chain_l0(lane) >> 2; // 4 elements per 16-byte granule
#pragma unroll
			for (int q = 0; q < 8 && !fast; ++q) {
				*reinterpret_cast<uint4*>(stage + (size_t)swz<T>(g0 + q) * 16) =
				    make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
			}
		} else if constexpr (T == 16) {
#pragma unroll
			for (int h = 0; h < 2; ++h) {
				uint32_t acc = (bw >> (16 * h)) & 0xffffu;
#pragma unroll
				for (int row = 0; row < 16; ++row) {
					acc += v[2 * row + h] + base;
					v[2 * row + h] = acc & 0xffffu;
				}
				const int g0 = chain_l0(2 * lane + h) >> 3; // 8 elements per granule
#pragma unroll
				for (int q = 0; q < 2 && !fast; ++q) {
					uint32_t w4[4];
#pragma unroll
					for (int m = 0; m < 4; ++m) {
						w4[m] = v[2 * (8 * q + 2 * m) + h] | (v[2 * (8 * q + 2 * m + 1) + h] << 16);
					}
					*reinterpret_cast<uint4*>(stage + (size_t)swz<T>(g0 + q) * 16) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
				}
			}
		} else {
#pragma unroll
			for (int b = 0; b < 4; ++b) {
				uint32_t acc = (bw >> (8 * b)) & 0xffu;
				uint32_t w2[2] = {0, 0};
#pragma unroll
				for (int row = 0; row < 8; ++row) {
					acc += v[4 * row + b] + base;
					w2[row >> 2] |= (acc & 0xffu) << (8 * (row & 3));
					if constexpr (AGG) {
						v[4 * row + b] = acc & 0xffu;
					}
				}
				const int l0 = chain_l0(4 * lane + b); // byte offset == element offset
				if (!fast) {
					*reinterpret_cast<uint2*>(stage + (size_t)swz<T>(l0 >> 4) * 16 + (l0 & 8)) = make_uint2(w2[0], w2[1]);
				}
			}
		}
		(void)st32;
		if constexpr (AGG) {
			if (fast) {
				part.template add32<T>(v, agg_signed, agg_float);
				return;
			}
		}
		__syncwarp();
		copy_out_linear<T, OB, true>(stage, out_vec, lane, conv);
		return;
	}
	// FFOR family / ALP / ALP_RD: packed blob starts at the slot
	unpack_small_dispatch<T>((int)W, reinterpret_cast<const uint32_t*>(slot), lane, v);
	uint32_t packed = 128u * W; // bytes of the packed part; the exception tail follows
	if constexpr (T == 32) {
		if ((FAM & kFamAlp) != 0 && (FAM == kFamAlp || enc == FLS_ENC_ALP)) { // alp.hpp:129-134: (float)digit * (float)FACT[f] * FRAC[e], two roundings, no FMA
			const float fa = c_fact_f[(d.misc2 >> 16) & 0xffu], fr = c_frac_f[(d.misc2 >> 8) & 0xffu];
#pragma unroll
			for (int j = 0; j < 32; ++j) {
				const float x = __int2float_rn((int32_t)(v[j] + base));
				v[j] = __float_as_uint(__fmul_rn(__fmul_rn(x, fa), fr));
			}
		} else if ((FAM & kFamAlpRd) != 0 && (FAM == kFamAlpRd || enc == FLS_ENC_ALP_RD)) { // alprd.hpp:222-226
			uint32_t left[32];
			const uint32_t W2 = d.misc2 & 0xffu;
			packed += 128u * W2;
			const uint4 dq = __ldg(reinterpret_cast<const uint4*>(d.ch->rd_dict)); // issued before the unpack: latency overlaps it
			unpack_small_dispatch<T>((int)W2, reinterpret_cast<const uint32_t*>(slot) + 32 * W, lane, left);
			const uint32_t dw[4] = {dq.x, dq.y, dq.z, dq.w};
#pragma unroll
			for (int j = 0; j < 32; ++j) {
				const uint32_t code = left[j] & 7u;
				const uint32_t pair = code < 4 ? (code < 2 ? dw[0] : dw[1]) : (code < 6 ? dw[2] : dw[3]);
				const uint32_t key = (code & 1u) ? (pair >> 16) : (pair & 0xffffu);
				v[j] |= key << W;
			}
		} else {
#pragma unroll
			for (int j = 0; j < 32; ++j) {
				v[j] += base;
			}
		}
	} else {
#pragma unroll
		for (int j = 0; j < 32; ++j) {
			v[j] = (v[j] + base) & TMASK;
		}
	}
	if constexpr (AGG) {
		if (fast) {
			part.template add32<T>(v, agg_signed, agg_float);
			return;
		}
	}
	store_direct<T, OB>(out_vec, lane, v, conv);
	patch_exceptions<T, OB>(out_vec, d, lane, conv, slot + packed);
}

// ---- T = 64 --------------------------------------------------------------------------------------------------------------
template <uint32_t FAM, bool AGG>
__device__ __forceinline__ void process_64(const VecDesc& d, const uint8_t* slot, uint8_t* stage, uint8_t* out_vec, int lane,
                                           AggPartial& part, bool agg_fast, uint32_t agg_signed, uint32_t agg_float) {
	const bool fast = AGG && agg_fast;
	const uint32_t enc = d.misc >> 24, W = (d.misc >> 16) & 0xffu;
	if constexpr ((FAM & kFamLight) != 0) {
		switch (enc) {
		case FLS_ENC_UNCOMPRESSED: copy_out_linear<64, 8, false>(slot, out_vec, lane, 0); return;
		case FLS_ENC_CONSTANT:
			fill_constant<64, 8>(out_vec, lane, (uint32_t)d.base, (uint32_t)(d.base >> 32), 0);
			return;
		case FLS_ENC_CROSS_RLE: decode_cross_rle<64, 8>(out_vec, d, lane, 0); return;
		default: break;
		}
	}
	uint32_t lo[32], hi[32];
	const uint32_t blo = (uint32_t)d.base, bhi = (uint32_t)(d.base >> 32);
	if ((FAM & kFamLight) != 0 && enc == FLS_ENC_DELTA) {
		unpack_64_dispatch((int)W, reinterpret_cast<const uint32_t*>(slot) + 32, lane, lo, hi);
		// thread (l = lane&15, h = lane>>4) owns rows 32h..32h+31 of lane l: local sums, one shuffle to stitch halves
		uint64_t sum = 0;
#pragma unroll
		for (int j = 0; j < 32; ++j) {
			sum += (((uint64_t)hi[j] << 32) | lo[j]) + d.base;
		}
		const uint64_t lane_base = reinterpret_cast<const uint64_t*>(slot)[lane & 15];
		const uint64_t first_half = __shfl_sync(0xffffffffu, sum, lane & 15);
		uint64_t acc = lane_base + ((lane >> 4) ? first_half : 0ull);
		const int g0 = (chain_l0(lane & 15) + 32 * (lane >> 4)) >> 1; // 2 elements per granule
#pragma unroll
		for (int q = 0; q < 16; ++q) {
			acc += (((uint64_t)hi[2 * q] << 32) | lo[2 * q]) + d.base;
			const uint64_t a0 = acc;
			acc += (((uint64_t)hi[2 * q + 1] << 32) | lo[2 * q + 1]) + d.base;
			if constexpr (AGG) {
				if (fast) { // keep the decoded values in the registers the deltas came from
					lo[2 * q] = (uint32_t)a0;
					hi[2 * q] = (uint32_t)(a0 >> 32);
					lo[2 * q + 1] = (uint32_t)acc;
					hi[2 * q + 1] = (uint32_t)(acc >> 32);
					continue;
				}
			}
			*reinterpret_cast<uint4*>(stage + (size_t)swz<64>(g0 + q) * 16) =
			    make_uint4((uint32_t)a0, (uint32_t)(a0 >> 32), (uint32_t)acc, (uint32_t)(acc >> 32));
		}
		if constexpr (AGG) {
			if (fast) {
				part.add64(lo, hi, agg_signed, agg_float);
				return;
			}
		}
		__syncwarp();
		copy_out_linear<64, 8, true>(stage, out_vec, lane, 0);
		return;
	}
	unpack_64_dispatch((int)W, reinterpret_cast<const uint32_t*>(slot), lane, lo, hi);
	uint32_t packed = 128u * W; // bytes of the packed part; the exception tail follows
	if ((FAM & kFamAlp) != 0 && (FAM == kFamAlp || enc == FLS_ENC_ALP)) { // (double)digit * (double)FACT[f] * FRAC[e]
		const double fa = c_fact_d[(d.misc2 >> 16) & 0xffu], fr = c_frac_d[(d.misc2 >> 8) & 0xffu];
#pragma unroll
		for (int j = 0; j < 32; ++j) {
			const uint64_t u = ((((uint64_t)hi[j]) << 32) | lo[j]) + d.base;
			const double x = __dmul_rn(__dmul_rn(__ll2double_rn((long long)u), fa), fr);
			const long long b = __double_as_longlong(x);
			lo[j] = (uint32_t)b;
			hi[j] = (uint32_t)(b >> 32);
		}
	} else if ((FAM & kFamAlpRd) != 0 && (FAM == kFamAlpRd || enc == FLS_ENC_ALP_RD)) {
		uint32_t llo[32], lhi[32];
		const uint32_t W2 = d.misc2 & 0xffu;
		packed += 128u * W2;
		const uint4 dq = __ldg(reinterpret_cast<const uint4*>(d.ch->rd_dict)); // issued before the unpack: latency overlaps it
		unpack_64_dispatch((int)W2, reinterpret_cast<const uint32_t*>(slot) + 32 * W, lane, llo, lhi);
		const uint32_t dw[4] = {dq.x, dq.y, dq.z, dq.w};
#pragma unroll
		for (int j = 0; j < 32; ++j) {
			const uint32_t code = llo[j] & 7u;
			const uint32_t pair = code < 4 ? (code < 2 ? dw[0] : dw[1]) : (code < 6 ? dw[2] : dw[3]);
			const uint64_t key = (code & 1u) ? (pair >> 16) : (pair & 0xffffu);
			const uint64_t u = ((((uint64_t)hi[j]) << 32) | lo[j]) | (key << W);
			lo[j] = (uint32_t)u;
			hi[j] = (uint32_t)(u >> 32);
		}
		(void)lhi;
	} else {
#pragma unroll
		for (int j = 0; j < 32; ++j) {
			const uint64_t u = ((((uint64_t)hi[j]) << 32) | lo[j]) + d.base;
			lo[j] = (uint32_t)u;
			hi[j] = (uint32_t)(u >> 32);
		}
	}
	(void)blo;
	(void)bhi;
	if constexpr (AGG) {
		if (fast) {
			part.add64(lo, hi, agg_signed, agg_float);
			return;
		}
	}
	store_direct64(out_vec, lane, lo, hi);
	patch_exceptions<64, 8>(out_vec, d, lane, 0, slot + packed);
}

// ---- T = 64, materialising kernels: two phases of 16 rows -------------------------------------------------------------
// A thread owns 32 rows of a half lane; decoding them in two phases of 16 keeps 16 instead of 32 (lo, hi) pairs (and
// half the packed words) live, so the kernel fits 128 registers and two CTAs (16 warps) stay resident per SM: the
// single-phase body needed ~200 registers = 8 warps per SM, and was latency bound (ncu: 2 warps per scheduler, stall
// "wait" + "short scoreboard" ~3.5 per issue).
template <int J0>
__device__ __forceinline__ void delta64_phase(uint32_t W, const uint32_t* pe, const uint32_t* po, uint64_t base, uint64_t& acc,
                                              uint8_t* stage, int g0) {
	uint32_t lo[16], hi[16];
	unpack_64_part_dispatch<J0>((int)W, pe, po, lo, hi);
#pragma unroll
	for (int q = 0; q < 8; ++q) {
#ifdef FLS_DELTA64_PAIRS
		// pair sums: the serial chain is 8 additions long instead of 16
		const uint64_t first = ((((uint64_t)hi[2 * q]) << 32) | lo[2 * q]) + base;
		const uint64_t pair = first + ((((uint64_t)hi[2 * q + 1]) << 32) | lo[2 * q + 1]) + base;
		const uint64_t a0 = acc + first;
		acc += pair;
#else
		acc += ((((uint64_t)hi[2 * q]) << 32) | lo[2 * q]) + base;
		const uint64_t a0 = acc;
		acc += ((((uint64_t)hi[2 * q + 1]) << 32) | lo[2 * q + 1]) + base;
#endif
		*reinterpret_cast<uint4*>(stage + (size_t)swz<64>(g0 + J0 / 2 + q) * 16) =
		    make_uint4((uint32_t)a0, (uint32_t)(a0 >> 32), (uint32_t)acc, (uint32_t)(acc >> 32));
	}
}

struct Ffor64Ctx {
	const uint32_t *pe, *po;   // packed values (right parts for ALP_RD)
	const uint32_t *pe2, *po2; // ALP_RD left codes
	uint64_t base;
	double fa, fr;             // ALP
	uint32_t dw[4];            // ALP_RD dictionary
	uint32_t W, W2;
};

template <uint32_t FAM, int J0>
__device__ __forceinline__ void ffor64_phase(const Ffor64Ctx& c, uint32_t enc, uint8_t* out_vec, int lane) {
	uint32_t lo[16], hi[16];
	unpack_64_part_dispatch<J0>((int)c.W, c.pe, c.po, lo, hi);
	if ((FAM & kFamAlp) != 0 && (FAM == kFamAlp || enc == FLS_ENC_ALP)) { // (double)digit * (double)FACT[f] * FRAC[e]
#pragma unroll
		for (int j = 0; j < 16; ++j) {
			const uint64_t u = ((((uint64_t)hi[j]) << 32) | lo[j]) + c.base;
			const double x = __dmul_rn(__dmul_rn(__ll2double_rn((long long)u), c.fa), c.fr);
			const long long b = __double_as_longlong(x);
			lo[j] = (uint32_t)b;
			hi[j] = (uint32_t)(b >> 32);
		}
	} else if ((FAM & kFamAlpRd) != 0 && (FAM == kFamAlpRd || enc == FLS_ENC_ALP_RD)) {
		uint32_t llo[16], lhi[16];
		unpack_64_part_dispatch<J0>((int)c.W2, c.pe2, c.po2, llo, lhi);
#pragma unroll
		for (int j = 0; j < 16; ++j) {
			const uint32_t code = llo[j] & 7u;
			const uint32_t pair = code < 4 ? (code < 2 ? c.dw[0] : c.dw[1]) : (code < 6 ? c.dw[2] : c.dw[3]);
			const uint64_t key = (code & 1u) ? (pair >> 16) : (pair & 0xffffu);
			const uint64_t u = ((((uint64_t)hi[j]) << 32) | lo[j]) | (key << c.W);
			lo[j] = (uint32_t)u;
			hi[j] = (uint32_t)(u >> 32);
		}
		(void)lhi;
	} else {
#pragma unroll
		for (int j = 0; j < 16; ++j) {
			const uint64_t u = ((((uint64_t)hi[j]) << 32) | lo[j]) + c.base;
			lo[j] = (uint32_t)u;
			hi[j] = (uint32_t)(u >> 32);
		}
	}
#pragma unroll
	for (int j = 0; j < 16; ++j) {
		st_cs_v2(out_vec + (size_t)value_pos<64>(lane, J0 + j) * 8, lo[j], hi[j]);
	}
}

template <uint32_t FAM>
__device__ __forceinline__ void process_64_2ph(const VecDesc& d, const uint8_t* slot, uint8_t* stage, uint8_t* out_vec, int lane) {
	const uint32_t enc = d.misc >> 24, W = (d.misc >> 16) & 0xffu;
	if constexpr ((FAM & kFamLight64) != 0) {
		switch (enc) {
		case FLS_ENC_UNCOMPRESSED: copy_out_linear<64, 8, false>(slot, out_vec, lane, 0); return;
		case FLS_ENC_CONSTANT:
			fill_constant<64, 8>(out_vec, lane, (uint32_t)d.base, (uint32_t)(d.base >> 32), 0);
			return;
		case FLS_ENC_CROSS_RLE: decode_cross_rle<64, 8>(out_vec, d, lane, 0); return;
		default: break;
		}
	}
	const uint32_t* s32 = reinterpret_cast<const uint32_t*>(slot);
	if ((FAM & kFamDelta) != 0 && (FAM == kFamDelta || enc == FLS_ENC_DELTA)) {
		// thread (l = lane&15, h = lane>>4) owns rows 32h..32h+31 of lane l = 32 logically consecutive values.  The second
		// half starts its running sum at 0; the first half's end value (totals[l]) is added while the vector is copied out
		const uint32_t *pe, *po;
		stream64_bases(s32 + 32, lane, (int)W, pe, po);
		const int h = lane >> 4;
		uint64_t acc = h ? 0ull : reinterpret_cast<const uint64_t*>(slot)[lane & 15];
		const int g0 = (chain_l0(lane & 15) + 32 * h) >> 1; // 2 elements per granule
		uint64_t* totals = reinterpret_cast<uint64_t*>(stage + 8192);
		delta64_phase<0>(W, pe, po, d.base, acc, stage, g0);
		delta64_phase<16>(W, pe, po, d.base, acc, stage, g0);
		if (!h) {
			totals[lane] = acc;
		}
		__syncwarp();
		// copy-out in two batches of 8 granules per thread: all loads of a batch are issued before its stores so that
		// eight 16-byte stores per thread are in flight (a load -> add -> store chain per granule recycles three
		// register quads and stalls on the store's source registers: ncu long-scoreboard on the copy-out)
#pragma unroll
		for (int half = 0; half < 2; ++half) {
			uint4 q[8];
			uint64_t add[8];
#pragma unroll
			for (int j = 0; j < 8; ++j) { // granule g = it*32 + lane: FastLanes lane `it`, half `h`
				const int it = half * 8 + j;
				q[j] = *reinterpret_cast<const uint4*>(stage + (size_t)swz<64>(it * 32 + lane) * 16);
				add[j] = h ? totals[it] : 0ull;
			}
#pragma unroll
			for (int j = 0; j < 8; ++j) {
				const int it = half * 8 + j;
				const uint64_t a = ((((uint64_t)q[j].y) << 32) | q[j].x) + add[j], b = ((((uint64_t)q[j].w) << 32) | q[j].z) + add[j];
				st_cs_v4(out_vec + (size_t)(it * 32 + lane) * 16, (uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32));
			}
		}
		return;
	}
	if constexpr (FAM != kFamDelta) {
	Ffor64Ctx c;
	c.W = W;
	c.W2 = d.misc2 & 0xffu;
	c.base = d.base;
	stream64_bases(s32, lane, (int)W, c.pe, c.po);
	uint32_t packed = 128u * W; // bytes of the packed part; the exception tail follows
	if ((FAM & kFamAlp) != 0 && (FAM == kFamAlp || enc == FLS_ENC_ALP)) {
		c.fa = c_fact_d[(d.misc2 >> 16) & 0xffu];
		c.fr = c_frac_d[(d.misc2 >> 8) & 0xffu];
	}
	if ((FAM & kFamAlpRd) != 0 && (FAM == kFamAlpRd || enc == FLS_ENC_ALP_RD)) {
		const uint4 dq = __ldg(reinterpret_cast<const uint4*>(d.ch->rd_dict));
		c.dw[0] = dq.x;
		c.dw[1] = dq.y;
		c.dw[2] = dq.z;
		c.dw[3] = dq.w;
		stream64_bases(s32 + 32 * W, lane, (int)c.W2, c.pe2, c.po2);
		packed += 128u * c.W2;
	}
	ffor64_phase<FAM, 0>(c, enc, out_vec, lane);
	ffor64_phase<FAM, 16>(c, enc, out_vec, lane);
	patch_exceptions<64, 8>(out_vec, d, lane, 0, slot + packed);
	}
}

// ---- the persistent kernel -----------------------------------------------------------------------------------------------
// MODE 0: main family (T = physical lane width = VT); MODE 1: DICT (T = code lane width, VT = key width);
// MODE 2: RLE (T = 16: run-index lanes, VT = value width).
// Register budget: two resident CTAs per SM (<= 128 registers) everywhere except the 64-bit fused-aggregate kernel, which
// keeps all 32 (lo, hi) pairs of a thread in registers (one CTA of 8 warps, no spills).
template <int MODE, int T, bool AGG, uint32_t FAM>
constexpr int min_ctas_per_sm() {
	// 8/16-bit materialising kernels: short vectors, the per-vector latencies (descriptor, TMA wait) want more resident warps
	return (MODE == 0 && T == 64 && AGG) ? 1 : FAM == kFamSwar8 ? 5 : (MODE == 0 && T <= 16 && !AGG) ? 3 : FLS_MIN_CTAS;
}

// AGG: the vector is decoded into the warp's shared-memory tile (FLS_SINK_SHARED translation unit) and folded into
// per-thread partial aggregates instead of being written to HBM.
template <int MODE, int T, int VT, int OB, bool AGG = false, uint32_t FAM = FLS_FAMILY_MAIN>
__global__ void __launch_bounds__(kWarpsPerCta * 32, min_ctas_per_sm<MODE, T, AGG, FAM>()) fls_decode_kernel(const FlsLaunch L) {
	extern __shared__ __align__(128) uint8_t smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t n_slots = L.n_slots;
	const uint32_t wpc = blockDim.x >> 5; // warps per CTA: kWarpsPerCta unless the launcher had to shrink the CTA to fit shared memory
	uint64_t* bars = reinterpret_cast<uint64_t*>(smem) + warp * n_slots;
	const uint32_t bar_bytes = (wpc * n_slots * 8 + 127u) & ~127u;
	constexpr uint32_t kTileBytes = AGG ? 1024u * OB : 0u;
	const uint32_t warp_bytes = kDescBytesPerWarp + n_slots * L.slot_bytes + L.stage_bytes + kTileBytes;
	uint8_t* descs = smem + bar_bytes + (size_t)warp * warp_bytes;
	uint8_t* slots = descs + kDescBytesPerWarp;
	uint8_t* stage = slots + (size_t)n_slots * L.slot_bytes;
	uint8_t* tile = stage + L.stage_bytes;
	AggPartial part;
	(void)tile;

	// Materialising launches of one stream are independent of each other (inputs: staged compressed bytes; outputs: the
	// caller's column buffers), so the next one may fill the SMs as this one's CTAs retire instead of waiting for the
	// slowest CTA and a launch latency: the tail of column k overlaps the ramp of column k+1.
	if (!AGG && L.pdl) {
		griddep_launch_dependents();
	}

	if (lane == 0) {
		for (uint32_t s = 0; s < n_slots; ++s) {
			mbar_init(bars + s, 1);
		}
		fence_mbar_init();
	}
	__syncwarp();

	const uint32_t n_warps = gridDim.x * wpc;
	const uint32_t gw = blockIdx.x * wpc + warp;
	uint32_t phases = 0; // bit s: parity to wait for on slot s
	uint32_t slot = 0;   // pipeline slot of the next vector this warp consumes (vectors processed % n_slots, kept incrementally)
	uint8_t* const out = static_cast<uint8_t*>(L.out);

	// Vector -> warp assignment: a warp takes runs of C = 2^run_shift consecutive vectors (same row group, same
	// dictionary, adjacent output), runs are dealt round-robin over all warps of the grid; 32 vectors per batch.
	const uint32_t cs = L.run_shift, cmask = (1u << cs) - 1u, runs_per_batch = 32u >> cs;
	for (uint64_t it = 0;; ++it) {
		const uint64_t run0 = it * runs_per_batch; // first run index (per warp) of this batch
		if (((run0 * n_warps + gw) << cs) >= L.total_vecs) {
			break;
		}
		const uint64_t my_v64 = (((run0 + ((uint32_t)lane >> cs)) * n_warps + gw) << cs) + ((uint32_t)lane & cmask);
		uint32_t todo;
		{
			const VecDesc mine = fetch_desc<T, VT>(L, my_v64 < L.total_vecs ? (uint32_t)my_v64 : 0xffffffffu);
			store_desc(descs, lane, mine);
			todo = __ballot_sync(0xffffffffu, mine.ch != nullptr);
			// The blobs of the NEXT batch (all 32 at once; for the first batch also its own) are requested into L2 now, in
			// one burst, instead of trickling out of HBM one pipeline slot at a time: a write stream with a trickle of
			// small reads in it loses ~12 % of its bandwidth to DRAM read/write turnarounds (tools/membw.cu: 6.2 -> 5.4
			// TB/s with 256 B read per 8 KiB written), and the warps of a launch run in near lock step, so the bursts of
			// all warps coincide.  The next batch's descriptor is fetched again when the batch starts (L1/L2 hits).
			if (L.prefetch) {
				if (it == 0 && mine.bytes) {
					tma_prefetch_l2(mine.blob, mine.bytes);
				}
				const uint64_t nx_v64 = my_v64 + (((uint64_t)runs_per_batch * n_warps) << cs);
				if (nx_v64 < L.total_vecs) {
					const VecDesc nxt = fetch_desc<T, VT>(L, (uint32_t)nx_v64);
					if (nxt.bytes) {
						tma_prefetch_l2(nxt.blob, nxt.bytes);
					}
				}
			}
		}
		__syncwarp();
		uint32_t to_issue = todo;
		// prologue: fill the pipeline
		if (lane == 0) {
			uint32_t slot_i = slot;
			for (uint32_t s = 0; s < n_slots && to_issue; ++s) {
				const int k = __ffs(to_issue) - 1;
				to_issue &= to_issue - 1;
				const uint4 q0 = *reinterpret_cast<const uint4*>(descs + (size_t)k * kDescBytes + 16u * desc_quad(k, 0));
				const uint32_t bytes = *reinterpret_cast<const uint32_t*>(descs + (size_t)k * kDescBytes + 16u * desc_quad(k, 1));
				if (bytes) {
					mbar_expect_tx(bars + slot_i, bytes);
					tma_load_1d(slots + (size_t)slot_i * L.slot_bytes, (const void*)(((unsigned long long)q0.y << 32) | q0.x), bytes,
					            bars + slot_i);
				}
				slot_i = slot_i + 1 == n_slots ? 0 : slot_i + 1;
			}
		}
		to_issue = __shfl_sync(0xffffffffu, to_issue, 0);
		const uint64_t vbase = (run0 * n_warps + gw) << cs;
		uint8_t* const out_batch = out + (size_t)vbase * (1024u * OB);
		while (todo) {
			const int k = __ffs(todo) - 1;
			todo &= todo - 1;
			const VecDesc d = load_desc(descs, k);
			if (d.bytes) {
				mbar_wait(bars + slot, (phases >> slot) & 1u);
				phases ^= 1u << slot;
			}
			// vector index = batch base + a 32-bit offset (<= 32 runs further on)
			const uint32_t voff = ((((uint32_t)k >> cs) * n_warps) << cs) + ((uint32_t)k & cmask);
			const uint64_t vk = vbase + voff;
			uint8_t* out_vec = AGG ? tile : out_batch + (size_t)voff * (1024u * OB);
			if constexpr (!AGG) {
				if (L.map_in_place) { // a family's vector list: the vector lands where it belongs in the column
					const uint32_t r = (uint32_t)(d.ch - L.chunks) / L.chunk_stride;
					out_vec = out + ((size_t)r * L.vec_per_rg + (d.vec_in_chunk & 0xffffu)) * (1024u * OB);
				}
			}
			(void)vk;
			const uint8_t* sp = slots + (size_t)slot * L.slot_bytes;
			// aggregate fast path: complete vector, no exceptions, an encoding whose values end up in registers
			const uint32_t enc_k = d.misc >> 24;
			const bool agg_fast = AGG && MODE == 0 && L.mask_mode == 0u && (d.vec_in_chunk >> 16) == 1024u && (d.misc & 0xffffu) == 0u &&
			                      enc_k != FLS_ENC_UNCOMPRESSED && enc_k != FLS_ENC_CONSTANT && enc_k != FLS_ENC_CROSS_RLE;
			if constexpr (MODE == 1) {
				process_dict<T, VT, OB>(d, sp, out_vec, lane, L.conv);
			} else if constexpr (MODE == 2) {
				process_rle<VT, OB>(d, sp, stage, out_vec, lane, L.conv);
			} else if constexpr (T == 64 && !AGG) {
				process_64_2ph<FAM>(d, sp, stage, out_vec, lane);
			} else if constexpr (T == 64) {
				process_64<FAM, AGG>(d, sp, stage, out_vec, lane, part, agg_fast, L.agg_signed, L.agg_float);
			} else {
				process_small<T, OB, FAM, AGG>(d, sp, stage, out_vec, lane, L.conv, part, agg_fast, L.agg_signed, L.agg_float);
			}
			__syncwarp(); // all lanes are done reading the slot (and the staging area)
			if constexpr (AGG) {
				if (L.mask_mode != 0u) { // filter pushdown: the tile holds the vector in natural order
					uint32_t* mask_vec = L.mask + (size_t)vk * 32u;
					if (L.mask_mode == 3u) {
						part.template fold_masked<VT>(tile, lane, d.vec_in_chunk >> 16, L.agg_signed, L.agg_float, mask_vec);
					} else {
						AggPartial::template build_mask<VT>(tile, lane, d.vec_in_chunk >> 16, L.agg_signed, L.agg_float, L.cmp_op, L.cmp_const,
						                                    L.cmp_op2, L.cmp_const2, L.mask_mode == 2u, mask_vec);
					}
					__syncwarp();
				} else if (!agg_fast) {
					part.template fold<VT>(tile, lane, d.vec_in_chunk >> 16, L.agg_signed, L.agg_float);
					__syncwarp();
				}
			}
			if (to_issue) {
				const int kn = __ffs(to_issue) - 1;
				to_issue &= to_issue - 1;
				if (lane == 0) {
					const uint4 q0 = *reinterpret_cast<const uint4*>(descs + (size_t)kn * kDescBytes + 16u * desc_quad(kn, 0));
					const uint32_t bytes = *reinterpret_cast<const uint32_t*>(descs + (size_t)kn * kDescBytes + 16u * desc_quad(kn, 1));
					if (bytes) {
						mbar_expect_tx(bars + slot, bytes);
						tma_load_1d(slots + (size_t)slot * L.slot_bytes, (const void*)(((unsigned long long)q0.y << 32) | q0.x), bytes,
						            bars + slot);
					}
				}
			}
			slot = slot + 1 == n_slots ? 0 : slot + 1;
		}
	}
	if constexpr (AGG) {
		if (L.mask_mode == 0u || L.mask_mode == 3u) {
			part.flush(L.agg, lane, L.agg_float);
		}
	} else if (L.pdl) {
		// keep completion transitive: this grid must not be seen complete before the grid it was allowed to overtake
		// (whatever follows in the stream waits for THIS grid only); at this point the predecessor has all but finished
		griddep_wait();
	}
}

} // namespace FLS_NS

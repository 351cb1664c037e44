// fls_decode_gather.cuh — DICT and RLE decode bodies (index lanes differ from the value width).
#pragma once
#include "fls_decode.cuh"

namespace FLS_NS {

// dec_dict_opr<KEY, IDX> (reference dictionary_kernel.hpp:11-22): out[i] = keys[index[i]].
// IT = lane width of the codes (8/16/32), VT = key width.  A thread's G = 32/IT adjacent codes give G adjacent
// outputs, stored with one vector store.  Keys come through the read-only path (dictionary is L1/L2 resident).
template <int IT, int VT, int OB>
__device__ __forceinline__ void process_dict(const VecDesc& d, const uint8_t* slot, uint8_t* out_vec, int lane, uint32_t conv) {
	const uint32_t W = (d.misc >> 16) & 0xffu;
	const uint32_t base = (uint32_t)d.base;
	const uint8_t* keys = d.aux0;
	const uint32_t n_keys = d.aux_n;
	if constexpr (IT == 8) {
		// 8-bit codes stay four to a register (SWAR unpack): byte b of x[row] is the code of position row*128 + 4*lane + b
		uint32_t x[8];
		unpack8_swar_dispatch((int)W, reinterpret_cast<const uint32_t*>(slot), lane, x);
		if constexpr (OB == 8) {
			// 8-byte outputs: a thread that kept its four positions would store 32 bytes at a 32-byte stride (half-filled
			// sectors per store instruction).  The CODES are exchanged instead — lane l takes positions 64*half + 2l, 2l+1
			// from lane 16*half + l/2 — so every store instruction writes 512 contiguous bytes.
#pragma unroll
			for (int row = 0; row < 8; ++row) {
#pragma unroll
				for (int half = 0; half < 2; ++half) {
					const uint32_t w = __shfl_sync(0xffffffffu, x[row], 16 * half + (lane >> 1));
					const uint32_t pair = (lane & 1) ? (w >> 16) : (w & 0xffffu);
					uint32_t lo[2], hi[2];
#pragma unroll
					for (int g = 0; g < 2; ++g) {
						const uint32_t code = min(((pair >> (8 * g)) & 0xffu) + base, n_keys - 1u);
						load_elem<VT>(keys, code, lo[g], hi[g]);
						convert_val<VT, OB>(lo[g], hi[g], conv);
					}
					store_adjacent<OB, 2>(out_vec, row * 128 + 64 * half + 2 * lane, lo, hi);
				}
			}
		} else {
#pragma unroll
			for (int row = 0; row < 8; ++row) {
				uint32_t lo[4], hi[4];
#pragma unroll
				for (int g = 0; g < 4; ++g) {
					const uint32_t code = min(((x[row] >> (8 * g)) & 0xffu) + base, n_keys - 1u); // clamp: never read outside the dictionary
					load_elem<VT>(keys, code, lo[g], hi[g]);
					convert_val<VT, OB>(lo[g], hi[g], conv);
				}
				store_adjacent<OB, 4>(out_vec, row * 128 + 4 * lane, lo, hi);
			}
		}
		return;
	}
	uint32_t v[32];
	unpack_small_dispatch<IT>((int)W, reinterpret_cast<const uint32_t*>(slot), lane, v);
	constexpr int G = 32 / IT;
#pragma unroll
	for (int r = 0; r < 32 / G; ++r) {
		uint32_t lo[G], hi[G];
#pragma unroll
		for (int g = 0; g < G; ++g) {
			const uint32_t code = min(v[r * G + g] + base, n_keys - 1u); // clamp: never read outside the dictionary
			load_elem<VT>(keys, code, lo[g], hi[g]);
			convert_val<VT, OB>(lo[g], hi[g], conv);
		}
		store_adjacent<OB, G>(out_vec, value_pos<IT>(lane, r * G), lo, hi);
	}
}

// dec_rle_map_opr<KEY, u16> (reference rle_map_kernel.hpp:15-23,45-58): the run index is a DELTA-coded u16 vector;
// decode it (per-lane running sums), transpose it through the staging area, then out[i] = rle_vals[idx[i]].
template <int VT, int OB>
__device__ __forceinline__ void process_rle(const VecDesc& d, const uint8_t* slot, uint8_t* stage, uint8_t* out_vec, int lane,
                                            uint32_t conv) {
	const uint32_t W = (d.misc >> 16) & 0xffu;
	const uint32_t* s32 = reinterpret_cast<const uint32_t*>(slot);
	uint32_t v[32];
	unpack_small_dispatch<16>((int)W, s32 + 32, lane, v);
	const uint32_t bw = s32[lane], dmin = (uint32_t)d.base;
#pragma unroll
	for (int h = 0; h < 2; ++h) {
		uint32_t acc = (bw >> (16 * h)) & 0xffffu;
#pragma unroll
		for (int row = 0; row < 16; ++row) {
			acc += v[2 * row + h] + dmin;
			v[2 * row + h] = acc & 0xffffu;
		}
		const int g0 = chain_l0(2 * lane + h) >> 3;
#pragma unroll
		for (int q = 0; q < 2; ++q) {
			uint32_t w4[4];
#pragma unroll
			for (int m = 0; m < 4; ++m) {
				w4[m] = v[2 * (8 * q + 2 * m) + h] | (v[2 * (8 * q + 2 * m + 1) + h] << 16);
			}
			*reinterpret_cast<uint4*>(stage + (size_t)swz<16>(g0 + q) * 16) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
		}
	}
	__syncwarp();
	const uint8_t* vals = d.aux0;
	const uint32_t first = d.aux_off, n_runs = d.misc & 0xffffu;
	if constexpr (OB >= 4) {
		// wide outputs: a thread takes A = 16 / OB adjacent indexes per store, so every store instruction of the warp writes
		// 512 contiguous bytes (8 indexes per thread would be 32- or 64-byte pieces at a stride: partly filled sectors)
		constexpr int A = 16 / OB;
#pragma unroll
		for (int it = 0; it < 1024 / (32 * A); ++it) {
			const int pos = (it * 32 + lane) * A;
			const uint8_t* src = stage + (size_t)swz<16>(pos >> 3) * 16 + (pos & 7) * 2;
			uint32_t w[A / 2];
			if constexpr (A == 2) {
				w[0] = *reinterpret_cast<const uint32_t*>(src);
			} else {
				const uint2 q = *reinterpret_cast<const uint2*>(src);
				w[0] = q.x;
				w[1] = q.y;
			}
			uint32_t lo[A], hi[A];
#pragma unroll
			for (int s = 0; s < A; ++s) {
				const uint32_t idx = min((s & 1) ? (w[s >> 1] >> 16) : (w[s >> 1] & 0xffffu), n_runs - 1u);
				load_elem<VT>(vals, first + idx, lo[s], hi[s]);
				convert_val<VT, OB>(lo[s], hi[s], conv);
			}
			store_adjacent<OB, A>(out_vec, pos, lo, hi);
		}
		return;
	}
#pragma unroll
	for (int it = 0; it < 4; ++it) { // 2 KB of u16 indexes = 128 granules of 8
		const int g = it * 32 + lane;
		const uint4 q = *reinterpret_cast<const uint4*>(stage + (size_t)swz<16>(g) * 16);
		const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
		for (int half = 0; half < 2; ++half) {
			uint32_t lo[4], hi[4];
#pragma unroll
			for (int s = 0; s < 4; ++s) {
				const uint32_t word = w[half * 2 + (s >> 1)];
				const uint32_t idx = min((s & 1) ? (word >> 16) : (word & 0xffffu), n_runs - 1u);
				load_elem<VT>(vals, first + idx, lo[s], hi[s]);
				convert_val<VT, OB>(lo[s], hi[s], conv);
			}
			store_adjacent<OB, 4>(out_vec, g * 8 + half * 4, lo, hi);
		}
	}
}

} // namespace FLS_NS

// fls_store.cuh — widening conversion + coalesced stores of the 32 values a thread holds after unpacking.
// Conversion semantics = the materializer's CastNumeric static_cast loop (reference fls_type_resolver.hpp:22-33) with
// signed columns viewed as signed first (unffor_kernel.hpp:16-25): conv 0 zero-extends, 1 sign-extends, 2 is
// float -> double.
#pragma once
#include "fls_ptx.cuh"
#include "fls_unpack.cuh"

namespace FLS_NS {

// (lo,hi) of one T-bit value -> OB-byte value
template <int T, int OB>
__device__ __forceinline__ void convert_val(uint32_t& lo, uint32_t& hi, uint32_t conv) {
	if constexpr (OB * 8 > T) {
		if (conv == 1) {
			if constexpr (T == 8) {
				lo = (uint32_t)(int32_t)(int8_t)lo;
			} else if constexpr (T == 16) {
				lo = (uint32_t)(int32_t)(int16_t)lo;
			}
			hi = (uint32_t)((int32_t)lo >> 31);
		} else if (conv == 2 && T == 32 && OB == 8) {
			const double d = (double)__uint_as_float(lo);
			const long long b = __double_as_longlong(d);
			lo = (uint32_t)b;
			hi = (uint32_t)(b >> 32);
		} else {
			hi = 0;
		}
	}
}

// Store G adjacent converted values (positions p..p+G-1, p multiple of G) with the widest store available.
template <int OB, int G>
__device__ __forceinline__ void store_adjacent(uint8_t* out_vec, int p, const uint32_t (&lo)[G], const uint32_t (&hi)[G]) {
	uint8_t* dst = out_vec + (size_t)p * OB;
	if constexpr (OB == 8) {
#pragma unroll
		for (int g = 0; g < G; g += 2) {
			if constexpr (G == 1) {
				st_cs_v2(dst, lo[0], hi[0]);
			} else {
				st_cs_v4(dst + g * 8, lo[g], hi[g], lo[g + 1], hi[g + 1]);
			}
		}
	} else if constexpr (OB == 4) {
		if constexpr (G == 1) {
			st_cs_u32(dst, lo[0]);
		} else if constexpr (G == 2) {
			st_cs_v2(dst, lo[0], lo[1]);
		} else {
			st_cs_v4(dst, lo[0], lo[1], lo[2], lo[3]);
		}
	} else if constexpr (OB == 2) {
		if constexpr (G == 1) {
			st_cs_u16(dst, lo[0]);
		} else if constexpr (G == 2) {
			st_cs_u32(dst, (lo[0] & 0xffffu) | (lo[1] << 16));
		} else {
			st_cs_v2(dst, (lo[0] & 0xffffu) | (lo[1] << 16), (lo[2] & 0xffffu) | (lo[3] << 16));
		}
	} else {
		if constexpr (G == 4) {
			st_cs_u32(dst, (lo[0] & 0xffu) | ((lo[1] & 0xffu) << 8) | ((lo[2] & 0xffu) << 16) | (lo[3] << 24));
		} else if constexpr (G == 2) {
			st_cs_u16(dst, (lo[0] & 0xffu) | ((lo[1] & 0xffu) << 8));
		} else {
			st_cs_u8(dst, lo[0]);
		}
	}
}

// Direct (non-transposed) store of a thread's 32 values at their FastLanes positions.  T <= 32.
template <int T, int OB>
__device__ __forceinline__ void store_direct(uint8_t* out_vec, int tid, const uint32_t (&v)[32], uint32_t conv) {
	constexpr int G = 32 / T;
#pragma unroll
	for (int r = 0; r < 32 / G; ++r) {
		uint32_t lo[G], hi[G];
#pragma unroll
		for (int g = 0; g < G; ++g) {
			lo[g] = v[r * G + g];
			hi[g] = 0;
			convert_val<T, OB>(lo[g], hi[g], conv);
		}
		store_adjacent<OB, G>(out_vec, value_pos<T>(tid, r * G), lo, hi);
	}
}

// T = 64
__device__ __forceinline__ void store_direct64(uint8_t* out_vec, int tid, const uint32_t (&lo)[32], const uint32_t (&hi)[32]) {
#pragma unroll
	for (int j = 0; j < 32; ++j) {
		st_cs_v2(out_vec + (size_t)value_pos<64>(tid, j) * 8, lo[j], hi[j]);
	}
}

// One converted value at logical position pos (exception patching, gathers).
template <int T, int OB>
__device__ __forceinline__ void store_one(uint8_t* out_vec, int pos, uint32_t lo, uint32_t hi, uint32_t conv) {
	convert_val<T, OB>(lo, hi, conv);
	const uint32_t l[1] = {lo}, h[1] = {hi};
	store_adjacent<OB, 1>(out_vec, pos, l, h);
}

// Swizzled 16-byte granule index of the per-warp staging area (conflict-free transposition):
// line = g >> 3; granule-in-line is XORed with (line >> SH) & 7, SH = 2 for T = 64 and 1 otherwise.
template <int T>
__device__ __forceinline__ int swz(int g) {
	constexpr int SH = T == 64 ? 2 : 1;
	return (g & ~7) | ((g ^ (g >> (3 + SH))) & 7);
}

// Copy a whole vector that sits in shared memory in LOGICAL order (T-bit elements) to global memory with
// conversion: 16-byte granules, thread t takes granule it*32 + t, so every warp store covers contiguous memory.
template <int T, int OB, bool SWZ>
__device__ __forceinline__ void copy_out_linear(const uint8_t* smem, uint8_t* out_vec, int tid, uint32_t conv) {
	constexpr int N_IT = T / 4;        // 1024*T/8 bytes / 16 / 32
	constexpr int PER = 128 / T;       // elements per granule
	if constexpr (T == 64 || OB * 8 == T) {
		// same-width copy: all loads of a batch (<= 8 granules per thread) are issued before its stores, so that the stores
		// do not wait one by one for a recycled register quad
		constexpr int B = N_IT < 8 ? N_IT : 8;
#pragma unroll
		for (int b0 = 0; b0 < N_IT; b0 += B) {
			uint4 q[B];
#pragma unroll
			for (int j = 0; j < B; ++j) {
				const int g = (b0 + j) * 32 + tid;
				q[j] = *reinterpret_cast<const uint4*>(smem + (size_t)(SWZ ? swz<T>(g) : g) * 16);
			}
#pragma unroll
			for (int j = 0; j < B; ++j) {
				const int g = (b0 + j) * 32 + tid;
				st_cs_v4(out_vec + (size_t)g * 16, q[j].x, q[j].y, q[j].z, q[j].w);
			}
		}
	} else {
#pragma unroll
		for (int it = 0; it < N_IT; ++it) {
			const int g = it * 32 + tid;
			const uint4 q = *reinterpret_cast<const uint4*>(smem + (size_t)(SWZ ? swz<T>(g) : g) * 16);
			const uint32_t w[4] = {q.x, q.y, q.z, q.w};
			const int e0 = g * PER;
			constexpr int PW = 32 / T; // elements per 32-bit word
#pragma unroll
			for (int k = 0; k < 4; ++k) {
				uint32_t lo[PW], hi[PW];
#pragma unroll
				for (int s = 0; s < PW; ++s) {
					lo[s] = T == 32 ? w[k] : (w[k] >> (s * T)) & ((1u << (T & 31)) - 1u);
					hi[s] = 0;
					convert_val<T, OB>(lo[s], hi[s], conv);
				}
				store_adjacent<OB, PW>(out_vec, e0 + k * PW, lo, hi);
			}
		}
	}
}

} // namespace FLS_NS

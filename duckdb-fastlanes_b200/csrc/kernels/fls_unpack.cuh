// fls_unpack.cuh — per-thread FastLanes bit-unpack, specialised on (T, W) at compile time.
//
// One warp decodes one 1024-value vector.  A thread owns one 32-bit "super lane": for T=32 that is FastLanes lane
// `tid`; for T=16 it holds lanes 2tid,2tid+1 side by side (SWAR); for T=8 lanes 4tid..4tid+3; for T=64 a thread
// owns HALF a lane (lane tid&15, rows 32*(tid>>4)..+31) and reads it as a stream of 32-bit words — the half-lane of
// rows 32..63 starts exactly W 32-bit words into the lane's bit stream, so both halves run the same code.
// Every thread therefore produces 32 values v[0..31]; value_pos<T>(tid, i) gives the logical position of v[i].
//
// The packed vector sits in shared memory exactly as in the file (TMA bulk copy): T-bit word k of lane l at
// index k*LANES + l, i.e. 32-bit super-lane word k of thread tid at s32[k*32 + tid] for T <= 32.
//
// Compiles for host too (FLS_HD) so that tests can run the exact index math without a GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define FLS_HD __host__ __device__ __forceinline__
#else
#define FLS_HD inline
#endif

// The aggregate translation unit re-instantiates every template with its stores aimed at shared memory
// (FLS_SINK_SHARED); it does so in its own namespace so that the two sets of instantiations never alias.
#ifndef FLS_NS
#define FLS_NS flsk
#endif

namespace FLS_NS {

FLS_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {
#if defined(__CUDA_ARCH__)
	return __funnelshift_r(lo, hi, sh);
#else
	return (uint32_t)(((((uint64_t)hi) << 32) | lo) >> (sh & 31));
#endif
}

FLS_HD constexpr int fl_order(int o) {
	return ((o & 1) << 2) | (o & 2) | ((o >> 2) & 1); // {0,4,2,6,1,5,3,7}: 3-bit reversal
}

// Logical position of value i (0..31) of thread tid.
template <int T>
FLS_HD int value_pos(int tid, int i) {
	if (T >= 32) { // row j = i (T=64: row 32*(tid>>4)+i of lane tid&15, which lands on the same formula)
		return fl_order(i >> 3) * 16 + (i & 7) * 128 + tid;
	} else if (T == 16) { // i = 2*row + h
		const int row = i >> 1, h = i & 1;
		return (row >> 3) * 64 + (row & 7) * 128 + 2 * tid + h;
	} else { // T == 8: i = 4*row + b
		const int row = i >> 2, b = i & 3;
		return row * 128 + 4 * tid + b;
	}
}

// Replicated field mask: low n bits of every T-bit field of a 32-bit word.
template <int T>
FLS_HD constexpr uint32_t field_mask(int n) {
	return T == 32 ? (n >= 32 ? 0xffffffffu : ((1u << n) - 1u))
	     : T == 16 ? (n >= 16 ? 0xffffffffu : (((1u << n) - 1u) * 0x00010001u))
	               : (n >= 8 ? 0xffffffffu : (((1u << n) - 1u) * 0x01010101u));
}

// ---- T in {8,16,32}: v[i] as separate 32-bit registers ---------------------------------------------------------------
template <int T, int W>
FLS_HD void unpack_small(const uint32_t* s32, int tid, uint32_t (&v)[32]) {
	constexpr int PER = 32 / T; // values per 32-bit word
	uint32_t w[W + 1];
#pragma unroll
	for (int k = 0; k < W; ++k) {
		w[k] = s32[k * 32 + tid];
	}
	w[W] = 0;
#pragma unroll
	for (int row = 0; row < T; ++row) {
		const int bit = row * W, k = bit / T, sh = bit % T;
		uint32_t x;
		if (W == 0) {
			x = 0;
		} else if (sh + W <= T) {
			x = (w[k] >> sh) & field_mask<T>(W);
		} else if (T == 32) {
			x = funnel_r(w[k], w[k + 1], sh) & field_mask<T>(W);
		} else {
			x = ((w[k] >> sh) & field_mask<T>(T - sh)) | ((w[k + 1] & field_mask<T>(sh + W - T)) << (T - sh));
		}
		if (T == 32) {
			v[row] = x;
		} else if (T == 16) {
			v[2 * row] = x & 0xffffu;
			v[2 * row + 1] = x >> 16;
		} else {
			v[4 * row] = x & 0xffu;
			v[4 * row + 1] = (x >> 8) & 0xffu;
			v[4 * row + 2] = (x >> 16) & 0xffu;
			v[4 * row + 3] = x >> 24;
		}
	}
	(void)PER;
}

// ---- T = 64: (lo, hi) register pairs -----------------------------------------------------------------------------------
// pe / po: base pointers for even / odd stream-word indices (see stream64_bases).
template <int W>
FLS_HD void unpack_64(const uint32_t* pe, const uint32_t* po, uint32_t (&lo)[32], uint32_t (&hi)[32]) {
	uint32_t w[W + 2];
#pragma unroll
	for (int k = 0; k < W; ++k) {
		w[k] = ((k & 1) ? po : pe)[(k >> 1) * 32 + (k & 1)];
	}
	w[W] = 0;
	w[W + 1] = 0;
#pragma unroll
	for (int j = 0; j < 32; ++j) {
		const int p = j * W, k = p >> 5, sh = p & 31;
		constexpr int WL = W < 32 ? W : 32; // bits of the value that live in `lo`
		uint32_t l, h;
		if (W == 0) {
			l = 0;
			h = 0;
		} else {
			if (sh == 0) {
				l = w[k];
			} else if (sh + WL <= 32) {
				l = w[k] >> sh;
			} else {
				l = funnel_r(w[k], w[k + 1], sh);
			}
			if (WL < 32) {
				l &= (1u << (WL & 31)) - 1u;
			}
			if (W <= 32) {
				h = 0;
			} else {
				constexpr int WH = W - 32; // 1..32 bits in `hi`, starting at stream bit p+32
				if (sh == 0) {
					h = w[k + 1];
				} else if (sh + WH <= 32) {
					h = w[k + 1] >> sh;
				} else {
					h = funnel_r(w[k + 1], w[k + 2], sh);
				}
				if (WH < 32) {
					h &= (1u << (WH & 31)) - 1u;
				}
			}
		}
		lo[j] = l;
		hi[j] = h;
	}
}

// Rows [J0, J0+NJ) of the thread's half lane only (NJ = 16: the materialising T = 64 kernels decode a vector in two
// phases so that 16 instead of 32 (lo, hi) pairs are live at a time: half the registers, twice the resident warps).
// Only the stream words those rows touch are loaded.
template <int W, int J0, int NJ>
FLS_HD void unpack_64_part(const uint32_t* pe, const uint32_t* po, uint32_t (&lo)[NJ], uint32_t (&hi)[NJ]) {
	constexpr int K0 = (J0 * W) >> 5;                      // first stream word touched
	constexpr int K1 = W == 0 ? 0 : (((J0 + NJ) * W + 31) >> 5); // one past the last
	uint32_t w[K1 - K0 + 2];
#pragma unroll
	for (int k = K0; k < K1; ++k) {
		w[k - K0] = ((k & 1) ? po : pe)[(k >> 1) * 32 + (k & 1)];
	}
	w[K1 - K0] = 0;
	w[K1 - K0 + 1] = 0;
#pragma unroll
	for (int jj = 0; jj < NJ; ++jj) {
		const int p = (J0 + jj) * W, k = (p >> 5) - K0, sh = p & 31;
		constexpr int WL = W < 32 ? W : 32;
		uint32_t l, h;
		if (W == 0) {
			l = 0;
			h = 0;
		} else {
			if (sh == 0) {
				l = w[k];
			} else if (sh + WL <= 32) {
				l = w[k] >> sh;
			} else {
				l = funnel_r(w[k], w[k + 1], sh);
			}
			if (WL < 32) {
				l &= (1u << (WL & 31)) - 1u;
			}
			if (W <= 32) {
				h = 0;
			} else {
				constexpr int WH = W - 32;
				if (sh == 0) {
					h = w[k + 1];
				} else if (sh + WH <= 32) {
					h = w[k + 1] >> sh;
				} else {
					h = funnel_r(w[k + 1], w[k + 2], sh);
				}
				if (WH < 32) {
					h &= (1u << (WH & 31)) - 1u;
				}
			}
		}
		lo[jj] = l;
		hi[jj] = h;
	}
}

// Base pointers of thread tid's 32-bit word stream inside a T=64 packed vector at s32 (see header comment).
FLS_HD void stream64_bases(const uint32_t* s32, int tid, int W, const uint32_t*& pe, const uint32_t*& po) {
	const int l = tid & 15, h = tid >> 4;
	const int s = h * W, odd = s & 1;
	const uint32_t* b = s32 + (s >> 1) * 32 + 2 * l;
	pe = b + odd;      // even k: f(k) + odd*1
	po = b + odd * 31; // odd  k: f(k) + odd*31
}

// ---- runtime-W dispatch -----------------------------------------------------------------------------------------------
// ---- T = 8, SWAR form: x[row] keeps the four 8-bit lanes 4tid..4tid+3 of a row side by side in one register -------------
// (what unpack_small<8, W> computes before it splits the word into four registers; the 8-bit kernels add the base / the
// running sum with byte-wise SWAR arithmetic and store the word as it is: 4 values per instruction instead of 1)
template <int W>
FLS_HD void unpack8_swar(const uint32_t* s32, int tid, uint32_t (&x)[8]) {
	uint32_t w[W + 1];
#pragma unroll
	for (int k = 0; k < W; ++k) {
		w[k] = s32[k * 32 + tid];
	}
	w[W] = 0;
#pragma unroll
	for (int row = 0; row < 8; ++row) {
		const int bit = row * W, k = bit / 8, sh = bit % 8;
		if (W == 0) {
			x[row] = 0;
		} else if (sh + W <= 8) {
			x[row] = (w[k] >> sh) & field_mask<8>(W);
		} else {
			x[row] = ((w[k] >> sh) & field_mask<8>(8 - sh)) | ((w[k + 1] & field_mask<8>(sh + W - 8)) << (8 - sh));
		}
	}
}

FLS_HD void unpack8_swar_dispatch(int W, const uint32_t* s32, int tid, uint32_t (&x)[8]) {
	switch (W) {
	case 0: unpack8_swar<0>(s32, tid, x); break;
	case 1: unpack8_swar<1>(s32, tid, x); break;
	case 2: unpack8_swar<2>(s32, tid, x); break;
	case 3: unpack8_swar<3>(s32, tid, x); break;
	case 4: unpack8_swar<4>(s32, tid, x); break;
	case 5: unpack8_swar<5>(s32, tid, x); break;
	case 6: unpack8_swar<6>(s32, tid, x); break;
	case 7: unpack8_swar<7>(s32, tid, x); break;
	default: unpack8_swar<8>(s32, tid, x); break;
	}
}

// byte-wise a + b (mod 256 per byte, no carries between bytes)
FLS_HD uint32_t add8x4(uint32_t a, uint32_t b) {
	return ((a & 0x7f7f7f7fu) + (b & 0x7f7f7f7fu)) ^ ((a ^ b) & 0x80808080u);
}

#define FLS_CASE_S(T, Wv) \
	case Wv: unpack_small<T, Wv>(s32, tid, v); break;
#define FLS_CASES8(M, T, b) M(T, b + 0) M(T, b + 1) M(T, b + 2) M(T, b + 3) M(T, b + 4) M(T, b + 5) M(T, b + 6) M(T, b + 7)

template <int T>
FLS_HD void unpack_small_dispatch(int W, const uint32_t* s32, int tid, uint32_t (&v)[32]) {
	if constexpr (T == 8) {
		switch (W) {
			FLS_CASES8(FLS_CASE_S, 8, 0)
			case 8: unpack_small<8, 8>(s32, tid, v); break;
		default: break;
		}
	} else if constexpr (T == 16) {
		switch (W) {
			FLS_CASES8(FLS_CASE_S, 16, 0)
			FLS_CASES8(FLS_CASE_S, 16, 8)
			case 16: unpack_small<16, 16>(s32, tid, v); break;
		default: break;
		}
	} else {
		switch (W) {
			FLS_CASES8(FLS_CASE_S, 32, 0)
			FLS_CASES8(FLS_CASE_S, 32, 8)
			FLS_CASES8(FLS_CASE_S, 32, 16)
			FLS_CASES8(FLS_CASE_S, 32, 24)
			case 32: unpack_small<32, 32>(s32, tid, v); break;
		default: break;
		}
	}
}

#define FLS_CASE_64(T, Wv) \
	case Wv: unpack_64<Wv>(pe, po, lo, hi); break;

FLS_HD void unpack_64_dispatch(int W, const uint32_t* s32, int tid, uint32_t (&lo)[32], uint32_t (&hi)[32]) {
	const uint32_t *pe, *po;
	stream64_bases(s32, tid, W, pe, po);
	switch (W) {
		FLS_CASES8(FLS_CASE_64, 64, 0)
		FLS_CASES8(FLS_CASE_64, 64, 8)
		FLS_CASES8(FLS_CASE_64, 64, 16)
		FLS_CASES8(FLS_CASE_64, 64, 24)
		FLS_CASES8(FLS_CASE_64, 64, 32)
		FLS_CASES8(FLS_CASE_64, 64, 40)
		FLS_CASES8(FLS_CASE_64, 64, 48)
		FLS_CASES8(FLS_CASE_64, 64, 56)
		case 64: unpack_64<64>(pe, po, lo, hi); break;
	default: break;
	}
}

#define FLS_CASE_64P(T, Wv) \
	case Wv: unpack_64_part<Wv, J0, 16>(pe, po, lo, hi); break;

// rows [J0, J0+16) of the half lane; pe/po from stream64_bases
template <int J0>
FLS_HD void unpack_64_part_dispatch(int W, const uint32_t* pe, const uint32_t* po, uint32_t (&lo)[16], uint32_t (&hi)[16]) {
	switch (W) {
		FLS_CASES8(FLS_CASE_64P, 64, 0)
		FLS_CASES8(FLS_CASE_64P, 64, 8)
		FLS_CASES8(FLS_CASE_64P, 64, 16)
		FLS_CASES8(FLS_CASE_64P, 64, 24)
		FLS_CASES8(FLS_CASE_64P, 64, 32)
		FLS_CASES8(FLS_CASE_64P, 64, 40)
		FLS_CASES8(FLS_CASE_64P, 64, 48)
		FLS_CASES8(FLS_CASE_64P, 64, 56)
		case 64: unpack_64_part<64, J0, 16>(pe, po, lo, hi); break;
	default: break;
	}
}

} // namespace FLS_NS

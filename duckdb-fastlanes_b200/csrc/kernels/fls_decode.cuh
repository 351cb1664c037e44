// fls_decode.cuh — the fused decode kernels (sm_100a).
//
// Work unit: one warp decodes one 1024-value vector end to end: TMA bulk copy of the packed blob into the warp's
// shared-memory slot -> register unpack specialised on the bit width -> encoding epilogue (FOR base, per-lane
// running sum + transposition, ALP multiply, ALP-RD glue, patches) -> coalesced / 128-bit stores.  This fuses what
// the reference does in two passes (FastLanes library decode into scratch, then the materializer's copy / cast /
// untranspose: reference src/reader/fls_reader.cpp:436-472, src/include/reader/materializer/kernels/*.hpp).
//
// Kernels are persistent: each warp walks vectors gw, gw + n_warps, ... in batches of 32; lane k of the warp
// fetches the descriptor (chunk + fls_vec_meta entry point) of the batch's k-th vector so that descriptor latency
// is paid once per 32 vectors, and the warp keeps n_slots TMA copies in flight ahead of the vector it is decoding.
#pragma once
#include "fls_device.h"
#include "fls_store.cuh"

namespace FLS_NS {

__constant__ double c_fact_d[19] = {1e0, 1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,
                                    1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18};
__constant__ double c_frac_d[21] = {1.0,
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
__constant__ float c_fact_f[11] = {1.0F,       10.0F,       100.0F,       1000.0F,       10000.0F,      100000.0F,
                                   1000000.0F, 10000000.0F, 100000000.0F, 1000000000.0F, 10000000000.0F};
__constant__ float c_frac_f[11] = {1.0F,      0.1F,       0.01F,       0.001F,       0.0001F,      0.00001F,
                                   0.000001F, 0.0000001F, 0.00000001F, 0.000000001F, 0.0000000001F};

constexpr int kWarpsPerCta = 8;
#ifndef FLS_MIN_CTAS
#define FLS_MIN_CTAS 1
#endif

// Everything a warp needs to decode one vector; filled per lane, broadcast with shuffles.
struct VecDesc {
	const uint8_t* blob;
	const FlsChunkDev* ch;
	uint64_t base;
	uint32_t aux_off;
	uint32_t bytes; // TMA bytes of the blob (0: nothing to copy)
	uint32_t misc;  // aux_cnt | width << 16 | encoding << 24
	uint32_t misc2; // width2 | e << 8 | f << 16
	uint32_t vec_in_chunk; // vector index inside its chunk | n_valid << 16
	// chunk-level pointers, copied at descriptor-fetch time (lane-parallel, once per 32 vectors) so that no warp ever
	// chases ch-> through global memory in the middle of a vector
	const uint8_t* aux0; // dictionary keys | run values | exception values | CROSS_RLE values
	const uint8_t* aux1; // exception positions | CROSS_RLE run starts
	uint32_t aux_n;      // ch->aux_count (#keys, #runs)
};

// A batch's 32 descriptors live in the warp's shared memory (64 B each): lane k stores the one it fetched, every lane
// then reads descriptor k with broadcast loads (same address across the warp = one wavefront) and only the fields the
// instantiation uses.  (Broadcasting through shuffles kept ~18 registers live per thread across the whole decode body;
// ptxas spilled them to local memory: 16 SHFL + ~12 LDL/STL per vector.)
constexpr uint32_t kDescBytes = 64, kDescBytesPerWarp = 32 * kDescBytes;
static_assert(kDescBytesPerWarp == FLS_DESC_BYTES_PER_WARP, "launcher and kernel disagree on the descriptor area");

// where quad j of descriptor k sits inside the descriptor's 64 bytes (see store_desc)
__device__ __forceinline__ uint32_t desc_quad(int k, int j) {
#ifdef FLS_DESC_PLAIN
	(void)k;
	return (uint32_t)j;
#else
	return (uint32_t)(j ^ ((k >> 1) & 3));
#endif
}

__device__ __forceinline__ void store_desc(uint8_t* descs, int lane, const VecDesc& d) {
	uint4* p = reinterpret_cast<uint4*>(descs + (size_t)lane * kDescBytes);
	const unsigned long long blob = (unsigned long long)d.blob, ch = (unsigned long long)d.ch;
	const unsigned long long a0 = (unsigned long long)d.aux0, a1 = (unsigned long long)d.aux1;
	// quad j of lane's descriptor sits at quad j ^ x, x = (lane >> 1) & 3: the 8 lanes of a quarter warp (the unit a 128-bit
	// shared store is checked in) then hit 8 different 16-byte bank groups instead of two (stride 64 B: 4-way conflicts, ncu
	// 12 excessive wavefronts per store)
	const int x = (int)desc_quad(lane, 0);
	p[0 ^ x] = make_uint4((uint32_t)blob, (uint32_t)(blob >> 32), (uint32_t)d.base, (uint32_t)(d.base >> 32));
	p[1 ^ x] = make_uint4(d.bytes, d.misc, d.misc2, d.vec_in_chunk);
	p[2 ^ x] = make_uint4((uint32_t)a0, (uint32_t)(a0 >> 32), (uint32_t)a1, (uint32_t)(a1 >> 32));
	p[3 ^ x] = make_uint4((uint32_t)ch, (uint32_t)(ch >> 32), d.aux_off, d.aux_n);
}

__device__ __forceinline__ VecDesc load_desc(const uint8_t* descs, int k) {
	const uint4* p = reinterpret_cast<const uint4*>(descs + (size_t)k * kDescBytes);
	const int x = (int)desc_quad(k, 0);
	const uint4 q0 = p[0 ^ x], q1 = p[1 ^ x], q2 = p[2 ^ x], q3 = p[3 ^ x];
	VecDesc r;
	r.blob = (const uint8_t*)(((unsigned long long)q0.y << 32) | q0.x);
	r.base = ((unsigned long long)q0.w << 32) | q0.z;
	r.bytes = q1.x;
	r.misc = q1.y;
	r.misc2 = q1.z;
	r.vec_in_chunk = q1.w;
	r.aux0 = (const uint8_t*)(((unsigned long long)q2.y << 32) | q2.x);
	r.aux1 = (const uint8_t*)(((unsigned long long)q2.w << 32) | q2.z);
	r.ch = (const FlsChunkDev*)(((unsigned long long)q3.y << 32) | q3.x);
	r.aux_off = q3.z;
	r.aux_n = q3.w;
	return r;
}

// Bytes of a vector's blob: packed part + exception tail (fls_format.h).  T = lane width of the packed data, VT = value width.
template <int T, int VT>
__device__ __forceinline__ uint32_t blob_bytes(uint32_t enc, uint32_t w, uint32_t w2, uint32_t n_exc) {
	switch (enc) {
	case FLS_ENC_UNCOMPRESSED: return 128u * T;
	case FLS_ENC_CONSTANT:
	case FLS_ENC_CROSS_RLE: return 0u;
	case FLS_ENC_DELTA:
	case FLS_ENC_RLE: return 128u + 128u * w;
	case FLS_ENC_DICT: return 128u * w;
	case FLS_ENC_ALP_RD: return 128u * (w + w2) + FLS_EXC_TAIL_BYTES(n_exc, VT / 8u);
	default: return 128u * w + FLS_EXC_TAIL_BYTES(n_exc, VT / 8u); // FFOR, SLPATCH, FREQUENCY, ALP
	}
}

// Lane-parallel descriptor fetch for vector v (v >= total or not of this kernel's family -> bytes = 0, ch = null).
template <int T, int VT>
__device__ __forceinline__ VecDesc fetch_desc(const FlsLaunch& L, uint32_t v) {
	VecDesc d;
	d.blob = nullptr;
	d.ch = nullptr;
	d.base = 0;
	d.aux_off = d.bytes = d.misc = d.misc2 = d.vec_in_chunk = d.aux_n = 0;
	d.aux0 = d.aux1 = nullptr;
	if (v < L.total_vecs) {
		if (L.vec_map != nullptr) {
			v = __ldg(L.vec_map + v);
		}
		const uint32_t r = v / L.vec_per_rg, vi = v - r * L.vec_per_rg;
		const FlsChunkDev* ch = L.chunks + (size_t)r * L.chunk_stride;
		const uint32_t enc = ch->encoding;
		if (vi < ch->n_vec && ((L.family_mask >> enc) & 1u) && (L.idx_bits == 0 || ch->idx_bits == L.idx_bits) &&
		    (L.rg_keep == nullptr || L.rg_keep[r])) {
			const uint4* mp = reinterpret_cast<const uint4*>(ch->meta + vi);
			const uint4 m0 = __ldg(mp), m1 = __ldg(mp + 1);
			// fls_vec_meta: base(8) data_off16(4) aux_off(4) | aux_cnt(2) width(1) width2(1) e(1) f(1) n_valid(2) reserved(8)
			d.base = ((uint64_t)m0.y << 32) | m0.x;
			d.aux_off = m0.w;
			const uint32_t aux_cnt = m1.x & 0xffffu, w = (m1.x >> 16) & 0xffu, w2 = m1.x >> 24;
			d.misc = aux_cnt | (w << 16) | (enc << 24);
			d.misc2 = w2 | ((m1.y & 0xffffu) << 8);
			d.vec_in_chunk = vi | (m1.y & 0xffff0000u);
			d.bytes = blob_bytes<T, VT>(enc, w, w2, aux_cnt);
			d.blob = ch->data + (size_t)m0.z * 16;
			d.ch = ch;
			d.aux0 = ch->aux0;
			d.aux1 = ch->aux1;
			d.aux_n = ch->aux_count;
		}
	}
	return d;
}

template <int T>
__device__ __forceinline__ void load_elem(const uint8_t* base, uint32_t idx, uint32_t& lo, uint32_t& hi) {
	hi = 0;
	if constexpr (T == 8) {
		lo = __ldg(base + idx);
	} else if constexpr (T == 16) {
		lo = __ldg(reinterpret_cast<const uint16_t*>(base) + idx);
	} else if constexpr (T == 32) {
		lo = __ldg(reinterpret_cast<const uint32_t*>(base) + idx);
	} else {
		const uint2 q = __ldg(reinterpret_cast<const uint2*>(base) + idx);
		lo = q.x;
		hi = q.y;
	}
}

// out[pos[k]] = exception[k]  (alp.hpp:401-404; slpatch / frequency patches).  The exception tail was staged into the
// vector's shared-memory slot by the same bulk copy as the packed data, so patching costs no global load; the scattered
// stores land after the vector's own stores (ordered by __syncwarp).
template <int T, int OB>
__device__ __forceinline__ void patch_exceptions(uint8_t* out_vec, const VecDesc& d, int lane, uint32_t conv, const uint8_t* tail) {
	const uint32_t n = d.misc & 0xffffu;
	if (n == 0) {
		return;
	}
	__syncwarp();
	const uint16_t* pos = reinterpret_cast<const uint16_t*>(tail + FLS_PAD16(n * (T / 8u)));
#pragma unroll 1
	for (uint32_t k = lane; k < n; k += 32) {
		uint32_t lo, hi = 0;
		if constexpr (T == 8) {
			lo = tail[k];
		} else if constexpr (T == 16) {
			lo = reinterpret_cast<const uint16_t*>(tail)[k];
		} else if constexpr (T == 32) {
			lo = reinterpret_cast<const uint32_t*>(tail)[k];
		} else {
			const uint2 q = reinterpret_cast<const uint2*>(tail)[k];
			lo = q.x;
			hi = q.y;
		}
		store_one<T, OB>(out_vec, pos[k], lo, hi, conv);
	}
}

// logical start of the chain that FastLanes lane `lane` decodes in a DELTA vector
__device__ __forceinline__ int chain_l0(int lane) {
	return (lane & 15) * 64 + fl_order((lane >> 4) & 7) * 8;
}

template <int T, int OB>
__device__ __forceinline__ void fill_constant(uint8_t* out_vec, int lane, uint32_t lo, uint32_t hi, uint32_t conv) {
	convert_val<T, OB>(lo, hi, conv);
	uint32_t w[4];
	if constexpr (OB == 8) {
		w[0] = w[2] = lo;
		w[1] = w[3] = hi;
	} else if constexpr (OB == 4) {
		w[0] = w[1] = w[2] = w[3] = lo;
	} else if constexpr (OB == 2) {
		w[0] = w[1] = w[2] = w[3] = (lo & 0xffffu) * 0x00010001u;
	} else {
		w[0] = w[1] = w[2] = w[3] = (lo & 0xffu) * 0x01010101u;
	}
#pragma unroll
	for (int it = 0; it < OB * 2; ++it) { // 1024*OB bytes / 16 / 32
		st_cs_v4(out_vec + (size_t)(it * 32 + lane) * 16, w[0], w[1], w[2], w[3]);
	}
}

// dec_cross_rle_opr: starts[] is the exclusive prefix sum of the run lengths (built at stage time).
template <int T, int OB>
__device__ __forceinline__ void decode_cross_rle(uint8_t* out_vec, const VecDesc& d, int lane, uint32_t conv) {
	const uint32_t* starts = reinterpret_cast<const uint32_t*>(d.aux1);
	const uint32_t n_runs = d.aux_n;
	const uint32_t g0 = (d.vec_in_chunk & 0xffffu) * 1024u;
	// first run of this vector (uniform binary search): largest r with starts[r] <= g0
	uint32_t lo_r = 0, hi_r = n_runs;
	while (hi_r - lo_r > 1) {
		const uint32_t mid = (lo_r + hi_r) >> 1;
		if (__ldg(starts + mid) <= g0) {
			lo_r = mid;
		} else {
			hi_r = mid;
		}
	}
	const uint32_t r_first = lo_r;
#pragma unroll 1
	for (int it = 0; it < 32; ++it) {
		const uint32_t gi = g0 + it * 32 + lane;
		uint32_t a = r_first, b = min(n_runs, r_first + 1025u);
		while (b - a > 1) {
			const uint32_t mid = (a + b) >> 1;
			if (__ldg(starts + mid) <= gi) {
				a = mid;
			} else {
				b = mid;
			}
		}
		uint32_t lo, hi;
		load_elem<T>(d.aux0, a, lo, hi);
		store_one<T, OB>(out_vec, it * 32 + lane, lo, hi, conv);
	}
}

// ---- fused aggregates (SUM as 128-bit, COUNT, MIN, MAX; reference semantics: DuckDB SUM(BIGINT) -> HUGEINT,
// SUM(DECIMAL(p,s)) -> DECIMAL(38,s), duckdb/extension/core_functions/aggregate/distributive/sum.cpp:161-210) -------------
struct AggPartial {
	unsigned long long lo = 0;   // low 64 bits of the running sum
	long long hi = 0;            // carries out of `lo` minus the number of negative addends
	unsigned long long cnt = 0;
	double fs = 0.0;

	// Fold the first n_valid values of the natural-order tile (VT-bit elements) into the partials.
	template <int VT>
	__device__ __forceinline__ void fold(const uint8_t* tile, int lane, uint32_t n_valid, uint32_t is_signed, uint32_t is_float) {
#pragma unroll 4
		for (uint32_t i = lane; i < 1024u; i += 32u) {
			if (i >= n_valid) {
				break;
			}
			unsigned long long u;
			if constexpr (VT == 8) {
				u = tile[i];
				u = is_signed ? (unsigned long long)(long long)(int8_t)u : u;
			} else if constexpr (VT == 16) {
				u = reinterpret_cast<const uint16_t*>(tile)[i];
				u = is_signed ? (unsigned long long)(long long)(int16_t)u : u;
			} else if constexpr (VT == 32) {
				const uint32_t w = reinterpret_cast<const uint32_t*>(tile)[i];
				if (is_float) {
					fs += (double)__uint_as_float(w);
					continue;
				}
				u = is_signed ? (unsigned long long)(long long)(int32_t)w : w;
			} else {
				u = reinterpret_cast<const unsigned long long*>(tile)[i];
				if (is_float) {
					fs += __longlong_as_double((long long)u);
					continue;
				}
			}
			lo += u;
			hi += (lo < u) ? 1 : 0;
			hi -= (is_signed && (long long)u < 0) ? 1 : 0;
		}
		cnt += n_valid > (uint32_t)lane ? (n_valid - lane + 31u) / 32u : 0u;
	}

	// Filter pushdown: value i of the natural-order tile as the 64-bit comparison domain of its column.
	template <int VT>
	static __device__ __forceinline__ unsigned long long tile_value(const uint8_t* tile, uint32_t i, uint32_t is_signed) {
		if constexpr (VT == 8) {
			const unsigned long long u = tile[i];
			return is_signed ? (unsigned long long)(long long)(int8_t)u : u;
		} else if constexpr (VT == 16) {
			const unsigned long long u = reinterpret_cast<const uint16_t*>(tile)[i];
			return is_signed ? (unsigned long long)(long long)(int16_t)u : u;
		} else if constexpr (VT == 32) {
			const uint32_t w = reinterpret_cast<const uint32_t*>(tile)[i];
			return is_signed ? (unsigned long long)(long long)(int32_t)w : w;
		} else {
			return reinterpret_cast<const unsigned long long*>(tile)[i];
		}
	}

	// DuckDB comparison semantics (ConstantFilter, duckdb/planner/filter/constant_filter.hpp): integers in their own
	// signedness; floating point with NaN equal to itself and greater than everything else.  A row's order against the
	// constant is 0 / 1 / 2 for < / == / >; a comparison operator is a 3-bit truth table indexed by that.
	static __device__ __forceinline__ uint32_t cmp_truth(uint32_t op) { // bit 0: passes when <, bit 1: when ==, bit 2: when >
		switch (op) {
		case 0: return 2u;      // EQ
		case 1: return 5u;      // NE
		case 2: return 1u;      // LT
		case 3: return 3u;      // LE
		case 4: return 4u;      // GT
		case 5: return 6u;      // GE
		default: return 7u;     // no second comparison
		}
	}

	// mask words of one vector: bit `lane` of word `it` = row it*32 + lane passes (rows >= n_valid never pass).
	// KIND 0: unsigned integers, 1: signed integers, 2: floating point - hoisted out of the row loop (the per-row version
	// with run-time type switches cost ~2300 instructions per vector, this one ~600).
	template <int VT, int KIND>
	static __device__ __forceinline__ uint32_t mask_rows(const uint8_t* tile, int lane, uint32_t n_valid, uint32_t t1, unsigned long long c,
	                                                     uint32_t t2, unsigned long long c2) {
		uint32_t mine = 0;
		// constants in the comparison domain of the loop
		double fc = 0.0, fc2 = 0.0;
		if constexpr (KIND == 2) {
			fc = VT == 32 ? (double)__uint_as_float((uint32_t)c) : __longlong_as_double((long long)c);
			fc2 = VT == 32 ? (double)__uint_as_float((uint32_t)c2) : __longlong_as_double((long long)c2);
		}
		const bool cn = fc != fc, cn2 = fc2 != fc2;
#pragma unroll 8
		for (uint32_t it = 0; it < 32u; ++it) {
			const uint32_t i = it * 32u + lane;
			uint32_t o1, o2; // 0: v < c, 1: v == c, 2: v > c
			if constexpr (KIND == 2) {
				double a;
				if constexpr (VT == 32) {
					a = (double)__uint_as_float(reinterpret_cast<const uint32_t*>(tile)[i]);
				} else {
					a = __longlong_as_double((long long)reinterpret_cast<const unsigned long long*>(tile)[i]);
				}
				const bool an = a != a; // DuckDB: NaN == NaN, NaN greater than everything else
				o1 = an ? (cn ? 1u : 2u) : (cn ? 0u : (a < fc ? 0u : (a == fc ? 1u : 2u)));
				o2 = an ? (cn2 ? 1u : 2u) : (cn2 ? 0u : (a < fc2 ? 0u : (a == fc2 ? 1u : 2u)));
			} else if constexpr (KIND == 1) {
				long long v;
				if constexpr (VT == 8) {
					v = (int8_t)tile[i];
				} else if constexpr (VT == 16) {
					v = reinterpret_cast<const int16_t*>(tile)[i];
				} else if constexpr (VT == 32) {
					v = reinterpret_cast<const int32_t*>(tile)[i];
				} else {
					v = reinterpret_cast<const long long*>(tile)[i];
				}
				o1 = v < (long long)c ? 0u : (v == (long long)c ? 1u : 2u);
				o2 = v < (long long)c2 ? 0u : (v == (long long)c2 ? 1u : 2u);
			} else {
				unsigned long long v;
				if constexpr (VT == 8) {
					v = tile[i];
				} else if constexpr (VT == 16) {
					v = reinterpret_cast<const uint16_t*>(tile)[i];
				} else if constexpr (VT == 32) {
					v = reinterpret_cast<const uint32_t*>(tile)[i];
				} else {
					v = reinterpret_cast<const unsigned long long*>(tile)[i];
				}
				o1 = v < c ? 0u : (v == c ? 1u : 2u);
				o2 = v < c2 ? 0u : (v == c2 ? 1u : 2u);
			}
			const bool pass = i < n_valid && ((t1 >> o1) & (t2 >> o2) & 1u);
			const uint32_t word = __ballot_sync(0xffffffffu, pass);
			if ((uint32_t)lane == it) {
				mine = word;
			}
		}
		return mine;
	}

	template <int VT>
	static __device__ __forceinline__ void build_mask(const uint8_t* tile, int lane, uint32_t n_valid, uint32_t is_signed, uint32_t is_float,
	                                                  uint32_t op, unsigned long long c, uint32_t op2, unsigned long long c2, bool and_into,
	                                                  uint32_t* mask_vec) {
		const uint32_t t1 = cmp_truth(op), t2 = cmp_truth(op2); // t2 = 7 (always true) without a second comparison
		uint32_t mine;
		if (VT >= 32 && is_float) {
			mine = mask_rows<VT, (VT >= 32 ? 2 : 0)>(tile, lane, n_valid, t1, c, t2, c2);
		} else if (is_signed) {
			mine = mask_rows<VT, 1>(tile, lane, n_valid, t1, c, t2, c2);
		} else {
			mine = mask_rows<VT, 0>(tile, lane, n_valid, t1, c, t2, c2);
		}
		mask_vec[lane] = and_into ? (mask_vec[lane] & mine) : mine;
	}

	// SUM / COUNT over the rows of the tile whose mask bit is set
	template <int VT>
	__device__ __forceinline__ void fold_masked(const uint8_t* tile, int lane, uint32_t n_valid, uint32_t is_signed, uint32_t is_float,
	                                            const uint32_t* mask_vec) {
		const uint32_t my_word = __ldg(mask_vec + lane);
		if (__ballot_sync(0xffffffffu, my_word != 0u) == 0u) {
			return;
		}
		(void)n_valid; // padding rows never have their bit set
		if (VT >= 32 && is_float) {
			double f = 0.0;
			uint32_t n = 0;
#pragma unroll 8
			for (uint32_t it = 0; it < 32u; ++it) {
				const uint32_t word = __shfl_sync(0xffffffffu, my_word, it);
				const uint32_t i = it * 32u + lane;
				if ((word >> lane) & 1u) {
					++n;
					if constexpr (VT == 32) {
						f += (double)__uint_as_float(reinterpret_cast<const uint32_t*>(tile)[i]);
					} else if constexpr (VT == 64) {
						f += __longlong_as_double((long long)reinterpret_cast<const unsigned long long*>(tile)[i]);
					}
				}
			}
			fs += f;
			cnt += n;
			return;
		}
		// integers: values of <= 32 bits cannot overflow a 64-bit partial over 32 rows; 64-bit values carry into `hi`
		unsigned long long s = 0;
		long long h = 0;
		uint32_t n = 0;
#pragma unroll 8
		for (uint32_t it = 0; it < 32u; ++it) {
			const uint32_t word = __shfl_sync(0xffffffffu, my_word, it);
			const uint32_t i = it * 32u + lane;
			const bool on = (word >> lane) & 1u;
			n += on ? 1u : 0u;
			if constexpr (VT == 64) {
				const unsigned long long u = on ? reinterpret_cast<const unsigned long long*>(tile)[i] : 0ull;
				s += u;
				h += (s < u) ? 1 : 0;
				h -= (is_signed && (long long)u < 0) ? 1 : 0;
			} else {
				const unsigned long long u = on ? tile_value<VT>(tile, i, is_signed) : 0ull; // sign-extended when signed
				s += u;
			}
		}
		cnt += n;
		if constexpr (VT == 64) {
			const unsigned long long before = lo;
			lo += s;
			hi += h + ((lo < before) ? 1 : 0);
		} else {
			add_u128(s, (is_signed && (long long)s < 0) ? -1ll : 0ll);
		}
	}

	// ---- register fast path: the 32 values a thread holds right after decoding (vector complete, no exceptions) -------
	// MIN/MAX are not computed here: flsgpu_aggregate_column folds the chunks' exact zone maps (fls_chunk_desc.min_val /
	// max_val, written from the data by the writer) into the result instead, like a scan answering from statistics.
	__device__ __forceinline__ void add_u128(unsigned long long u, long long sign_word) {
		const unsigned long long before = lo;
		lo += u;
		hi += sign_word + ((lo < before) ? 1 : 0);
	}
	// T <= 32: values are T-bit fields held in 32-bit registers; 32 of them cannot overflow a 64-bit partial
	template <int T>
	__device__ __forceinline__ void add32(const uint32_t (&v)[32], uint32_t is_signed, uint32_t is_float) {
		cnt += 32;
		if (T == 32 && is_float) {
			double f0 = 0.0, f1 = 0.0;
#pragma unroll
			for (int j = 0; j < 32; j += 2) {
				f0 += (double)__uint_as_float(v[j]);
				f1 += (double)__uint_as_float(v[j + 1]);
			}
			fs += f0 + f1;
			return;
		}
		if (is_signed) {
			long long s0 = 0, s1 = 0;
#pragma unroll
			for (int j = 0; j < 32; j += 2) {
				s0 += T == 32 ? (int)v[j] : T == 16 ? (int)(short)v[j] : (int)(signed char)v[j];
				s1 += T == 32 ? (int)v[j + 1] : T == 16 ? (int)(short)v[j + 1] : (int)(signed char)v[j + 1];
			}
			const long long s = s0 + s1;
			add_u128((unsigned long long)s, s < 0 ? -1ll : 0ll);
		} else {
			unsigned long long s0 = 0, s1 = 0;
#pragma unroll
			for (int j = 0; j < 32; j += 2) {
				s0 += v[j];
				s1 += v[j + 1];
			}
			add_u128(s0 + s1, 0);
		}
	}
	// 64-bit values as (lo, hi) halves: low and high words are summed separately (4 independent chains), negatives are
	// counted, and the vector's total enters the 128-bit partial once: sum = S_lo + (S_hi << 32) - (n_neg << 64).
	__device__ __forceinline__ void add64(const uint32_t (&l)[32], const uint32_t (&h)[32], uint32_t is_signed, uint32_t is_float) {
		cnt += 32;
		if (is_float) {
			double f0 = 0.0, f1 = 0.0;
#pragma unroll
			for (int j = 0; j < 32; j += 2) {
				f0 += __longlong_as_double((long long)(((unsigned long long)h[j] << 32) | l[j]));
				f1 += __longlong_as_double((long long)(((unsigned long long)h[j + 1] << 32) | l[j + 1]));
			}
			fs += f0 + f1;
			return;
		}
		unsigned long long sl0 = 0, sl1 = 0, sh0 = 0, sh1 = 0;
		uint32_t neg = 0;
#pragma unroll
		for (int j = 0; j < 32; j += 2) {
			sl0 += l[j];
			sl1 += l[j + 1];
			sh0 += h[j];
			sh1 += h[j + 1];
			neg += (h[j] >> 31) + (h[j + 1] >> 31);
		}
		const unsigned long long sl = sl0 + sl1, sh = sh0 + sh1; // each < 2^37
		add_u128(sl, 0);
		add_u128(sh << 32, (long long)(sh >> 32) - (is_signed ? (long long)neg : 0ll));
	}

	// Warp reduction, then one set of atomics per warp.
	__device__ __forceinline__ void flush(flsgpu_agg_acc* acc, int lane, uint32_t is_float) {
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			const unsigned long long olo = __shfl_down_sync(0xffffffffu, lo, o);
			const long long ohi = __shfl_down_sync(0xffffffffu, hi, o);
			lo += olo;
			hi += ohi + ((lo < olo) ? 1 : 0);
			cnt += __shfl_down_sync(0xffffffffu, cnt, o);
			fs += __shfl_down_sync(0xffffffffu, fs, o);
		}
		if (lane == 0 && cnt) {
			atomicAdd(&acc->count, cnt);
			if (is_float) {
				atomicAdd(&acc->fsum, fs);
			} else {
				const unsigned long long old = atomicAdd(&acc->sum_lo, lo);
				const unsigned long long carry = (old + lo < old) ? 1ull : 0ull;
				atomicAdd(&acc->sum_hi, (unsigned long long)hi + carry);
			}
		}
	}
};

} // namespace FLS_NS

// flsgpu_kernels_select.cu — selected-rows-only scans (sm_100a): the rows a filter mask keeps are packed together on the
// device, so that only they cross PCIe.
//
// The reference evaluates pushed-down filters after decode and hands DuckDB the whole chunk sliced by a selection vector
// (FilterExecutor::Apply, reference src/reader/filter_executor.cpp:19-60: chunk.Slice(sel, count)); every decoded value has
// been produced in host memory by then.  Here the decoded columns of a slab sit in HBM next to the slab's selection mask
// (flsgpu_filter_mask), and two small kernels compact them before the device-to-host copy:
//   fls_select_plan_kernel    per 1024-row vector: popcount of its 32 mask words, exclusive prefix sum over the slab's vectors
//                             (one CTA; a slab has a few thousand vectors) -> where each vector's selected rows go, and how many
//                             rows the slab selects
//   fls_select_compact_kernel one warp per vector: lane L owns mask word L, a warp scan of the 32 popcounts gives every word's
//                             offset inside the vector, then for each of the 32 row groups of 32 rows the lanes whose bit is
//                             set copy their row to out[base + word offset + rank of the bit]: reads are fully coalesced,
//                             writes land in one contiguous run per vector.  Element widths 1 / 2 / 4 / 8 / 16 bytes
//                             (16 = DuckDB string_t); IOTA: the row's index inside the range is written instead of a value
//                             (file_row_number of a compacted scan).
// Both kernels are bandwidth trivial next to the decode (they re-read what the decode just wrote, mostly from L2).
#include <cuda_runtime.h>

#include <stdint.h>

namespace {

constexpr int kPlanThreads = 1024;

__global__ void __launch_bounds__(kPlanThreads) fls_select_plan_kernel(const uint32_t* __restrict__ mask, uint32_t n_vec,
                                                                       uint32_t* __restrict__ vec_offsets /* n_vec + 1 */) {
	__shared__ uint32_t s_warp[kPlanThreads / 32];
	__shared__ uint32_t s_carry;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	if (threadIdx.x == 0) {
		s_carry = 0;
	}
	__syncthreads();
	for (uint32_t v0 = 0; v0 < n_vec; v0 += kPlanThreads) {
		const uint32_t v = v0 + threadIdx.x;
		uint32_t cnt = 0;
		if (v < n_vec) {
			const uint4* w = reinterpret_cast<const uint4*>(mask + (size_t)v * 32u);
#pragma unroll
			for (int k = 0; k < 8; ++k) {
				const uint4 q = __ldg(w + k);
				cnt += __popc(q.x) + __popc(q.y) + __popc(q.z) + __popc(q.w);
			}
		}
		uint32_t incl = cnt; // inclusive scan inside the warp
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
			if ((int)lane >= d) {
				incl += up;
			}
		}
		if (lane == 31) {
			s_warp[warp] = incl;
		}
		__syncthreads();
		if (warp == 0) {
			const uint32_t t = s_warp[lane];
			uint32_t winc = t;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t up = __shfl_up_sync(0xffffffffu, winc, d);
				if ((int)lane >= d) {
					winc += up;
				}
			}
			s_warp[lane] = winc - t; // exclusive offset of every warp
		}
		__syncthreads();
		const uint32_t carry = s_carry;
		if (v < n_vec) {
			vec_offsets[v] = carry + s_warp[warp] + incl - cnt;
		}
		__syncthreads();
		if (threadIdx.x == kPlanThreads - 1) {
			s_carry = carry + s_warp[warp] + incl;
		}
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		vec_offsets[n_vec] = s_carry;
	}
}

// dictionary codes: the codes of a row group count from its own dictionary; packed rows of several row groups share ONE
// dictionary (all keys of the range back to back), so every code is moved up by the number of keys in front of its row group
struct FlsKeyBase {
	const unsigned long long* key_prefix; // per chunk of the column (stride entries apart), or nullptr
	unsigned long long key_prefix0;       // keys in front of the range
	uint32_t stride, vec_per_rg, first_rg; // first_rg: row group of vector 0 inside the range
};

template <typename T, bool IOTA, bool CODES>
__global__ void __launch_bounds__(256) fls_select_compact_kernel(const T* __restrict__ in, const uint32_t* __restrict__ mask,
                                                                 const uint32_t* __restrict__ vec_offsets, T* __restrict__ out, uint32_t n_vec, uint32_t iota_base,
                                                                 const FlsKeyBase kb) {
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t v = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (v >= n_vec) {
		return;
	}
	const uint32_t word = __ldg(mask + (size_t)v * 32u + lane);
	const uint32_t cnt = __popc(word);
	uint32_t incl = cnt;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
		if ((int)lane >= d) {
			incl += up;
		}
	}
	const uint32_t word_off = incl - cnt;
	if (__shfl_sync(0xffffffffu, incl, 31) == 0) {
		return; // nothing selected in this vector
	}
	const uint32_t base = __ldg(vec_offsets + v);
	const T* src = in + (size_t)v * 1024u;
	const uint32_t below = (1u << lane) - 1u;
	uint32_t key_add = 0;
	if constexpr (CODES) {
		key_add = (uint32_t)(__ldg(kb.key_prefix + (size_t)(kb.first_rg + v / kb.vec_per_rg) * kb.stride) - kb.key_prefix0);
	}
#pragma unroll 4
	for (int it = 0; it < 32; ++it) {
		const uint32_t w = __shfl_sync(0xffffffffu, word, it);
		const uint32_t off = __shfl_sync(0xffffffffu, word_off, it);
		if ((w >> lane) & 1u) {
			const uint32_t pos = base + off + __popc(w & below);
			if constexpr (IOTA) {
				out[pos] = (T)(iota_base + v * 1024u + (uint32_t)it * 32u + lane);
			} else if constexpr (CODES) {
				out[pos] = src[it * 32 + lane] + key_add;
			} else {
				out[pos] = src[it * 32 + lane];
			}
		}
	}
}

} // namespace

cudaError_t fls_select_launch_plan(const uint32_t* mask, uint32_t n_vec, uint32_t* vec_offsets, cudaStream_t stream) {
	fls_select_plan_kernel<<<1, kPlanThreads, 0, stream>>>(mask, n_vec, vec_offsets);
	return cudaGetLastError();
}

// elem_bytes 1 / 2 / 4 / 8 / 16; in == nullptr: the rows' indexes inside the range (u32) are written instead of values
cudaError_t fls_select_launch_compact_codes(const uint32_t* in, const uint32_t* mask, const uint32_t* vec_offsets, uint32_t* out, uint32_t n_vec,
                                            const unsigned long long* key_prefix, unsigned long long key_prefix0, uint32_t stride,
                                            uint32_t vec_per_rg, uint32_t first_rg, cudaStream_t stream) {
	if (!n_vec) {
		return cudaSuccess;
	}
	const FlsKeyBase kb {key_prefix, key_prefix0, stride, vec_per_rg, first_rg};
	fls_select_compact_kernel<uint32_t, false, true><<<(n_vec + 7u) / 8u, 256, 0, stream>>>(in, mask, vec_offsets, out, n_vec, 0u, kb);
	return cudaGetLastError();
}

cudaError_t fls_select_launch_compact(const void* in, uint32_t elem_bytes, const uint32_t* mask, const uint32_t* vec_offsets, void* out,
                                      uint32_t n_vec, uint32_t iota_base, cudaStream_t stream) {
	const uint32_t grid = (n_vec + 7u) / 8u;
	if (!n_vec) {
		return cudaSuccess;
	}
	const FlsKeyBase kb {nullptr, 0ull, 0u, 1u, 0u};
	if (!in) {
		fls_select_compact_kernel<uint32_t, true, false><<<grid, 256, 0, stream>>>(nullptr, mask, vec_offsets, static_cast<uint32_t*>(out), n_vec, iota_base, kb);
		return cudaGetLastError();
	}
	switch (elem_bytes) {
	case 1:
		fls_select_compact_kernel<uint8_t, false, false><<<grid, 256, 0, stream>>>(static_cast<const uint8_t*>(in), mask, vec_offsets, static_cast<uint8_t*>(out), n_vec, 0u, kb);
		break;
	case 2:
		fls_select_compact_kernel<uint16_t, false, false><<<grid, 256, 0, stream>>>(static_cast<const uint16_t*>(in), mask, vec_offsets, static_cast<uint16_t*>(out), n_vec, 0u, kb);
		break;
	case 4:
		fls_select_compact_kernel<uint32_t, false, false><<<grid, 256, 0, stream>>>(static_cast<const uint32_t*>(in), mask, vec_offsets, static_cast<uint32_t*>(out), n_vec, 0u, kb);
		break;
	case 8:
		fls_select_compact_kernel<unsigned long long, false, false><<<grid, 256, 0, stream>>>(static_cast<const unsigned long long*>(in), mask, vec_offsets, static_cast<unsigned long long*>(out), n_vec, 0u, kb);
		break;
	case 16:
		fls_select_compact_kernel<uint4, false, false><<<grid, 256, 0, stream>>>(static_cast<const uint4*>(in), mask, vec_offsets, static_cast<uint4*>(out), n_vec, 0u, kb);
		break;
	default: return cudaErrorInvalidValue;
	}
	return cudaGetLastError();
}

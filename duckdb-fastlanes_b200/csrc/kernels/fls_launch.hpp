// fls_launch.hpp — kernel lookup shared by the kernel translation units and the host-side launcher.
// Each family is compiled in its own .cu (compile time); the launcher picks the instantiation by
// (lane width, value width, output bytes) and launches it through cudaLaunchKernel.
#pragma once
#include "fls_device.h"

// MODE 0 (T-lane encodings): T = column physical bits, OB = output bytes per value (>= T/8).
// fam: 0 = light (integer encodings), 1 = ALP, 2 = ALP_RD, 3 = DELTA of 64-bit columns (own instantiation)
const void* fls_kernel_main(int T, int OB, int fam);
// MODE 1 (DICT): IT = code lane bits (8/16/32), VT = key bits, OB = output bytes (>= VT/8).
const void* fls_kernel_dict(int IT, int VT, int OB);
// MODE 2 (RLE, u16 run index): VT = run value bits.
const void* fls_kernel_rle(int VT, int OB);
// Fused decode + aggregate (SUM/COUNT/MIN/MAX): same families, output replaced by a reduction.
const void* fls_kernel_agg(int mode, int T, int VT, int fam);

#ifdef __CUDACC__
#include <cuda_runtime.h>
#endif
#if defined(__CUDACC__) || defined(CUDART_VERSION)
cudaError_t fls_agg_finalize(flsgpu_agg_acc* acc, uint32_t is_signed, unsigned long long key_min, unsigned long long key_max,
                             cudaStream_t stream);
cudaError_t fls_agg_reset(flsgpu_agg_acc* acc, cudaStream_t stream);
// String columns (flsgpu_kernels_str.cu): heap fill (plain copy / parallel FSST decode), string_t rows or u32 codes of
// whole vectors, string_t of the keys of one dictionary chunk.
cudaError_t fls_str_launch_heap(FlsStrLaunch L, cudaStream_t stream);
cudaError_t fls_str_launch_rows(FlsStrLaunch L, cudaStream_t stream);
cudaError_t fls_str_launch_keys(FlsStrLaunch L, uint32_t max_keys_per_chunk, cudaStream_t stream);
// Constant comparison over decoded string_t rows -> selection mask words (one warp per vector).
cudaError_t fls_str_launch_filter(FlsStrFilterLaunch L, cudaStream_t stream);
// The same comparison over the keys of dictionary chunks -> one flag byte per key (then fls_str_launch_rows in filter mode).
cudaError_t fls_str_launch_key_flags(FlsStrFilterLaunch L, cudaStream_t stream);
// Selected-rows-only scans (flsgpu_kernels_select.cu): where every vector's selected rows go (exclusive prefix sums of the mask
// popcounts over n_vec vectors, total at [n_vec]), then the compaction of one decoded column (in == nullptr: row indexes).
cudaError_t fls_select_launch_plan(const uint32_t* mask, uint32_t n_vec, uint32_t* vec_offsets, cudaStream_t stream);
cudaError_t fls_select_launch_compact(const void* in, uint32_t elem_bytes, const uint32_t* mask, const uint32_t* vec_offsets, void* out,
                                      uint32_t n_vec, uint32_t iota_base, cudaStream_t stream);
// dictionary codes: every code moved up by the keys in front of its row group (key_prefix as in FlsStrLaunch)
cudaError_t fls_select_launch_compact_codes(const uint32_t* in, const uint32_t* mask, const uint32_t* vec_offsets, uint32_t* out, uint32_t n_vec,
                                            const unsigned long long* key_prefix, unsigned long long key_prefix0, uint32_t stride,
                                            uint32_t vec_per_rg, uint32_t first_rg, cudaStream_t stream);
#endif

constexpr int kFlsWarpsPerCta = 8;

// flsgpu_kernels_agg.cu — fused decode + aggregate kernels (SUM/COUNT/MIN/MAX over a scan): the same
// decode bodies as the materialising kernels, re-instantiated with their stores aimed at the warp's shared-memory tile
// (FLS_SINK_SHARED) and followed by a fold of the tile into per-thread partials.  No decoded byte reaches HBM.
#define FLS_NS flsk_agg
#define FLS_SINK_SHARED 1
#include "fls_decode_main.cuh"
#include "fls_launch.hpp"

using namespace flsk_agg;

const void* fls_kernel_agg(int mode, int T, int VT, int fam) {
	if (mode == 0 && fam == 0) {
#define FLS_KL(Tv) if (T == Tv) return reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, Tv / 8, true, kFamLight>);
		FLS_KL(8) FLS_KL(16) FLS_KL(32) FLS_KL(64)
#undef FLS_KL
	}
	if (mode == 0 && fam != 0) {
#define FLS_KL(Tv) if (T == Tv) return fam == 1 ? reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, Tv / 8, true, kFamAlp>) : reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, Tv / 8, true, kFamAlpRd>);
		FLS_KL(32) FLS_KL(64)
#undef FLS_KL
	}
#define FLS_K(M, Tv, VTv) \
	if (mode == M && T == Tv && VT == VTv) return reinterpret_cast<const void*>(&fls_decode_kernel<M, Tv, VTv, VTv / 8, true>);
	FLS_K(1, 8, 8) FLS_K(1, 8, 16) FLS_K(1, 8, 32) FLS_K(1, 8, 64)
	FLS_K(1, 16, 8) FLS_K(1, 16, 16) FLS_K(1, 16, 32) FLS_K(1, 16, 64)
	FLS_K(1, 32, 8) FLS_K(1, 32, 16) FLS_K(1, 32, 32) FLS_K(1, 32, 64)
	FLS_K(2, 16, 8) FLS_K(2, 16, 16) FLS_K(2, 16, 32) FLS_K(2, 16, 64)
#undef FLS_K
	return nullptr;
}

// min_val / max_val of the public result are written from the order keys after every aggregate launch.
__global__ void fls_agg_finalize_kernel(flsgpu_agg_acc* acc, uint32_t is_signed, unsigned long long key_min, unsigned long long key_max) {
	const unsigned long long flip = is_signed ? 0x8000000000000000ull : 0ull;
	// merge the zone-map extremes of the chunks this call covered (order keys: value ^ sign bit)
	acc->key_min = key_min < acc->key_min ? key_min : acc->key_min;
	acc->key_max = key_max > acc->key_max ? key_max : acc->key_max;
	if (acc->count && acc->key_min <= acc->key_max) {
		acc->min_val = (long long)(acc->key_min ^ flip);
		acc->max_val = (long long)(acc->key_max ^ flip);
	}
}

__global__ void fls_agg_reset_kernel(flsgpu_agg_acc* acc) {
	acc->sum_lo = acc->sum_hi = acc->count = 0;
	acc->min_val = acc->max_val = 0;
	acc->fsum = 0.0;
	acc->key_min = ~0ull;
	acc->key_max = 0ull;
}

cudaError_t fls_agg_finalize(flsgpu_agg_acc* acc, uint32_t is_signed, unsigned long long key_min, unsigned long long key_max,
                             cudaStream_t stream) {
	fls_agg_finalize_kernel<<<1, 1, 0, stream>>>(acc, is_signed, key_min, key_max);
	return cudaGetLastError();
}

cudaError_t fls_agg_reset(flsgpu_agg_acc* acc, cudaStream_t stream) {
	fls_agg_reset_kernel<<<1, 1, 0, stream>>>(acc);
	return cudaGetLastError();
}

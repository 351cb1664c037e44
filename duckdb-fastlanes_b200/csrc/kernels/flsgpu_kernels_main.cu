// flsgpu_kernels_main.cu — instantiations of the main decode kernel family (UNCOMPRESSED, CONSTANT, FFOR, DELTA, ALP,
// ALP_RD, CROSS_RLE, FREQUENCY, SLPATCH) for every (physical width, output width) pair the materializer allows
// (reference fls_type_resolver.hpp:42-83,120-131: same width copy or widening static_cast).
#include "fls_decode_main.cuh"
#include "fls_launch.hpp"

using namespace flsk;

const void* fls_kernel_main(int T, int OB) {
#define FLS_K(Tv, OBv) \
	if (T == Tv && OB == OBv) return reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, OBv>);
	FLS_K(8, 1) FLS_K(8, 2) FLS_K(8, 4) FLS_K(8, 8)
	FLS_K(16, 2) FLS_K(16, 4) FLS_K(16, 8)
	FLS_K(32, 4) FLS_K(32, 8)
	FLS_K(64, 8)
#undef FLS_K
	return nullptr;
}

// flsgpu_kernels_rle.cu — RLE decode kernels (dec_rle_map_opr<KEY, u16>, reference rle_map_kernel.hpp:7-24).
#include "fls_decode_main.cuh"
#include "fls_launch.hpp"

using namespace flsk;

const void* fls_kernel_rle(int VT, int OB) {
#define FLS_K(VTv, OBv) \
	if (VT == VTv && OB == OBv) return reinterpret_cast<const void*>(&fls_decode_kernel<2, 16, VTv, OBv>);
	FLS_K(8, 1) FLS_K(8, 2) FLS_K(8, 4) FLS_K(8, 8)
	FLS_K(16, 2) FLS_K(16, 4) FLS_K(16, 8)
	FLS_K(32, 4) FLS_K(32, 8)
	FLS_K(64, 8)
#undef FLS_K
	return nullptr;
}

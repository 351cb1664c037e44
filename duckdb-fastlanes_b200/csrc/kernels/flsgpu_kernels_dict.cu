// flsgpu_kernels_dict.cu — DICT decode kernels (dec_dict_opr<KEY, IDX>, reference dictionary_kernel.hpp:11-50).
#include "fls_decode_main.cuh"
#include "fls_launch.hpp"

using namespace flsk;

const void* fls_kernel_dict(int IT, int VT, int OB) {
#define FLS_K(ITv, VTv, OBv) \
	if (IT == ITv && VT == VTv && OB == OBv) return reinterpret_cast<const void*>(&fls_decode_kernel<1, ITv, VTv, OBv>);
#define FLS_KI(ITv) \
	FLS_K(ITv, 8, 1) FLS_K(ITv, 8, 2) FLS_K(ITv, 8, 4) FLS_K(ITv, 8, 8) \
	FLS_K(ITv, 16, 2) FLS_K(ITv, 16, 4) FLS_K(ITv, 16, 8) \
	FLS_K(ITv, 32, 4) FLS_K(ITv, 32, 8) \
	FLS_K(ITv, 64, 8)
	FLS_KI(8) FLS_KI(16) FLS_KI(32)
#undef FLS_KI
#undef FLS_K
	return nullptr;
}

// flsgpu_kernels_main.cu — instantiations of the main decode kernel family (UNCOMPRESSED, CONSTANT, FFOR, DELTA,
// CROSS_RLE, FREQUENCY, SLPATCH = "light"; ALP; ALP_RD) for every (physical width, output width) pair the materializer
// allows (reference fls_type_resolver.hpp:42-83,120-131: same width copy or widening static_cast).
#include "fls_decode_main.cuh"
#include "fls_launch.hpp"

using namespace flsk;

// fam: 0 = light, 1 = ALP, 2 = ALP_RD, 3 = DELTA (T = 64 only; narrower types keep DELTA in the light kernel),
// 4 = FFOR + DELTA of 8-bit columns written as 8 bits (SWAR)
const void* fls_kernel_main(int T, int OB, int fam) {
	if (fam == 4) {
		return T == 8 && OB == 1 ? reinterpret_cast<const void*>(&fls_decode_kernel<0, 8, 8, 1, false, kFamSwar8>) : nullptr;
	}
#define FLS_K(Tv, OBv) \
	if (fam == 0 && T == Tv && OB == OBv) return reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, OBv, false, kFamLight>);
	FLS_K(8, 1) FLS_K(8, 2) FLS_K(8, 4) FLS_K(8, 8)
	FLS_K(16, 2) FLS_K(16, 4) FLS_K(16, 8)
	FLS_K(32, 4) FLS_K(32, 8)
#undef FLS_K
	if (fam == 0 && T == 64 && OB == 8) return reinterpret_cast<const void*>(&fls_decode_kernel<0, 64, 64, 8, false, kFamLight64>);
	if (fam == 3 && T == 64 && OB == 8) return reinterpret_cast<const void*>(&fls_decode_kernel<0, 64, 64, 8, false, kFamDelta>);
#define FLS_K(Tv, OBv) \
	if (fam == 1 && T == Tv && OB == OBv) return reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, OBv, false, kFamAlp>); \
	if (fam == 2 && T == Tv && OB == OBv) return reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, OBv, false, kFamAlpRd>);
	FLS_K(32, 4) FLS_K(32, 8) FLS_K(64, 8)
#undef FLS_K
	// fam 5: ALP and ALP_RD chunks of one column in one launch (columns that mix the two)
#define FLS_K(Tv, OBv) \
	if (fam == 5 && T == Tv && OB == OBv) return reinterpret_cast<const void*>(&fls_decode_kernel<0, Tv, Tv, OBv, false, kFamAlp | kFamAlpRd>);
	FLS_K(32, 4) FLS_K(32, 8) // (T = 64 fused would spill ~600 bytes: the two 64-bit float paths stay separate launches)
#undef FLS_K
	return nullptr;
}

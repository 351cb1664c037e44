// fls_ptx.cuh — sm_100a PTX wrappers: mbarrier + 1-D TMA bulk copy (cp.async.bulk), cache-hinted stores.
#pragma once
#include <stdint.h>

#ifndef FLS_NS
#define FLS_NS flsk // see fls_unpack.cuh
#endif

namespace FLS_NS {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
	return (uint32_t)__cvta_generic_to_shared(p);
}

// Programmatic dependent launch (sm_90+): launch_dependents lets the next kernel of the stream (if it was launched with
// programmatic stream serialization) begin once every CTA of this grid has executed it or exited; wait blocks until the
// preceding grid has completed and its memory is visible.  Both are no-ops for ordinary launches.
__device__ __forceinline__ void griddep_launch_dependents() {
	asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void griddep_wait() {
	asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

// make mbarrier.init visible to the async (TMA) proxy
__device__ __forceinline__ void fence_mbar_init() {
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void fence_proxy_async() {
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
	asm volatile("{\n\t"
	             ".reg .pred p;\n\t"
	             "WAIT_%=:\n\t"
	             "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
	             "@p bra DONE_%=;\n\t"
	             "bra WAIT_%=;\n\t"
	             "DONE_%=:\n\t"
	             "}" ::"r"(smem_u32(bar)),
	             "r"(parity)
	             : "memory");
}

// TMA bulk copy global -> shared (SASS: UBLKCP), completion signalled on `bar` as transaction bytes.
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
	                 smem_u32(smem_dst)),
	             "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
	             : "memory");
}

// TMA prefetch of a byte range into L2 (no shared-memory destination, nothing to wait for).
__device__ __forceinline__ void tma_prefetch_l2(const void* gmem_src, uint32_t bytes) {
	asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem_src), "r"(bytes) : "memory");
}

#if defined(FLS_SINK_SHARED)
// Aggregate kernels: the "output vector" is the warp's shared-memory tile (generic pointer into shared memory).
__device__ __forceinline__ void st_cs_u32(void* p, uint32_t a) {
	asm volatile("st.shared.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(a) : "memory");
}
__device__ __forceinline__ void st_cs_v2(void* p, uint32_t a, uint32_t b) {
	asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(smem_u32(p)), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void st_cs_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
	asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_cs_u16(void* p, uint32_t a) {
	asm volatile("st.shared.u16 [%0], %1;" ::"r"(smem_u32(p)), "h"((uint16_t)a) : "memory");
}
__device__ __forceinline__ void st_cs_u8(void* p, uint32_t a) {
	*reinterpret_cast<uint8_t*>(p) = (uint8_t)a;
}
#else
// streaming stores: decoded output is written once and not re-read by this kernel
__device__ __forceinline__ void st_cs_u32(void* p, uint32_t a) {
	asm volatile("st.global.cs.u32 [%0], %1;" ::"l"(p), "r"(a) : "memory");
}
__device__ __forceinline__ void st_cs_v2(void* p, uint32_t a, uint32_t b) {
	asm volatile("st.global.cs.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void st_cs_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
	asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_cs_u16(void* p, uint32_t a) {
	asm volatile("st.global.cs.u16 [%0], %1;" ::"l"(p), "h"((uint16_t)a) : "memory");
}
__device__ __forceinline__ void st_cs_u8(void* p, uint32_t a) {
	*reinterpret_cast<uint8_t*>(p) = (uint8_t)a;
}

#endif

} // namespace FLS_NS

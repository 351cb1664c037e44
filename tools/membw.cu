// membw.cu — what a write-dominated streaming kernel can reach on this GPU (context for the decode roofline: a DELTA/FFOR
// scan of a 64-bit column writes ~60x more bytes than it reads, while MEASURED_PEAKS.json's hbm_gbs is a 1:1 copy).
// Variants: register stores (st.global / st.global.cs, 16 B per thread), TMA bulk stores from shared memory
// (cp.async.bulk.global.shared::cta), a 1:1 copy, and a read-only sum.  Build: make -C tools membw.  Run: tools/membw [MB]
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                                    \
	do {                                                                                         \
		cudaError_t e_ = (x);                                                                    \
		if (e_ != cudaSuccess) {                                                                 \
			std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
			std::exit(1);                                                                        \
		}                                                                                        \
	} while (0)

template <int MODE> // 0: st.global, 1: st.global.cs, 2: st.global.wt? (not used)
__global__ void __launch_bounds__(256) k_write(uint4* out, size_t n16, uint32_t seed) {
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
		const uint32_t v = seed + (uint32_t)i;
		if (MODE == 0) {
			out[i] = make_uint4(v, v, v, v);
		} else {
			asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(out + i), "r"(v), "r"(v), "r"(v), "r"(v) : "memory");
		}
	}
}

// each warp writes contiguous 8 KiB tiles (like one decoded 64-bit vector), tiles dealt round-robin to warps
__global__ void __launch_bounds__(256) k_write_tiles(uint4* out, size_t n_tiles, uint32_t seed) {
	const int lane = threadIdx.x & 31;
	const size_t n_warps = (size_t)gridDim.x * (blockDim.x >> 5), gw = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	for (size_t t = gw; t < n_tiles; t += n_warps) {
		uint4* dst = out + t * 512;
#pragma unroll
		for (int it = 0; it < 16; ++it) {
			const uint32_t v = seed + it;
			asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst + it * 32 + lane), "r"(v), "r"(v), "r"(v), "r"(v) : "memory");
		}
	}
}

// 8 KiB tiles per warp + a small read per tile (rd_bytes from a side buffer, like a vector's packed blob): does a trickle
// of reads inside a write stream cost DRAM bus turnarounds?
__global__ void __launch_bounds__(256) k_write_tiles_rd(uint4* out, size_t n_tiles, const uint4* side, uint32_t rd16, uint32_t ahead,
                                                        uint32_t* sink) {
	const int lane = threadIdx.x & 31;
	const size_t n_warps = (size_t)gridDim.x * (blockDim.x >> 5), gw = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	uint32_t acc = 0;
	for (size_t t = gw; t < n_tiles; t += n_warps) {
		uint4 r = make_uint4(0, 0, 0, 0);
		if ((uint32_t)lane < rd16) {
			const size_t tt = t + ahead * n_warps < n_tiles ? t + ahead * n_warps : t;
			asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(side + tt * rd16 + lane));
		}
		acc ^= r.x ^ r.y ^ r.z ^ r.w;
		uint4* dst = out + t * 512;
#pragma unroll
		for (int it = 0; it < 16; ++it) {
			const uint32_t v = acc + it;
			asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst + it * 32 + lane), "r"(v), "r"(v), "r"(v), "r"(v) : "memory");
		}
	}
	if (acc == 0x12345678u) {
		*sink = acc;
	}
}

// same, but the reads of 32 tiles are issued together every 32nd tile (per-warp batching), or the whole side buffer is
// pulled into L2 by a prologue (mode 2) so that the per-tile reads never reach DRAM
__global__ void __launch_bounds__(256) k_write_tiles_rd_batched(uint4* out, size_t n_tiles, const uint4* side, uint32_t rd16, int mode,
                                                                uint32_t* sink) {
	const int lane = threadIdx.x & 31;
	const size_t n_warps = (size_t)gridDim.x * (blockDim.x >> 5), gw = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	uint32_t acc = 0;
	if (mode == 2) { // prologue: every warp prefetches its share of the side buffer into L2
		const size_t total16 = n_tiles * rd16;
		for (size_t i = (gw * 32 + lane) * 8; i < total16; i += n_warps * 32 * 8) { // 128 B per lane
			asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(side + i));
		}
	}
	size_t k = 0;
	for (size_t t = gw; t < n_tiles; t += n_warps, ++k) {
		if (mode == 1) {
			if ((k & 31) == 0) { // reads of the next 32 tiles of this warp, all at once
				for (uint32_t j = 0; j < 32; ++j) {
					const size_t tt = t + j * n_warps;
					if (tt < n_tiles && (uint32_t)lane < rd16) {
						uint4 r;
						asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(side + tt * rd16 + lane));
						acc ^= r.x ^ r.y ^ r.z ^ r.w;
					}
				}
			}
		} else if ((uint32_t)lane < rd16) {
			uint4 r;
			asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(side + t * rd16 + lane));
			acc ^= r.x ^ r.y ^ r.z ^ r.w;
		}
		uint4* dst = out + t * 512;
#pragma unroll
		for (int it = 0; it < 16; ++it) {
			const uint32_t v = acc + it;
			asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst + it * 32 + lane), "r"(v), "r"(v), "r"(v), "r"(v) : "memory");
		}
	}
	if (acc == 0x12345678u) {
		*sink = acc;
	}
}

// TMA bulk store: each warp fills an 8 KiB shared tile once, then streams it out tile after tile (two tiles in flight)
__global__ void __launch_bounds__(256) k_write_tma(uint8_t* out, size_t n_tiles, uint32_t tile_bytes, int depth) {
	extern __shared__ __align__(128) uint8_t smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint8_t* tile = smem + (size_t)warp * tile_bytes;
	for (uint32_t i = lane * 16; i < tile_bytes; i += 32 * 16) {
		*reinterpret_cast<uint4*>(tile + i) = make_uint4(i, i, i, i);
	}
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	__syncwarp();
	const size_t n_warps = (size_t)gridDim.x * (blockDim.x >> 5), gw = (size_t)blockIdx.x * (blockDim.x >> 5) + warp;
	for (size_t t = gw; t < n_tiles; t += n_warps) {
		if (lane == 0) {
			asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + t * tile_bytes),
			             "r"((uint32_t)__cvta_generic_to_shared(tile)), "r"(tile_bytes)
			             : "memory");
			asm volatile("cp.async.bulk.commit_group;" ::: "memory");
			if (depth == 1) {
				asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
			} else {
				asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
			}
		}
		__syncwarp();
	}
	if (lane == 0) {
		asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
	}
}

__global__ void __launch_bounds__(256) k_copy(const uint4* in, uint4* out, size_t n16) {
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
		uint4 v;
		asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(in + i));
		asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(out + i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
	}
}

__global__ void __launch_bounds__(256) k_read(const uint4* in, size_t n16, uint32_t* sink) {
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	uint32_t acc = 0;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
		uint4 v;
		asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(in + i));
		acc += v.x ^ v.y ^ v.z ^ v.w;
	}
	if (acc == 0x12345678u) {
		*sink = acc;
	}
}

template <class F>
static float timed(F f, int iters = 10) {
	for (int i = 0; i < 3; ++i) {
		f();
	}
	cudaEvent_t e0, e1;
	CK(cudaEventCreate(&e0));
	CK(cudaEventCreate(&e1));
	CK(cudaEventRecord(e0));
	for (int i = 0; i < iters; ++i) {
		f();
	}
	CK(cudaEventRecord(e1));
	CK(cudaEventSynchronize(e1));
	CK(cudaGetLastError());
	float ms = 0;
	CK(cudaEventElapsedTime(&ms, e0, e1));
	return ms / iters;
}

int main(int argc, char** argv) {
	const size_t mb = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 800;
	const size_t bytes = mb * 1000000ull / 8192 * 8192, n16 = bytes / 16;
	uint8_t *a, *b, *c;
	uint32_t* sink;
	// three buffers used in rotation so that consecutive launches never find their lines in L2
	CK(cudaMalloc(&a, bytes));
	CK(cudaMalloc(&b, bytes));
	CK(cudaMalloc(&c, bytes));
	CK(cudaMalloc(&sink, 4));
	CK(cudaMemset(a, 1, bytes));
	CK(cudaMemset(b, 2, bytes));
	CK(cudaMemset(c, 3, bytes));
	int sms = 0;
	CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
	uint8_t* bufs[3] = {a, b, c};
	int rot = 0;
	auto next = [&]() { return bufs[rot++ % 3]; };
	std::printf("buffer %zu MB, %d SMs\n", bytes / 1000000, sms);
	for (int cps : {2, 4, 8}) {
		const int grid = sms * cps;
		float ms = timed([&] { k_write<0><<<grid, 256>>>((uint4*)next(), n16, 7); });
		std::printf("write st.global       grid %4d: %.4f ms %7.1f GB/s\n", grid, ms, bytes / ms / 1e6);
		ms = timed([&] { k_write<1><<<grid, 256>>>((uint4*)next(), n16, 7); });
		std::printf("write st.global.cs    grid %4d: %.4f ms %7.1f GB/s\n", grid, ms, bytes / ms / 1e6);
		ms = timed([&] { k_write_tiles<<<grid, 256>>>((uint4*)next(), bytes / 8192, 7); });
		std::printf("write 8K tiles/warp   grid %4d: %.4f ms %7.1f GB/s\n", grid, ms, bytes / ms / 1e6);
	}
	for (int cps : {2, 8}) {
		for (uint32_t rd16 : {0u, 8u, 16u, 32u}) {
			const int grid = sms * cps;
			uint8_t* side = c;
			float ms = timed([&] {
				uint8_t* d = bufs[rot++ % 2];
				k_write_tiles_rd<<<grid, 256>>>((uint4*)d, bytes / 8192, (const uint4*)side, rd16, 0, sink);
			});
			std::printf("write 8K tiles + %3u B read/tile grid %4d: %.4f ms %7.1f GB/s written\n", rd16 * 16, grid, ms, bytes / ms / 1e6);
		}
	}
	for (int mode : {1, 2}) {
		for (uint32_t rd16 : {8u, 16u, 32u}) {
			const int grid = sms * 2;
			uint8_t* side = c;
			float ms = timed([&] {
				uint8_t* d = bufs[rot++ % 2];
				k_write_tiles_rd_batched<<<grid, 256>>>((uint4*)d, bytes / 8192, (const uint4*)side, rd16, mode, sink);
			});
			std::printf("write 8K tiles + %3u B read/tile %s grid %4d: %.4f ms %7.1f GB/s written\n", rd16 * 16,
			            mode == 1 ? "batched x32 per warp" : "L2 prefetch prologue", grid, ms, bytes / ms / 1e6);
		}
	}
	for (uint32_t tile : {4096u, 8192u}) {
		for (int cps : {1, 2}) {
			for (int depth : {1, 3}) {
				const int grid = sms * cps;
				CK(cudaFuncSetAttribute(k_write_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 8192));
				float ms = timed([&] { k_write_tma<<<grid, 256, 8 * tile>>>(next(), bytes / tile, tile, depth); });
				std::printf("write TMA bulk %5u B grid %4d depth %d: %.4f ms %7.1f GB/s\n", tile, grid, depth, ms, bytes / ms / 1e6);
			}
		}
	}
	for (int cps : {4, 8}) {
		const int grid = sms * cps;
		float ms = timed([&] {
			uint8_t* s = next();
			uint8_t* d = next();
			k_copy<<<grid, 256>>>((const uint4*)s, (uint4*)d, n16);
		});
		std::printf("copy 1:1              grid %4d: %.4f ms %7.1f GB/s (read+write)\n", grid, ms, 2.0 * bytes / ms / 1e6);
		ms = timed([&] { k_read<<<grid, 256>>>((const uint4*)next(), n16, sink); });
		std::printf("read only             grid %4d: %.4f ms %7.1f GB/s\n", grid, ms, bytes / ms / 1e6);
	}
	float ms = timed([&] { CK(cudaMemsetAsync(next(), 5, bytes)); });
	std::printf("cudaMemset                       : %.4f ms %7.1f GB/s\n", ms, bytes / ms / 1e6);
	ms = timed([&] {
		uint8_t* s = next();
		uint8_t* d = next();
		CK(cudaMemcpyAsync(d, s, bytes, cudaMemcpyDeviceToDevice));
	});
	std::printf("cudaMemcpy D2D                   : %.4f ms %7.1f GB/s (read+write)\n", ms, 2.0 * bytes / ms / 1e6);
	return 0;
}

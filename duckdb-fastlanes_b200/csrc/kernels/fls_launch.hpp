// fls_launch.hpp — kernel lookup shared by the kernel translation units and the host-side launcher.
// Each family is compiled in its own .cu (compile time); the launcher picks the instantiation by
// (lane width, value width, output bytes) and launches it through cudaLaunchKernel.
#pragma once
#include "fls_device.h"

// MODE 0 (T-lane encodings): T = column physical bits, OB = output bytes per value (>= T/8).
const void* fls_kernel_main(int T, int OB);
// MODE 1 (DICT): IT = code lane bits (8/16/32), VT = key bits, OB = output bytes (>= VT/8).
const void* fls_kernel_dict(int IT, int VT, int OB);
// MODE 2 (RLE, u16 run index): VT = run value bits.
const void* fls_kernel_rle(int VT, int OB);
// Fused decode + aggregate (SUM/COUNT/MIN/MAX): same families, output replaced by a reduction.
const void* fls_kernel_agg(int mode, int T, int VT);

constexpr int kFlsWarpsPerCta = 8;

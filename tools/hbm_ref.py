#!/usr/bin/env python
"""Reference HBM rates on this GPU for context: write-only (memset), read+write copy, for a given buffer size."""
import sys
import torch

n = int(sys.argv[1]) if len(sys.argv) > 1 else 800_000_000
a = torch.empty(n, dtype=torch.uint8, device="cuda")
b = torch.empty(n, dtype=torch.uint8, device="cuda")


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


ms = timed(lambda: a.zero_())
print(f"memset {n/1e6:.0f} MB: {ms:.4f} ms  {n/ms/1e6:.1f} GB/s written")
ms = timed(lambda: a.fill_(7))
print(f"fill   {n/1e6:.0f} MB: {ms:.4f} ms  {n/ms/1e6:.1f} GB/s written")
ms = timed(lambda: b.copy_(a))
print(f"copy   {n/1e6:.0f} MB: {ms:.4f} ms  {2*n/ms/1e6:.1f} GB/s read+write")
a64 = a.view(torch.int64)
ms = timed(lambda: a64.sum())
print(f"read   {n/1e6:.0f} MB: {ms:.4f} ms  {n/ms/1e6:.1f} GB/s read (torch sum)")

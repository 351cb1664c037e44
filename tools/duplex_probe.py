"""PCIe probe: pinned H2D and D2H alone and at the same time (two streams), 1 GiB each."""
import torch
n = 1 << 30
dev_a, dev_b = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.empty(n, dtype=torch.uint8, device="cuda")
host_a, host_b = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=4):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def h2d():
    with torch.cuda.stream(s1):
        dev_a.copy_(host_a, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        host_b.copy_(dev_b, non_blocking=True)


def both():
    h2d()
    d2h()


import time
for name, fn, b in (("h2d", h2d, n), ("d2h", d2h, n), ("both", both, 2 * n)):
    t0 = time.perf_counter()
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 4
    print(name, round(b / dt / 1e9, 1), "GB/s total")

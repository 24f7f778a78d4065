#!/usr/bin/env python
"""What HBM bandwidth does this B200 give for the convert kernel's traffic MIX (1 byte read : 2 bytes
written)?  MEASURED_PEAKS.json's copy figure is a 1:1 mix.  Plain torch element-wise kernels, CUDA events."""
import torch

def bw(fn, nbytes, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    return nbytes * iters / (a.elapsed_time(b) * 1e-3) / 1e9

n = 327_680_000                      # samples of one cfg3 step
x16 = torch.randint(-30000, 30000, (n,), dtype=torch.int16, device="cuda")
y32 = torch.empty(n, dtype=torch.float32, device="cuda")
z16 = torch.empty_like(x16)
big = torch.empty(n, dtype=torch.float32, device="cuda")
print("copy int16->int16 (1:1)        %.0f GB/s" % bw(lambda: z16.copy_(x16), 4 * n))
print("copy f32->f32 (1:1)            %.0f GB/s" % bw(lambda: big.copy_(y32), 8 * n))
print("convert int16->f32 (1:2)       %.0f GB/s" % bw(lambda: y32.copy_(x16), 6 * n))
print("fill f32 (write only)          %.0f GB/s" % bw(lambda: y32.fill_(1.0), 4 * n))
print("sum int16 (read only)          %.0f GB/s" % bw(lambda: x16.sum(), 2 * n))

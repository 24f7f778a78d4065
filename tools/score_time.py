#!/usr/bin/env python
"""Time score_kernel (ga_k_signal_strength) for 4,096 window members with 4,800-sample wake-word buffers
and check the scores against numpy's sequential float64 sum of float32 squares."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import __graft_entry__ as entry  # noqa: E402
entry.build()
from loqa_hub_b200 import api  # noqa: E402

n, w = 4096, 4800
dev = torch.device("cuda", 0)
rng = np.random.default_rng(3)
x = (rng.integers(-32768, 32768, size=(n, w)).astype(np.float32) / np.float32(32767.0))
x[5, 100:200] = 0.0
x[6] = 0.0
x[7, :10] = np.float32(1e-30)          # squares underflow to zero / subnormal: the predicated fallback
x[8, :4] = [np.float32(3e-20), np.float32(1e-19), np.float32(2e19), np.float32(1e-22)]
d_x = torch.from_numpy(x).to(dev)
off = torch.arange(n, dtype=torch.int64, device=dev) * w
cnt = torch.full((n,), w, dtype=torch.int32, device=dev)
sc = torch.zeros(n, dtype=torch.float64, device=dev)
run = lambda: api.k_signal_strength(d_x.data_ptr(), off.data_ptr(), cnt.data_ptr(), 0, n, sc.data_ptr(), 0, 0)
for _ in range(3):
    run()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20):
    run()
b.record()
torch.cuda.synchronize()
got = sc.cpu().numpy()
bad = 0
for m in list(range(0, n, 97)) + [5, 6, 7, 8]:
    s = 0.0
    for q in (x[m] * x[m]).astype(np.float64):
        s += q
    bad += got[m] != np.sqrt(s / w)
print("score_kernel: %.2f us per launch, %d mismatches" % (a.elapsed_time(b) / 20 * 1e3, bad))

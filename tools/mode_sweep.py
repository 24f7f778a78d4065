#!/usr/bin/env python
"""Time ga_k_ingest_uniform for every mode combination on the configs[2] batch (4,096 streams x 250 x 640 B):
what each part (convert = 1, features = 2, SHA-256 = 4) adds to the launch.  Prints one line per mode."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import __graft_entry__ as entry  # noqa: E402
entry.build()
from loqa_hub_b200 import _lib as L  # noqa: E402
from loqa_hub_b200 import api, synth  # noqa: E402

S, M, ms = 4096, 250, 320
dev = torch.device("cuda", 0)
n = S * M
pcm = torch.empty(n * ms, dtype=torch.int16, device=dev)
g = torch.arange(S, device=dev, dtype=torch.int32) // 8
r = torch.arange(S, device=dev, dtype=torch.int32) % 8
api.synth_batch_device_map(pcm.data_ptr(), synth.SEED, g.data_ptr(), r.data_ptr(), S, M, ms, 0)
k = np.arange(n, dtype=np.uint64)
descs = api.make_descs(k * 640, np.full(n, 640, np.uint32), (k % S) * (M * ms) + (k // S) * ms, (k % S).astype(np.uint32),
                       np.zeros(n, np.uint32), frame_off=np.zeros(n, np.uint32))
d_desc = torch.from_numpy(descs.view(np.uint8)).to(dev)
samples = torch.empty(S * M * ms + 4, dtype=torch.float32, device=dev)
results = torch.empty(n * 64, dtype=torch.uint8, device=dev)
for mode in (4, 2, 6, 1, 5, 3, 7):
    def run():
        api.k_ingest_uniform(pcm.data_ptr(), d_desc.data_ptr(), n, 0, 640, 640, samples.data_ptr() if mode & 1 else 0,
                             results.data_ptr(), 0, mode, 0.01, 0)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        run()
    b.record()
    torch.cuda.synchronize()
    print("mode %d (%s): %.4f ms" % (mode, "+".join(x for x, bit in (("convert", 1), ("features", 2), ("sha256", 4)) if mode & bit),
                                     a.elapsed_time(b) / 10))

#!/usr/bin/env python
"""Window-close winner path: time ga_fetch_wavs (WAV + SHA-256 of the samples) for N winners at once.
Each winner holds a 300 ms wake-word buffer (4,800 samples) plus `speech` chunks of 20 ms.  Prints JSON."""
import ctypes as C
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import __graft_entry__ as entry  # noqa: E402
entry.build()
from loqa_hub_b200 import _lib as L  # noqa: E402
from loqa_hub_b200 import api  # noqa: E402
import hashlib  # noqa: E402

out = []
for n_win, speech in ((1024, 0), (1024, 15), (8192, 0), (1024, 235)):
    S = n_win
    with api.Context(device=0, max_streams=S, max_batch_msgs=S * 25, ring_bytes=S * 640 * 32, arena_samples=80_000,
                     wake_samples_max=4_800, mode=L.DO_ALL) as c:
        rng = np.random.default_rng(1)
        chunk = rng.integers(-20000, 20000, size=(S, 320)).astype("<i2")
        ids = np.arange(S, dtype=np.uint32)
        total = 15 + speech
        for t0 in range(0, total, 25):
            T = min(25, total - t0)
            view, off = c.reserve(T * S * 640)
            view[:T * S * 640] = np.tile(chunk.view(np.uint8).reshape(-1), T)
            fl = np.zeros((T, S), np.uint8)
            for t in range(T):
                if t0 + t < 15:
                    fl[t] |= L.MSG_WAKE
                if t0 + t == 0:
                    fl[t] |= L.MSG_FIRST
            c.wait(c.submit_ticks(off, T, ids, 640, fl.reshape(-1)))
        lib = L.load()
        base, offs, dg = C.c_void_p(), np.zeros(S + 1, np.uint64), np.zeros((S, 32), np.uint8)
        ts = []
        for _ in range(12):
            a = time.perf_counter()
            L.check(lib.ga_fetch_wavs(c._h, L.ptr(ids), S, 16000, C.byref(base), offs.ctypes.data_as(C.c_void_p), L.ptr(dg)))
            ts.append((time.perf_counter() - a) * 1e3)
        want = hashlib.sha256((np.tile(chunk[7], total).astype(np.float32) / np.float32(32767.0)).astype("<f4").tobytes()).digest()
        assert dg[7].tobytes() == want, "digest mismatch"
        ts = sorted(ts[2:])
        out.append({"winners": n_win, "samples_each": 320 * total, "wav_mbytes": float(offs[S] - offs[0]) / 1e6,
                    "ms_p50": ts[len(ts) // 2], "ms_max": ts[-1]})
print(json.dumps({"tool": "winner_latency", "cases": out}))

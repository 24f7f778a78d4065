#!/usr/bin/env python
"""Window-close latency of the single-process multi-GPU path: one context per visible GPU, 8-relay
windows, every 16th window with half of its relays on the next GPU; times ga_cluster_arbitrate
(host call to results on the host) for ALL windows at once and checks the winners."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402
from loqa_hub_b200 import _lib as L  # noqa: E402
from loqa_hub_b200 import api, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--streams", type=int, default=8192, help="per GPU")
ap.add_argument("--calls", type=int, default=30)
a = ap.parse_args()
entry.build()
n = L.load().ga_device_count()
S, G, wake, ms = a.streams, 8, 15, 320
groups = S // G
ctxs = [api.Context(device=d, max_streams=S, max_batch_msgs=S * wake, ring_bytes=S * wake * ms * 2, arena_samples=8000,
                    wake_samples_max=wake * ms, mode=L.DO_ALL) for d in range(n)]
ids = np.arange(S, dtype=np.uint32)
for d, c in enumerate(ctxs):                       # wake-word audio of every stream: 15 ticks in one regular batch
    pcm = api.synth_batch_host(synth.SEED, d * groups, S, G, wake, ms, n_threads=8)
    view, off = c.reserve(pcm.nbytes)
    view[:pcm.nbytes] = pcm.view(np.uint8).reshape(-1)
    flags = np.full(wake * S, L.MSG_WAKE, np.uint8); flags[:S] |= L.MSG_FIRST
    c.wait_view(c.submit_ticks(off, wake, ids, ms * 2, flags))
seg, shard, stream = [0], [], []
for d in range(n):
    for g in range(groups):
        for r in range(G):
            far = n > 1 and g % 16 == 0 and r >= G // 2
            shard.append((d + 1) % n if far else d); stream.append(g * G + r)
        seg.append(len(shard))
seg, shard, stream = np.array(seg, np.uint32), np.array(shard, np.uint32), np.array(stream, np.uint32)
with api.Cluster(ctxs, max_guests=S) as cl:
    times = []
    for _ in range(a.calls + 3):
        t0 = time.perf_counter()
        winner, scores, status = cl.arbitrate(seg, shard, stream)
        times.append(time.perf_counter() - t0)
    times = np.array(times[3:]) * 1e3
    # relay 0 of a group hears it loudest; for a split window the far half are relays 4..7 of the NEXT GPU's group g
    local = [w for w in range(len(seg) - 1) if len(set(shard[seg[w]:seg[w + 1]])) == 1]
    ok = all(int(winner[w]) == int(seg[w]) for w in local)
    print(json.dumps({"gpus": n, "streams_per_gpu": S, "windows": int(len(seg) - 1),
                      "spanning_windows": int(len(seg) - 1 - len(local)),
                      "cluster_arbitrate_ms": {"median": float(np.median(times)), "p99": float(np.percentile(times, 99)),
                                               "max": float(times.max())},
                      "local_winners_are_relay0": bool(ok)}))
for c in ctxs:
    c.close()

#!/usr/bin/env python
"""Tick mode: one 20 ms chunk per stream per batch tick through the host C ABI —
per-tick latency from "chunks are in the pinned ring" to "result records are on the host", and the
latency of arbitrating all windows at window close.  The real-time requirement is << 20 ms."""
import argparse
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from loqa_hub_b200 import _lib as L  # noqa: E402
from loqa_hub_b200 import api, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--streams", type=int, default=8192)
ap.add_argument("--ticks", type=int, default=250)
ap.add_argument("--utterance", action="store_true", help="also keep the per-utterance running SHA-256 (GA_DO_UTTERANCE)")
ap.add_argument("--submit", default="ticks", choices=["ticks", "msgs"],
                help="ga_submit_ticks (descriptors built on the GPU) or ga_submit")
args = ap.parse_args()
use_ticks = args.submit == "ticks"
S, ms, wake, group = args.streams, 320, 15, 8
pcm = api.synth_batch_host(synth.SEED, 0, S, group, args.ticks, ms, n_threads=8).reshape(args.ticks, S * ms)
ctx = api.Context(device=0, max_streams=S, max_batch_msgs=S, max_inflight=2, ring_bytes=S * ms * 2 * 8,
                  arena_samples=args.ticks * ms, wake_samples_max=wake * ms,
                  mode=L.DO_ALL | (L.DO_UTTERANCE if args.utterance else 0))
msgs = np.zeros(S, dtype=L.MSG_DTYPE)
msgs["nbytes"] = ms * 2
msgs["stream_id"] = np.arange(S)
rel = np.arange(S, dtype=np.uint64) * (ms * 2)
lat, arb = [], None
seg = np.arange(0, S + 1, group, dtype=np.uint32)
sid = np.arange(S, dtype=np.uint32)
for t in range(args.ticks):
    view, off = ctx.reserve(S * ms * 2)
    view[:S * ms * 2] = pcm[t].view(np.uint8)                 # the relays' chunks land in the ring (not timed)
    msgs["ring_offset"] = rel + np.uint64(off)
    msgs["flags"] = (L.MSG_WAKE if t < wake else 0) | (L.MSG_FIRST if t == 0 else 0)
    tick_flags = np.full(S, int(msgs["flags"][0]), np.uint8)
    t0 = time.perf_counter()
    if use_ticks:
        res, _ = ctx.wait_view(ctx.submit_ticks(off, 1, sid, ms * 2, tick_flags))
    else:
        res, _ = ctx.wait_view(ctx.submit(msgs))
    lat.append(time.perf_counter() - t0)
    if t == wake - 1:                                         # the 300 ms windows close
        t0 = time.perf_counter()
        winner, scores, status = ctx.arbitrate(seg, sid)
        arb_first = time.perf_counter() - t0                  # includes growing the context's tables
        assert (winner == seg[:-1]).all()
        t0 = time.perf_counter()
        winner, scores, status = ctx.arbitrate(seg, sid)
        arb = time.perf_counter() - t0
        assert (winner == seg[:-1]).all()
lat = np.array(lat[5:]) * 1e3
if args.utterance:
    import hashlib
    whole = pcm.reshape(args.ticks, S, ms)[:, 5, :].tobytes()
    assert ctx.utterance_digest(5) == (hashlib.sha256(whole).digest(), len(whole))
print(json.dumps({"streams": S, "ticks": args.ticks, "submit": "ga_submit_ticks" if use_ticks else "ga_submit", "utterance_digest": bool(args.utterance), "tick_latency_ms": {"median": float(np.median(lat)),
      "p99": float(np.percentile(lat, 99)), "max": float(lat.max())},
      "arbitration_ms_all_windows": arb * 1e3, "arbitration_ms_first_call": arb_first * 1e3, "windows": S // group,
      "real_time_budget_ms": 20.0, "h2d_bytes_per_tick": S * ms * 2 + (S * 13 if use_ticks else S * 32), "d2h_bytes_per_tick": S * 64}))
ctx.close()

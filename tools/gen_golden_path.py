#!/usr/bin/env python
"""Generate tests/golden/path_golden.json: the parts of the path ingest_golden.json does not hold —
arbitration windows, the 48 kHz decimator, the WAV payload for winners and the running utterance
digest.

Like tools/gen_golden.py it uses implementations independent of BOTH the C oracle and the CUDA
path: oracle/numpy_mirror.py (numpy float32 / sequential float64; the decimator with every fmaf
evaluated as an exact rational and rounded once; struct for the RIFF header) and hashlib.  The
filter taps are read from include/gpuaudio_taps.h.  Reference inputs restated here:
TestSignalStrengthArbitration (internal/grpc/collision_detection_test.go:226-293),
the single-relay / empty-signal requests (:322-459) and float32ToWAV (internal/llm/stt_client.go:365-398).
"""
import hashlib
import json
import os
import re
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import numpy_mirror as npm  # noqa: E402

CONTENDING, CONNECTED, WINNER, CANCELLED = 1, 0, 2, 3


def f32_bits(a):
    return [int(b) for b in np.asarray(a, dtype=np.float32).view(np.uint32)]


def f64_hex(x):
    return struct.pack("<d", float(x)).hex()


def taps_from_header():
    with open(os.path.join(ROOT, "include", "gpuaudio_taps.h")) as f:
        body = f.read()
    body = body[body.index("GA_RESAMPLE_TAP_BITS"):]
    bits = [int(h, 16) for h in re.findall(r"0x([0-9a-fA-F]{8})u", body)]
    assert len(bits) == 48, len(bits)
    return np.asarray(bits, dtype=np.uint32).view(np.float32)


rng = np.random.default_rng(0x10CA0B20)

# ------------------------------------------------------------------ arbitration windows
def window(name, signals, status=None):
    """signals: float32 arrays, or bytes = int16 PCM as received (converted by A1)."""
    status = status or [CONTENDING] * len(signals)
    members = [{"pcm_hex": s.hex(), "status": st} if isinstance(s, bytes) else {"f32_bits": f32_bits(s), "status": st}
               for s, st in zip(signals, status)]
    signals = [npm.bytes_to_float32(s) if isinstance(s, bytes) else s for s in signals]
    scores = [npm.signal_strength(s)[0] if st == CONTENDING else None for s, st in zip(signals, status)]
    w = npm.select_winner([0.0 if sc is None else sc for sc in scores], [st == CONTENDING for st in status])
    return {"name": name, "members": members,
            "scores_f64_hex": [None if sc is None else f64_hex(sc) for sc in scores], "winner": w,
            "status_out": [WINNER if i == w else CANCELLED for i in range(len(signals))]}


f = np.float32
relay8 = []
base = rng.integers(-12000, 12000, size=1600)
for r in range(8):                                              # one utterance at 8 gains + noise, as SURVEY §8d
    x = np.clip((base * int(32768 * 0.9 ** r) >> 15) + rng.integers(-64 * (r + 1), 64 * (r + 1) + 1, size=1600), -32768, 32767)
    relay8.append(x.astype("<i2").tobytes())
windows = [
    window("reference_three_relays", [np.full(4, f(0.1)), np.full(4, f(1.0)), np.full(4, f(0.5))]),
    window("single_relay", [np.full(3, f(0.5))]),
    window("empty_wake_signal_single_relay", [np.zeros(0, np.float32)]),
    window("all_zero_falls_back_to_first", [np.zeros(8, np.float32), np.zeros(3, np.float32), np.zeros(0, np.float32)]),
    window("tie_first_in_join_order_wins", [np.full(4, f(0.5)), np.full(4, f(0.5)), np.full(4, f(0.25))]),
    window("loudest_is_not_contending", [np.full(4, f(0.9)), np.full(4, f(0.5)), np.full(4, f(0.6))],
           [CANCELLED, CONTENDING, CONTENDING]),
    window("nobody_contending_first_member", [np.full(4, f(0.9)), np.full(4, f(0.5))], [CONNECTED, WINNER]),
    window("eight_relays_100ms", list(reversed(relay8))),      # loudest joined last
    window("empty_window", []),
]

# ------------------------------------------------------------------ decimator
taps = taps_from_header()
streams = []
for name, sizes in (("ragged", [4 * 7, 4 * 1 + 2, 4 * 100, 4 * 46, 4 * 200 + 3, 4 * 300, 0, 4 * 47, 4 * 48]),
                    ("two_full_chunks", [3840, 3840])):
    total = sum(s - s % 4 for s in sizes) // 4
    if name == "ragged":
        pcm = rng.integers(-30000, 30000, size=(total, 2)).astype("<i2")
    else:                                                      # a 1 kHz tone, R = L - L/8
        left = (9000 * np.sin(2 * np.pi * 1000 * np.arange(total) / 48000)).astype(np.int16)
        pcm = np.stack([left, left - (left >> 3)], axis=1).astype("<i2")
    raw, pos, r = pcm.tobytes(), 0, npm.Resampler(taps)
    chunks = []
    for n in sizes:
        data = raw[pos:pos + n]
        if n % 4:                                              # a trailing partial frame is dropped, not carried
            data = data[: n - n % 4] + bytes(rng.integers(0, 256, size=n % 4, dtype=np.uint8))
        pos += n - n % 4
        y = r.process(data)
        rms, sumsq = npm.signal_strength(y)
        chunks.append({"pcm_hex": data.hex(), "out_f32_bits": f32_bits(y), "sha256": hashlib.sha256(data).hexdigest(),
                       "sumsq_f64_hex": f64_hex(sumsq), "rms_f64_hex": f64_hex(rms)})
    streams.append({"name": name, "chunks": chunks})

# ------------------------------------------------------------------ WAV for winners + utterance digest
wav_samples = npm.bytes_to_float32(rng.integers(-20000, 20000, size=37).astype("<i2").tobytes())
wav = {"f32_bits": f32_bits(wav_samples), "sample_rate": 16000, "wav_hex": npm.float32_to_wav(wav_samples, 16000).hex(),
       "empty_wav_hex": npm.float32_to_wav(np.zeros(0, np.float32), 16000).hex()}
utt_chunks = [bytes(rng.integers(0, 256, size=n, dtype=np.uint8)) for n in (640, 640, 1, 63, 64, 65, 0, 640, 3)]
h, running = hashlib.sha256(), []
for c in utt_chunks:
    h.update(c)
    running.append(h.hexdigest())
utterance = {"chunks_hex": [c.hex() for c in utt_chunks], "running_sha256": running,
             "samples_sha256": hashlib.sha256(np.concatenate([npm.bytes_to_float32(c) for c in utt_chunks]).tobytes()).hexdigest()}

# ------------------------------------------------------------------ 20 ms framing (N1): A2 per 320-sample frame
frame_msgs = []
for n in (1000, 320, 321, 5):
    x = rng.integers(-25000, 25000, size=n)
    x[rng.random(n) < 0.2] = 0
    data = x.astype("<i2").tobytes()
    s = npm.bytes_to_float32(data)
    frs = []
    for off in range(0, n, 320):
        fr = s[off:off + 320]
        rms, sumsq = npm.signal_strength(fr)
        frs.append({"n_samples": int(fr.size), "sumsq_f64_hex": f64_hex(sumsq), "rms_f64_hex": f64_hex(rms),
                    "nonzero": int((fr != 0).sum()), "peak_bits": int(np.abs(fr).max().view(np.uint32)),
                    "vad": int(rms > 0.01)})
    frame_msgs.append({"pcm_hex": data.hex(), "frames": frs})

out = os.path.join(ROOT, "tests", "golden", "path_golden.json")
with open(out, "w") as fh:
    json.dump({"generator": "tools/gen_golden_path.py", "tap_bits": [int(b) for b in taps.view(np.uint32)],
               "arbitration": windows, "resample": streams, "wav": wav, "utterance": utterance,
               "frames": frame_msgs}, fh, indent=0)
print("wrote", out, os.path.getsize(out), "bytes")

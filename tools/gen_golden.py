#!/usr/bin/env python
"""Generate tests/golden/ingest_golden.json.

The reference is Go and cannot run in this image (no Go toolchain), so the fixture
is produced by implementations that are independent of BOTH the C oracle and the
CUDA path: numpy float32 division / strictly-sequential cumsum (oracle/numpy_mirror.py
restating internal/grpc/audio_service.go:1048-1101 and :818-852) and hashlib (OpenSSL)
for SHA-256.  It includes the reference's own unit-test inputs
(internal/grpc/audio_service_test.go:258-313) and the 440 Hz int16 sine of
tests/integration/audio_pipeline_test.go:497-507.
"""
import json
import math
import os
import struct
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import numpy_mirror as npm  # noqa: E402


def sine_bytes(n):  # tests/integration/audio_pipeline_test.go:497-507
    out = bytearray()
    for i in range(n):
        s = int(1000.0 * math.sin(2 * math.pi * float(i) * 440.0 / 16000.0))  # Go int16() truncates
        out += struct.pack("<h", s)
    return bytes(out)


rng = np.random.default_rng(0x10CA)
cases = {
    "ref_empty": b"",
    "ref_single": bytes([0x00, 0x10]),
    "ref_multiple": bytes([0x00, 0x00, 0xFF, 0x7F]),
    "ref_odd": bytes([0x00, 0x10, 0xFF]),
    "one_byte": b"\x7f",
    "int16_min_run": struct.pack("<8h", *([-32768] * 8)),
    "int16_max_run": struct.pack("<8h", *([32767] * 8)),
    "plus_minus_one": struct.pack("<6h", 1, -1, 1, -1, 0, 1),
    "sine440_20ms": sine_bytes(320),
    "silent_20ms": b"\x00" * 640,
}
for n in (55, 56, 63, 64, 65, 119, 120, 127, 128, 129, 639, 641):
    cases["rand_%dB" % n] = rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()
quiet = rng.integers(-20000, 20000, size=4800).astype("<i2")
quiet[::5] = rng.integers(-9, 10, size=quiet[::5].size)
cases["wake_window_4800"] = quiet.tobytes()

msgs = []
for name, data in cases.items():
    s = npm.bytes_to_float32(data)
    rms, sumsq = npm.signal_strength(s)
    msgs.append({
        "name": name,
        "pcm_hex": data.hex(),
        "f32_bits": [int(b) for b in s.view(np.uint32)],
        "sha256": npm.sha256(data).hex(),
        "rms_f64_hex": struct.pack("<d", rms).hex(),
        "sumsq_f64_hex": struct.pack("<d", sumsq).hex(),
        "nonzero": int((s != 0).sum()),
        "peak_bits": int(np.abs(s).max().view(np.uint32)) if s.size else 0,
    })
out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                   "ingest_golden.json")
with open(out, "w") as f:
    json.dump({"generator": "tools/gen_golden.py", "messages": msgs}, f, indent=0)
print("wrote", out, len(msgs), "cases", os.path.getsize(out), "bytes")

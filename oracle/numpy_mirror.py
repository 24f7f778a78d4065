"""Independent numpy restatement of A1/A2/A3 — TEST INFRASTRUCTURE ONLY.

A second, differently-written statement of the same reference semantics
(internal/grpc/audio_service.go:1048-1101, :818-852, :517-551) used to
cross-check the C oracle; hashlib (OpenSSL) plays the same role for SHA-256.
"""
import hashlib

import numpy as np


def bytes_to_float32(data):
    b = np.frombuffer(bytes(data), dtype=np.uint8)
    b = b[: b.size - (b.size % 2)]                       # :1056-1063
    v = b.view("<i2")                                     # :1076
    return v.astype(np.float32) / np.float32(32767.0)     # :1078 (f32 / f32)


def signal_strength(samples):
    s = np.asarray(samples, dtype=np.float32)
    if s.size == 0:
        return 0.0, 0.0                                   # :819-822
    sq = (s * s).astype(np.float32).astype(np.float64)    # :828 product rounded to f32 first
    total = float(np.cumsum(sq)[-1])                      # cumsum is strictly sequential
    return float(np.sqrt(total / s.size)), total          # :841


def select_winner(scores, contending):
    winner, best = -1, 0.0                                # :519
    for i, (sc, c) in enumerate(zip(scores, contending)):
        if not c:
            continue                                      # :522-524
        if sc > best:                                     # :535
            best, winner = sc, i
    if winner < 0 and len(scores) > 0:
        winner = 0                                        # :542-547
    return winner


def sha256(data):
    return hashlib.sha256(bytes(data)).digest()


# ---------------------------------------------------------------- N2 decimator, exact arithmetic
# A third statement of the 48 kHz stereo -> 16 kHz mono decimator, independent of the C oracle's
# compiler and libm: every fused multiply-add is evaluated as an exact rational and rounded once
# to binary32 (round-to-nearest-even), which is what fmaf() is specified to do.
from fractions import Fraction  # noqa: E402


def round_to_f32(x):
    """Exact rational -> nearest binary32 (ties to even), as a Python float."""
    if x == 0:
        return 0.0
    sign = -1 if x < 0 else 1
    a = abs(x)
    e = a.numerator.bit_length() - a.denominator.bit_length()      # 2^(e-1) <= a < 2^(e+1)
    if Fraction(2) ** e > a:
        e -= 1                                                      # now 2^e <= a < 2^(e+1)
    q = Fraction(2) ** (max(e, -126) - 23)                          # spacing of binary32 at a (subnormals: 2^-149)
    n = a / q
    lo = n.numerator // n.denominator
    rem = n - lo
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and lo % 2 == 1):
        lo += 1
    return sign * float(lo * q)                                     # lo*q has <= 24 significant bits: exact in a double


def fma_f32(a, b, c):
    return round_to_f32(Fraction(float(a)) * Fraction(float(b)) + Fraction(float(c)))


class Resampler:
    """mono = (L/32767 + R/32767) * 0.5 in float32, then y[M] = sum over k = 0..47 (in this order,
    one fmaf each) of taps[k] * mono[3M + 2 - k]; zero history before the stream; a partial
    trailing frame of a chunk is dropped."""

    def __init__(self, taps):
        self.taps = [float(t) for t in np.asarray(taps, dtype=np.float32)]
        self.hist = [0.0] * (len(self.taps) - 1)
        self.n_frames = 0

    def process(self, data):
        b = np.frombuffer(bytes(data), dtype=np.uint8)
        nfr = b.size // 4
        v = b[: nfr * 4].view("<i2").reshape(nfr, 2).astype(np.float32) / np.float32(32767.0)
        mono = ((v[:, 0] + v[:, 1]).astype(np.float32) * np.float32(0.5)).astype(np.float32)
        x = self.hist + [float(m) for m in mono]
        t, n0, n1 = len(self.taps), self.n_frames, self.n_frames + nfr
        out = []
        m = n0 // 3
        while 3 * m + 2 < n0:
            m += 1
        while 3 * m + 2 < n1:
            top = 3 * m + 2 - n0 + (t - 1)
            acc = 0.0
            for k in range(t):
                acc = fma_f32(self.taps[k], x[top - k], acc)
            out.append(acc)
            m += 1
        self.hist = x[len(x) - (t - 1):]
        self.n_frames = n1
        return np.asarray(out, dtype=np.float32)


def float32_to_wav(samples, sample_rate):
    """internal/llm/stt_client.go:365-398."""
    import struct
    s = np.asarray(samples, dtype="<f4")
    data = s.tobytes()
    return (b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE" + b"fmt " + struct.pack("<I", 16) +
            struct.pack("<HHIIHH", 3, 1, sample_rate, sample_rate * 4, 4, 32) + b"data" +
            struct.pack("<I", len(data)) + data)

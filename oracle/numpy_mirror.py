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

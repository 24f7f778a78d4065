"""ctypes binding of oracle/_build/liboracle.so — TEST INFRASTRUCTURE ONLY.

Mirrors the C restatement one-to-one; see gpuaudio_oracle.h for the reference
citations of every function.  build() compiles the library with the committed
Makefile (gcc only).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

RELAY_CONNECTED, RELAY_CONTENDING, RELAY_WINNER, RELAY_CANCELLED = 0, 1, 2, 3
RESAMPLE_TAPS = 48


class Stats(C.Structure):
    _fields_ = [("n_samples", C.c_uint32), ("nonzero", C.c_uint32), ("peak", C.c_float),
                ("_pad", C.c_uint32), ("sumsq", C.c_double), ("rms", C.c_double)]


class Sha256Ctx(C.Structure):
    _fields_ = [("h", C.c_uint32 * 8), ("tail", C.c_uint8 * 64), ("tail_len", C.c_uint32),
                ("_pad", C.c_uint32), ("total_len", C.c_uint64)]


class ResampleState(C.Structure):
    _fields_ = [("hist", C.c_float * (RESAMPLE_TAPS - 1)), ("_pad", C.c_uint32),
                ("n_frames", C.c_uint64)]


class BaselineCfg(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("n_streams", "n_msgs", "msg_bytes", "wake_msgs", "group_size",
                                          "format", "layout", "sha_impl", "n_threads")]


FMT_MONO16K, FMT_STEREO48K = 0, 1
LAYOUT_TIME_MAJOR, LAYOUT_STREAM_MAJOR = 0, 1
SHA_PORT, SHA_OPENSSL_EVP, SHA_OPENSSL_LOW = 0, 1, 2
SYNTH_SEED = 0x10CA0B2000000001

FRAME_DTYPE = np.dtype([("sumsq", "<f8"), ("rms", "<f8"), ("n_samples", "<u4"),
                        ("nonzero", "<u4"), ("peak", "<f4"), ("vad", "<u4")])

_lib = None


def build(force=False):
    if force or not os.path.exists(_SO) or any(
            os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_SO)
            for f in ("gpuaudio_oracle.c", "cpu_baseline.c", "synth_oracle.c", "gpuaudio_oracle.h", "Makefile")):
        subprocess.check_call(["make", "-C", _HERE, "-B"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p, f32p = C.POINTER(C.c_uint8), C.POINTER(C.c_float)
        L.ora_bytes_to_float32.restype = C.c_size_t
        L.ora_bytes_to_float32.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(Stats)]
        L.ora_signal_strength.restype = C.c_double
        L.ora_signal_strength.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Stats)]
        L.ora_arbitrate.restype = C.c_int
        L.ora_arbitrate.argtypes = [C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t),
                                    C.c_void_p, C.c_void_p, C.c_void_p]
        L.ora_select_winner.restype = C.c_int
        L.ora_select_winner.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ora_sha256.restype = None
        L.ora_sha256.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.ora_sha256_init.argtypes = [C.POINTER(Sha256Ctx)]
        L.ora_sha256_update.argtypes = [C.POINTER(Sha256Ctx), C.c_void_p, C.c_size_t]
        L.ora_sha256_final.argtypes = [C.POINTER(Sha256Ctx), C.c_void_p]
        L.ora_frame_features.restype = C.c_size_t
        L.ora_frame_features.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_double, C.c_void_p]
        L.ora_resample_init.argtypes = [C.POINTER(ResampleState)]
        L.ora_resample_48k_stereo.restype = C.c_size_t
        L.ora_resample_48k_stereo.argtypes = [C.POINTER(ResampleState), C.c_void_p, C.c_size_t,
                                              C.c_void_p]
        L.ora_resample_taps.restype = f32p
        L.ora_float32_to_wav.restype = C.c_size_t
        L.ora_float32_to_wav.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p]
        L.ora_have_openssl.restype = C.c_int
        L.ora_cpu_baseline.restype = C.c_double
        L.ora_cpu_baseline.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                       C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_void_p]
        L.ora_cpu_baseline_run.restype = C.c_double
        L.ora_cpu_baseline_run.argtypes = [C.POINTER(BaselineCfg), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ora_memcpy_gbs.restype = C.c_double
        L.ora_memcpy_gbs.argtypes = [C.c_uint32, C.c_size_t, C.c_uint32]
        L.ora_synth_fill.restype = None
        L.ora_synth_fill.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                     C.c_uint32, C.c_int, C.c_uint32, C.c_uint32]
        L.ora_synth_sample.restype = C.c_int16
        L.ora_synth_sample.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        _ = u8p
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def bytes_to_float32(data):
    """A1.  data: bytes-like / uint8 array.  Returns (float32 array, Stats)."""
    buf = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else \
        np.ascontiguousarray(data, dtype=np.uint8)
    out = np.empty(buf.size // 2, dtype=np.float32)
    st = Stats()
    n = lib().ora_bytes_to_float32(_ptr(buf), buf.size, _ptr(out), C.byref(st))
    return out[:n], st


def signal_strength(samples):
    """A2.  Returns (rms, Stats)."""
    s = np.ascontiguousarray(samples, dtype=np.float32)
    st = Stats()
    rms = lib().ora_signal_strength(_ptr(s), s.size, C.byref(st))
    return rms, st


def arbitrate(wake_signals, status_in=None):
    """A3 selection.  wake_signals: list of float32 arrays (join order).
    Returns (winner_index, scores float64[n], status_out uint8[n])."""
    n = len(wake_signals)
    arrs = [np.ascontiguousarray(w, dtype=np.float32) for w in wake_signals]
    ptrs = (C.c_void_p * max(n, 1))(*[a.ctypes.data if a.size else None for a in arrs])
    lens = (C.c_size_t * max(n, 1))(*[a.size for a in arrs])
    st_in = np.full(n, RELAY_CONTENDING, dtype=np.uint8) if status_in is None else \
        np.ascontiguousarray(status_in, dtype=np.uint8)
    scores = np.zeros(n, dtype=np.float64)
    st_out = np.zeros(n, dtype=np.uint8)
    w = lib().ora_arbitrate(n, ptrs, lens, _ptr(st_in), _ptr(scores), _ptr(st_out))
    return w, scores, st_out


def select_winner(scores, status_in=None):
    sc = np.ascontiguousarray(scores, dtype=np.float64)
    n = sc.size
    st_in = np.full(n, RELAY_CONTENDING, dtype=np.uint8) if status_in is None else \
        np.ascontiguousarray(status_in, dtype=np.uint8)
    st_out = np.zeros(n, dtype=np.uint8)
    w = lib().ora_select_winner(n, _ptr(sc), _ptr(st_in), _ptr(st_out))
    return w, st_out


def sha256(data):
    buf = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else \
        np.ascontiguousarray(data, dtype=np.uint8)
    out = np.empty(32, dtype=np.uint8)
    lib().ora_sha256(_ptr(buf), buf.size, _ptr(out))
    return out.tobytes()


class Sha256:
    """Streaming SHA-256 (the per-utterance running digest)."""

    def __init__(self):
        self.ctx = Sha256Ctx()
        lib().ora_sha256_init(C.byref(self.ctx))

    def update(self, data):
        buf = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else \
            np.ascontiguousarray(data, dtype=np.uint8)
        lib().ora_sha256_update(C.byref(self.ctx), _ptr(buf), buf.size)
        return self

    def digest(self):
        out = np.empty(32, dtype=np.uint8)
        lib().ora_sha256_final(C.byref(self.ctx), _ptr(out))
        return out.tobytes()


def frame_features(samples, frame_len=320, vad_threshold=0.01):
    s = np.ascontiguousarray(samples, dtype=np.float32)
    nf = (s.size + frame_len - 1) // frame_len
    out = np.zeros(nf, dtype=FRAME_DTYPE)
    n = lib().ora_frame_features(_ptr(s), s.size, frame_len, vad_threshold, _ptr(out))
    return out[:n]


class Resampler:
    """N2: stateful 48 kHz stereo int16 -> 16 kHz mono f32 decimator for one stream."""

    def __init__(self):
        self.state = ResampleState()
        lib().ora_resample_init(C.byref(self.state))

    def process(self, data):
        buf = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else \
            np.ascontiguousarray(data, dtype=np.uint8)
        out = np.empty(buf.size // 4 // 3 + 2, dtype=np.float32)
        n = lib().ora_resample_48k_stereo(C.byref(self.state), _ptr(buf), buf.size, _ptr(out))
        return out[:n].copy()


def resample_taps():
    p = lib().ora_resample_taps()
    return np.ctypeslib.as_array(p, shape=(RESAMPLE_TAPS,)).copy()


def float32_to_wav(samples, sample_rate=16000):
    s = np.ascontiguousarray(samples, dtype=np.float32)
    out = np.empty(44 + 4 * s.size, dtype=np.uint8)
    n = lib().ora_float32_to_wav(_ptr(s), s.size, sample_rate, _ptr(out))
    return out[:n].tobytes()


def have_openssl():
    return bool(lib().ora_have_openssl())


def cpu_baseline(pcm, n_streams, n_msgs, msg_bytes, wake_msgs, group_size, n_threads,
                 use_openssl=True, want_outputs=False):
    """Timed CPU baseline over a time-major batch (see gpuaudio_oracle.h).
    Returns (seconds, digests|None, winners|None, scores|None)."""
    buf = np.ascontiguousarray(pcm, dtype=np.uint8)
    assert buf.size >= n_streams * n_msgs * msg_bytes
    dg = np.empty((n_msgs, n_streams, 32), dtype=np.uint8) if want_outputs else None
    wn = np.empty(n_streams // max(group_size, 1), dtype=np.int32) if want_outputs else None
    sc = np.empty(n_streams, dtype=np.float64) if want_outputs else None
    t = lib().ora_cpu_baseline(_ptr(buf), n_streams, n_msgs, msg_bytes, wake_msgs, group_size,
                               n_threads, 1 if use_openssl else 0, _ptr(dg), _ptr(wn), _ptr(sc))
    return t, dg, wn, sc


def cpu_baseline_run(pcm, n_streams, n_msgs, msg_bytes, wake_msgs, group_size, n_threads,
                     fmt=FMT_MONO16K, layout=LAYOUT_STREAM_MAJOR, sha_impl=SHA_OPENSSL_LOW,
                     want_outputs=False):
    """General form of the timed CPU baseline (gpuaudio_oracle.h: ora_cpu_baseline_run).
    Returns (seconds, digests [n_msgs][n_streams][32] | None, winners | None, scores | None)."""
    buf = np.ascontiguousarray(pcm).view(np.uint8).reshape(-1)
    assert buf.size >= n_streams * n_msgs * msg_bytes
    cfg = BaselineCfg(n_streams, n_msgs, msg_bytes, wake_msgs, group_size, fmt, layout, sha_impl, n_threads)
    dg = np.empty((n_msgs, n_streams, 32), dtype=np.uint8) if want_outputs else None
    wn = np.empty(n_streams // max(group_size, 1), dtype=np.int32) if want_outputs else None
    sc = np.empty(n_streams, dtype=np.float64) if want_outputs else None
    t = lib().ora_cpu_baseline_run(C.byref(cfg), _ptr(buf), _ptr(dg), _ptr(wn), _ptr(sc))
    return t, dg, wn, sc


def memcpy_gbs(n_threads, bytes_per_thread=64 << 20, reps=8):
    return lib().ora_memcpy_gbs(n_threads, bytes_per_thread, reps)


def synth_fill(group_base, n_streams, group_size, n_msgs, msg_samples, stereo48=False,
               layout=LAYOUT_STREAM_MAJOR, n_threads=1, seed=SYNTH_SEED):
    """int16 array of the synthetic batch (gpuaudio_oracle.h: ora_synth_fill); shape
    [n_streams][n_msgs][k] (stream-major) or [n_msgs][n_streams][k] (time-major), k = int16 per message."""
    k = msg_samples * (6 if stereo48 else 1)
    shape = (n_streams, n_msgs, k) if layout == LAYOUT_STREAM_MAJOR else (n_msgs, n_streams, k)
    out = np.empty(shape, dtype=np.int16)
    lib().ora_synth_fill(_ptr(out), seed, group_base, n_streams, group_size, n_msgs, msg_samples,
                         1 if stereo48 else 0, layout, n_threads)
    return out


def synth_sample(group, relay, n, utt_samples, seed=SYNTH_SEED):
    return int(lib().ora_synth_sample(seed, group, relay, n, utt_samples))

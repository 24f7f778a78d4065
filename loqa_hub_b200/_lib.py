"""ctypes binding of libgpuaudio.so (include/gpuaudio.h).

The library is the product; this module only declares its C ABI.  Loading fails
loudly when the shared object is missing — there is no Python or CPU fallback.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libgpuaudio.so")

GA_OK = 0
GA_ERR_ARG, GA_ERR_CUDA, GA_ERR_NOMEM, GA_ERR_RING_FULL = -1, -2, -3, -4
GA_ERR_OVERFLOW, GA_ERR_STATE, GA_ERR_COMM = -5, -6, -7
RELAY_CONNECTED, RELAY_CONTENDING, RELAY_WINNER, RELAY_CANCELLED = 0, 1, 2, 3
MSG_WAKE, MSG_EOS, MSG_FIRST = 1, 2, 4
DO_CONVERT, DO_FEATURES, DO_SHA256, DO_ALL = 1, 2, 4, 7
DO_UTTERANCE = 8
FRAME_SAMPLES = 320
NO_WINNER = 0xFFFFFFFF

RESULT_DTYPE = np.dtype([("digest", "u1", 32), ("sumsq", "<f8"), ("rms", "<f8"),
                         ("n_samples", "<u4"), ("nonzero", "<u4"), ("peak", "<f4"),
                         ("flags", "<u4")])
FRAME_DTYPE = np.dtype([("sumsq", "<f8"), ("rms", "<f8"), ("n_samples", "<u4"),
                        ("nonzero", "<u4"), ("peak", "<f4"), ("vad", "<u4")])
DESC_DTYPE = np.dtype([("in_off", "<u8"), ("out_off", "<u8"), ("nbytes", "<u4"),
                       ("stream_id", "<u4"), ("flags", "<u4"), ("frame_off", "<u4")])
MSG_DTYPE = np.dtype([("ring_offset", "<u8"), ("nbytes", "<u4"), ("stream_id", "<u4"),
                      ("flags", "<u4"), ("reserved", "<u4")])
assert RESULT_DTYPE.itemsize == 64 and FRAME_DTYPE.itemsize == 32
RDESC_DTYPE = np.dtype([("in_off", "<u8"), ("out_off", "<u8"), ("hist_off", "<u8"), ("n0", "<u8"),
                        ("nbytes", "<u4"), ("stream_id", "<u4"), ("flags", "<u4"), ("frame_off", "<u4"),
                        ("hist_kind", "<u4"), ("reserved", "<u4", 3)])
HIST_ZERO, HIST_SLAB, HIST_STATE = 0, 1, 2
RS_SAVE_HIST = 0x10000
FMT_MONO16K, FMT_STEREO48K = 0, 1
STAGING_AUTO, STAGING_BULK = 0, 1
assert DESC_DTYPE.itemsize == 32 and MSG_DTYPE.itemsize == 24 and RDESC_DTYPE.itemsize == 64


class P2PArgs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("d_samples", "d_wake_off", "d_wake_n", "d_status_in", "d_seg_offsets",
                                          "d_scores", "d_winner", "d_status_out", "d_guest_dst", "d_notify",
                                          "d_wait", "d_inbox", "d_route_dst", "d_counter", "d_error")] + \
               [("inbox_parity_stride", C.c_uint64), ("epoch", C.c_uint64), ("timeout_cycles", C.c_longlong)] + \
               [(n, C.c_uint32) for n in ("n_windows", "n_home_members", "n_guests", "n_notify", "n_wait", "n_in")]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_streams", C.c_uint32), ("max_batch_msgs", C.c_uint32),
                ("max_inflight", C.c_uint32), ("ring_bytes", C.c_uint64), ("arena_samples", C.c_uint64),
                ("wake_samples_max", C.c_uint64), ("mode", C.c_uint32), ("want_frames", C.c_uint32),
                ("vad_threshold", C.c_double), ("input_format", C.c_uint32), ("staging", C.c_uint32)]


# every symbol include/gpuaudio.h declares: name -> (restype, argtypes)
_vp, _u32, _u64, _i32, _sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_size_t
SYMBOLS = {
    "ga_k_ingest": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp, _u32, C.c_double, _vp]),
    "ga_k_arbitrate_p2p": (_i32, [C.POINTER(P2PArgs), _vp]),
    "ga_p2p_alloc": (_i32, [_u64, C.POINTER(_vp), _vp]),
    "ga_p2p_open": (_i32, [_vp, C.POINTER(_vp)]),
    "ga_p2p_close": (_i32, [_vp]),
    "ga_p2p_free": (_i32, [_vp]),
    "ga_k_sha256_chain": (_i32, [_vp, _vp, _vp, _vp, _u32, _vp, _vp]),
    "ga_k_sha256_final": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp]),
    "ga_k_sha256_reset": (_i32, [_vp, _u32, _vp]),
    "ga_samples_digest": (_i32, [_vp, _u32, _vp]),
    "ga_utterance_digest": (_i32, [_vp, _u32, _vp, C.POINTER(_u64)]),
    "ga_k_resample48": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp, _vp, C.c_double, _vp]),
    "ga_k_ingest_uniform": (_i32, [_vp, _vp, _u32, _u64, _u32, _u64, _vp, _vp, _vp, _u32, C.c_double, _vp]),
    "ga_k_arbitrate": (_i32, [_vp, _vp, _vp, _vp, _vp, _u32, _u32, _vp, _vp, _vp, _vp]),
    "ga_k_arbitrate_noalloc": (_i32, [_vp, _vp, _vp, _vp, _vp, _u32, _u32, _vp, _vp, _vp]),
    "ga_k_signal_strength": (_i32, [_vp, _vp, _vp, _vp, _u32, _vp, _vp, _vp]),
    "ga_k_select": (_i32, [_vp, _vp, _vp, _u32, _vp, _vp, _vp]),
    "ga_k_route_scores": (_i32, [_vp, _vp, _vp, _u32, _vp, _vp]),
    "ga_init": (_i32, [C.POINTER(Config), C.POINTER(_vp)]),
    "ga_shutdown": (None, [_vp]),
    "ga_last_error": (C.c_char_p, []),
    "ga_ring_reserve": (_i32, [_vp, _u64, C.POINTER(_vp), C.POINTER(_u64)]),
    "ga_submit": (_i32, [_vp, _vp, _u32, C.POINTER(_u64)]),
    "ga_set_audio_metadata": (_i32, [_vp, _vp, C.c_size_t, _i32, _i32, C.POINTER(C.c_double), C.c_char_p]),
    "ga_stream_audio_metadata": (_i32, [_vp, _u32, _i32, _i32, C.POINTER(C.c_double), C.c_char_p]),
    "ga_format_audio_hash": (None, [_vp, C.c_char_p]),
    "ga_cluster_size": (_u32, [_vp]),
    "ga_cluster_context": (_vp, [_vp, _u32]),
    "ga_hub_create_cluster": (_i32, [_vp, _u32, _u64, C.POINTER(_vp)]),
    "ga_hub_connect_on": (_i32, [_vp, C.c_char_p, _u32, _u32, C.POINTER(_u32)]),
    "ga_hub_relay_shard": (_i32, [_vp, C.c_char_p, C.POINTER(_u32)]),
    "ga_hub_fetch_wav": (_i32, [_vp, C.c_char_p, _u32, _vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "ga_cluster_create": (_i32, [_vp, _u32, _u32, C.POINTER(_vp)]),
    "ga_cluster_destroy": (None, [_vp]),
    "ga_cluster_arbitrate": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ga_submit_ticks": (_i32, [_vp, _u64, _u32, _u32, _u32, _vp, _vp, C.POINTER(_u64)]),
    "ga_wait": (_i32, [_vp, _u64, _vp, _vp]),
    "ga_wait_view": (_i32, [_vp, _u64, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_u32), C.POINTER(_u32)]),
    "ga_ticket_frames": (_i32, [_vp, _u64, C.POINTER(_u32)]),
    "ga_fetch_samples": (_i32, [_vp, _u32, _u64, _u64, _vp]),
    "ga_stream_info": (_i32, [_vp, _u32, C.POINTER(_u64), C.POINTER(_u64)]),
    "ga_stream_reset": (_i32, [_vp, _u32]),
    "ga_device_arena": (_i32, [_vp, C.POINTER(_vp)]),
    "ga_arbitrate": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp]),
    "ga_fetch_wav": (_i32, [_vp, _u32, _u32, _vp, _sz, C.POINTER(_sz)]),
    "ga_hub_create": (_i32, [_vp, _u32, _u64, C.POINTER(_vp)]),
    "ga_hub_destroy": (None, [_vp]),
    "ga_hub_connect": (_i32, [_vp, C.c_char_p, _u32, C.POINTER(_u32)]),
    "ga_hub_start_window": (_i32, [_vp, C.c_char_p, _u64, C.POINTER(_u64)]),
    "ga_hub_join_window": (_i32, [_vp, C.c_char_p, _u64, C.POINTER(_i32)]),
    "ga_hub_on_chunk": (_i32, [_vp, C.c_char_p, _vp, _u32, _i32, _i32, _u64, C.POINTER(_i32)]),
    "ga_fetch_wavs": (_i32, [_vp, _vp, _u32, _u32, C.POINTER(_vp), _vp, _vp]),
    "ga_samples_digests": (_i32, [_vp, _vp, _u32, _vp]),
    "ga_hub_poll_windows": (_i32, [_vp, _u64, _vp, _u32, C.POINTER(_u32)]),
    "ga_hub_fetch_winner_wavs": (_i32, [_vp, _vp, _u32, _u32, _vp, _vp, _vp]),
    "ga_hub_relay_handle": (_i32, [_vp, C.c_char_p, C.POINTER(_u64)]),
    "ga_hub_on_chunk_h": (_i32, [_vp, _u64, _vp, _u32, _i32, _i32, _u64, C.POINTER(_i32)]),
    "ga_hub_flush": (_i32, [_vp, C.POINTER(_u32)]),
    "ga_hub_perform_arbitration": (_i32, [_vp, _u64]),
    "ga_hub_poll": (_i32, [_vp, _u64, C.POINTER(_u32)]),
    "ga_hub_window_info": (_i32, [_vp, _u64, C.POINTER(_i32), C.c_char_p, _sz, C.POINTER(_u32)]),
    "ga_hub_active_window": (_i32, [_vp, _u32, C.POINTER(_u64)]),
    "ga_hub_relay_info": (_i32, [_vp, C.c_char_p, C.POINTER(_i32), C.POINTER(C.c_double), C.POINTER(_i32), C.POINTER(_u32)]),
    "ga_hub_is_relay_active": (_i32, [_vp, C.c_char_p]),
    "ga_hub_set_relay_status": (_i32, [_vp, C.c_char_p, _i32]),
    "ga_hub_complete_request": (_i32, [_vp, C.c_char_p]),
    "ga_hub_cleanup_relay": (_i32, [_vp, C.c_char_p]),
    "ga_hub_active_streams": (_i32, [_vp, C.POINTER(_u32)]),
    "ga_bytes_to_float32": (_i32, [_vp, _vp, _sz, _vp, C.POINTER(_sz), _vp]),
    "ga_signal_strength": (_i32, [_vp, _vp, _sz, C.POINTER(C.c_double), _vp]),
    "ga_sha256": (_i32, [_vp, _vp, _sz, _vp]),
    "ga_synth_fill_host": (_i32, [_vp, _u64, _u32, _u32, _u32, _u32, _u32]),
    "ga_synth_fill_batch_device": (_i32, [_vp, _u64, _u32, _u32, _u32, _u32, _u32, _vp]),
    "ga_synth_fill_batch_device_map": (_i32, [_vp, _u64, _vp, _vp, _u32, _u32, _u32, _vp]),
    "ga_synth_fill_batch48_device_map": (_i32, [_vp, _u64, _vp, _vp, _u32, _u32, _u32, _vp]),
    "ga_synth_fill_batch_host": (_i32, [_vp, _u64, _u32, _u32, _u32, _u32, _u32, _u32]),
    "ga_launch_count": (_u64, []),
    "ga_abi_version": (_i32, []),
    "ga_abi_sizes": (None, [C.POINTER(_u32 * 7)]),
    "ga_device_count": (_i32, []),
}

_lib = None


class GpuAudioError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libgpuaudio error %d: %s" % (code, msg))
        self.code = code


def load():
    """Load libgpuaudio.so and bind every declared symbol.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). libgpuaudio has no CPU fallback." % SO_PATH)
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)      # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    sizes = (_u32 * 7)()
    lib.ga_abi_sizes(C.byref(sizes))
    want = [RESULT_DTYPE.itemsize, FRAME_DTYPE.itemsize, DESC_DTYPE.itemsize, MSG_DTYPE.itemsize,
            C.sizeof(Config), RDESC_DTYPE.itemsize, 128]
    if list(sizes) != want:
        raise ImportError("libgpuaudio ABI struct sizes %s != binding %s" % (list(sizes), want))
    _lib = lib
    return lib


def check(rc):
    if rc != GA_OK:
        raise GpuAudioError(rc, load().ga_last_error().decode("utf-8", "replace"))


def ptr(a):
    """void* of a numpy array (None for empty / None)."""
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p) if a.size else None

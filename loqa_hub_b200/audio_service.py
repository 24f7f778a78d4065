"""Python face of the C++ hub (csrc/hub.cu): the reference's AudioService for the ingest path.

Method names follow internal/grpc/audio_service.go so the tests in tests/test_gpu_hub.py read
like internal/grpc/collision_detection_test.go.  Every method is one ga_hub_* call."""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib as L
from . import api

RelayStatusConnected, RelayStatusContending, RelayStatusWinner, RelayStatusCancelled = 0, 1, 2, 3


@dataclass
class WindowInfo:
    is_active: bool
    winner_id: str
    n_relays: int


@dataclass
class RelayInfo:
    status: int
    signal_strength: float
    cancel_closed: bool
    slot: int


class AudioService:
    def __init__(self, ctx, max_streams, window_duration_us=0):
        """ctx: an api.Context, None for the bookkeeping-only hub the CPU tests use, or an
        api.Cluster — one hub over every GPU of the box (max_streams is then per GPU)."""
        self._h = C.c_void_p()
        if isinstance(ctx, api.Cluster):
            self.cluster, self.ctx = ctx, None
            L.check(L.load().ga_hub_create_cluster(ctx._h, max_streams, window_duration_us, C.byref(self._h)))
            return
        self.cluster, self.ctx = None, ctx
        L.check(L.load().ga_hub_create(ctx._h if ctx is not None else None, max_streams, window_duration_us,
                                       C.byref(self._h)))

    def connect_on(self, relay_id, group, shard):
        """ga_hub_connect_on: the relay's audio lives on GPU `shard` whatever its group."""
        slot = C.c_uint32()
        L.check(L.load().ga_hub_connect_on(self._h, relay_id.encode(), group, shard, C.byref(slot)))
        return slot.value

    def relay_shard(self, relay_id):
        sh = C.c_uint32()
        L.check(L.load().ga_hub_relay_shard(self._h, relay_id.encode(), C.byref(sh)))
        return sh.value

    def relay_context(self, relay_id):
        return self.cluster.contexts[self.relay_shard(relay_id)] if self.cluster else self.ctx

    def close(self):
        if self._h:
            L.load().ga_hub_destroy(self._h)
            self._h = C.c_void_p()

    def connect(self, relay_id, group=0):
        slot = C.c_uint32()
        L.check(L.load().ga_hub_connect(self._h, relay_id.encode(), group, C.byref(slot)))
        return slot.value

    def start_arbitration_window(self, relay_id, now_us):
        w = C.c_uint64()
        L.check(L.load().ga_hub_start_window(self._h, relay_id.encode(), now_us, C.byref(w)))
        return w.value

    def join_arbitration_window(self, relay_id, now_us):
        j = C.c_int()
        L.check(L.load().ga_hub_join_window(self._h, relay_id.encode(), now_us, C.byref(j)))
        return bool(j.value)

    def on_chunk(self, relay_id, audio, is_wake_word, is_end_of_speech, now_us):
        b = np.frombuffer(bytes(audio), dtype=np.uint8)
        keep = C.c_int()
        L.check(L.load().ga_hub_on_chunk(self._h, relay_id.encode(), L.ptr(b), b.size, int(is_wake_word),
                                         int(is_end_of_speech), now_us, C.byref(keep)))
        return bool(keep.value)

    def relay_handle(self, relay_id):
        hd = C.c_uint64()
        L.check(L.load().ga_hub_relay_handle(self._h, relay_id.encode(), C.byref(hd)))
        return hd.value

    def on_chunk_h(self, handle, audio, is_wake_word, is_end_of_speech, now_us):
        """ga_hub_on_chunk_h: the per-chunk path by handle (no name lookup, no hub lock for speech)."""
        b = audio if isinstance(audio, np.ndarray) else np.frombuffer(bytes(audio), dtype=np.uint8)
        keep = C.c_int()
        L.check(L.load().ga_hub_on_chunk_h(self._h, handle, L.ptr(b), b.size, int(is_wake_word),
                                           int(is_end_of_speech), now_us, C.byref(keep)))
        return bool(keep.value)

    def flush(self):
        n = C.c_uint32()
        L.check(L.load().ga_hub_flush(self._h, C.byref(n)))
        return n.value

    def perform_arbitration(self, window_id):
        L.check(L.load().ga_hub_perform_arbitration(self._h, window_id))

    def poll(self, now_us):
        n = C.c_uint32()
        L.check(L.load().ga_hub_poll(self._h, now_us, C.byref(n)))
        return n.value

    def poll_windows(self, now_us, cap=65536):
        """ga_hub_poll_windows: ids of the windows this poll closed."""
        ids = np.zeros(cap, dtype=np.uint64)
        n = C.c_uint32()
        L.check(L.load().ga_hub_poll_windows(self._h, now_us, L.ptr(ids), cap, C.byref(n)))
        return [int(x) for x in ids[:min(n.value, cap)]]

    def fetch_winner_wavs(self, window_ids, sample_rate=16000, hashes=True):
        """ga_hub_fetch_winner_wavs: [(wav bytes | None, "sha256:<hex>" | "")] per window."""
        ids = np.ascontiguousarray(window_ids, dtype=np.uint64)
        n = ids.size
        ptrs = (C.c_void_p * max(n, 1))()
        sizes = np.zeros(max(n, 1), dtype=np.uint64)
        hs = C.create_string_buffer(72 * max(n, 1)) if hashes else None
        L.check(L.load().ga_hub_fetch_winner_wavs(self._h, L.ptr(ids), n, sample_rate, ptrs, L.ptr(sizes), hs))
        out = []
        for k in range(n):
            wav = C.string_at(ptrs[k], int(sizes[k])) if ptrs[k] else None
            out.append((wav, hs.raw[72 * k:72 * (k + 1)].split(b"\0")[0].decode() if hashes else ""))
        return out

    def window(self, window_id):
        a, n = C.c_int(), C.c_uint32()
        buf = C.create_string_buffer(256)
        L.check(L.load().ga_hub_window_info(self._h, window_id, C.byref(a), buf, 256, C.byref(n)))
        return WindowInfo(bool(a.value), buf.value.decode(), n.value)

    def arbitration_window(self, group=0):
        w = C.c_uint64()
        L.check(L.load().ga_hub_active_window(self._h, group, C.byref(w)))
        return w.value or None

    def relay(self, relay_id):
        st, cc, slot, sig = C.c_int(), C.c_int(), C.c_uint32(), C.c_double()
        L.check(L.load().ga_hub_relay_info(self._h, relay_id.encode(), C.byref(st), C.byref(sig), C.byref(cc), C.byref(slot)))
        return RelayInfo(st.value, sig.value, bool(cc.value), slot.value)

    def is_relay_active(self, relay_id):
        r = L.load().ga_hub_is_relay_active(self._h, relay_id.encode())
        if r < 0:
            L.check(r)
        return bool(r)

    def set_relay_status(self, relay_id, status):
        L.check(L.load().ga_hub_set_relay_status(self._h, relay_id.encode(), status))

    def complete_request(self, relay_id):
        L.check(L.load().ga_hub_complete_request(self._h, relay_id.encode()))

    def cleanup_relay(self, relay_id):
        L.check(L.load().ga_hub_cleanup_relay(self._h, relay_id.encode()))

    def active_streams(self):
        n = C.c_uint32()
        L.check(L.load().ga_hub_active_streams(self._h, C.byref(n)))
        return n.value

    def fetch_wav(self, relay_id, sample_rate=16000):
        n = C.c_size_t()
        L.check(L.load().ga_hub_fetch_wav(self._h, relay_id.encode(), sample_rate, None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=np.uint8)
        L.check(L.load().ga_hub_fetch_wav(self._h, relay_id.encode(), sample_rate, L.ptr(out), out.size, C.byref(n)))
        return out.tobytes()

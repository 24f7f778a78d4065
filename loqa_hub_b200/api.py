"""numpy-facing wrappers over the libgpuaudio C ABI (tests, bench and smoke call these).

Nothing here computes: every function marshals arrays and calls through ctypes.
`Context` wraps the host-buffer layer (ga_init / ga_ring_reserve / ga_submit / ga_wait /
ga_arbitrate and the one-shot mirrors); the `k_*` functions call the device-pointer
layer with raw CUDA addresses (e.g. torch tensors' data_ptr()).
"""
import ctypes as C

import numpy as np

from . import _lib as L
from ._lib import (DESC_DTYPE, FRAME_DTYPE, MSG_DTYPE, RESULT_DTYPE, Config, GpuAudioError,  # noqa: F401
                   check, load, ptr)


def make_descs(in_off, nbytes, out_off, stream_id=None, flags=None, frame_off=None):
    n = len(nbytes)
    d = np.zeros(n, dtype=DESC_DTYPE)
    d["in_off"] = in_off
    d["nbytes"] = nbytes
    d["out_off"] = out_off
    if stream_id is not None:
        d["stream_id"] = stream_id
    if flags is not None:
        d["flags"] = flags
    if frame_off is None:
        ns = np.asarray(nbytes, dtype=np.uint64) // 2
        nf = (ns + L.FRAME_SAMPLES - 1) // L.FRAME_SAMPLES
        frame_off = np.concatenate([[0], np.cumsum(nf)[:-1]]) if n else np.zeros(0)
    d["frame_off"] = frame_off
    return d


def k_ingest(d_pcm, d_desc, n_msgs, d_samples, d_results, d_frames=0, mode=L.DO_ALL,
             vad_threshold=0.01, stream=0):
    check(load().ga_k_ingest(d_pcm, d_desc, n_msgs, d_samples or None, d_results, d_frames or None,
                             mode, vad_threshold, stream or None))


def k_sha256_chain(d_pcm, d_desc, d_run_first, d_next, n_runs, d_state, stream=0):
    check(load().ga_k_sha256_chain(d_pcm, d_desc, d_run_first, d_next, n_runs, d_state, stream or None))


def k_sha256_final(d_state, d_stream_ids, n, d_digests, d_lengths=0, stream=0):
    check(load().ga_k_sha256_final(d_state, d_stream_ids or None, n, d_digests, d_lengths or None, stream or None))


def k_sha256_reset(d_state, n, stream=0):
    check(load().ga_k_sha256_reset(d_state, n, stream or None))


def k_resample48(d_pcm, d_rdesc, n_msgs, d_hist_state, d_samples, d_results, d_frames=0,
                 vad_threshold=0.01, stream=0):
    check(load().ga_k_resample48(d_pcm, d_rdesc, n_msgs, d_hist_state or None, d_samples, d_results,
                                 d_frames or None, vad_threshold, stream or None))


def k_ingest_uniform(d_pcm, d_desc, n_msgs, first_in_off, msg_bytes, msg_stride, d_samples, d_results,
                     d_frames=0, mode=L.DO_ALL, vad_threshold=0.01, stream=0):
    check(load().ga_k_ingest_uniform(d_pcm, d_desc, n_msgs, first_in_off, msg_bytes, msg_stride,
                                     d_samples or None, d_results, d_frames or None, mode, vad_threshold,
                                     stream or None))


def k_arbitrate(d_samples, d_wake_off, d_wake_n, d_status_in, d_seg, n_windows, n_members,
                d_scores, d_winner, d_status_out, stream=0):
    check(load().ga_k_arbitrate(d_samples, d_wake_off, d_wake_n, d_status_in, d_seg, n_windows,
                                n_members, d_scores, d_winner, d_status_out, stream or None))


def k_arbitrate_noalloc(d_samples, d_wake_off, d_wake_n, d_status_in, d_seg, n_windows, n_members,
                        d_winner, d_status_out, stream=0):
    check(load().ga_k_arbitrate_noalloc(d_samples, d_wake_off, d_wake_n, d_status_in, d_seg, n_windows,
                                        n_members, d_winner, d_status_out, stream or None))


def k_signal_strength(d_samples, d_wake_off, d_wake_n, d_status_in, n_members, d_scores,
                      d_stats=0, stream=0):
    check(load().ga_k_signal_strength(d_samples, d_wake_off, d_wake_n, d_status_in or None,
                                      n_members, d_scores, d_stats or None, stream or None))


def k_select(d_scores, d_status_in, d_seg, n_windows, d_winner, d_status_out, stream=0):
    check(load().ga_k_select(d_scores, d_status_in, d_seg, n_windows, d_winner, d_status_out,
                             stream or None))


def k_route_scores(d_gathered, d_src_idx, d_dst_slot, n, d_scores, stream=0):
    check(load().ga_k_route_scores(d_gathered, d_src_idx, d_dst_slot, n, d_scores, stream or None))


def synth_batch_device_map(d_pcm, seed, d_group, d_relay, n_streams, n_msgs, msg_samples, stream=0):
    check(load().ga_synth_fill_batch_device_map(d_pcm, seed, d_group, d_relay, n_streams, n_msgs,
                                                msg_samples, stream or None))


def synth_batch48_device_map(d_pcm, seed, d_group, d_relay, n_streams, n_msgs, msg_frames, stream=0):
    check(load().ga_synth_fill_batch48_device_map(d_pcm, seed, d_group, d_relay, n_streams, n_msgs,
                                                  msg_frames, stream or None))


def launch_count():
    return int(load().ga_launch_count())


def device_count():
    return int(load().ga_device_count())


def synth_host(seed, group, relay, first, n_samples, utt_samples):
    out = np.empty(n_samples, dtype=np.int16)
    check(load().ga_synth_fill_host(ptr(out), seed, group, relay, first, n_samples, utt_samples))
    return out


def synth_batch_host(seed, group_base, n_streams, group_size, n_msgs, msg_samples, n_threads=1,
                     out=None):
    """Time-major batch: message m of stream s at [(m*n_streams + s) * msg_samples]."""
    if out is None:
        out = np.empty(n_msgs * n_streams * msg_samples, dtype=np.int16)
    check(load().ga_synth_fill_batch_host(ptr(out), seed, group_base, n_streams, group_size,
                                          n_msgs, msg_samples, n_threads))
    return out


def synth_batch_device(d_pcm, seed, group_base, n_streams, group_size, n_msgs, msg_samples,
                       stream=0):
    check(load().ga_synth_fill_batch_device(d_pcm, seed, group_base, n_streams, group_size,
                                            n_msgs, msg_samples, stream or None))


class Context:
    """One GPU's ingest context (ga_ctx)."""

    def __init__(self, device=0, max_streams=1024, max_batch_msgs=65536, max_inflight=4,
                 ring_bytes=64 << 20, arena_samples=80_000, wake_samples_max=8_000,
                 mode=L.DO_ALL, want_frames=False, vad_threshold=0.01, input_format=L.FMT_MONO16K,
                 staging=L.STAGING_AUTO):
        self.cfg = Config(device, max_streams, max_batch_msgs, max_inflight, ring_bytes,
                          arena_samples, wake_samples_max, mode, 1 if want_frames else 0,
                          vad_threshold, input_format, staging)
        self._h = C.c_void_p()
        check(load().ga_init(C.byref(self.cfg), C.byref(self._h)))
        self._pending = {}

    def close(self):
        if self._h:
            load().ga_shutdown(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- pinned ring
    def reserve(self, nbytes):
        """Returns (uint8 view of the pinned ring slot, virtual ring offset)."""
        p, off = C.c_void_p(), C.c_uint64()
        check(load().ga_ring_reserve(self._h, nbytes, C.byref(p), C.byref(off)))
        padded = max(16, (nbytes + 15) & ~15)
        view = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(padded,))
        return view, off.value

    def push(self, data):
        """Reserve + copy one chunk; returns its ring offset (what the Go shim does per chunk)."""
        b = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else \
            data.view(np.uint8).reshape(-1)
        view, off = self.reserve(b.size)
        view[:b.size] = b
        return off

    # ---- batches
    def submit(self, msgs):
        m = np.ascontiguousarray(msgs, dtype=MSG_DTYPE)
        t = C.c_uint64()
        check(load().ga_submit(self._h, ptr(m), m.size, C.byref(t)))
        self._pending[t.value] = m.size
        return t.value

    def submit_ticks(self, ring_offset, n_ticks, stream_ids, chunk_bytes, flags=None):
        """ga_submit_ticks: n_ticks periods x len(stream_ids) relays, message t*S+j at
        ring_offset + (t*S+j)*round16(chunk_bytes); flags is a u8 array [n_ticks*S] or None."""
        ids = np.ascontiguousarray(stream_ids, dtype=np.uint32)
        n = int(n_ticks) * ids.size
        fl = None
        if flags is not None:
            fl = np.ascontiguousarray(flags, dtype=np.uint8)
            if fl.size != n:
                raise ValueError("flags must hold n_ticks * n_streams bytes")
        t = C.c_uint64()
        check(load().ga_submit_ticks(self._h, int(ring_offset), int(n_ticks), ids.size, int(chunk_bytes),
                                     ptr(ids), ptr(fl) if fl is not None else None, C.byref(t)))
        self._pending[t.value] = n
        return t.value

    def wait(self, ticket):
        n = self._pending.pop(ticket)
        nf = C.c_uint32()
        check(load().ga_ticket_frames(self._h, ticket, C.byref(nf)))
        res = np.zeros(n, dtype=RESULT_DTYPE)
        frames = np.zeros(nf.value, dtype=FRAME_DTYPE)
        check(load().ga_wait(self._h, ticket, ptr(res), ptr(frames)))
        return res, frames

    def wait_view(self, ticket):
        """Zero-copy wait: numpy views of the pinned result buffers (ticket t's records are
        overwritten by the submit that returns ticket t + max_inflight)."""
        self._pending.pop(ticket)
        pr, pf, nr, nf = C.c_void_p(), C.c_void_p(), C.c_uint32(), C.c_uint32()
        check(load().ga_wait_view(self._h, ticket, C.byref(pr), C.byref(pf), C.byref(nr), C.byref(nf)))
        res = np.ctypeslib.as_array(C.cast(pr, C.POINTER(C.c_uint8)), shape=(nr.value * 64,)).view(RESULT_DTYPE) \
            if nr.value else np.zeros(0, RESULT_DTYPE)
        frames = np.ctypeslib.as_array(C.cast(pf, C.POINTER(C.c_uint8)), shape=(nf.value * 32,)).view(FRAME_DTYPE) \
            if nf.value else np.zeros(0, FRAME_DTYPE)
        return res, frames

    def ingest(self, chunks, stream_ids, flags=None):
        """Convenience: push + submit + wait a list of byte chunks."""
        n = len(chunks)
        msgs = np.zeros(n, dtype=MSG_DTYPE)
        for i, c in enumerate(chunks):
            msgs["ring_offset"][i] = self.push(c)
            msgs["nbytes"][i] = len(c)
        msgs["stream_id"] = stream_ids
        if flags is not None:
            msgs["flags"] = flags
        return self.wait(self.submit(msgs))

    # ---- streams
    def stream_info(self, stream_id):
        a, b = C.c_uint64(), C.c_uint64()
        check(load().ga_stream_info(self._h, stream_id, C.byref(a), C.byref(b)))
        return a.value, b.value

    def utterance_digest(self, stream_id):
        """(digest bytes, bytes hashed) of the stream's running per-utterance SHA-256."""
        out = np.zeros(32, dtype=np.uint8)
        n = C.c_uint64()
        check(load().ga_utterance_digest(self._h, stream_id, ptr(out), C.byref(n)))
        return out.tobytes(), n.value

    def samples_digest(self, stream_id):
        out = np.zeros(32, dtype=np.uint8)
        check(load().ga_samples_digest(self._h, stream_id, ptr(out)))
        return out.tobytes()

    def samples_digests(self, stream_ids):
        ids = np.ascontiguousarray(stream_ids, dtype=np.uint32)
        out = np.zeros((ids.size, 32), dtype=np.uint8)
        check(load().ga_samples_digests(self._h, ptr(ids), ids.size, ptr(out)))
        return [out[i].tobytes() for i in range(ids.size)]

    def fetch_wavs(self, stream_ids, sample_rate=16000, digests=True):
        """ga_fetch_wavs: (list of WAV byte strings, list of 32-byte digests or None) for many streams."""
        ids = np.ascontiguousarray(stream_ids, dtype=np.uint32)
        base = C.c_void_p()
        off = np.zeros(ids.size + 1, dtype=np.uint64)
        dg = np.zeros((ids.size, 32), dtype=np.uint8) if digests else None
        check(load().ga_fetch_wavs(self._h, ptr(ids), ids.size, sample_rate, C.byref(base), off.ctypes.data_as(C.c_void_p),
                                   ptr(dg) if digests else None))
        wavs = []
        for k in range(ids.size):
            hdr = C.string_at(base.value + int(off[k]), 8)
            size = 8 + int.from_bytes(hdr[4:8], "little")
            wavs.append(C.string_at(base.value + int(off[k]), size))
        return wavs, ([dg[i].tobytes() for i in range(ids.size)] if digests else None)

    def stream_reset(self, stream_id):
        check(load().ga_stream_reset(self._h, stream_id))

    def fetch_samples(self, stream_id, first=0, n=None):
        if n is None:
            n = self.stream_info(stream_id)[0] - first
        out = np.empty(n, dtype=np.float32)
        check(load().ga_fetch_samples(self._h, stream_id, first, n, ptr(out)))
        return out

    # ---- arbitration
    def arbitrate(self, seg_offsets, stream_ids, status_in=None, scores_in=None):
        seg = np.ascontiguousarray(seg_offsets, dtype=np.uint32)
        nw = seg.size - 1
        sid = np.ascontiguousarray(stream_ids, dtype=np.uint32)
        nm = sid.size
        st = np.full(nm, L.RELAY_CONTENDING, np.uint8) if status_in is None else \
            np.ascontiguousarray(status_in, dtype=np.uint8)
        scores = np.zeros(nm, np.float64) if scores_in is None else \
            np.array(scores_in, dtype=np.float64)
        winner = np.zeros(max(nw, 0), np.uint32)
        st_out = np.zeros(nm, np.uint8)
        check(load().ga_arbitrate(self._h, ptr(seg), max(nw, 0), ptr(sid), ptr(st), ptr(winner),
                                  ptr(scores), ptr(st_out)))
        return winner, scores, st_out

    # ---- one-shot mirrors of the Go functions
    def bytes_to_float32(self, data):
        b = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else \
            np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        out = np.empty(b.size // 2, dtype=np.float32)
        n = C.c_size_t()
        res = np.zeros(1, dtype=RESULT_DTYPE)
        check(load().ga_bytes_to_float32(self._h, ptr(b), b.size, ptr(out), C.byref(n), ptr(res)))
        return out[:n.value], res[0]

    def signal_strength(self, samples):
        s = np.ascontiguousarray(samples, dtype=np.float32)
        rms = C.c_double()
        res = np.zeros(1, dtype=RESULT_DTYPE)
        check(load().ga_signal_strength(self._h, ptr(s), s.size, C.byref(rms), ptr(res)))
        return rms.value, res[0]

    def sha256(self, data):
        b = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else \
            np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        out = np.zeros(32, dtype=np.uint8)
        check(load().ga_sha256(self._h, ptr(b), b.size, ptr(out)))
        return out.tobytes()


class Cluster:
    """ga_cluster: several contexts of this process, one per GPU; arbitration of windows whose
    relays sit on different GPUs in one fused launch per GPU (peer stores over NVLink)."""

    def __init__(self, contexts, max_guests=4096):
        self.contexts = list(contexts)
        arr = (C.c_void_p * len(self.contexts))(*[c._h for c in self.contexts])
        self._h = C.c_void_p()
        check(load().ga_cluster_create(arr, len(self.contexts), max_guests, C.byref(self._h)))

    def close(self):
        if self._h:
            load().ga_cluster_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def arbitrate(self, seg_offsets, shard_ids, stream_ids, status_in=None, scores_in=None):
        seg = np.ascontiguousarray(seg_offsets, dtype=np.uint32)
        nw = seg.size - 1
        sh = np.ascontiguousarray(shard_ids, dtype=np.uint32)
        sid = np.ascontiguousarray(stream_ids, dtype=np.uint32)
        nm = sid.size
        st = np.full(nm, L.RELAY_CONTENDING, np.uint8) if status_in is None else \
            np.ascontiguousarray(status_in, dtype=np.uint8)
        scores = np.zeros(nm, np.float64) if scores_in is None else np.array(scores_in, dtype=np.float64)
        winner = np.zeros(max(nw, 0), np.uint32)
        st_out = np.zeros(nm, np.uint8)
        check(load().ga_cluster_arbitrate(self._h, ptr(seg), max(nw, 0), ptr(sh), ptr(sid), ptr(st), ptr(winner),
                                          ptr(scores), ptr(st_out)))
        return winner, scores, st_out

"""ga_submit_ticks — the regular tick batch whose per-message descriptors are built on the GPU —
must be indistinguishable from ga_submit of the same messages: same result records, same samples
in the stream arenas, same stream cursors; and a refused call must leave the context untouched."""
import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


def make(**kw):
    import __graft_entry__ as entry
    entry.build()
    args = dict(device=0, max_streams=96, max_batch_msgs=4096, max_inflight=2, ring_bytes=8 << 20,
                arena_samples=40_000, wake_samples_max=8_000, mode=L.DO_ALL, want_frames=True)
    args.update(kw)
    return api.Context(**args)


def lay_out(c, pcm, n, chunk):
    """Copy n chunks into one reservation at a stride of round16(chunk)."""
    stride = (chunk + 15) & ~15
    view, off = c.reserve(max(stride * n, 1))
    for m in range(n):
        view[m * stride:m * stride + chunk] = pcm[m]
    return off, stride


def per_message(off, stride, chunk, ids, flags, n_ticks):
    S = len(ids)
    msgs = np.zeros(n_ticks * S, dtype=L.MSG_DTYPE)
    k = np.arange(n_ticks * S, dtype=np.uint64)
    msgs["ring_offset"] = off + k * np.uint64(stride)
    msgs["nbytes"] = chunk
    msgs["stream_id"] = np.tile(np.asarray(ids, dtype=np.uint32), n_ticks)
    msgs["flags"] = 0 if flags is None else flags
    return msgs


def same_records(a, b):
    for f in ("digest", "sumsq", "rms", "n_samples", "nonzero", "peak", "flags"):
        assert np.array_equal(a[f], b[f]), f


@pytest.mark.parametrize("chunk,n_ticks,with_flags", [(640, 5, True), (1001, 3, True), (6400, 4, False),
                                                      (2, 7, True), (0, 2, True), (3838, 1, True)])
def test_matches_per_message_submit(chunk, n_ticks, with_flags):
    rng = np.random.default_rng(chunk * 31 + n_ticks)
    ids = rng.permutation(96)[:37].astype(np.uint32)           # a scattered subset, not 0..S-1
    S, n = ids.size, n_ticks * ids.size
    pcm = rng.integers(0, 256, size=(n, chunk), dtype=np.uint8)
    flags = None
    if with_flags:
        flags = np.zeros((n_ticks, S), dtype=np.uint8)
        flags[:2] |= L.MSG_WAKE                                 # wake-word chunks first ...
        flags[0] |= L.MSG_FIRST
        flags[-1] |= L.MSG_EOS
        odd = rng.random(S) < 0.3                               # ... and some relays keep sending wake audio
        if n_ticks > 2:
            flags[2, odd] |= L.MSG_WAKE
        flags = flags.reshape(-1)
    with make() as a, make() as b:
        for round_ in range(2):                                 # second round: cursors carried over
            fl = flags
            if round_ == 1 and flags is not None:
                fl = flags & ~np.uint8(L.MSG_FIRST)             # keep appending
            off_a, stride = lay_out(a, pcm, n, chunk)
            off_b, _ = lay_out(b, pcm, n, chunk)
            ra, fa = a.wait(a.submit_ticks(off_a, n_ticks, ids, chunk, fl))
            rb, fb = b.wait(b.submit(per_message(off_b, stride, chunk, ids, fl, n_ticks)))
            same_records(ra, rb)
            assert fa.tobytes() == fb.tobytes()
            for m in (0, n // 2, n - 1):                        # and against the CPU checker directly
                assert ra["digest"][m].tobytes() == ora.sha256(pcm[m].tobytes())
                samples, _ = ora.bytes_to_float32(pcm[m].tobytes())
                rms, st = ora.signal_strength(samples)
                assert (ra["sumsq"][m], ra["rms"][m]) == (st.sumsq, rms)
        for sid in ids:
            assert a.stream_info(int(sid)) == b.stream_info(int(sid))
            assert a.fetch_samples(int(sid)).tobytes() == b.fetch_samples(int(sid)).tobytes()


def test_refused_calls_leave_the_context_as_it_was():
    with make(max_streams=8, arena_samples=2_400, wake_samples_max=400, want_frames=False) as c:
        ids = np.arange(4, dtype=np.uint32)
        pcm = np.random.default_rng(3).integers(0, 256, size=(40, 400), dtype=np.uint8)
        off, _ = lay_out(c, pcm, 8, 400)
        c.wait(c.submit_ticks(off, 2, ids, 400))                # 400 speech samples per stream
        before = [c.stream_info(i) for i in range(8)]
        off, _ = lay_out(c, pcm, 40, 400)
        with pytest.raises(L.GpuAudioError) as e:               # 10 more periods overflow the speech arena
            c.submit_ticks(off, 10, ids, 400)
        assert e.value.code == L.GA_ERR_OVERFLOW
        with pytest.raises(L.GpuAudioError) as e:               # the wake-word region holds two chunks
            c.submit_ticks(off, 3, ids, 400, np.full(12, L.MSG_WAKE, np.uint8))
        assert e.value.code == L.GA_ERR_OVERFLOW
        with pytest.raises(L.GpuAudioError) as e:
            c.submit_ticks(off, 2, np.array([0, 1, 1, 2], np.uint32), 400)
        assert e.value.code == L.GA_ERR_ARG and "twice" in str(e.value)
        with pytest.raises(L.GpuAudioError) as e:
            c.submit_ticks(off, 2, np.array([0, 1, 8, 2], np.uint32), 400)
        assert e.value.code == L.GA_ERR_ARG
        with pytest.raises(L.GpuAudioError) as e:               # beyond what was reserved
            c.submit_ticks(off, 11, ids, 400)
        assert e.value.code == L.GA_ERR_ARG
        with pytest.raises(L.GpuAudioError) as e:
            c.submit_ticks(off + 8, 2, ids, 400)
        assert e.value.code == L.GA_ERR_ARG
        assert [c.stream_info(i) for i in range(8)] == before
        res, _ = c.wait(c.submit_ticks(off, 8, ids, 400))       # the same reservation is still usable
        assert res["digest"][5].tobytes() == ora.sha256(pcm[5].tobytes())
        assert c.stream_info(0) == (2000, 0)
        assert np.array_equal(c.fetch_samples(0, 400, 200), ora.bytes_to_float32(pcm[0].tobytes())[0])


@pytest.mark.parametrize("mode", [L.DO_ALL | L.DO_UTTERANCE, L.DO_UTTERANCE])
def test_running_utterance_digests_through_tick_batches(mode):
    """GA_DO_UTTERANCE contexts: the per-stream lists the digest kernel walks are laid out on the GPU
    too; digests after several tick batches (a restart in between) equal hashlib over the bytes."""
    import hashlib
    rng = np.random.default_rng(77)
    ids = np.array([9, 2, 40, 17, 5], dtype=np.uint32)
    S, chunk = ids.size, 650                                    # 650 B: the carried tail is exercised
    got = {int(s): b"" for s in ids}
    with make(mode=mode, want_frames=False) as c:
        for rnd, n_ticks in enumerate((3, 1, 4, 2)):
            n = n_ticks * S
            pcm = rng.integers(0, 256, size=(n, chunk), dtype=np.uint8)
            flags = np.zeros((n_ticks, S), np.uint8)
            if rnd == 0:
                flags[0] = L.MSG_FIRST
            if rnd == 2:
                flags[0, 1] = L.MSG_FIRST                        # stream 2 starts a new utterance
                got[2] = b""
            off, _ = lay_out(c, pcm, n, chunk)
            res, _ = c.wait(c.submit_ticks(off, n_ticks, ids, chunk, flags.reshape(-1)))
            for t in range(n_ticks):
                for j, sid in enumerate(ids):
                    got[int(sid)] += pcm[t * S + j].tobytes()
            if mode & L.DO_SHA256:
                assert res["digest"][n - 1].tobytes() == hashlib.sha256(pcm[n - 1].tobytes()).digest()
            for sid in ids:
                d, nbytes = c.utterance_digest(int(sid))
                assert (bytes(d), nbytes) == (hashlib.sha256(got[int(sid)]).digest(), len(got[int(sid)])), (rnd, sid)


@pytest.mark.parametrize("chunk,utter", [(3840, False), (1920, True), (192, False)])
def test_48k_tick_batches_match_per_message_submit(chunk, utter):
    """48 kHz stereo contexts: the decimator's descriptors (position on the stream's frame timeline,
    where the 47 frames of history come from, who saves it) are built on the GPU as well.  Two tick
    batches, then a ragged chunk through ga_submit on both contexts: the carried state must agree."""
    rng = np.random.default_rng(chunk)
    ids = np.array([7, 1, 30, 12], dtype=np.uint32)
    S = ids.size
    mode = L.DO_ALL | (L.DO_UTTERANCE if utter else 0)
    kw = dict(mode=mode, input_format=L.FMT_STEREO48K, arena_samples=40_000, wake_samples_max=8_000)
    with make(**kw) as a, make(**kw) as b:
        for rnd, n_ticks in enumerate((3, 2)):
            n = n_ticks * S
            pcm = rng.integers(0, 256, size=(n, chunk), dtype=np.uint8)
            flags = np.zeros((n_ticks, S), np.uint8)
            flags[:1] |= L.MSG_WAKE
            if rnd == 0:
                flags[0] |= L.MSG_FIRST
            off_a, stride = lay_out(a, pcm, n, chunk)
            off_b, _ = lay_out(b, pcm, n, chunk)
            ra, fa = a.wait(a.submit_ticks(off_a, n_ticks, ids, chunk, flags.reshape(-1)))
            rb, fb = b.wait(b.submit(per_message(off_b, stride, chunk, ids, flags.reshape(-1), n_ticks)))
            same_records(ra, rb)
            assert fa.tobytes() == fb.tobytes()
        tail = rng.integers(0, 256, size=4 * 101 + 3, dtype=np.uint8).tobytes()      # ragged, history from the saved state
        ra, _ = a.ingest([tail] * S, ids, [0] * S)
        rb, _ = b.ingest([tail] * S, ids, [0] * S)
        same_records(ra, rb)
        for sid in ids:
            assert a.stream_info(int(sid)) == b.stream_info(int(sid))
            assert a.fetch_samples(int(sid)).tobytes() == b.fetch_samples(int(sid)).tobytes()
            if utter:
                assert a.utterance_digest(int(sid)) == b.utterance_digest(int(sid))
        assert a.stream_info(7)[0] == 5 * (chunk // 12) + 33        # 5 tick chunks + the 101-frame tail


def test_48k_tick_batches_need_whole_groups_of_three_frames():
    with make(input_format=L.FMT_STEREO48K, arena_samples=40_000) as c:
        _, off = c.reserve(64_000)
        for bad in (640, 3838, 188, 3852, 0):
            with pytest.raises(L.GpuAudioError) as e:
                c.submit_ticks(off, 1, np.arange(10, dtype=np.uint32), bad)
            assert e.value.code == L.GA_ERR_ARG and "multiple of 12" in str(e.value)
        assert [c.stream_info(i) for i in range(10)] == [(0, 0)] * 10


def test_restarting_an_utterance_inside_a_batch():
    """All messages of a batch are processed at once, so a GA_MSG_FIRST after chunks of the same
    stream would write two utterances over each other: converting contexts refuse it in both
    submit calls; without conversion there is nothing to overwrite and both accept it alike."""
    rng = np.random.default_rng(8)
    ids = np.arange(5, dtype=np.uint32)
    pcm = rng.integers(0, 256, size=(15, 640), dtype=np.uint8)
    flags = np.zeros((3, 5), dtype=np.uint8)
    flags[0] = L.MSG_FIRST
    flags[2, 3] = L.MSG_FIRST
    with make(want_frames=False) as c:
        off, stride = lay_out(c, pcm, 15, 640)
        for call in (lambda: c.submit_ticks(off, 3, ids, 640, flags.reshape(-1)),
                     lambda: c.submit(per_message(off, stride, 640, ids, flags.reshape(-1), 3))):
            with pytest.raises(L.GpuAudioError) as e:
                call()
            assert e.value.code == L.GA_ERR_ARG and "next batch" in str(e.value)
            assert [c.stream_info(i) for i in range(5)] == [(0, 0)] * 5
        # split where the restart is, it goes through: stream 3 holds only the new utterance
        c.wait(c.submit_ticks(off, 2, ids, 640, flags[:2].reshape(-1)))
        c.wait(c.submit_ticks(off + 10 * stride, 1, ids, 640, flags[2]))
        assert c.stream_info(3) == (320, 0) and c.stream_info(2) == (960, 0)
        assert np.array_equal(c.fetch_samples(3), ora.bytes_to_float32(pcm[13].tobytes())[0])
    with make(mode=L.DO_SHA256 | L.DO_FEATURES, want_frames=False) as a, \
            make(mode=L.DO_SHA256 | L.DO_FEATURES, want_frames=False) as b:
        off_a, stride = lay_out(a, pcm, 15, 640)
        off_b, _ = lay_out(b, pcm, 15, 640)
        ra, _ = a.wait(a.submit_ticks(off_a, 3, ids, 640, flags.reshape(-1)))
        rb, _ = b.wait(b.submit(per_message(off_b, stride, 640, ids, flags.reshape(-1), 3)))
        same_records(ra, rb)
        assert [a.stream_info(i) for i in range(5)] == [b.stream_info(i) for i in range(5)]
        assert a.stream_info(3) == (320, 0) and a.stream_info(2) == (960, 0)


def test_a_restart_in_the_next_batch_waits_for_the_batch_in_flight():
    """Tickets run on their own CUDA streams; a stream restarted while its previous chunks are in
    a ticket that has not been waited must still end up holding the new utterance only."""
    rng = np.random.default_rng(10)
    S, chunk, T = 64, 6400, 8
    ids = np.arange(S, dtype=np.uint32)
    old = rng.integers(0, 256, size=(T * S, chunk), dtype=np.uint8)
    new = rng.integers(0, 256, size=(S, chunk), dtype=np.uint8)
    with make(max_streams=S, max_batch_msgs=T * S, ring_bytes=16 << 20, arena_samples=40_000,
              wake_samples_max=400, want_frames=False) as c:
        for _ in range(5):
            off1, _ = lay_out(c, old, T * S, chunk)
            off2, _ = lay_out(c, new, S, chunk)
            t1 = c.submit_ticks(off1, T, ids, chunk, np.r_[np.full(S, L.MSG_FIRST, np.uint8), np.zeros((T - 1) * S, np.uint8)])
            t2 = c.submit_ticks(off2, 1, ids, chunk, np.full(S, L.MSG_FIRST, np.uint8))    # not waited in between
            c.wait(t1); c.wait(t2)
            for sid in (0, 17, S - 1):
                assert c.stream_info(sid) == (chunk // 2, 0)
                assert np.array_equal(c.fetch_samples(sid), ora.bytes_to_float32(new[sid].tobytes())[0])

"""The batched winner path (ga_fetch_wavs / ga_samples_digests / ga_hub_fetch_winner_wavs): what the
reference does once per closed window — processWinningRelayAudio (internal/grpc/audio_service.go:590-609),
float32ToWAV (internal/llm/stt_client.go:365-398) and the fingerprint of the samples
(internal/events/voice_event.go:84-88) — for many winners in one call.  WAV bytes against the
oracle's float32ToWAV restatement, digests against hashlib."""
import hashlib
import threading

import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api
from loqa_hub_b200.audio_service import AudioService, RelayStatusWinner
from oracle import oracle as ora

pytestmark = pytest.mark.gpu
MS = 1000


def make(S=32, **kw):
    import __graft_entry__ as entry
    entry.build()
    args = dict(device=0, max_streams=S, max_batch_msgs=4096, ring_bytes=32 << 20, arena_samples=20_000,
                wake_samples_max=4_800, mode=L.DO_ALL)
    args.update(kw)
    return api.Context(**args)


def fill(ctx, S, rng, lens):
    """Stream s gets lens[s] = (wake chunks, speech chunks) of 20 ms; returns its samples (wake ++ speech)."""
    chunks, sid, flags, per = [], [], [], {}
    for s in range(S):
        nw, ns = lens[s]
        pcm = rng.integers(-30000, 30000, size=(nw + ns) * 320).astype("<i2")
        if s % 5 == 0:
            pcm[-7:] = 0
        per[s] = ora.bytes_to_float32(pcm.tobytes())[0] if pcm.size else np.zeros(0, np.float32)
        for m in range(nw + ns):
            chunks.append(pcm[m * 320:(m + 1) * 320].tobytes()); sid.append(s)
            flags.append((L.MSG_FIRST if m == 0 else 0) | (L.MSG_WAKE if m < nw else 0))
    ctx.ingest(chunks, np.array(sid, np.uint32), np.array(flags, np.uint32))
    return per


def test_many_wavs_and_digests_in_one_call():
    S = 24
    rng = np.random.default_rng(9)
    lens = {s: (int(rng.integers(0, 4)), int(rng.integers(0, 9))) for s in range(S)}
    lens[3] = (0, 0); lens[4] = (3, 0); lens[5] = (0, 2)              # nothing at all / wake word only / speech only
    with make(S) as c:
        per = fill(c, S, rng, lens)
        ids = [5, 3, 4, 0, 23, 11, 11, 7]                              # any order, repeats allowed
        wavs, dgs = c.fetch_wavs(ids, 16000)
        for k, s in enumerate(ids):
            assert wavs[k] == ora.float32_to_wav(per[s], 16000), s
            assert dgs[k] == hashlib.sha256(per[s].astype("<f4").tobytes()).digest(), s
        assert c.fetch_wavs([], 16000) == ([], [])
        # the one-stream entry points agree with the batch (same kernels, n = 1)
        assert c.samples_digest(7) == dgs[ids.index(7)]
        assert c.samples_digests(list(range(S))) == [hashlib.sha256(per[s].astype("<f4").tobytes()).digest() for s in range(S)]
        wav8k, none = c.fetch_wavs([0], 8000, digests=False)
        assert none is None and wav8k[0] == ora.float32_to_wav(per[0], 8000)
        with pytest.raises(L.GpuAudioError) as e:
            c.fetch_wavs([S], 16000)
        assert e.value.code == L.GA_ERR_ARG


def test_winner_fetch_does_not_hold_up_the_batching_thread():
    """The winner path waits for the GPU without the context lock: submits from another thread go on
    while 256 winners are fetched again and again, and both sides stay correct."""
    S = 256
    rng = np.random.default_rng(10)
    with make(S, arena_samples=12_000, wake_samples_max=3_200, max_inflight=2) as c:
        per = fill(c, S, rng, {s: (2, 6) for s in range(S)})
        want = [hashlib.sha256(per[s].astype("<f4").tobytes()).digest() for s in range(S)]
        extra = rng.integers(-9000, 9000, size=(8, 320)).astype("<i2")
        errors, stop = [], threading.Event()

        def batcher():                                   # keeps appending speech to streams 0..7 only
            try:
                n = 0
                while not stop.is_set() and n < 20:
                    c.ingest([extra[s].tobytes() for s in range(8)], np.arange(8, dtype=np.uint32), np.zeros(8, np.uint32))
                    n += 1
            except Exception as e:                       # noqa: BLE001
                errors.append(e)
        t = threading.Thread(target=batcher)
        t.start()
        for _ in range(10):
            wavs, dgs = c.fetch_wavs(list(range(8, S)), 16000)
            assert dgs == want[8:]
            assert wavs[100] == ora.float32_to_wav(per[108], 16000)
        stop.set()
        t.join()
        assert not errors, errors
        n_extra = (c.stream_info(0)[0] - 8 * 320) // 320
        tail = np.concatenate([per[0]] + [ora.bytes_to_float32(extra[0].tobytes())[0]] * n_extra)
        assert np.array_equal(c.fetch_samples(0).view(np.uint32), tail.view(np.uint32))


def test_hub_closes_many_windows_and_hands_out_all_winner_wavs_at_once():
    G, R, wake, speech = 12, 4, 3, 4
    import __graft_entry__ as entry
    entry.build()
    ctx = make(G * R)
    svc = AudioService(ctx, max_streams=G * R, window_duration_us=100 * MS)
    try:
        data = ora.synth_fill(70, G * R, R, wake + speech, 320)              # relay 0 of each group is the loudest
        for s in range(G * R):
            svc.connect("g%d-r%d" % (s // R, s % R), group=s // R)
        for m in range(wake + speech):
            for s in range(G * R):
                svc.on_chunk("g%d-r%d" % (s // R, s % R), data[s, m].tobytes(), m < wake, False, m * MS)
        assert svc.flush() == G * R * wake + G * R * speech
        assert svc.poll_windows(50 * MS) == []
        closed = svc.poll_windows(150 * MS)
        assert len(closed) == G
        out = svc.fetch_winner_wavs(closed)
        for w, (wav, h) in zip(closed, out):
            info = svc.window(w)
            g = int(info.winner_id[1:].split("-")[0])
            assert info.winner_id == "g%d-r0" % g and svc.relay(info.winner_id).status == RelayStatusWinner
            full, _ = ora.bytes_to_float32(data[g * R].tobytes())
            assert wav == ora.float32_to_wav(full, 16000)
            assert h == "sha256:" + hashlib.sha256(full.astype("<f4").tobytes()).hexdigest()
        assert svc.fetch_winner_wavs([], hashes=False) == []
        # a window whose winner has been cleaned up meanwhile yields nothing, not an error
        svc.cleanup_relay(svc.window(closed[0]).winner_id)
        assert svc.fetch_winner_wavs(closed[:2])[0] == (None, "")
    finally:
        svc.close(); ctx.close()

"""ga_cluster — one process, one context per GPU, windows whose relays sit on different GPUs
(SURVEY §8e).  performArbitration (audio_service.go:505-571) must not depend on which GPU a relay's
samples live on: winners, scores and statuses are compared with the oracle over the same wake-word
signals, for windows that are local, spanning, empty, all-silent, and with members that are not
contending.  The spanning cases need two GPUs; the single-shard cases run anywhere."""
import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


def n_devices():
    return L.load().ga_device_count()


def make_ctx(dev, streams=64):
    import __graft_entry__ as entry
    entry.build()
    return api.Context(device=dev, max_streams=streams, max_batch_msgs=512, max_inflight=2, ring_bytes=4 << 20,
                       arena_samples=40_000, wake_samples_max=8_000, mode=L.DO_ALL)


def fill(ctxs, rng, per_shard):
    """Wake-word audio for `per_shard` streams of every context; returns sig[shard][stream]."""
    sig = []
    for c in ctxs:
        chunks, mine = [], []
        for s in range(per_shard):
            n = int(rng.integers(0, 3000))
            kind = rng.random()
            x = np.zeros(n, "<i2") if kind < 0.15 else (rng.integers(-30000, 30000, size=n) * rng.random()).astype("<i2")
            chunks.append(x.tobytes())
            mine.append(ora.bytes_to_float32(x.tobytes())[0])
        c.ingest(chunks, list(range(per_shard)), [L.MSG_WAKE | L.MSG_FIRST] * per_shard)
        sig.append(mine)
    return sig


def check(cl, sig, rng, n_windows, n_shards, per_shard, spanning):
    seg, shard, stream, status = [0], [], [], []
    for w in range(n_windows):
        k = int(rng.integers(0, 7))
        home = int(rng.integers(0, n_shards))
        for _ in range(k):
            sh = int(rng.integers(0, n_shards)) if (spanning and rng.random() < 0.5) else home
            shard.append(sh); stream.append(int(rng.integers(0, per_shard)))
            status.append(L.RELAY_CONTENDING if rng.random() < 0.8 else int(rng.choice([0, 2, 3])))
        seg.append(len(shard))
    winner, scores, st_out = cl.arbitrate(seg, shard, stream, status, np.full(len(shard), -2.0))
    for w in range(n_windows):
        lo, hi = seg[w], seg[w + 1]
        want_w, want_sc, want_st = ora.arbitrate([sig[shard[m]][stream[m]] for m in range(lo, hi)],
                                                 np.array(status[lo:hi], np.uint8))
        assert int(winner[w]) == (lo + want_w if want_w >= 0 else L.NO_WINNER), (w, lo, hi)
        assert list(st_out[lo:hi]) == list(want_st)
        for m in range(lo, hi):
            assert scores[m] == (want_sc[m - lo] if status[m] == L.RELAY_CONTENDING else -2.0)
    return sum(1 for w in range(n_windows) if len({shard[m] for m in range(seg[w], seg[w + 1])}) > 1)


def test_single_shard_cluster_equals_the_context_call():
    rng = np.random.default_rng(1)
    with make_ctx(0) as c, api.Cluster([c]) as cl:
        sig = fill([c], rng, 40)
        for _ in range(3):
            check(cl, sig, rng, 25, 1, 40, spanning=False)
        # and the plain context entry point on the same state
        w, sc, st = c.arbitrate([0, 3], [5, 6, 7])
        w2, sc2, st2 = cl.arbitrate([0, 3], [0, 0, 0], [5, 6, 7])
        assert (list(w), list(sc), list(st)) == (list(w2), list(sc2), list(st2))


def test_two_contexts_on_one_device_are_refused():
    with make_ctx(0) as a, make_ctx(0) as b:
        with pytest.raises(L.GpuAudioError) as e:
            api.Cluster([a, b])
        assert e.value.code == L.GA_ERR_ARG and "share device" in str(e.value)


def test_more_remote_relays_than_the_cluster_was_sized_for():
    if n_devices() < 2:
        pytest.skip("needs 2 GPUs")
    rng = np.random.default_rng(2)
    with make_ctx(0) as a, make_ctx(1) as b, api.Cluster([a, b], max_guests=2) as cl:
        fill([a, b], rng, 8)
        with pytest.raises(L.GpuAudioError) as e:
            cl.arbitrate([0, 4], [0, 1, 1, 1], [0, 1, 2, 3])
        assert e.value.code == L.GA_ERR_OVERFLOW
        w, _, _ = cl.arbitrate([0, 3], [0, 1, 1], [0, 1, 2])       # two guests fit
        assert int(w[0]) in (0, 1, 2)


def test_spanning_windows_across_gpus_match_the_oracle():
    n = min(n_devices(), 8)
    if n < 2:
        pytest.skip("needs 2 GPUs")
    rng = np.random.default_rng(3)
    ctxs = [make_ctx(d) for d in range(n)]
    try:
        with api.Cluster(ctxs) as cl:
            sig = fill(ctxs, rng, 48)
            spanning = 0
            for _ in range(6):                                    # consecutive calls: inbox parity and flags alternate
                spanning += check(cl, sig, rng, 60, n, 48, spanning=True)
            assert spanning > 50
            # every window homed on shard 1, every relay on shard 0: shard 0 only sends, shard 1 only receives
            seg = list(range(0, 41, 4))
            shard = [1 if m % 4 == 0 else 0 for m in range(40)]
            stream = [int(x) for x in rng.integers(0, 48, size=40)]
            winner, scores, _ = cl.arbitrate(seg, shard, stream)
            for w in range(10):
                lo = seg[w]
                want_w, want_sc, _ = ora.arbitrate([sig[shard[m]][stream[m]] for m in range(lo, lo + 4)])
                assert int(winner[w]) == lo + want_w and list(scores[lo:lo + 4]) == list(want_sc)
            # new wake-word audio, not waited separately: the cluster orders itself behind the ingest tickets
            sig = fill(ctxs, rng, 48)
            check(cl, sig, rng, 60, n, 48, spanning=True)
    finally:
        for c in ctxs:
            c.close()


# ---------------------------------------------------------------- the hub over several GPUs
def test_hub_over_two_gpus_spanning_windows_and_winner_audio():
    """The StreamAudio flow (audio_service.go:939-1041) with relays on two GPUs: groups homed by
    group % nGPU, two relays of every group forced onto the other GPU so that every window spans;
    wake-word chunks, window close by the timer, winners, cancelled relays, and the winner's WAV
    (stt_client.go:365-398) from whichever GPU holds it — all against the oracle."""
    if n_devices() < 2:
        pytest.skip("needs 2 GPUs")
    from loqa_hub_b200 import synth
    from loqa_hub_b200.audio_service import AudioService, RelayStatusCancelled, RelayStatusWinner
    MS = 1000
    ctxs = [make_ctx(d, streams=64) for d in range(2)]
    try:
        with api.Cluster(ctxs) as cl:
            svc = AudioService(cl, max_streams=64, window_duration_us=100 * MS)
            n_groups, n_relays, wake_msgs, ms = 6, 4, 5, 320
            pcm, name = {}, lambda g, r: "g%d-relay%d" % (g, r)
            for g in range(n_groups):
                for r in range(n_relays):
                    if r < 2:
                        svc.connect(name(g, r), group=g)                     # home shard g % 2
                    else:
                        svc.connect_on(name(g, r), g, (g + 1) % 2)           # the other GPU
                    assert svc.relay_shard(name(g, r)) == (g % 2 if r < 2 else (g + 1) % 2)
                    pcm[g, r] = synth.relay_view(g, r, wake_msgs * ms + 3 * ms).astype("<i2")
            # the quietest relay speaks first, the loudest (relay 0) joins last: order must not matter
            for m in range(wake_msgs):
                for g in range(n_groups):
                    for r in reversed(range(n_relays)):
                        assert svc.on_chunk(name(g, r), pcm[g, r][m * ms:(m + 1) * ms].tobytes(), True, False, (10 * m + n_relays - 1 - r) * MS)
            assert svc.flush() == n_groups * n_relays * wake_msgs
            assert svc.poll(now_us=50 * MS) == 0
            closed = svc.poll_windows(200 * MS)                              # ga_hub_poll that also names the windows it closed
            assert len(closed) == n_groups
            for g in range(n_groups):
                sig = [ora.bytes_to_float32(pcm[g, r][:wake_msgs * ms].tobytes())[0] for r in reversed(range(n_relays))]
                w, sc, _ = ora.arbitrate(sig)
                winner = name(g, n_relays - 1 - w)
                assert winner == name(g, 0)                                   # relay 0 has gain 1.0
                for k, r in enumerate(reversed(range(n_relays))):
                    info = svc.relay(name(g, r))
                    assert info.signal_strength == sc[k]
                    assert info.status == (RelayStatusWinner if name(g, r) == winner else RelayStatusCancelled)
                assert svc.arbitration_window(g) is None
            # speech for the winners only (cancelled relays' handlers have returned), then their WAVs
            for m in range(wake_msgs, wake_msgs + 3):
                for g in range(n_groups):
                    svc.on_chunk(name(g, 0), pcm[g, 0][m * ms:(m + 1) * ms].tobytes(), False, m == wake_msgs + 2, 300 * MS)
            svc.flush()
            for g in range(n_groups):
                full = ora.bytes_to_float32(pcm[g, 0].tobytes())[0]
                assert svc.fetch_wav(name(g, 0)) == ora.float32_to_wav(full, 16000)
            # ... and all winners at once, one gather + hash + copy per GPU (winners sit on both GPUs)
            import hashlib
            for w, (wav, audio_hash) in zip(closed, svc.fetch_winner_wavs(closed)):
                g = int(svc.window(w).winner_id[1:].split("-")[0])
                full = ora.bytes_to_float32(pcm[g, 0].tobytes())[0]
                assert wav == ora.float32_to_wav(full, 16000)
                assert audio_hash == "sha256:" + hashlib.sha256(full.astype("<f4").tobytes()).hexdigest()
            svc.close()
    finally:
        for c in ctxs:
            c.close()

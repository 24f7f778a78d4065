"""Property test of the context layer: ANY batch the boundary accepts — ragged lengths, shuffled
streams, wake/speech flags, every mode and both staging paths — must come back bit-identical to the
CPU oracle, record by record, frame by frame, and sample by sample in the stream arenas.
Hypothesis draws the batches (derandomised, so the run is the same everywhere)."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api
from oracle import oracle as ora

pytestmark = pytest.mark.gpu

N_STREAMS = 12
_ctx = {}


def context(mode, staging, frames):
    key = (mode, staging, frames)
    if key not in _ctx:
        import __graft_entry__ as entry
        entry.build()
        _ctx[key] = api.Context(device=0, max_streams=N_STREAMS, max_batch_msgs=256, max_inflight=2,
                                ring_bytes=4 << 20, arena_samples=400_000, wake_samples_max=100_000,
                                mode=mode, want_frames=frames, staging=staging)
    return _ctx[key]


@pytest.fixture(scope="module", autouse=True)
def _close_contexts():
    yield
    for c in _ctx.values():
        c.close()
    _ctx.clear()


lengths = st.one_of(st.integers(0, 200), st.integers(630, 650), st.integers(1270, 1290),
                    st.integers(0, 4000), st.sampled_from([55, 56, 63, 64, 65, 119, 120, 127, 128, 129, 640]))
kinds = st.sampled_from(["random", "zeros", "max", "min", "alternating", "small"])


def payload(kind, n, seed):
    rng = np.random.default_rng(seed)
    if kind == "random":
        return rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()
    if kind == "zeros":
        return bytes(n)
    v = {"max": [32767], "min": [-32768], "alternating": [32767, -32768, 0], "small": [1, -1, 0, 0]}[kind]
    return np.resize(np.asarray(v, dtype="<i2"), (n + 1) // 2).tobytes()[:n]


batches = st.lists(st.tuples(lengths, kinds, st.integers(0, N_STREAMS - 1), st.booleans()), min_size=1, max_size=48)


@settings(max_examples=200, deadline=None, derandomize=True, database=None,
          suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(batch=batches, uniform_len=st.one_of(st.none(), st.sampled_from([2, 64, 640, 642, 1000])),
       mode=st.sampled_from([L.DO_ALL, L.DO_CONVERT | L.DO_FEATURES, L.DO_SHA256, L.DO_SHA256 | L.DO_FEATURES,
                             L.DO_CONVERT | L.DO_SHA256, L.DO_FEATURES, L.DO_CONVERT]),
       staging=st.sampled_from([L.STAGING_AUTO, L.STAGING_BULK]), frames=st.booleans(), seed=st.integers(0, 2 ** 31))
def test_any_accepted_batch_matches_the_oracle(batch, uniform_len, mode, staging, frames, seed):
    c = context(mode, staging, frames)
    chunks, sids, flags, seen = [], [], [], set()
    for i, (n, kind, sid, wake) in enumerate(batch):
        if uniform_len is not None:
            n = uniform_len                       # equal sizes pushed back to back: the tensor-map path
        chunks.append(payload(kind, n, seed + i))
        sids.append(sid)
        flags.append((L.MSG_WAKE if wake else 0) | (0 if sid in seen else L.MSG_FIRST))   # restart each stream once
        seen.add(sid)
    res, fr = c.ingest(chunks, sids, flags)
    want_wake = {s: [] for s in seen}
    want_speech = {s: [] for s in seen}
    fpos = 0
    for i, data in enumerate(chunks):
        x, cs = ora.bytes_to_float32(data)
        rms, ss = ora.signal_strength(x)
        r = res[i]
        if mode & L.DO_SHA256:
            assert r["digest"].tobytes() == ora.sha256(data), (i, len(data))
        else:
            assert not r["digest"].any()
        assert r["n_samples"] == x.size and (r["flags"] >> 8) == flags[i]
        if mode & L.DO_FEATURES:
            assert (r["sumsq"], r["rms"], r["nonzero"]) == (ss.sumsq, rms, ss.nonzero), (i, len(data))
            assert np.float32(r["peak"]).view(np.uint32) == np.float32(cs.peak).view(np.uint32)
            assert (r["flags"] & 1) == (1 if rms > 0.01 else 0)
            if frames:
                of = ora.frame_features(x, L.FRAME_SAMPLES, 0.01)
                for a, b in zip(fr[fpos:fpos + len(of)], of):
                    assert (a["sumsq"], a["rms"], a["n_samples"], a["nonzero"], a["vad"]) == \
                           (b["sumsq"], b["rms"], b["n_samples"], b["nonzero"], b["vad"])
        if frames:
            fpos += (x.size + L.FRAME_SAMPLES - 1) // L.FRAME_SAMPLES
        (want_wake if flags[i] & L.MSG_WAKE else want_speech)[sids[i]].append(x)
    if frames:
        assert fpos == len(fr)
    for s in seen:
        w = np.concatenate(want_wake[s]) if want_wake[s] else np.zeros(0, np.float32)
        sp = np.concatenate(want_speech[s]) if want_speech[s] else np.zeros(0, np.float32)
        assert c.stream_info(s) == (w.size + sp.size, w.size)
        if mode & L.DO_CONVERT:
            got = c.fetch_samples(s)
            assert got.view(np.uint32).tobytes() == np.concatenate([w, sp]).view(np.uint32).tobytes(), s


# ---------------------------------------------------------------- 48 kHz stereo + running digest
chunk48 = st.one_of(st.integers(0, 40), st.integers(180, 200), st.just(3840), st.integers(0, 3840),
                    st.sampled_from([4 * 46, 4 * 47, 4 * 48, 4 * 959 + 1, 3839]))


@settings(max_examples=150, deadline=None, derandomize=True, database=None,
          suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(plan=st.lists(st.tuples(st.integers(0, 3), chunk48, st.booleans()), min_size=1, max_size=40),
       seed=st.integers(0, 2 ** 31))
def test_48k_streams_any_chunking_any_batching(plan, seed):
    """Four 48 kHz stereo streams cut into arbitrary chunks (partial frames, chunks shorter than the
    filter history, empty ones) and arbitrary batches: samples, statistics and the running
    per-utterance digest must not depend on where the cuts fall.  `new_batch` closes the current
    batch; one is also closed where the boundary demands it (a chunk whose stream's previous
    chunk in the batch is shorter than the 47-frame history)."""
    import hashlib
    key = ("48k",)
    if key not in _ctx:
        import __graft_entry__ as entry
        entry.build()
        _ctx[key] = api.Context(device=0, max_streams=4, max_batch_msgs=64, max_inflight=2, ring_bytes=2 << 20,
                                arena_samples=60_000, wake_samples_max=20_000, mode=L.DO_ALL | L.DO_UTTERANCE,
                                want_frames=True, input_format=L.FMT_STEREO48K)
    c = _ctx[key]
    rng = np.random.default_rng(seed)
    oracle = [ora.Resampler() for _ in range(4)]
    received = [b"" for _ in range(4)]
    want = [[] for _ in range(4)]
    started = [False] * 4
    batch, last_len = [], {}

    def flush():
        if not batch:
            return
        res, fr = c.ingest([b[1] for b in batch], [b[0] for b in batch], [b[2] for b in batch])
        fpos = 0
        for r, (sid, data, _fl, y) in zip(res, batch):
            rms, ss = ora.signal_strength(y)
            assert r["digest"].tobytes() == hashlib.sha256(data).digest()
            assert (r["n_samples"], r["nonzero"], r["sumsq"], r["rms"]) == (y.size, ss.nonzero, ss.sumsq, rms)
            of = ora.frame_features(y, L.FRAME_SAMPLES, 0.01)
            for a, b in zip(fr[fpos:fpos + len(of)], of):
                assert (a["sumsq"], a["n_samples"], a["nonzero"]) == (b["sumsq"], b["n_samples"], b["nonzero"])
            fpos += len(of)
        assert fpos == len(fr)
        batch.clear()
        last_len.clear()

    for sid, nbytes, new_batch in plan:
        if new_batch or (sid in last_len and last_len[sid] < 47 * 4):
            flush()
        data = rng.integers(0, 256, size=nbytes, dtype=np.uint8).tobytes()
        y = oracle[sid].process(data)
        fl = 0 if started[sid] else L.MSG_FIRST
        started[sid] = True
        received[sid] += data
        want[sid].append(y)
        batch.append((sid, data, fl, y))
        last_len[sid] = nbytes
    flush()
    for sid in range(4):
        if not started[sid]:
            continue
        y = np.concatenate(want[sid])
        assert c.stream_info(sid) == (y.size, 0)
        assert c.fetch_samples(sid).view(np.uint32).tobytes() == y.view(np.uint32).tobytes()
        digest, n = c.utterance_digest(sid)
        assert (bytes(digest), n) == (hashlib.sha256(received[sid]).digest(), len(received[sid]))


# ---------------------------------------------------------------- arbitration
member = st.tuples(st.integers(0, 2600), st.sampled_from(["random", "zeros", "max", "small", "quiet"]),
                   st.sampled_from([L.RELAY_CONTENDING] * 4 + [L.RELAY_CONNECTED, L.RELAY_WINNER, L.RELAY_CANCELLED]))


@settings(max_examples=120, deadline=None, derandomize=True, database=None,
          suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(windows=st.lists(st.lists(member, min_size=0, max_size=9), min_size=1, max_size=7), seed=st.integers(0, 2 ** 31),
       twin=st.booleans())
def test_any_set_of_windows_arbitrates_like_the_reference(windows, seed, twin):
    """performArbitration (audio_service.go:505-571) for arbitrary windows — empty ones, empty and
    all-zero wake-word signals (first-member fallback), members that are not contending (skipped,
    :522-524), exact ties (`twin`: two members hear identical audio; the first in join order keeps
    the win because the comparison is a strict `>`)."""
    c = context(L.DO_ALL, L.STAGING_AUTO, False)
    rng = np.random.default_rng(seed)
    sid, chunks, flags, sig, status, seg = 0, [], [], [], [], [0]
    for w in windows:
        w = w[:N_STREAMS - sid]                   # the context has N_STREAMS stream slots
        for k, (n, kind, st_in) in enumerate(w):
            if kind == "quiet":
                data = rng.integers(-3, 4, size=n).astype("<i2").tobytes()
            else:
                data = payload(kind, 2 * n, seed + sid)
            if twin and k == 1:
                data = prev                        # same bytes as the window's first member
            prev = data
            chunks.append(data); flags.append(L.MSG_WAKE | L.MSG_FIRST)
            sig.append(ora.bytes_to_float32(data)[0]); status.append(st_in)
            sid += 1
        seg.append(sid)
    if sid:
        c.ingest(chunks, list(range(sid)), flags)
    winner, scores, st_out = c.arbitrate(np.array(seg, np.uint32), np.arange(sid, dtype=np.uint32),
                                         np.array(status, np.uint8), np.full(sid, -1.0))
    for k in range(len(seg) - 1):
        lo, hi = seg[k], seg[k + 1]
        w, sc, so = ora.arbitrate(sig[lo:hi], np.array(status[lo:hi], np.uint8))
        assert int(winner[k]) == (lo + w if w >= 0 else 0xffffffff), (k, lo, hi)
        for m in range(lo, hi):
            if status[m] == L.RELAY_CONTENDING:
                assert scores[m] == sc[m - lo]
            else:
                assert scores[m] == -1.0           # relay.SignalStrength untouched for skipped relays
        assert list(st_out[lo:hi]) == list(so)


# ---------------------------------------------------------------- tick batches vs per-message batches
_pairs = {}


@settings(max_examples=80, deadline=None, derandomize=True, database=None,
          suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(fmt48=st.booleans(), utter=st.booleans(), n_streams=st.integers(1, 12),
       ticks=st.lists(st.integers(1, 5), min_size=1, max_size=3), wake_ticks=st.integers(0, 3),
       chunk16=st.sampled_from([0, 2, 64, 640, 642, 1000, 1281]), chunk48=st.sampled_from([192, 960, 1920, 3840]),
       seed=st.integers(0, 2 ** 31))
def test_tick_batches_equal_per_message_batches(fmt48, utter, n_streams, ticks, wake_ticks, chunk16, chunk48, seed):
    """ga_submit_ticks (descriptors, decimator descriptors and digest lists built on the GPU) and
    ga_submit (built on the host) on the same chunks, over consecutive batches of one utterance:
    records, arenas, cursors and running digests must be identical."""
    key = (fmt48, utter)
    if key not in _pairs:
        import __graft_entry__ as entry
        entry.build()
        kw = dict(device=0, max_streams=N_STREAMS, max_batch_msgs=64, max_inflight=2, ring_bytes=1 << 20,
                  arena_samples=40_000, wake_samples_max=16_000, mode=L.DO_ALL | (L.DO_UTTERANCE if utter else 0),
                  want_frames=True, input_format=L.FMT_STEREO48K if fmt48 else L.FMT_MONO16K)
        _pairs[key] = (api.Context(**kw), api.Context(**kw))
        _ctx[("pair-a",) + key], _ctx[("pair-b",) + key] = _pairs[key]      # closed with the module
    a, b = _pairs[key]
    rng = np.random.default_rng(seed)
    chunk = chunk48 if fmt48 else chunk16
    stride = (chunk + 15) & ~15
    ids = rng.permutation(N_STREAMS)[:n_streams].astype(np.uint32)
    done = 0
    for rnd, T in enumerate(ticks):
        n = T * n_streams
        pcm = rng.integers(0, 256, size=(n, chunk), dtype=np.uint8)
        flags = np.zeros((T, n_streams), np.uint8)
        for t in range(T):
            if done + t < wake_ticks:
                flags[t] |= L.MSG_WAKE
        if rnd == 0:
            flags[0] |= L.MSG_FIRST
        done += T
        offs = []
        for c in (a, b):
            view, off = c.reserve(max(stride * n, 1))
            for m in range(n):
                view[m * stride:m * stride + chunk] = pcm[m]
            offs.append(off)
        ra, fa = a.wait(a.submit_ticks(offs[0], T, ids, chunk, flags.reshape(-1)))
        msgs = np.zeros(n, dtype=L.MSG_DTYPE)
        msgs["ring_offset"] = offs[1] + np.arange(n, dtype=np.uint64) * np.uint64(stride)
        msgs["nbytes"] = chunk
        msgs["stream_id"] = np.tile(ids, T)
        msgs["flags"] = flags.reshape(-1)
        rb, fb = b.wait(b.submit(msgs))
        assert ra.tobytes() == rb.tobytes() and fa.tobytes() == fb.tobytes()
    for sid in ids:
        assert a.stream_info(int(sid)) == b.stream_info(int(sid))
        assert a.fetch_samples(int(sid)).tobytes() == b.fetch_samples(int(sid)).tobytes()
        if utter:
            assert a.utterance_digest(int(sid)) == b.utterance_digest(int(sid))

"""48 kHz stereo -> 16 kHz mono decimator (BASELINE configs[3]) through the C ABI against the
oracle's ora_resample_48k_stereo.  The filter is new spec (the reference ignores
AudioChunk.SampleRate, internal/grpc/audio_service.go:926-1043), so the oracle is the definition
("parity unpinned" at the reference level); what IS pinned is A1's conversion inside it and the
statistics (calculateSignalStrength, :818-852) of the resampled samples.  Fixed tap order +
explicit FMAs make the output reproducible, so everything here is compared bit for bit."""
import struct

import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api, synth
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


def bits(x):
    return np.asarray(x, dtype=np.float32).view(np.uint32)


def stereo48(group, relay, n_frames):
    """SURVEY §8d cfg4 input: 3x zero-order hold of the relay's 16 kHz view, R = L - (L >> 3)."""
    x = synth.relay_view(group, relay, (n_frames + 2) // 3).astype(np.int32)
    left = np.repeat(x, 3)[:n_frames]
    right = left - (left >> 3)
    return np.stack([left, right], axis=1).astype("<i2")


@pytest.fixture(scope="module")
def ctx48():
    import __graft_entry__ as entry
    entry.build()
    c = api.Context(device=0, max_streams=64, max_batch_msgs=8192, max_inflight=4, ring_bytes=64 << 20,
                    arena_samples=40_000, wake_samples_max=8_000, mode=L.DO_ALL, want_frames=True,
                    input_format=L.FMT_STEREO48K)
    yield c
    c.close()


def oracle_stream(chunks):
    r, parts = ora.Resampler(), []
    for c in chunks:
        parts.append(r.process(c))
    return parts


def check(res, frames, chunks, parts):
    fpos = 0
    for i, (data, y) in enumerate(zip(chunks, parts)):
        assert res["digest"][i].tobytes() == ora.sha256(data)          # fingerprint of the RECEIVED bytes
        rms, st = ora.signal_strength(y)
        assert res["n_samples"][i] == y.size
        assert struct.pack("<d", res["sumsq"][i]) == struct.pack("<d", st.sumsq)
        assert struct.pack("<d", res["rms"][i]) == struct.pack("<d", rms)
        assert res["nonzero"][i] == st.nonzero and bits(res["peak"][i]) == bits(st.peak)
        of = ora.frame_features(y, L.FRAME_SAMPLES, 0.01)
        for a, b in zip(frames[fpos:fpos + len(of)], of):
            assert (a["sumsq"], a["rms"], a["n_samples"], a["nonzero"], a["vad"]) == \
                   (b["sumsq"], b["rms"], b["n_samples"], b["nonzero"], b["vad"])
            assert bits(a["peak"]) == bits(b["peak"])
        fpos += len(of)
    assert fpos == len(frames)


def test_cfg4_streams_whole_utterance_in_one_batch(ctx48):
    """8 streams x 25 chunks of 3,840 B (20 ms of 48 kHz stereo), time-major, ONE submit: every
    chunk after the first takes its 47-frame history from its predecessor in the slab."""
    n_streams, n_msgs, fr = 8, 25, 960
    pcm = [stereo48(3 + s // 4, s % 4, n_msgs * fr) for s in range(n_streams)]
    chunks, sid, flags = [], [], []
    for m in range(n_msgs):
        for s in range(n_streams):
            chunks.append(pcm[s][m * fr:(m + 1) * fr].tobytes())
            sid.append(s)
            flags.append((L.MSG_FIRST if m == 0 else 0) | (L.MSG_WAKE if m < 15 else 0))
    res, frames = ctx48.ingest(chunks, np.array(sid, np.uint32), np.array(flags, np.uint32))
    for s in range(n_streams):
        mine = [i for i in range(len(chunks)) if sid[i] == s]
        parts = oracle_stream([chunks[i] for i in mine])
        got = ctx48.fetch_samples(s)
        assert np.array_equal(bits(got), bits(np.concatenate(parts))), s
        assert ctx48.stream_info(s) == (n_msgs * 320, 15 * 320)
        check(res[mine], frames[mine], [chunks[i] for i in mine], parts)     # 1 frame per 320-sample message
    # arbitration on the resampled wake-word buffers: relay 0 of each group is the loudest
    winner, scores, _ = ctx48.arbitrate(np.array([0, 4, 8], np.uint32), np.arange(8, dtype=np.uint32))
    assert list(winner) == [0, 4]
    y0 = np.concatenate(oracle_stream([pcm[0][m * fr:(m + 1) * fr].tobytes() for m in range(15)]))
    assert scores[0] == ora.signal_strength(y0)[0]


def test_history_carries_across_batches_and_odd_chunk_sizes(ctx48):
    """Ragged chunking (frame counts that are not multiples of 3, short chunks, a trailing partial
    frame), one chunk per submit so the history comes from the saved per-stream state, mixed with
    batches that hold several chunks of the stream."""
    rng = np.random.default_rng(5)
    total = 6000
    pcm = np.stack([rng.integers(-30000, 30000, total), rng.integers(-30000, 30000, total)], axis=1).astype("<i2")
    raw = pcm.tobytes()
    sizes = [4 * 7, 4 * 1, 3840, 4 * 50, 4 * 2 + 3, 4 * 959, 4 * 100, 4 * 47, 3840, 3840, 4 * 333]
    chunks, pos = [], 0
    for n in sizes:
        chunks.append(raw[pos:pos + n]); pos += n - (n % 4)        # a partial frame is dropped, not carried
    parts = oracle_stream(chunks)
    plan = [[0], [1], [2], [3], [4], [5, 6], [7], [8, 9, 10]]       # chunks per submit (predecessors >= 47 frames)
    out_res = []
    for k, group in enumerate(plan):
        fl = [L.MSG_FIRST if (k == 0 and j == 0) else 0 for j in range(len(group))]
        res, frames = ctx48.ingest([chunks[i] for i in group], np.full(len(group), 9, np.uint32), np.array(fl, np.uint32))
        out_res.append(res)
    res = np.concatenate(out_res)
    got = ctx48.fetch_samples(9)
    assert np.array_equal(bits(got), bits(np.concatenate(parts)))
    for i, (data, y) in enumerate(zip(chunks, parts)):
        assert res["digest"][i].tobytes() == ora.sha256(data)
        assert res["n_samples"][i] == y.size
        assert struct.pack("<d", res["rms"][i]) == struct.pack("<d", ora.signal_strength(y)[0])


def test_saved_history_is_not_overwritten_while_an_earlier_chunk_still_needs_it(ctx48):
    """A batch of more than 32 messages in which streams have their first chunk (history from the
    state saved by the PREVIOUS batch, n0 != 0) in one 32-message group and their last chunk (which
    saves the state for the NEXT batch) in another: the save must not reach the state before every
    read of it is done.  Messages 31 / 32 of one stream is the tightest case (end of group 0, start
    of group 1).  Repeated over three batches so the saved state is also consumed again."""
    rng = np.random.default_rng(17)
    n_streams, fr = 40, 960
    per = {s: [] for s in range(n_streams)}
    pcm = {s: np.stack([rng.integers(-30000, 30000, 12 * fr), rng.integers(-30000, 30000, 12 * fr)], axis=1).astype("<i2")
           for s in range(n_streams)}
    taken = {s: 0 for s in range(n_streams)}

    def next_chunk(s, frames=fr):
        c = pcm[s][taken[s]:taken[s] + frames].tobytes()
        taken[s] += frames
        per[s].append(c)
        return c

    got = {s: [] for s in range(n_streams)}

    def run(order, first=False, frames=None):
        chunks = [next_chunk(s, (frames or {}).get(i, fr)) for i, s in enumerate(order)]
        seen = set()
        fl = []
        for s in order:
            fl.append(L.MSG_FIRST if first and s not in seen else 0)
            seen.add(s)
        res, _ = ctx48.ingest(chunks, np.array(order, np.uint32), np.array(fl, np.uint32))
        for i, s in enumerate(order):
            got[s].append(res[i])

    run(list(range(n_streams)), first=True, frames={i: 700 + 3 * i + (i % 3) for i in range(n_streams)})   # n0 != 0 afterwards
    # batch 2: stream 7 sits at messages 31 and 32; 31 other streams fill the rest of group 0 and
    # come back (their saving chunk) in groups 1-3
    front = [s for s in range(32) if s != 7]
    order = front + [7, 7] + front + list(range(32, 40)) + front[::2]
    assert len(front) == 31 and order[31] == 7 and order[32] == 7 and 7 not in front and len(order) > 64
    run(order)
    # batch 3: every stream again, first chunks spread over two groups, second chunks behind them
    run(list(range(n_streams)) + list(range(n_streams - 1, -1, -1)))
    for s in range(n_streams):
        parts = oracle_stream(per[s])
        assert np.array_equal(bits(ctx48.fetch_samples(s)), bits(np.concatenate(parts))), s
        for r, data, y in zip(got[s], per[s], parts):
            assert r["digest"].tobytes() == ora.sha256(data)
            rms, st = ora.signal_strength(y)
            assert r["n_samples"] == y.size and r["nonzero"] == st.nonzero
            assert struct.pack("<d", r["sumsq"]) == struct.pack("<d", st.sumsq)
            assert struct.pack("<d", r["rms"]) == struct.pack("<d", rms)
            assert bits(r["peak"]) == bits(st.peak)


def test_every_int16_value_in_both_channels(ctx48):
    """The kernel takes A1's division (audio_service.go:1078) as one fused multiply-add per channel
    (csrc/resample.cu mono_of); that is exact only because it was checked for every input, so every
    int16 value goes through both channels here, against the oracle's plain division."""
    rng = np.random.default_rng(33)
    left = rng.permutation(np.arange(-32768, 32768)).astype("<i2")
    right = np.concatenate([left[::-1][1:], left[:1]]).astype("<i2")       # a different value beside each
    pcm = np.stack([left, right], axis=1)
    chunks = [pcm[i:i + 960].tobytes() for i in range(0, len(pcm), 960)]
    fl = np.zeros(len(chunks), np.uint32); fl[0] = L.MSG_FIRST
    ctx48.ingest(chunks, np.full(len(chunks), 21, np.uint32), fl)
    want = np.concatenate(oracle_stream(chunks))
    assert np.array_equal(bits(ctx48.fetch_samples(21)), bits(want))


def test_48k_boundary_rules(ctx48):
    with pytest.raises(L.GpuAudioError) as e:
        ctx48.ingest([b"\x00" * 3844], [1], [L.MSG_FIRST])
    assert e.value.code == L.GA_ERR_ARG and "3840" in str(e.value)
    before = ctx48.stream_info(2)
    with pytest.raises(L.GpuAudioError) as e:                       # predecessor in the same batch too short
        ctx48.ingest([b"\x01" * 40, b"\x01" * 400], [2, 2], [L.MSG_FIRST, 0])
    assert e.value.code == L.GA_ERR_ARG and "47 frames" in str(e.value)
    assert ctx48.stream_info(2) == before                           # the failed submit left no trace
    res, _ = ctx48.ingest([b"\x01" * 400], [2], [L.MSG_FIRST])
    assert res["n_samples"][0] == 33


def test_48k_context_with_utterance_digest_and_partial_fetches():
    """All context options together on 48 kHz input: per-message digests of the received bytes, the
    per-utterance running digest, frames, and ga_fetch_samples ranges that straddle the boundary
    between the wake-word region and the speech region of the arena."""
    import hashlib
    import __graft_entry__ as entry
    entry.build()
    with api.Context(device=0, max_streams=8, max_batch_msgs=512, ring_bytes=8 << 20, arena_samples=20_000,
                     wake_samples_max=4_000, mode=L.DO_ALL | L.DO_UTTERANCE, want_frames=True,
                     input_format=L.FMT_STEREO48K) as c:
        pcm = stereo48(11, 1, 960 * 20)
        chunks = [pcm[i * 960:(i + 1) * 960].tobytes() for i in range(20)]
        fl = np.zeros(20, np.uint32)
        fl[:6] |= L.MSG_WAKE; fl[0] |= L.MSG_FIRST
        r1, f1 = c.ingest(chunks[:9], np.full(9, 3, np.uint32), fl[:9])          # 6 wake + 3 speech chunks
        r2, f2 = c.ingest(chunks[9:], np.full(11, 3, np.uint32), fl[9:])
        res = np.concatenate([r1, r2])
        parts = oracle_stream(chunks)
        whole = np.concatenate(parts)
        assert c.stream_info(3) == (20 * 320, 6 * 320)
        assert np.array_equal(bits(c.fetch_samples(3)), bits(whole))
        for first, n in ((0, 1), (1900, 40), (1919, 2), (1920, 5), (1000, 3000), (6399, 1), (6400, 0)):
            assert np.array_equal(bits(c.fetch_samples(3, first, n)), bits(whole[first:first + n])), (first, n)
        with pytest.raises(L.GpuAudioError):
            c.fetch_samples(3, 6000, 401)
        for i, data in enumerate(chunks):
            assert res["digest"][i].tobytes() == hashlib.sha256(data).digest()
        assert c.utterance_digest(3) == (hashlib.sha256(b"".join(chunks)).digest(), 20 * 3840)
        assert len(f1) + len(f2) == 20
        assert c.samples_digest(3) == hashlib.sha256(whole.astype("<f4").tobytes()).digest()

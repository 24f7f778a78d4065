"""Per-utterance running SHA-256 (one digest per voice event: API.md:170, hook
internal/events/voice_event.go:84-88) through the C ABI against the oracle's streaming SHA-256
and hashlib.  Bit-exact."""
import hashlib

import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api, synth
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import __graft_entry__ as entry
    entry.build()
    c = api.Context(device=0, max_streams=128, max_batch_msgs=1 << 15, max_inflight=4, ring_bytes=64 << 20,
                    arena_samples=120_000, wake_samples_max=8_000, mode=L.DO_ALL | L.DO_UTTERANCE)
    yield c
    c.close()


def test_fresh_stream_digest_is_sha256_of_nothing(ctx):
    d, n = ctx.utterance_digest(5)
    assert (d.hex(), n) == ("e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855", 0)


def test_utterance_digest_over_many_batches_and_many_messages_per_batch(ctx):
    """16 streams x 250 chunks of 640 B: first 40 ticks in ONE submit (each lane walks 40 messages of
    its stream), then tick by tick, then the rest at once; the digest is checked along the way."""
    n_streams, n_msgs, ms = 16, 250, 320
    pcm = api.synth_batch_host(synth.SEED, 40, n_streams, 8, n_msgs, ms).reshape(n_msgs, n_streams, ms)
    ref = [ora.Sha256() for _ in range(n_streams)]

    def submit(m0, m1):
        chunks = [pcm[m, s].tobytes() for m in range(m0, m1) for s in range(n_streams)]
        sid = np.tile(np.arange(n_streams, dtype=np.uint32), m1 - m0)
        fl = np.array([(L.MSG_FIRST if m == 0 else 0) | (L.MSG_WAKE if m < 15 else 0) | (L.MSG_EOS if m == n_msgs - 1 else 0)
                       for m in range(m0, m1) for _ in range(n_streams)], np.uint32)
        res, _ = ctx.ingest(chunks, sid, fl)
        for m in range(m0, m1):
            for s in range(n_streams):
                ref[s].update(pcm[m, s].tobytes())
        return res

    submit(0, 40)
    for m in range(40, 60):
        submit(m, m + 1)
        if m % 7 == 0:
            for s in (0, 3, 15):
                assert ctx.utterance_digest(s) == (ref[s].digest(), (m + 1) * 640)
    res = submit(60, n_msgs)
    for s in range(n_streams):
        whole = pcm[:, s, :].tobytes()
        assert ctx.utterance_digest(s) == (hashlib.sha256(whole).digest(), len(whole))
        assert ctx.utterance_digest(s)[0] == ref[s].digest()
    # the per-message digests are still produced alongside
    assert res["digest"][-1].tobytes() == hashlib.sha256(pcm[-1, -1].tobytes()).digest()


def test_ragged_chunks_carry_a_partial_block_between_messages_and_batches(ctx):
    rng = np.random.default_rng(8)
    sizes = [1, 63, 64, 65, 640, 3, 119, 120, 0, 5000, 55, 56, 7, 641, 640, 640, 129]
    data = [rng.integers(0, 256, size=n, dtype=np.uint8).tobytes() for n in sizes]
    plan = [[0, 1, 2], [3], [4, 5, 6, 7, 8], [9], [10, 11], [12, 13, 14, 15, 16]]
    h = hashlib.sha256()
    for k, group in enumerate(plan):
        fl = [L.MSG_FIRST if (k == 0 and j == 0) else 0 for j in range(len(group))]
        ctx.ingest([data[i] for i in group], np.full(len(group), 77, np.uint32), np.array(fl, np.uint32))
        for i in group:
            h.update(data[i])
        assert ctx.utterance_digest(77) == (h.digest(), sum(sizes[:group[-1] + 1]))
    # GA_MSG_FIRST starts a new utterance; ga_stream_reset (cleanupRelay) empties the state
    ctx.ingest([b"abc"], [77], [L.MSG_FIRST])
    assert ctx.utterance_digest(77)[0].hex() == "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"
    ctx.stream_reset(77)
    assert ctx.utterance_digest(77)[1] == 0


def test_utterance_only_context():
    with api.Context(max_streams=4, max_batch_msgs=64, ring_bytes=1 << 20, arena_samples=4096,
                     wake_samples_max=64, mode=L.DO_UTTERANCE) as c:
        res, _ = c.ingest([b"a" * 1000, b"b" * 24], [1, 1], [L.MSG_FIRST, 0])
        assert not res["digest"].any()
        assert c.utterance_digest(1) == (hashlib.sha256(b"a" * 1000 + b"b" * 24).digest(), 1024)


def test_fingerprint_of_the_converted_samples(ctx):
    """SURVEY §8c secondary mode: SHA-256 over the little-endian f32 bytes of WakeWordSignal ++
    SpeechAudio (the serialisation of stt_client.go:390-395), hashed on the device."""
    pcm = synth.relay_view(9, 2, 7000)
    chunks = [pcm[i:i + 320].tobytes() for i in range(0, 7000, 320)]          # the last chunk is short
    fl = np.zeros(len(chunks), np.uint32)
    fl[:15] |= L.MSG_WAKE; fl[0] |= L.MSG_FIRST
    ctx.ingest(chunks, np.full(len(chunks), 33, np.uint32), fl)
    samples = ctx.fetch_samples(33)
    assert samples.size == 7000
    assert ctx.samples_digest(33) == hashlib.sha256(samples.astype("<f4").tobytes()).digest()
    want, _ = ora.bytes_to_float32(pcm.tobytes())
    assert ctx.samples_digest(33) == ora.sha256(want.astype("<f4").tobytes())
    ctx.stream_reset(33)
    assert ctx.samples_digest(33) == hashlib.sha256(b"").digest()

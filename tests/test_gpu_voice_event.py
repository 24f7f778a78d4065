"""The voice-event audio hook (events/voice_event.go:84-88) with the documented fingerprint
(API.md:170) restored: duration / rate / flag exactly as SetAudioMetadata computes them, the hash
against hashlib (an independent SHA-256) over the bytes the hub serialises (stt_client.go:390-395)."""
import hashlib
import math

import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api
from loqa_hub_b200.voice_event import VoiceEvent

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import __graft_entry__ as entry
    entry.build()
    with api.Context(device=0, max_streams=8, max_batch_msgs=256, ring_bytes=4 << 20,
                     arena_samples=100_000, wake_samples_max=8_000, mode=L.DO_ALL) as c:
        yield c


@pytest.mark.parametrize("n,rate", [(80_000, 16000), (1, 16000), (0, 16000), (15, 44100), (48_123, 48000)])
def test_set_audio_metadata(ctx, n, rate):
    audio = np.random.default_rng(n).standard_normal(n).astype(np.float32)
    ev = VoiceEvent(relay_id="relay-1", request_id="req-1").set_audio_metadata(ctx, audio, rate, True)
    assert ev.audio_duration == float(n) / float(rate)           # voice_event.go:85
    assert (ev.sample_rate, ev.wake_word_detected) == (rate, True)
    assert ev.audio_hash == "sha256:" + hashlib.sha256(audio.astype("<f4").tobytes()).hexdigest()
    assert ev.to_json()["audio_hash"].startswith("sha256:") and len(ev.audio_hash) == 71


def test_rest_handler_dummy_audio_and_zero_rate(ctx):
    """api/voice_events.go:216-219 calls the hook with one zero sample; a zero rate is not an error in Go."""
    ev = VoiceEvent().set_audio_metadata(ctx, np.zeros(1, np.float32), 16000, False)
    assert ev.audio_duration == 1 / 16000 and ev.audio_hash == "sha256:" + hashlib.sha256(b"\0\0\0\0").hexdigest()
    ev = VoiceEvent().set_audio_metadata(ctx, np.zeros(1, np.float32), 0, False)
    assert math.isinf(ev.audio_duration) and ev.audio_duration > 0
    assert math.isnan(VoiceEvent().set_audio_metadata(ctx, np.zeros(0, np.float32), 0, False).audio_duration)


def test_zero_persistence_keeps_no_fingerprint(ctx):
    ev = VoiceEvent().set_audio_metadata(ctx, np.ones(100, np.float32), 16000, True, zero_persistence=True)
    assert ev.audio_hash == "" and "audio_hash" not in ev.to_json()
    assert ev.audio_duration == 100 / 16000 and ev.wake_word_detected


def test_winner_stream_metadata_without_moving_the_samples(ctx):
    """The hub calls the hook with WakeWordSignal ++ SpeechAudio of the winner; from a stream
    already in HBM only the digest crosses PCIe, and it equals the host-array call."""
    rng = np.random.default_rng(5)
    chunks = [rng.integers(-20000, 20000, size=320, dtype=np.int16).astype("<i2").tobytes() for _ in range(30)]
    flags = [(L.MSG_FIRST if m == 0 else 0) | (L.MSG_WAKE if m < 15 else 0) for m in range(30)]
    ctx.ingest(chunks, [3] * 30, flags)
    samples = ctx.fetch_samples(3)
    a = VoiceEvent().set_audio_metadata_from_stream(ctx, 3, 16000, True)
    b = VoiceEvent().set_audio_metadata(ctx, samples, 16000, True)
    assert samples.size == 9600 and a == b and a.audio_duration == 0.6
    assert a.audio_hash == "sha256:" + hashlib.sha256(samples.tobytes()).hexdigest()
    assert VoiceEvent().set_audio_metadata_from_stream(ctx, 3, 16000, True, zero_persistence=True).audio_hash == ""

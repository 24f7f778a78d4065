"""Host-side mirror of the voice-event audio hook — `VoiceEvent.SetAudioMetadata`
(internal/events/voice_event.go:84-88) — with the fingerprint the docs still describe
(`"audio_hash": "sha256:…"`, API.md:170) computed on the GPU.  Only the audio half of the event
is mirrored; transcription / intent / storage stay in the hub.

The field the reference dropped (internal/storage/migrations/001_remove_audio_hash.sql) is additive
and privacy-gated: with `zero_persistence` (PrivacyConfig.ZeroPersistence, config/config.go:102-103)
no fingerprint is computed and `audio_hash` stays "".
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib as L


@dataclass
class VoiceEvent:
    relay_id: str = ""
    request_id: str = ""
    audio_duration: float = 0.0          # json:"audio_duration"
    sample_rate: int = 0                 # json:"sample_rate"
    wake_word_detected: bool = False     # json:"wake_word_detected"
    audio_hash: str = ""                 # json:"audio_hash" — "sha256:" + hex, or "" (zero persistence)

    def set_audio_metadata(self, ctx, audio_data, sample_rate, is_wake_word, zero_persistence=False):
        """SetAudioMetadata(audioData []float32, sampleRate int, isWakeWord bool)."""
        a = np.ascontiguousarray(audio_data, dtype=np.float32)
        dur, buf = C.c_double(), C.create_string_buffer(72)
        L.check(L.load().ga_set_audio_metadata(ctx._h, L.ptr(a) if a.size else None, a.size, int(sample_rate),
                                               int(bool(zero_persistence)), C.byref(dur), buf))
        self._fill(dur.value, sample_rate, is_wake_word, buf)
        return self

    def set_audio_metadata_from_stream(self, ctx, stream_id, sample_rate, is_wake_word, zero_persistence=False):
        """The same for a stream whose WakeWordSignal ++ SpeechAudio already sits in HBM
        (an arbitration winner, audio_service.go:592-594): only the digest crosses PCIe."""
        dur, buf = C.c_double(), C.create_string_buffer(72)
        L.check(L.load().ga_stream_audio_metadata(ctx._h, int(stream_id), int(sample_rate),
                                                  int(bool(zero_persistence)), C.byref(dur), buf))
        self._fill(dur.value, sample_rate, is_wake_word, buf)
        return self

    def _fill(self, duration, sample_rate, is_wake_word, buf):
        self.audio_duration = duration
        self.sample_rate = int(sample_rate)
        self.wake_word_detected = bool(is_wake_word)
        self.audio_hash = buf.value.decode()

    def to_json(self):
        out = {"relay_id": self.relay_id, "request_id": self.request_id, "audio_duration": self.audio_duration,
               "sample_rate": self.sample_rate, "wake_word_detected": self.wake_word_detected}
        if self.audio_hash:
            out["audio_hash"] = self.audio_hash
        return out


def format_audio_hash(digest):
    d = np.frombuffer(bytes(digest), dtype=np.uint8)
    if d.size != 32:
        raise ValueError("a SHA-256 digest has 32 bytes")
    buf = C.create_string_buffer(72)
    L.load().ga_format_audio_hash(L.ptr(d), buf)
    return buf.value.decode()

"""Pin the CPU oracle against the reference's own test vectors and the FIPS 180-4 vectors.

Reference tests restated here (relative to /root/reference):
  internal/grpc/audio_service_test.go:258-313   TestBytesToFloat32Array
  internal/grpc/audio_service_test.go:315-364   TestAudioService_CalculateSignalStrength
  internal/grpc/collision_detection_test.go:48-92    TestCalculateSignalStrength
  internal/grpc/collision_detection_test.go:226-293  TestSignalStrengthArbitration
  internal/grpc/collision_detection_test.go:395-459  TestConsecutiveRequestsWithSingleRelay
  internal/grpc/collision_detection_test.go:322-391  TestRequestScopedArbitrationLifecycle (empty wake signal)
"""
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from oracle import numpy_mirror as npm
from oracle import oracle as ora

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def f32bits(x):
    return np.asarray(x, dtype=np.float32).view(np.uint32)


# ---------------------------------------------------------------- A1
CONVERT_CASES = [
    # name, input bytes, expected values (reference test), expected f32 bit patterns
    ("empty input", b"", [], []),
    ("single sample", bytes([0x00, 0x10]), [4096.0 / 32767.0], [0x3E000100]),
    ("multiple samples", bytes([0x00, 0x00, 0xFF, 0x7F]), [0.0, 1.0], [0x00000000, 0x3F800000]),
    ("odd byte count", bytes([0x00, 0x10, 0xFF]), [4096.0 / 32767.0], [0x3E000100]),
    ("int16 min (extra)", bytes([0x00, 0x80]), [-32768.0 / 32767.0], [0xBF800100]),
]


@pytest.mark.parametrize("name,data,expected,bits", CONVERT_CASES, ids=[c[0] for c in CONVERT_CASES])
def test_bytes_to_float32_reference_vectors(name, data, expected, bits):
    out, st = ora.bytes_to_float32(data)
    assert len(out) == len(expected)                       # wantLen
    for got, want in zip(out, expected):
        assert abs(float(got) - want) <= 1e-6              # the reference's epsilon
    assert list(f32bits(out)) == bits                      # and the exact bit patterns
    assert st.n_samples == len(expected)


def test_bytes_to_float32_all_int16_values_match_numpy_mirror():
    v = np.arange(-32768, 32768, dtype=np.int32).astype("<i2")
    out, st = ora.bytes_to_float32(v.tobytes())
    ref = npm.bytes_to_float32(v.tobytes())
    assert np.array_equal(f32bits(out), f32bits(ref))
    assert st.nonzero == 65535
    assert f32bits(st.peak) == 0x3F800100                  # |-32768/32767|
    # multiply-by-reciprocal is NOT the reference arithmetic (SURVEY Appendix B rule 2)
    recip = v.astype(np.float32) * (np.float32(1.0) / np.float32(32767.0))
    assert int((f32bits(recip) != f32bits(out)).sum()) == 1536


# ---------------------------------------------------------------- A2
RMS_CASES = [
    ("empty", [], 0.0, 1e-7),
    ("single 0.5", [0.5], 0.5, 1e-7),
    ("0 1 0 -1", [0.0, 1.0, 0.0, -1.0], 0.7071067811865476, 1e-7),
    ("ones", [1.0, 1.0, 1.0, 1.0], 1.0, 1e-7),
    ("0.1 x4", [0.1, 0.1, 0.1, 0.1], 0.1, 1e-7),
    ("zeros", [0.0, 0.0, 0.0], 0.0, 1e-6),
    ("single 1.0", [1.0], 1.0, 1e-6),
    ("mixed", [0.5, -0.5, 0.3, -0.3], 0.412, 1e-2),
]
RMS_EXACT = {"0.1 x4": 0.10000000353902572, "mixed": 0.4123105647302035,
             "0 1 0 -1": 0.7071067811865476}


@pytest.mark.parametrize("name,samples,expected,eps", RMS_CASES, ids=[c[0] for c in RMS_CASES])
def test_signal_strength_reference_vectors(name, samples, expected, eps):
    rms, st = ora.signal_strength(np.array(samples, dtype=np.float32))
    assert abs(rms - expected) <= eps
    assert rms >= 0
    if name in RMS_EXACT:
        assert rms == RMS_EXACT[name]                      # exact under reference semantics
    mirror, total = npm.signal_strength(np.array(samples, dtype=np.float32))
    assert rms == mirror
    assert st.sumsq == total


def test_signal_strength_sequential_order_matters_and_matches_mirror():
    rng = np.random.default_rng(7)
    pcm = rng.integers(-32768, 32768, size=4800, dtype=np.int64).astype("<i2")
    pcm[::7] = rng.integers(-12, 12, size=pcm[::7].size)   # quiet samples force f64 rounding
    s, _ = ora.bytes_to_float32(pcm.tobytes())
    rms, st = ora.signal_strength(s)
    mirror, total = npm.signal_strength(s)
    assert (rms, st.sumsq) == (mirror, total)
    assert st.nonzero == int((pcm != 0).sum())
    assert f32bits(st.peak) == f32bits(np.abs(s).max())


# ---------------------------------------------------------------- A3
def test_arbitration_reference_three_relays():
    wake = [np.full(4, 0.1, np.float32), np.full(4, 1.0, np.float32), np.full(4, 0.5, np.float32)]
    w, scores, status = ora.arbitrate(wake)
    assert w == 1                                          # relay2 wins
    assert scores[0] < scores[1] and scores[2] < scores[1]
    assert list(status) == [ora.RELAY_CANCELLED, ora.RELAY_WINNER, ora.RELAY_CANCELLED]
    assert npm.select_winner(scores, [True] * 3) == 1


def test_arbitration_single_relay_wins():
    w, scores, status = ora.arbitrate([np.array([0.5, 0.5, 0.5], np.float32)])
    assert w == 0 and status[0] == ora.RELAY_WINNER and scores[0] == 0.5


def test_arbitration_empty_wake_signal_falls_back_to_first_member():
    # audio_service.go:542-547 — no relay exceeded 0.0 -> first member of the window
    w, scores, status = ora.arbitrate([np.zeros(0, np.float32)])
    assert w == 0 and scores[0] == 0.0 and status[0] == ora.RELAY_WINNER
    w, scores, status = ora.arbitrate([np.zeros(8, np.float32), np.zeros(3, np.float32)])
    assert w == 0 and list(status) == [ora.RELAY_WINNER, ora.RELAY_CANCELLED]


def test_arbitration_skips_non_contending_and_first_index_wins_ties():
    wake = [np.full(4, 0.9, np.float32), np.full(4, 0.5, np.float32), np.full(4, 0.5, np.float32)]
    st_in = np.array([ora.RELAY_CANCELLED, ora.RELAY_CONTENDING, ora.RELAY_CONTENDING], np.uint8)
    w, scores, status = ora.arbitrate(wake, st_in)
    assert w == 1                                          # strict > keeps the first of equals
    assert scores[0] == 0.0                                # skipped relay's score untouched
    assert list(status) == [ora.RELAY_CANCELLED, ora.RELAY_WINNER, ora.RELAY_CANCELLED]
    assert ora.arbitrate([])[0] == -1                      # empty window


# ---------------------------------------------------------------- SHA-256
NIST = [
    (b"", "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"),
    (b"abc", "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"),
    (b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq",
     "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1"),
    (b"abcdefghbcdefghicdefghijdefghijkefghijklfghijklmghijklmnhijklmnoijklmnopjklmnopqklmnopqrlmnopqrsmnopqrstnopqrstu",
     "cf5b16a778af8380036ce59e7b0492370b249b11e8f07a51afac45037afee9d1"),
    (b"\x00" * 640, "9e132485d5107211de325a45e7917cbe3e4b5b9cde3e4ee91d7d2102317759ee"),
]


@pytest.mark.parametrize("msg,hexd", NIST, ids=["empty", "abc", "448bit", "896bit", "silent20ms"])
def test_sha256_known_answers(msg, hexd):
    assert ora.sha256(msg).hex() == hexd


def test_sha256_million_a():
    assert ora.sha256(b"a" * 1_000_000).hex() == \
        "cdc76e5c9914fb9281a1c7e284d73e67f1809a48a497200e046d39ccc7112cd0"


def test_sha256_matches_openssl_on_padding_boundaries_and_chunked_updates():
    rng = np.random.default_rng(1)
    for n in list(range(0, 201)) + [640, 3840, 160000]:
        m = rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()
        assert ora.sha256(m) == hashlib.sha256(m).digest(), n
    m = rng.integers(0, 256, size=10_000, dtype=np.uint8).tobytes()
    ctx, pos = ora.Sha256(), 0
    for step in [1, 63, 64, 65, 640, 3, 119, 120, 5000]:
        ctx.update(m[pos:pos + step]); pos += step
        assert ctx.digest() == hashlib.sha256(m[:pos]).digest()


# ---------------------------------------------------------------- N1 / N2 / WAV
def test_frame_features_are_a2_per_frame():
    rng = np.random.default_rng(3)
    pcm = rng.integers(-20000, 20000, size=1000).astype("<i2")
    s, _ = ora.bytes_to_float32(pcm.tobytes())
    fr = ora.frame_features(s, 320, 0.01)
    assert len(fr) == 4 and list(fr["n_samples"]) == [320, 320, 320, 40]
    for i, f in enumerate(fr):
        rms, st = ora.signal_strength(s[i * 320:(i + 1) * 320])
        assert (f["rms"], f["sumsq"], f["nonzero"]) == (rms, st.sumsq, st.nonzero)
        assert f32bits(f["peak"]) == f32bits(st.peak) and f["vad"] == (1 if rms > 0.01 else 0)


def test_resampler_matches_scipy_and_is_chunking_invariant():
    from scipy.signal import upfirdn
    rng = np.random.default_rng(5)
    n = 960 * 5
    left = (8000 * np.sin(2 * np.pi * 440 * np.arange(n) / 48000)).astype(np.int16)
    right = (left - (left >> 3)).astype(np.int16)
    inter = np.stack([left, right], axis=1).astype("<i2").tobytes()
    whole = ora.Resampler().process(inter)
    assert whole.size == n // 3
    mono = (npm.bytes_to_float32(left.astype("<i2").tobytes()) +
            npm.bytes_to_float32(right.astype("<i2").tobytes())) * np.float32(0.5)
    taps = ora.resample_taps()
    full = upfirdn(taps.astype(np.float64), mono.astype(np.float64), 1, 1)
    np.testing.assert_allclose(whole, full[2::3][: whole.size], rtol=0, atol=2e-6)
    # arbitrary (also non-multiple-of-3-frames) chunking gives bit-identical output
    r, pos, parts = ora.Resampler(), 0, []
    for step in [4 * 7, 4 * 1, 3840, 4 * 50, 4 * 2, 3840 * 2, 10 ** 9]:
        parts.append(r.process(inter[pos:pos + step])); pos += step
    assert np.array_equal(f32bits(np.concatenate(parts)), f32bits(whole))
    _ = rng


def test_float32_to_wav_header_and_payload():
    s = np.array([0.0, 1.0, -1.0000305, 0.12500381], dtype=np.float32)
    wav = ora.float32_to_wav(s, 16000)
    assert len(wav) == 44 + 16
    assert wav[:4] == b"RIFF" and struct.unpack("<I", wav[4:8])[0] == 36 + 16
    assert wav[8:16] == b"WAVEfmt " and struct.unpack("<IHHIIHH", wav[16:36]) == (16, 3, 1, 16000, 64000, 4, 32)
    assert wav[36:40] == b"data" and struct.unpack("<I", wav[40:44])[0] == 16
    assert wav[44:] == s.astype("<f4").tobytes()


# ---------------------------------------------------------------- committed fixtures
def test_committed_golden_fixture():
    with open(os.path.join(GOLDEN, "ingest_golden.json")) as f:
        g = json.load(f)
    for case in g["messages"]:
        data = bytes.fromhex(case["pcm_hex"])
        out, st = ora.bytes_to_float32(data)
        assert [int(b) for b in f32bits(out)] == case["f32_bits"]
        assert ora.sha256(data).hex() == case["sha256"]
        rms, st2 = ora.signal_strength(out)
        assert struct.pack("<d", rms).hex() == case["rms_f64_hex"]
        assert struct.pack("<d", st2.sumsq).hex() == case["sumsq_f64_hex"]
        assert (st.nonzero, int(f32bits(st.peak))) == (case["nonzero"], case["peak_bits"])


# ---------------------------------------------------------------- the timed CPU baseline and its input
def test_oracle_synth_matches_the_numpy_statement_in_every_layout():
    """oracle/synth_oracle.c (what `bench.py --impl reference` feeds the CPU arm with) against the
    independent numpy statement of the same integer generator, mono and 48 kHz stereo."""
    from loqa_hub_b200 import synth
    S, gs, M, ms = 6, 3, 4, 320
    sm = ora.synth_fill(5, S, gs, M, ms, layout=ora.LAYOUT_STREAM_MAJOR, n_threads=3)
    tm = ora.synth_fill(5, S, gs, M, ms, layout=ora.LAYOUT_TIME_MAJOR, n_threads=2)
    assert np.array_equal(tm, synth.batch_time_major(5, S, gs, M, ms))
    assert np.array_equal(sm, tm.transpose(1, 0, 2))
    assert ora.synth_sample(7, 2, 12345, 80000) == int(synth.relay_view(7, 2, 1, 80000, 12345)[0])
    st = ora.synth_fill(5, S, gs, M, ms, stereo48=True, n_threads=2)
    assert st.shape == (S, M, ms * 6)
    left = np.repeat(sm.reshape(S, M * ms).astype(np.int32), 3, axis=1)
    want = np.stack([left, left - (left >> 3)], axis=2).astype(np.int16).reshape(S, M, ms * 6)
    assert np.array_equal(st, want)


def test_cpu_baseline_outputs_do_not_depend_on_layout_threads_or_sha_implementation():
    S, gs, M, ms, wake = 16, 4, 12, 320, 5
    sm = ora.synth_fill(0, S, gs, M, ms, layout=ora.LAYOUT_STREAM_MAJOR)
    tm = np.ascontiguousarray(sm.transpose(1, 0, 2))
    _, dg0, wn0, sc0 = ora.cpu_baseline(tm.view(np.uint8).reshape(-1), S, M, ms * 2, wake, gs, 1,
                                        use_openssl=False, want_outputs=True)
    for impl in (ora.SHA_PORT, ora.SHA_OPENSSL_EVP, ora.SHA_OPENSSL_LOW):
        for th in (1, 3):
            _, dg, wn, sc = ora.cpu_baseline_run(sm, S, M, ms * 2, wake, gs, th, sha_impl=impl, want_outputs=True)
            assert np.array_equal(dg, dg0) and np.array_equal(wn, wn0) and np.array_equal(sc, sc0)
    assert dg0[3, 5].tobytes() == hashlib.sha256(sm[5, 3].tobytes()).digest()
    assert list(wn0) == [0] * (S // gs)           # relay 0 has the highest gain in every group
    assert sc0[5] == ora.signal_strength(ora.bytes_to_float32(sm[5, :wake].tobytes())[0])[0]


def test_cpu_baseline_48k_leg_is_the_oracle_decimator_bit_for_bit():
    """The 48 kHz leg of the timed baseline restates ora_resample_48k_stereo for speed (no malloc per
    chunk, interleaved FMA chains); its scores must be those of the oracle functions."""
    S, gs, M, ms, wake = 4, 2, 9, 320, 4
    st = ora.synth_fill(3, S, gs, M, ms, stereo48=True)
    _, dg, wn, sc = ora.cpu_baseline_run(st, S, M, ms * 12, wake, gs, 2, fmt=ora.FMT_STEREO48K, want_outputs=True)
    for s in range(S):
        r = ora.Resampler()
        y = np.concatenate([r.process(st[s, m].tobytes()) for m in range(wake)])
        assert struct.pack("<d", sc[s]) == struct.pack("<d", ora.signal_strength(y)[0])
        assert dg[2, s].tobytes() == hashlib.sha256(st[s, 2].tobytes()).digest()
    assert list(wn) == [0, 0]

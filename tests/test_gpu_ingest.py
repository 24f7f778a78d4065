"""GPU parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle.

Bar: bit-exact for samples, digests, counts, peaks, sumsq/rms (the kernels keep the
reference's summation order, so even the float energies are compared bit for bit — the
1e-5 relative tolerance north_star allows is not used) and for arbitration winners.
The reference tests these read like: internal/grpc/audio_service_test.go:258-364 and
internal/grpc/collision_detection_test.go:48-92, :226-293, :395-459.
"""
import json
import os
import struct

import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api, synth
from oracle import oracle as ora

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def bits(x):
    return np.asarray(x, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def ctx():
    import __graft_entry__ as entry
    entry.build()
    c = api.Context(device=0, max_streams=256, max_batch_msgs=1 << 16, max_inflight=4,
                    ring_bytes=64 << 20, arena_samples=200_000, wake_samples_max=20_000,
                    mode=L.DO_ALL, want_frames=True)
    yield c
    c.close()


def check_message(res, data, frames=None):
    """One ga_result (+ its frame records) against the oracle, everything bit-exact."""
    want, st = ora.bytes_to_float32(data)
    rms, st2 = ora.signal_strength(want)
    assert res["digest"].tobytes() == ora.sha256(data)
    assert res["n_samples"] == want.size
    assert res["nonzero"] == st.nonzero == st2.nonzero
    assert bits(res["peak"]) == bits(st.peak)
    assert struct.pack("<d", res["sumsq"]) == struct.pack("<d", st2.sumsq)
    assert struct.pack("<d", res["rms"]) == struct.pack("<d", rms)
    assert (res["flags"] & 1) == (1 if rms > 0.01 else 0)
    if frames is not None:
        of = ora.frame_features(want, L.FRAME_SAMPLES, 0.01)
        assert len(frames) == len(of)
        for a, b in zip(frames, of):
            assert (a["sumsq"], a["rms"], a["n_samples"], a["nonzero"], a["vad"]) == \
                   (b["sumsq"], b["rms"], b["n_samples"], b["nonzero"], b["vad"])
            assert bits(a["peak"]) == bits(b["peak"])
    return want


# ------------------------------------------------------------------ reference unit tests, restated
CONVERT_CASES = [
    ("empty input", b"", []),
    ("single sample", bytes([0x00, 0x10]), [4096.0 / 32767.0]),
    ("multiple samples", bytes([0x00, 0x00, 0xFF, 0x7F]), [0.0, 1.0]),
    ("odd byte count", bytes([0x00, 0x10, 0xFF]), [4096.0 / 32767.0]),
]


@pytest.mark.parametrize("name,data,expected", CONVERT_CASES, ids=[c[0] for c in CONVERT_CASES])
def test_bytes_to_float32_array(ctx, name, data, expected):
    """TestBytesToFloat32Array, audio_service_test.go:258-313."""
    result, res = ctx.bytes_to_float32(data)
    assert len(result) == len(expected)
    for got, want in zip(result, expected):
        assert abs(float(got) - want) <= 0.000001
    want, st = ora.bytes_to_float32(data)
    assert np.array_equal(bits(result), bits(want))
    assert (res["nonzero"], res["n_samples"]) == (st.nonzero, st.n_samples)


def test_convert_all_65536_int16_values_bit_exact(ctx):
    v = np.arange(-32768, 32768, dtype=np.int32).astype("<i2")
    got, res = ctx.bytes_to_float32(v.tobytes())
    want, st = ora.bytes_to_float32(v.tobytes())
    assert np.array_equal(bits(got), bits(want))
    assert bits(res["peak"]) == bits(st.peak) == 0x3F800100
    assert res["nonzero"] == 65535


RMS_CASES = [
    ("Empty samples", [], 0.0, 1e-7), ("Single sample", [0.5], 0.5, 1e-7),
    ("Multiple samples", [0.0, 1.0, 0.0, -1.0], 0.7071067811865476, 1e-7),
    ("High amplitude", [1.0, 1.0, 1.0, 1.0], 1.0, 1e-7),
    ("Low amplitude", [0.1, 0.1, 0.1, 0.1], 0.1, 1e-7),
    ("zero samples", [0.0, 0.0, 0.0], 0.0, 1e-6), ("single positive sample", [1.0], 1.0, 1e-6),
    ("mixed samples", [0.5, -0.5, 0.3, -0.3], 0.412, 1e-2),
]


@pytest.mark.parametrize("name,samples,expected,eps", RMS_CASES, ids=[c[0] for c in RMS_CASES])
def test_calculate_signal_strength(ctx, name, samples, expected, eps):
    """TestCalculateSignalStrength (collision_detection_test.go:48-92) and
    TestAudioService_CalculateSignalStrength (audio_service_test.go:315-364)."""
    s = np.array(samples, dtype=np.float32)
    result, res = ctx.signal_strength(s)
    assert abs(result - expected) <= eps and result >= 0
    rms, st = ora.signal_strength(s)
    assert result == rms
    if s.size:
        assert (res["sumsq"], res["nonzero"], res["n_samples"]) == (st.sumsq, st.nonzero, st.n_samples)
        assert bits(res["peak"]) == bits(st.peak)


def test_signal_strength_long_buffers_keep_reference_order(ctx):
    rng = np.random.default_rng(11)
    for n in (1, 3, 4, 5, 127, 128, 129, 4800, 80_000):
        pcm = rng.integers(-32768, 32768, size=n).astype("<i2")
        pcm[::3] = rng.integers(-9, 10, size=pcm[::3].size)      # quiet samples force f64 rounding
        s, _ = ora.bytes_to_float32(pcm.tobytes())
        got, res = ctx.signal_strength(s)
        want, st = ora.signal_strength(s)
        assert struct.pack("<d", got) == struct.pack("<d", want), n
        assert res["sumsq"] == st.sumsq and res["nonzero"] == st.nonzero


def _window(ctx, wake_signals, status_in=None):
    """Load wake-word buffers into streams 0..n-1 (as int16 chunks) and arbitrate one window."""
    n = len(wake_signals)
    chunks = [np.asarray(w, dtype="<i2").tobytes() for w in wake_signals]
    ctx.ingest(chunks, np.arange(n, dtype=np.uint32),
               np.full(n, L.MSG_WAKE | L.MSG_FIRST, dtype=np.uint32))
    return ctx.arbitrate(np.array([0, n], np.uint32), np.arange(n, dtype=np.uint32), status_in)


def test_signal_strength_arbitration(ctx):
    """TestSignalStrengthArbitration, collision_detection_test.go:226-293: low / high / medium
    amplitude relays -> the high one wins, the others are cancelled."""
    lo, hi, mid = [3277] * 4, [32767] * 4, [16384] * 4      # ~0.1, 1.0, ~0.5 as int16
    winner, scores, status = _window(ctx, [lo, hi, mid])
    assert winner[0] == 1
    assert scores[0] < scores[1] and scores[2] < scores[1]
    assert list(status) == [L.RELAY_CANCELLED, L.RELAY_WINNER, L.RELAY_CANCELLED]
    wake = [ora.bytes_to_float32(np.asarray(w, "<i2").tobytes())[0] for w in (lo, hi, mid)]
    w, sc, st = ora.arbitrate(wake)
    assert (winner[0], list(scores), list(status)) == (w, list(sc), list(st))


def test_single_relay_and_empty_wake_signal_fall_back_to_first_member(ctx):
    """TestConsecutiveRequestsWithSingleRelay (:395-459) and the empty-WakeWordSignal relay of
    TestRequestScopedArbitrationLifecycle (:341-347; audio_service.go:542-547)."""
    for _ in range(3):
        winner, scores, status = _window(ctx, [[16384, 16384, 16384]])
        assert winner[0] == 0 and status[0] == L.RELAY_WINNER and scores[0] > 0.5
    winner, scores, status = _window(ctx, [[]])
    assert winner[0] == 0 and scores[0] == 0.0 and status[0] == L.RELAY_WINNER
    winner, scores, status = _window(ctx, [[0] * 8, [0] * 3])        # all-zero window
    assert winner[0] == 0 and list(status) == [L.RELAY_WINNER, L.RELAY_CANCELLED]


def test_non_contending_members_are_skipped_and_ties_go_to_first(ctx):
    st_in = np.array([L.RELAY_CANCELLED, L.RELAY_CONTENDING, L.RELAY_CONTENDING], np.uint8)
    winner, scores, status = _window(ctx, [[30000] * 4, [16384] * 4, [16384] * 4], st_in)
    assert winner[0] == 1 and scores[0] == 0.0
    assert list(status) == [L.RELAY_CANCELLED, L.RELAY_WINNER, L.RELAY_CANCELLED]
    # several windows at once, one of them empty
    seg = np.array([0, 2, 2, 3], np.uint32)
    winner, _, status = ctx.arbitrate(seg, np.array([1, 0, 2], np.uint32))
    assert list(winner) == [1, L.NO_WINNER, 2]      # window 0: stream 0 (30000) beats stream 1
    assert list(status) == [L.RELAY_CANCELLED, L.RELAY_WINNER, L.RELAY_WINNER]


# ------------------------------------------------------------------ committed golden fixture
def test_committed_golden_fixture_through_a_batch(ctx):
    with open(os.path.join(GOLDEN, "ingest_golden.json")) as f:
        cases = json.load(f)["messages"]
    chunks = [bytes.fromhex(c["pcm_hex"]) for c in cases]
    sid = np.arange(len(chunks), dtype=np.uint32)
    res, frames = ctx.ingest(chunks, sid, np.full(len(chunks), L.MSG_FIRST, np.uint32))
    for i, c in enumerate(cases):
        assert res["digest"][i].tobytes().hex() == c["sha256"], c["name"]
        assert struct.pack("<d", res["rms"][i]).hex() == c["rms_f64_hex"], c["name"]
        assert struct.pack("<d", res["sumsq"][i]).hex() == c["sumsq_f64_hex"], c["name"]
        assert (int(res["nonzero"][i]), int(bits(res["peak"][i]))) == (c["nonzero"], c["peak_bits"])
        got = ctx.fetch_samples(i)
        assert [int(b) for b in bits(got)] == c["f32_bits"], c["name"]


# ------------------------------------------------------------------ ragged / adversarial batches
def test_ragged_batch_every_length_and_alignment(ctx):
    rng = np.random.default_rng(2)
    lengths = list(range(0, 140)) + [191, 192, 193, 255, 256, 257, 639, 640, 641, 1279, 1280, 1281,
                                     3840, 3841, 9999] + list(rng.integers(0, 5000, size=60))
    rng.shuffle(lengths)
    chunks = [rng.integers(0, 256, size=int(n), dtype=np.uint8).tobytes() for n in lengths]
    n_streams = 37                                    # several messages per stream, odd sample counts
    sid = np.arange(len(chunks), dtype=np.uint32) % n_streams
    flags = np.where(np.arange(len(chunks)) < n_streams, L.MSG_FIRST, 0).astype(np.uint32)
    res, frames = ctx.ingest(chunks, sid, flags)
    fpos, per_stream = 0, {}
    for i, data in enumerate(chunks):
        nf = (len(data) // 2 + L.FRAME_SAMPLES - 1) // L.FRAME_SAMPLES
        want = check_message(res[i], data, frames[fpos:fpos + nf])
        fpos += nf
        per_stream.setdefault(int(sid[i]), []).append(want)
    assert fpos == len(frames)
    for s, parts in per_stream.items():
        got = ctx.fetch_samples(s)
        assert np.array_equal(bits(got), bits(np.concatenate(parts))), s


def test_extreme_sample_values_and_silence(ctx):
    runs = [np.full(640, 32767, "<i2"), np.full(640, -32768, "<i2"), np.zeros(640, "<i2"),
            np.tile(np.array([32767, -32768], "<i2"), 320), np.tile(np.array([1, -1, 0], "<i2"), 100)]
    chunks = [r.tobytes() for r in runs]
    res, frames = ctx.ingest(chunks, np.arange(len(chunks), dtype=np.uint32),
                             np.full(len(chunks), L.MSG_FIRST, np.uint32))
    fpos = 0
    for i, data in enumerate(chunks):
        nf = (len(data) // 2 + 319) // 320
        check_message(res[i], data, frames[fpos:fpos + nf]); fpos += nf
    assert res["digest"][2].tobytes().hex()[:16] != "0" * 16
    assert res["rms"][2] == 0.0 and res["nonzero"][2] == 0 and (res["flags"][2] & 1) == 0


def test_peak_of_negative_samples_in_either_half(ctx):
    """The peak must be found wherever the loudest sample sits and whatever its sign (this
    caught an nvcc VIMNMX3 miscompile of the packed min/max statistics)."""
    for val in (-32555, 32555, -32768):
        for n in (320, 321, 37):
            chunks = []
            for pos in range(n):
                v = np.full(n, 5, dtype="<i2"); v[pos] = val
                chunks.append(v.tobytes())
            sid = np.arange(n, dtype=np.uint32) % 256
            res, _ = ctx.ingest(chunks, sid, np.where(np.arange(n) < 256, L.MSG_FIRST, 0).astype(np.uint32))
            want = np.abs(np.float32(val) / np.float32(32767.0))
            assert (bits(res["peak"]) == bits(want)).all(), (val, n)
            assert (res["nonzero"] == n).all()


# ------------------------------------------------------------------ synthetic streams end to end
def test_cfg1_single_stream_5s_utterance_as_250_chunks(ctx):
    """BASELINE.json configs[0]: one relay stream, 5 s of 16 kHz mono int16 in 20 ms chunks."""
    pcm = synth.relay_view(0, 0, 80_000)
    assert np.array_equal(pcm, api.synth_host(synth.SEED, 0, 0, 0, 80_000, 80_000))
    chunks = [pcm[i:i + 320].tobytes() for i in range(0, 80_000, 320)]
    flags = np.zeros(250, np.uint32)
    flags[:15] |= L.MSG_WAKE; flags[0] |= L.MSG_FIRST; flags[-1] |= L.MSG_EOS
    res, frames = ctx.ingest(chunks, np.zeros(250, np.uint32), flags)
    for i, data in enumerate(chunks):
        check_message(res[i], data, frames[i:i + 1])
    want, _ = ora.bytes_to_float32(pcm.tobytes())
    assert np.array_equal(bits(ctx.fetch_samples(0)), bits(want))
    assert ctx.stream_info(0) == (80_000, 4_800)
    winner, scores, status = ctx.arbitrate(np.array([0, 1], np.uint32), np.array([0], np.uint32))
    assert winner[0] == 0 and scores[0] == ora.signal_strength(want[:4800])[0]
    # one-shot fingerprint of the whole utterance (2,501 chained compressions in one lane)
    assert ctx.sha256(pcm.tobytes()) == ora.sha256(pcm.tobytes())


def test_cfg3_collision_groups_winners_match_cpu_baseline(ctx):
    """32 collision groups x 8 relays, 15 wake-word + 10 speech chunks each, time-major batch:
    digests, scores and winners against the multi-threaded CPU baseline."""
    n_streams, group, n_msgs, wake, ms = 256, 8, 25, 15, 320
    pcm = api.synth_batch_host(synth.SEED, 100, n_streams, group, n_msgs, ms, n_threads=4)
    _, dg, wn, sc = ora.cpu_baseline(pcm.view(np.uint8), n_streams, n_msgs, ms * 2, wake, group, 4,
                                     use_openssl=True, want_outputs=True)
    msgs = np.zeros(n_msgs * n_streams, dtype=L.MSG_DTYPE)
    view, off0 = ctx.reserve(pcm.nbytes)
    view[:pcm.nbytes] = pcm.view(np.uint8)
    k = np.arange(n_msgs * n_streams)
    msgs["ring_offset"] = off0 + k * ms * 2
    msgs["nbytes"] = ms * 2
    msgs["stream_id"] = k % n_streams
    msgs["flags"] = np.where(k // n_streams < wake, L.MSG_WAKE, 0) | np.where(k < n_streams, L.MSG_FIRST, 0)
    res, _ = ctx.wait(ctx.submit(msgs))
    assert np.array_equal(res["digest"].reshape(n_msgs, n_streams, 32), dg)
    seg = np.arange(0, n_streams + 1, group, dtype=np.uint32)
    winner, scores, status = ctx.arbitrate(seg, np.arange(n_streams, dtype=np.uint32))
    assert np.array_equal(scores, sc)                       # float64, bit for bit
    assert np.array_equal(winner, np.arange(0, n_streams, group) + wn)
    assert (wn == 0).all()                                  # relay 0 has the highest gain
    assert (status[winner] == L.RELAY_WINNER).all() and (status == L.RELAY_WINNER).sum() == len(winner)
    # top-2 gap of every window is far above rounding (the generator guarantees tie-free input)
    s2 = np.sort(scores.reshape(-1, group), axis=1)
    assert ((s2[:, -1] - s2[:, -2]) / s2[:, -1] > 1e-3).all()


@pytest.mark.parametrize("msg_bytes", [640, 128, 64, 16, 656, 3840, 992, 641, 3, 130])
def test_uniform_batches_through_the_tensor_map_path(msg_bytes):
    """Equal-sized chunks back to back in the ring are staged with one 2-D TMA copy per 32 messages
    (ga_k_ingest_uniform, picked automatically by ga_submit); a context forced to per-message bulk
    copies must give byte-identical records and samples, and both must match the oracle."""
    rng = np.random.default_rng(msg_bytes)
    n = 333                                               # not a multiple of 32: partial last tile
    chunks = [rng.integers(0, 256, size=msg_bytes, dtype=np.uint8).tobytes() for _ in range(n)]
    sid = np.arange(n, dtype=np.uint32) % 50
    fl = np.where(np.arange(n) < 50, L.MSG_FIRST, 0).astype(np.uint32)
    out = {}
    for staging in (L.STAGING_AUTO, L.STAGING_BULK):
        with api.Context(max_streams=64, max_batch_msgs=512, ring_bytes=4 << 20, arena_samples=32_000,
                         wake_samples_max=64, mode=L.DO_ALL, want_frames=True, staging=staging) as c:
            res, frames = c.ingest(chunks, sid, fl)
            out[staging] = (res.copy(), frames.copy(), [c.fetch_samples(s) for s in range(50)])
    a, b = out[L.STAGING_AUTO], out[L.STAGING_BULK]
    assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()
    for x, y in zip(a[2], b[2]):
        assert np.array_equal(bits(x), bits(y))
    fpos = 0
    for i in range(0, n, 7):
        nf = (msg_bytes // 2 + 319) // 320
        check_message(a[0][i], chunks[i], a[1][i * nf:(i + 1) * nf])
    want0 = np.concatenate([ora.bytes_to_float32(chunks[i])[0] for i in range(n) if sid[i] == 0])
    assert np.array_equal(bits(a[2][0]), bits(want0))


def test_full_size_cfg3_batch_matches_the_cpu_baseline_exactly():
    """BASELINE.json configs[2] at FULL size — 4,096 streams x 250 chunks = 1,024,000 messages,
    655 MB of PCM — in one launch of the fused kernel: every digest, every wake-word score and every
    winner against the multi-threaded CPU baseline, plus a checksum of the converted samples."""
    import torch
    dev = torch.device("cuda", 0)
    S, gs, M, ms, wake = 4096, 8, 250, 320, 15
    pcm_host = api.synth_batch_host(synth.SEED, 0, S, gs, M, ms, n_threads=16)
    _, dg, wn, sc = ora.cpu_baseline(pcm_host.view(np.uint8), S, M, ms * 2, wake, gs, 16, want_outputs=True)
    pcm = torch.from_numpy(pcm_host).to(dev)
    n = S * M
    k = np.arange(n, dtype=np.uint64)
    arena = M * ms
    descs = api.make_descs(k * (ms * 2), np.full(n, ms * 2, np.uint32), (k % S) * arena + (k // S) * ms,
                           (k % S).astype(np.uint32), frame_off=np.zeros(n, np.uint32))
    d_desc = torch.from_numpy(descs.view(np.uint8)).to(dev)
    samples = torch.empty(S * arena + 4, dtype=torch.float32, device=dev)
    results = torch.empty(n * 64, dtype=torch.uint8, device=dev)
    api.k_ingest_uniform(pcm.data_ptr(), d_desc.data_ptr(), n, 0, ms * 2, ms * 2, samples.data_ptr(), results.data_ptr())
    off = torch.arange(S, dtype=torch.int64, device=dev) * arena
    cnt = torch.full((S,), wake * ms, dtype=torch.int32, device=dev)
    st = torch.full((S,), L.RELAY_CONTENDING, dtype=torch.uint8, device=dev)
    seg = torch.arange(0, S + 1, gs, dtype=torch.int32, device=dev)
    scores = torch.zeros(S, dtype=torch.float64, device=dev)
    winner = torch.zeros(S // gs, dtype=torch.int32, device=dev)
    st_out = torch.zeros(S, dtype=torch.uint8, device=dev)
    api.k_arbitrate(samples.data_ptr(), off.data_ptr(), cnt.data_ptr(), st.data_ptr(), seg.data_ptr(), S // gs, S,
                    scores.data_ptr(), winner.data_ptr(), st_out.data_ptr())
    torch.cuda.synchronize()
    res = results.cpu().numpy().view(L.RESULT_DTYPE).reshape(M, S)
    assert np.array_equal(res["digest"], dg)                                  # 1,024,000 digests
    assert np.array_equal(scores.cpu().numpy(), sc)                           # float64, bit for bit
    assert np.array_equal(winner.cpu().numpy().astype(np.int64), np.arange(0, S, gs) + wn)
    # samples: conversion is a pure function of the int16 value, so a per-value table gives the exact answer
    table = ora.bytes_to_float32(np.arange(-32768, 32768, dtype=np.int32).astype("<i2").tobytes())[0]
    for s_id in (0, 1234, 4095):
        want = table[pcm_host.reshape(M, S, ms)[:, s_id, :].reshape(-1).astype(np.int32) + 32768]
        got = samples[s_id * arena:(s_id + 1) * arena].cpu().numpy()
        assert np.array_equal(bits(got), bits(want))
    total = samples[:S * arena].double().sum().item()
    want_total = float(table.astype(np.float64)[pcm_host.astype(np.int32) + 32768].sum())
    assert abs(total - want_total) <= 1e-9 * max(1.0, abs(want_total)) + 1e-3   # order-independent checksum


def test_mode_subsets_agree_with_the_fused_kernel():
    rng = np.random.default_rng(9)
    chunks = [rng.integers(0, 256, size=int(n), dtype=np.uint8).tobytes()
              for n in [640] * 40 + [0, 1, 77, 641, 1000]]
    sid = np.arange(len(chunks), dtype=np.uint32)
    fl = np.full(len(chunks), L.MSG_FIRST, np.uint32)
    out = {}
    for mode in (L.DO_ALL, L.DO_SHA256, L.DO_FEATURES, L.DO_CONVERT | L.DO_FEATURES, L.DO_CONVERT,
                 L.DO_CONVERT | L.DO_SHA256, L.DO_FEATURES | L.DO_SHA256):
        with api.Context(max_streams=64, max_batch_msgs=256, ring_bytes=1 << 20, arena_samples=4096,
                         wake_samples_max=64, mode=mode) as c:
            res, _ = c.ingest(chunks, sid, fl)
            smp = [c.fetch_samples(i) for i in range(len(chunks))] if mode & L.DO_CONVERT else None
            out[mode] = (res, smp)
    full, full_smp = out[L.DO_ALL]
    for mode, (res, smp) in out.items():
        assert np.array_equal(res["n_samples"], full["n_samples"])
        if mode & L.DO_SHA256:
            assert np.array_equal(res["digest"], full["digest"]), mode
        else:
            assert not res["digest"].any()
        if mode & L.DO_FEATURES:
            for f in ("sumsq", "rms", "nonzero", "peak", "flags"):
                assert np.array_equal(res[f], full[f]), (mode, f)
        if smp is not None:
            for a, b in zip(smp, full_smp):
                assert np.array_equal(bits(a), bits(b))


def test_boundary_misuse_is_reported_not_fatal(ctx):
    with pytest.raises(L.GpuAudioError) as e:
        ctx.submit(np.array([(12345 * 16, 640, 0, 0, 0)], dtype=L.MSG_DTYPE) if False else
                   np.array([(1 << 50, 640, 0, 0, 0)], dtype=L.MSG_DTYPE))
    assert e.value.code == L.GA_ERR_ARG
    off = ctx.push(b"\x00" * 64)
    with pytest.raises(L.GpuAudioError) as e:
        ctx.submit(np.array([(off, 64, 10_000, 0, 0)], dtype=L.MSG_DTYPE))
    assert e.value.code == L.GA_ERR_ARG and "stream_id" in str(e.value)
    big = np.zeros(1, dtype=L.MSG_DTYPE)
    big["ring_offset"] = ctx.push(b"\x01" * 50_000); big["nbytes"] = 50_000; big["flags"] = L.MSG_WAKE | L.MSG_FIRST
    with pytest.raises(L.GpuAudioError) as e:       # 25,000 wake samples > wake_samples_max 20,000
        ctx.submit(big)
    assert e.value.code == L.GA_ERR_OVERFLOW
    # the context is still usable afterwards
    res, _ = ctx.ingest([b"abc"], [0], [L.MSG_FIRST])
    assert res["digest"][0].tobytes() == ora.sha256(b"abc")

"""tests/golden/path_golden.json — arbitration windows, the 48 kHz decimator, the WAV payload and
the running utterance digest, produced by tools/gen_golden_path.py from implementations that share
no code with the C oracle or the CUDA path (numpy / exact-rational fmaf / struct / hashlib).
CPU half: the C oracle against the fixture.  GPU half: the CUDA path, through the C ABI, against
the same fixture — no oracle involved."""
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from oracle import oracle as ora

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "path_golden.json")


@pytest.fixture(scope="module")
def g():
    with open(GOLDEN) as f:
        return json.load(f)


def f32(bits):
    return np.asarray(bits, dtype=np.uint32).view(np.float32)


def bits(a):
    return [int(b) for b in np.asarray(a, dtype=np.float32).view(np.uint32)]


def d64(hexs):
    return struct.unpack("<d", bytes.fromhex(hexs))[0]


def member_signal(m):
    return ora.bytes_to_float32(bytes.fromhex(m["pcm_hex"]))[0] if "pcm_hex" in m else f32(m["f32_bits"])


# ---------------------------------------------------------------- CPU: the oracle is held to the fixture
def test_oracle_taps_are_the_committed_header(g):
    assert bits(ora.resample_taps()) == g["tap_bits"]


def test_oracle_arbitration(g):
    for w in g["arbitration"]:
        sig = [member_signal(m) for m in w["members"]]
        st_in = np.array([m["status"] for m in w["members"]], np.uint8)
        winner, scores, st_out = ora.arbitrate(sig, st_in)
        assert winner == w["winner"], w["name"]
        assert list(st_out) == w["status_out"], w["name"]
        for sc, want in zip(scores, w["scores_f64_hex"]):
            if want is not None:
                assert struct.pack("<d", sc).hex() == want, w["name"]


def test_oracle_decimator(g):
    for s in g["resample"]:
        r = ora.Resampler()
        for c in s["chunks"]:
            y = r.process(bytes.fromhex(c["pcm_hex"]))
            assert bits(y) == c["out_f32_bits"], s["name"]
            rms, st = ora.signal_strength(y)
            assert (struct.pack("<d", st.sumsq).hex(), struct.pack("<d", rms).hex()) == (c["sumsq_f64_hex"], c["rms_f64_hex"])


def test_oracle_wav_and_utterance(g):
    assert ora.float32_to_wav(f32(g["wav"]["f32_bits"]), g["wav"]["sample_rate"]).hex() == g["wav"]["wav_hex"]
    assert ora.float32_to_wav(np.zeros(0, np.float32), 16000).hex() == g["wav"]["empty_wav_hex"]
    u, acc = g["utterance"], b""
    for c, want in zip(u["chunks_hex"], u["running_sha256"]):
        acc += bytes.fromhex(c)
        assert ora.sha256(acc).hex() == want


def _frame_tuple(a):
    return (int(a["n_samples"]), struct.pack("<d", a["sumsq"]).hex(), struct.pack("<d", a["rms"]).hex(),
            int(a["nonzero"]), int(np.float32(a["peak"]).view(np.uint32)), int(a["vad"]))


def _want_frames(m):
    return [(f["n_samples"], f["sumsq_f64_hex"], f["rms_f64_hex"], f["nonzero"], f["peak_bits"], f["vad"])
            for f in m["frames"]]


def test_oracle_frame_features(g):
    for m in g["frames"]:
        x, _ = ora.bytes_to_float32(bytes.fromhex(m["pcm_hex"]))
        got = ora.frame_features(x, 320, 0.01)
        assert [_frame_tuple(a) for a in got] == _want_frames(m)


# ---------------------------------------------------------------- GPU: the CUDA path is held to the fixture
@pytest.mark.gpu
def test_gpu_frame_features(g):
    import __graft_entry__ as entry
    from loqa_hub_b200 import _lib as L
    from loqa_hub_b200 import api
    entry.build()
    with api.Context(device=0, max_streams=8, max_batch_msgs=16, ring_bytes=1 << 20, arena_samples=4_000,
                     wake_samples_max=400, mode=L.DO_ALL, want_frames=True) as c:
        msgs = g["frames"]
        res, fr = c.ingest([bytes.fromhex(m["pcm_hex"]) for m in msgs], list(range(len(msgs))),
                           [L.MSG_FIRST] * len(msgs))
        pos = 0
        for m in msgs:
            want = _want_frames(m)
            assert [_frame_tuple(a) for a in fr[pos:pos + len(want)]] == want
            pos += len(want)
        assert pos == len(fr)


@pytest.mark.gpu
def test_gpu_arbitration_kernels(g):
    import torch
    import __graft_entry__ as entry
    from loqa_hub_b200 import _lib as L
    from loqa_hub_b200 import api
    entry.build()
    dev = torch.device("cuda", 0)
    sig, status, seg = [], [], [0]
    for w in g["arbitration"]:
        for m in w["members"]:
            sig.append(member_signal(m)); status.append(m["status"])
        seg.append(len(sig))
    n, stride = len(sig), 1600
    flat = np.zeros((n, stride), np.float32)
    for i, s in enumerate(sig):
        flat[i, :s.size] = s
    t_s = torch.from_numpy(flat).to(dev)
    t_off = torch.from_numpy(np.arange(n, dtype=np.int64) * stride).to(dev)
    t_cnt = torch.from_numpy(np.array([s.size for s in sig], np.int32)).to(dev)
    t_st = torch.from_numpy(np.array(status, np.uint8)).to(dev)
    t_seg = torch.from_numpy(np.array(seg, np.int32)).to(dev)
    scores = torch.full((n,), -1.0, dtype=torch.float64, device=dev)
    winner = torch.full((len(seg) - 1,), -7, dtype=torch.int32, device=dev)
    st_out = torch.full((n,), 9, dtype=torch.uint8, device=dev)
    api.k_arbitrate(t_s.data_ptr(), t_off.data_ptr(), t_cnt.data_ptr(), t_st.data_ptr(), t_seg.data_ptr(),
                    len(seg) - 1, n, scores.data_ptr(), winner.data_ptr(), st_out.data_ptr())
    torch.cuda.synchronize()
    winner, scores, st_out = winner.cpu().numpy(), scores.cpu().numpy(), st_out.cpu().numpy()
    for k, w in enumerate(g["arbitration"]):
        lo, hi = seg[k], seg[k + 1]
        assert int(winner[k]) == (lo + w["winner"] if w["winner"] >= 0 else -1), w["name"]
        assert list(st_out[lo:hi]) == w["status_out"], w["name"]
        for m, want in zip(range(lo, hi), w["scores_f64_hex"]):
            assert struct.pack("<d", scores[m]).hex() == (want if want is not None else struct.pack("<d", -1.0).hex()), w["name"]
    assert L.NO_WINNER == 0xFFFFFFFF


@pytest.mark.gpu
def test_gpu_decimator_context(g):
    import __graft_entry__ as entry
    from loqa_hub_b200 import _lib as L
    from loqa_hub_b200 import api
    entry.build()
    with api.Context(device=0, max_streams=4, max_batch_msgs=64, max_inflight=2, ring_bytes=1 << 20,
                     arena_samples=8_000, wake_samples_max=400, mode=L.DO_ALL, want_frames=False,
                     input_format=L.FMT_STEREO48K) as c:
        for sid, s in enumerate(g["resample"]):
            want = []
            for k, ch in enumerate(s["chunks"]):                 # one chunk per batch: history from the saved state
                data = bytes.fromhex(ch["pcm_hex"])
                res, _ = c.ingest([data], [sid], [L.MSG_FIRST if k == 0 else 0])
                assert res["digest"][0].tobytes().hex() == ch["sha256"]
                assert res["n_samples"][0] == len(ch["out_f32_bits"])
                assert (struct.pack("<d", res["sumsq"][0]).hex(), struct.pack("<d", res["rms"][0]).hex()) == \
                       (ch["sumsq_f64_hex"], ch["rms_f64_hex"]), (s["name"], k)
                want += ch["out_f32_bits"]
            assert bits(c.fetch_samples(sid)) == want, s["name"]
        # the same stream again with both chunks in ONE batch: history taken from the slab
        s = g["resample"][1]
        chunks = [bytes.fromhex(ch["pcm_hex"]) for ch in s["chunks"]]
        c.ingest(chunks, [3] * len(chunks), [L.MSG_FIRST] + [0] * (len(chunks) - 1))
        assert bits(c.fetch_samples(3)) == [b for ch in s["chunks"] for b in ch["out_f32_bits"]]


@pytest.mark.gpu
def test_gpu_wav_and_utterance_digest(g):
    import __graft_entry__ as entry
    from loqa_hub_b200 import _lib as L
    from loqa_hub_b200 import api
    import ctypes as C
    entry.build()
    u = g["utterance"]
    with api.Context(device=0, max_streams=2, max_batch_msgs=16, max_inflight=2, ring_bytes=1 << 20,
                     arena_samples=8_000, wake_samples_max=400, mode=L.DO_ALL | L.DO_UTTERANCE) as c:
        for k, (ch, want) in enumerate(zip(u["chunks_hex"], u["running_sha256"])):
            c.ingest([bytes.fromhex(ch)], [1], [L.MSG_FIRST if k == 0 else 0])
            digest, _ = c.utterance_digest(1)
            assert bytes(digest).hex() == want, k
        assert c.samples_digest(1).hex() == u["samples_sha256"]
        # WAV of the stream = header + the samples as they sit in HBM
        samples = c.fetch_samples(1)
        n = C.c_size_t()
        L.check(L.load().ga_fetch_wav(c._h, 1, 16000, None, 0, C.byref(n)))
        out = np.zeros(n.value, np.uint8)
        L.check(L.load().ga_fetch_wav(c._h, 1, 16000, L.ptr(out), out.size, C.byref(n)))
        assert out.tobytes()[:4] == b"RIFF" and out.tobytes()[44:] == samples.tobytes()
        assert out.tobytes()[:44].hex()[:8] == g["wav"]["wav_hex"][:8]
        assert hashlib.sha256(samples.tobytes()).hexdigest() == u["samples_sha256"]
    # the fixture's own WAV: push its samples through a stream as int16 is not possible (arbitrary
    # floats), so compare the 44-byte header layout for the same sample count instead
    with api.Context(device=0, max_streams=1, max_batch_msgs=4, ring_bytes=1 << 16, arena_samples=400,
                     wake_samples_max=40, mode=L.DO_ALL) as c:
        nwav = len(g["wav"]["f32_bits"])
        c.ingest([bytes(2 * nwav)], [0], [L.MSG_FIRST])
        n = C.c_size_t()
        out = np.zeros(44 + 4 * nwav, np.uint8)
        L.check(L.load().ga_fetch_wav(c._h, 0, g["wav"]["sample_rate"], L.ptr(out), out.size, C.byref(n)))
        assert out.tobytes()[:44].hex() == g["wav"]["wav_hex"][:88]

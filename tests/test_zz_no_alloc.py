"""ga_k_arbitrate_noalloc — the arbitration entry point for callers that must not allocate GPU
memory (a hub that gave all of its HBM to its own pool at start-up, or a window-close path that
cannot afford the device-wide synchronisation of cudaMalloc while ingest streams are running).

It keeps the per-member scores in an array built into the library instead of taking a score
buffer, so this test calls it through the C ABI next to the ordinary ga_k_arbitrate on the same
device data and expects identical winners and statuses (the decision is exact, no tolerance),
on two calls in a row: the second call uses fewer, different members, so anything left in the
shared array by the first call would show.  The winners are also held to the CPU oracle's
performArbitration window by window, so the check does not rest on one CUDA path agreeing with
another.  On a box with a second GPU the same runs on cuda:1 — the array's address is looked up
per device, and a hub with one context per GPU calls this on each of them.  Runs last (file name)
and also passes on its own.
"""
import numpy as np
import pytest
import torch

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api
from oracle import oracle as ora

pytestmark = pytest.mark.gpu
MAX_MEMBERS = 65536


def _case(dev, rng, n_windows, max_members, n_samples):
    sizes = rng.integers(0, max_members + 1, size=n_windows)
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
    n = int(seg[-1])
    lens = rng.integers(0, n_samples + 1, size=n).astype(np.uint32)
    lens[rng.random(n) < 0.1] = 0                                  # empty wake-word buffers
    stride = (n_samples + 3) // 4 * 4
    samples = (rng.standard_normal((max(n, 1), stride)) * rng.random((max(n, 1), 1))).astype(np.float32)
    samples[rng.random(max(n, 1)) < 0.1] = 0.0                      # silent relays (all-zero windows happen)
    status = np.where(rng.random(n) < 0.85, L.RELAY_CONTENDING, L.RELAY_CANCELLED).astype(np.uint8)
    t = dict(
        host=(samples, lens, status, seg),
        samples=torch.from_numpy(samples).to(dev),
        off=torch.from_numpy((np.arange(n, dtype=np.int64) * stride)).to(dev),
        cnt=torch.from_numpy(lens.astype(np.int32)).to(dev),
        status=torch.from_numpy(status).to(dev),
        seg=torch.from_numpy(seg.astype(np.int32)).to(dev),
    )
    return t, n_windows, n


def _run_both(dev, t, n_windows, n):
    out = {}
    for name in ("plain", "noalloc"):
        winner = torch.full((n_windows,), -7, dtype=torch.int32, device=dev)
        st_out = torch.full((max(n, 1),), 9, dtype=torch.uint8, device=dev)
        args = (t["samples"].data_ptr(), t["off"].data_ptr(), t["cnt"].data_ptr(), t["status"].data_ptr(),
                t["seg"].data_ptr(), n_windows, n)
        if name == "plain":
            scores = torch.zeros(max(n, 1), dtype=torch.float64, device=dev)
            api.k_arbitrate(*args, scores.data_ptr(), winner.data_ptr(), st_out.data_ptr())
        else:
            api.k_arbitrate_noalloc(*args, winner.data_ptr(), st_out.data_ptr())
        torch.cuda.synchronize()
        out[name] = (winner.cpu().numpy(), st_out.cpu().numpy()[:n])
    return out


def _oracle_winners(t, n_windows):
    """performArbitration (audio_service.go:505-571) per window on the host copies of the same data."""
    samples, lens, status, seg = t["host"]
    win = np.full(n_windows, L.NO_WINNER, dtype=np.int64)
    st_out = np.zeros(int(seg[-1]), dtype=np.uint8)
    for w in range(n_windows):
        lo, hi = int(seg[w]), int(seg[w + 1])
        if hi == lo:
            continue
        k, _, so = ora.arbitrate([samples[m, :lens[m]] for m in range(lo, hi)], status[lo:hi])
        win[w] = lo + k
        st_out[lo:hi] = so
    return win, st_out


def _twice_in_a_row(dev):
    rng = np.random.default_rng(20261106 + dev.index)
    # a warm-up call that does only the per-device set-up (what a host does at start-up)
    with torch.cuda.device(dev):
        api.k_arbitrate_noalloc(0, 0, 0, 0, 0, 0, 0, 0, 0)
        # call 1: many members; call 2: fewer and different ones, reusing the same built-in scratch
        for n_windows, max_members, n_samples in ((700, 12, 1500), (90, 5, 333)):
            t, nw, n = _case(dev, rng, n_windows, max_members, n_samples)
            out = _run_both(dev, t, nw, n)
            assert np.array_equal(out["noalloc"][0], out["plain"][0])          # winners
            assert np.array_equal(out["noalloc"][1], out["plain"][1])          # statuses
            w = out["noalloc"][0].astype(np.int64) & 0xFFFFFFFF
            assert ((w == L.NO_WINNER) | (w < n)).all()
            assert (w != L.NO_WINNER).sum() > 0
            want_w, want_st = _oracle_winners(t, nw)                           # ... and against the CPU oracle
            assert np.array_equal(w, want_w)
            assert np.array_equal(out["noalloc"][1], want_st)


def test_noalloc_arbitration_matches_the_ordinary_entry_point_twice_in_a_row():
    import __graft_entry__ as entry
    entry.build()
    _twice_in_a_row(torch.device("cuda", 0))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs a second GPU")
def test_noalloc_arbitration_on_a_second_device():
    """Each GPU has its own copy of the built-in score array at its own address: a call made with
    cuda:1 current must use that one (looked up with cuda:1 current), not cuda:0's."""
    import __graft_entry__ as entry
    entry.build()
    _twice_in_a_row(torch.device("cuda", 1))
    _twice_in_a_row(torch.device("cuda", 0))


def test_noalloc_refuses_more_members_than_its_scratch_holds():
    import __graft_entry__ as entry
    entry.build()
    dev = torch.device("cuda", 0)
    dummy = torch.zeros(64, dtype=torch.float32, device=dev)
    seg = torch.zeros(2, dtype=torch.int32, device=dev)
    with pytest.raises(L.GpuAudioError) as e:
        api.k_arbitrate_noalloc(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                seg.data_ptr(), 1, MAX_MEMBERS + 1, dummy.data_ptr(), dummy.data_ptr())
    assert e.value.code == L.GA_ERR_OVERFLOW and "65536" in str(e.value)

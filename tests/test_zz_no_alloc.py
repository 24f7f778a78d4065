"""ga_k_arbitrate_noalloc — the arbitration entry point for callers that must not allocate GPU
memory (a hub that gave all of its HBM to its own pool at start-up, or a window-close path that
cannot afford the device-wide synchronisation of cudaMalloc while ingest streams are running).

It keeps the per-member scores in an array built into the library instead of taking a score
buffer, so this test calls it through the C ABI next to the ordinary ga_k_arbitrate on the same
device data and expects identical winners and statuses (the decision is exact, no tolerance),
on two calls in a row: the second call uses fewer, different members, so anything left in the
shared array by the first call would show.  Runs last (file name) and also passes on its own.
"""
import numpy as np
import pytest
import torch

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api

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


def test_noalloc_arbitration_matches_the_ordinary_entry_point_twice_in_a_row():
    import __graft_entry__ as entry
    entry.build()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(20261106)
    # call 1: many members; call 2: fewer and different ones, reusing the same built-in scratch
    for n_windows, max_members, n_samples in ((700, 12, 1500), (90, 5, 333)):
        t, nw, n = _case(dev, rng, n_windows, max_members, n_samples)
        out = _run_both(dev, t, nw, n)
        assert np.array_equal(out["noalloc"][0], out["plain"][0])          # winners
        assert np.array_equal(out["noalloc"][1], out["plain"][1])          # statuses
        w = out["plain"][0].astype(np.int64) & 0xFFFFFFFF
        assert ((w == L.NO_WINNER) | (w < n)).all()
        assert (w != L.NO_WINNER).sum() > 0


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

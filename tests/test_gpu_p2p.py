"""Fused cross-shard arbitration (ga_k_arbitrate_p2p) on ONE GPU.

Two shards of a world-size-2 plan live on cuda:0 and use each other's inbox through plain device
pointers (over NVLink the same addresses come from ga_p2p_open).  Kernels of different shards
must not wait on one another on a single GPU (nothing guarantees they are co-resident), so the
exchange is exercised in two passes at the same epoch: pass 1 delivers scores and flags without
waiting, pass 2 repeats the step with the wait enabled — every flag is already there — and its
winners are checked against the un-sharded CPU baseline.  The real two-GPU run (with concurrent
waiting) is `bench.py --gpus 2 --exchange p2p`, which verifies the winners itself."""
import numpy as np
import pytest
import torch

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api, p2p, sharding, synth
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


def test_fused_exchange_two_shards_on_one_gpu():
    import __graft_entry__ as entry
    entry.build()
    dev = torch.device("cuda", 0)
    world, gpr, gs, n_msgs, ms, wake = 2, 16, 8, 18, 320, 15
    plans = [sharding.make_plan(r, world, gpr, gs, split_every=4) for r in range(world)]
    regions = [p2p.Region(pl.n_guest_max, world) for pl in plans]
    bases = [reg.ptr.value for reg in regions]
    to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    shards = []
    for pl, reg in zip(plans, regions):
        S, arena = pl.n_streams, n_msgs * ms
        pcm = torch.empty(n_msgs * S * ms, dtype=torch.int16, device=dev)
        g, r = to_dev(pl.stream_group.astype(np.int32)), to_dev(pl.stream_relay.astype(np.int32))
        api.synth_batch_device_map(pcm.data_ptr(), synth.SEED, g.data_ptr(), r.data_ptr(), S, n_msgs, ms)
        k = np.arange(n_msgs * S, dtype=np.uint64)
        descs = api.make_descs(k * (ms * 2), np.full(k.size, ms * 2, np.uint32), (k % S) * arena + (k // S) * ms)
        d_desc = to_dev(descs.view(np.uint8))
        samples = torch.zeros(S * arena + 4, dtype=torch.float32, device=dev)
        results = torch.zeros(k.size * 64, dtype=torch.uint8, device=dev)
        api.k_ingest(pcm.data_ptr(), d_desc.data_ptr(), k.size, samples.data_ptr(), results.data_ptr(), 0, L.DO_CONVERT)
        members, n_slots = sharding.score_members(pl)
        local = members >= 0
        t = dict(samples=samples,
                 wake_off=to_dev(np.where(local, members, 0).astype(np.int64) * arena),
                 wake_n=to_dev(np.where(local, wake * ms, 0).astype(np.int32)),
                 status_in=torch.full((len(members),), L.RELAY_CONTENDING, dtype=torch.uint8, device=dev),
                 seg=to_dev(pl.seg_offsets.astype(np.int32)),
                 scores=torch.zeros(len(members), dtype=torch.float64, device=dev),
                 winner=torch.full((pl.n_windows,), -5, dtype=torch.int32, device=dev),
                 status_out=torch.zeros(n_slots, dtype=torch.uint8, device=dev))
        ex = p2p.Exchange(pl, reg, bases, {n: v.data_ptr() for n, v in t.items()}, to_dev, timeout_cycles=200_000_000)
        shards.append((pl, t, ex, n_slots))
    torch.cuda.synchronize()
    for epoch in (1, 2, 3):                      # several steps: both inbox parities, growing flags
        for _, _, ex, _ in shards:
            ex.launch(epoch=epoch, wait=False)   # pass 1: deliver, do not wait
        torch.cuda.synchronize()
        for _, t, ex, _ in shards:
            t["winner"].fill_(-5)
            ex.launch(epoch=epoch, wait=True)    # pass 2: flags are already >= epoch
        torch.cuda.synchronize()
        assert all(ex.error_flag() == 0 for _, _, ex, _ in shards)
    # un-sharded reference
    n_groups = gpr * world
    pcm_all = api.synth_batch_host(synth.SEED, 0, n_groups * gs, gs, n_msgs, ms, n_threads=4)
    _, _, wn, sc = ora.cpu_baseline(pcm_all.view(np.uint8), n_groups * gs, n_msgs, ms * 2, wake, gs, 4,
                                    want_outputs=True)
    for pl, t, _, n_slots in shards:
        got_scores = t["scores"][:n_slots].cpu().numpy()
        got_winner = t["winner"].cpu().numpy().astype(np.int64)
        status = t["status_out"].cpu().numpy()
        for w in range(pl.n_windows):
            a, b = int(pl.seg_offsets[w]), int(pl.seg_offsets[w + 1])
            g = int(pl.member_global[a, 0])
            assert np.array_equal(got_scores[a:b], sc[g * gs:(g + 1) * gs]), (pl.rank, w)
            assert got_winner[w] - a == wn[g]
            assert status[got_winner[w]] == L.RELAY_WINNER and (status[a:b] == L.RELAY_WINNER).sum() == 1
    for reg in regions:
        reg.free()


def test_wait_times_out_instead_of_hanging():
    """A shard whose neighbour never delivers must come back with the error word set."""
    import __graft_entry__ as entry
    entry.build()
    dev = torch.device("cuda", 0)
    pl = sharding.make_plan(0, 2, 4, 8, split_every=2)
    reg = p2p.Region(pl.n_guest_max, 2)
    other = p2p.Region(pl.n_guest_max, 2)
    to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    members, n_slots = sharding.score_members(pl)
    samples = torch.ones(pl.n_streams * 64 + 4, dtype=torch.float32, device=dev)
    local = members >= 0
    t = dict(samples=samples, wake_off=to_dev(np.where(local, members, 0).astype(np.int64) * 64),
             wake_n=to_dev(np.where(local, 64, 0).astype(np.int32)),
             status_in=torch.full((len(members),), L.RELAY_CONTENDING, dtype=torch.uint8, device=dev),
             seg=to_dev(pl.seg_offsets.astype(np.int32)),
             scores=torch.zeros(len(members), dtype=torch.float64, device=dev),
             winner=torch.zeros(pl.n_windows, dtype=torch.int32, device=dev),
             status_out=torch.zeros(n_slots, dtype=torch.uint8, device=dev))
    ex = p2p.Exchange(pl, reg, [reg.ptr.value, other.ptr.value], {n: v.data_ptr() for n, v in t.items()}, to_dev,
                      timeout_cycles=20_000_000)     # ~10 ms
    ex.launch(epoch=1, wait=True)
    torch.cuda.synchronize()
    assert ex.error_flag() == 1
    reg.free(); other.free()

"""The N > 1 arbitration path on ONE GPU: both ranks' shards are processed in turn on
cuda:0 and the all-gather is a concatenation on the device, so ga_k_signal_strength ->
ga_k_route_scores -> ga_k_select are exercised exactly as bench.py drives them under
torchrun (whose NCCL all-gather only moves the guest scores between the two steps).
Checked against the CPU baseline run over the un-sharded collision groups."""
import numpy as np
import pytest
import torch

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api, sharding, synth
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


def run_rank(plan, n_msgs, ms, wake, dev):
    S = plan.n_streams
    arena = n_msgs * ms
    pcm = torch.empty(n_msgs * S * ms, dtype=torch.int16, device=dev)
    g = torch.from_numpy(plan.stream_group.astype(np.int64)).to(dev).to(torch.int32)
    r = torch.from_numpy(plan.stream_relay.astype(np.int64)).to(dev).to(torch.int32)
    api.synth_batch_device_map(pcm.data_ptr(), synth.SEED, g.data_ptr(), r.data_ptr(), S, n_msgs, ms)
    k = np.arange(n_msgs * S, dtype=np.uint64)
    m_idx, s_idx = k // S, k % S
    descs = api.make_descs(k * (ms * 2), np.full(k.size, ms * 2, np.uint32), s_idx * arena + m_idx * ms,
                           s_idx.astype(np.uint32), np.where(m_idx < wake, L.MSG_WAKE, 0).astype(np.uint32))
    d_desc = torch.from_numpy(descs.view(np.uint8)).to(dev)
    samples = torch.zeros(S * arena + 4, dtype=torch.float32, device=dev)
    results = torch.zeros(k.size * 64, dtype=torch.uint8, device=dev)
    api.k_ingest(pcm.data_ptr(), d_desc.data_ptr(), k.size, samples.data_ptr(), results.data_ptr())
    members, n_slots = sharding.score_members(plan)
    local = members >= 0
    off = torch.from_numpy(np.where(local, members, 0).astype(np.int64) * arena).to(dev)
    cnt = torch.from_numpy(np.where(local, wake * ms, 0).astype(np.int32)).to(dev)
    st = torch.full((len(members),), L.RELAY_CONTENDING, dtype=torch.uint8, device=dev)
    scores = torch.zeros(len(members), dtype=torch.float64, device=dev)
    api.k_signal_strength(samples.data_ptr(), off.data_ptr(), cnt.data_ptr(), st.data_ptr(), len(members),
                          scores.data_ptr())
    torch.cuda.synchronize()
    return dict(scores=scores, status=st, n_slots=n_slots, results=results, pcm=pcm, keep=(samples, off, cnt))


@pytest.mark.parametrize("world", [2, 4])
def test_cross_shard_windows_match_unsharded_cpu_baseline(world):
    dev = torch.device("cuda", 0)
    gpr, gs, n_msgs, ms, wake = 16, 8, 20, 320, 15
    plans = [sharding.make_plan(r, world, gpr, gs, split_every=4) for r in range(world)]
    ranks = [run_rank(p, n_msgs, ms, wake, dev) for p in plans]
    # the exchange step: all-gather of every rank's guest scores (here: a device concatenation)
    gathered = torch.cat([x["scores"][x["n_slots"]:] for x in ranks]).contiguous()
    assert gathered.numel() == world * plans[0].n_guest_max
    # un-sharded reference: all groups, relay-major inside a group
    n_groups = gpr * world
    pcm_all = api.synth_batch_host(synth.SEED, 0, n_groups * gs, gs, n_msgs, ms, n_threads=4)
    _, dg, wn, sc = ora.cpu_baseline(pcm_all.view(np.uint8), n_groups * gs, n_msgs, ms * 2, wake, gs, 4,
                                     want_outputs=True)
    for p, x in zip(plans, ranks):
        src = torch.from_numpy(p.route_src.astype(np.int32)).to(dev)
        dst = torch.from_numpy(p.route_dst.astype(np.int32)).to(dev)
        api.k_route_scores(gathered.data_ptr(), src.data_ptr(), dst.data_ptr(), len(p.route_dst),
                           x["scores"].data_ptr())
        seg = torch.from_numpy(p.seg_offsets.astype(np.int32)).to(dev)
        winner = torch.zeros(p.n_windows, dtype=torch.int32, device=dev)
        st_out = torch.zeros(x["n_slots"], dtype=torch.uint8, device=dev)
        api.k_select(x["scores"].data_ptr(), x["status"].data_ptr(), seg.data_ptr(), p.n_windows,
                     winner.data_ptr(), st_out.data_ptr())
        torch.cuda.synchronize()
        got_scores = x["scores"][:x["n_slots"]].cpu().numpy()
        got_winner = winner.cpu().numpy().astype(np.int64)
        for w in range(p.n_windows):
            a, b = int(p.seg_offsets[w]), int(p.seg_offsets[w + 1])
            g = int(p.member_global[a, 0])
            assert np.array_equal(got_scores[a:b], sc[g * gs:(g + 1) * gs]), (p.rank, w)   # bit for bit
            assert got_winner[w] - a == wn[g]
        # per-message digests of this shard's streams against the un-sharded run
        res = x["results"].cpu().numpy().view(L.RESULT_DTYPE).reshape(n_msgs, p.n_streams)
        gl = p.stream_group.astype(np.int64) * gs + p.stream_relay.astype(np.int64)
        assert np.array_equal(res["digest"], dg[:, gl, :])

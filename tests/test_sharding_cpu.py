"""Host-side logic of the N > 1 path, on CPU: the shard plan, and (world_size 2, gloo) the
exchange of guest scores + routing into home-window member slots + the selection, checked
against the oracle's select_winner.  No CUDA here — the kernels that do the same routing and
selection on the GPU are covered by tests/test_gpu_multi.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from loqa_hub_b200 import sharding
from oracle import oracle as ora


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_plan_covers_every_relay_exactly_once_and_is_balanced(world):
    plans = [sharding.make_plan(r, world, groups_per_rank=64, group_size=8, split_every=16)
             for r in range(world)]
    seen = set()
    for p in plans:
        pairs = list(zip(p.stream_group.tolist(), p.stream_relay.tolist()))
        assert len(pairs) == len(set(pairs))
        assert not (seen & set(pairs))
        seen |= set(pairs)
        assert p.n_streams == 64 * 8                                   # balanced shards
        assert p.n_windows == 64 and p.seg_offsets[-1] == 64 * 8
        home = p.stream_group[:p.n_home_streams]
        assert (home % world == p.rank).all()                          # home shard = group % world
        if world > 1:
            assert p.n_guest_streams == 4 * 4 and p.n_spanning_windows == 4
            guests = p.stream_group[p.n_home_streams:]
            assert (guests % world == (p.rank - 1) % world).all()
            assert (p.member_stream[p.route_dst] == -1).all()          # routed slots are the remote ones
        else:
            assert p.n_guest_streams == 0 and len(p.route_dst) == 0
    assert seen == {(g, r) for g in range(64 * world) for r in range(8)}


def fake_score(group, relay):
    # strictly decreasing in relay within a group, distinct across groups
    return 1.0 / (1.0 + relay) + 1e-3 * (group % 97)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = sharding.make_plan(rank, world, groups_per_rank=32, group_size=8, split_every=4)
        members, n_slots = sharding.score_members(plan)
        scores = np.zeros(len(members))
        for m, s in enumerate(members):                                # "score kernel": local members only
            if s >= 0:
                scores[m] = fake_score(int(plan.stream_group[s]), int(plan.stream_relay[s]))
        mine = torch.from_numpy(scores[n_slots:].copy())               # this rank's guest scores
        gathered = torch.zeros(world * plan.n_guest_max, dtype=torch.float64)
        dist.all_gather_into_tensor(gathered, mine)                    # the one exchange step
        scores[plan.route_dst] = gathered.numpy()[plan.route_src]      # "route kernel"
        ok = True
        for w in range(plan.n_windows):                                # "select kernel" vs the oracle
            a, b = plan.seg_offsets[w], plan.seg_offsets[w + 1]
            want = [fake_score(int(g), int(r)) for g, r in plan.member_global[a:b]]
            ok &= np.array_equal(scores[a:b], np.array(want))
            win, status = ora.select_winner(scores[a:b])
            ok &= (win == 0 and status[0] == ora.RELAY_WINNER and (status[1:] == ora.RELAY_CANCELLED).all())
        q.put((rank, bool(ok), plan.n_spanning_windows))
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_routes_guest_scores_to_home_windows():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out == [(0, True, 8), (1, True, 8)]

"""Host-side shard plan: which relay streams live on which GPU, which collision windows
span shards, and how exchanged scores are routed back to their window's home shard.

Streams are independent for convert / features / fingerprint, so they shard with no
data-path communication; arbitration couples only the relays of one collision window
(performArbitration, internal/grpc/audio_service.go:505-571).  A window is *homed* on
shard `group % world`.  To exercise windows that span shards (BASELINE.json configs[4])
every `split_every`-th home group of a shard keeps relays [0, split_at) at home and places
relays [split_at, group_size) on the next shard as *guests*.  The only exchange step is
an all-gather of the guests' float64 scores (one value per guest relay); the routing
table that scatters gathered scores into home-window member slots is static per plan —
membership is control-plane state (startArbitrationWindow / joinArbitrationWindow,
:408-491), scores are the data plane.

Pure Python / numpy: no CUDA, no collectives here (bench.py and the tests drive those).
"""
from dataclasses import dataclass, field

import numpy as np


@dataclass
class ShardPlan:
    rank: int
    world: int
    groups_per_rank: int
    group_size: int
    split_every: int
    split_at: int
    # local streams, in local stream-slot order: home relays first, then guests
    stream_group: np.ndarray = field(default=None)    # global collision-group id per local stream
    stream_relay: np.ndarray = field(default=None)    # relay index within its group
    n_home_streams: int = 0
    n_guest_streams: int = 0
    # home windows (one per home group), members in join order = relay order
    seg_offsets: np.ndarray = field(default=None)     # [n_windows + 1] into the member arrays
    member_stream: np.ndarray = field(default=None)   # local stream slot, or -1 if remote
    member_global: np.ndarray = field(default=None)   # (group, relay) pairs, for checking
    # exchange: every rank contributes `n_guest_max` scores (padded); gathered index ->
    # home member slot on THIS rank
    n_guest_max: int = 0
    route_src: np.ndarray = field(default=None)       # index into the flat gathered array
    route_dst: np.ndarray = field(default=None)       # member slot on this rank

    @property
    def n_streams(self):
        return self.n_home_streams + self.n_guest_streams

    @property
    def n_windows(self):
        return self.groups_per_rank

    @property
    def n_spanning_windows(self):
        return int(len(self.route_dst) // max(self.group_size - self.split_at, 1)) if self.world > 1 else 0


def global_group(rank, world, local_group):
    return local_group * world + rank


def is_split(world, local_group, split_every):
    return world > 1 and split_every > 0 and local_group % split_every == 0


def make_plan(rank, world, groups_per_rank=512, group_size=8, split_every=16, split_at=None):
    if split_at is None:
        split_at = group_size // 2
    sg, sr, seg, mem_stream, mem_global = [], [], [0], [], []
    # home relays
    for lg in range(groups_per_rank):
        g = global_group(rank, world, lg)
        split = is_split(world, lg, split_every)
        for r in range(group_size):
            if split and r >= split_at:
                mem_stream.append(-1)                      # lives on shard (rank + 1) % world
            else:
                mem_stream.append(len(sg))
                sg.append(g); sr.append(r)
            mem_global.append((g, r))
        seg.append(len(mem_stream))
    n_home = len(sg)
    # guests: relays [split_at, group_size) of the previous shard's split groups
    prev = (rank - 1) % world
    guests = []
    for lg in range(groups_per_rank):
        if is_split(world, lg, split_every):
            g = global_group(prev, world, lg)
            for r in range(split_at, group_size):
                guests.append((g, r))
                sg.append(g); sr.append(r)
    n_guest = len(guests)
    # routing: this rank's remote members are guests of shard `nxt`, in the same order
    nxt = (rank + 1) % world
    route_src, route_dst = [], []
    k = 0
    for lg in range(groups_per_rank):
        if is_split(world, lg, split_every):
            for r in range(split_at, group_size):
                route_src.append(nxt * n_guest + k)          # every rank pads to the same n_guest
                route_dst.append(seg[lg] + r)
                k += 1
    return ShardPlan(rank, world, groups_per_rank, group_size, split_every, split_at,
                     np.array(sg, np.uint32), np.array(sr, np.uint32), n_home, n_guest,
                     np.array(seg, np.uint32), np.array(mem_stream, np.int64),
                     np.array(mem_global, np.uint32).reshape(-1, 2), n_guest,
                     np.array(route_src, np.uint32), np.array(route_dst, np.uint32))


def score_members(plan):
    """Member list handed to the score kernel on this rank: the home-window slots (remote
    ones get an empty buffer and are filled by the routed scores) followed by the guest
    streams.  Returns (stream slot per scored member with -1 for remote, n_home_slots)."""
    guest_streams = np.arange(plan.n_home_streams, plan.n_streams, dtype=np.int64)
    return np.concatenate([plan.member_stream, guest_streams]), len(plan.member_stream)

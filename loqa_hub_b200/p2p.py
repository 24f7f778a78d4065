"""Harness-side set-up of the fused cross-shard arbitration (ga_k_arbitrate_p2p).

Allocates this rank's peer-visible inbox + flag words with ga_p2p_alloc, swaps the CUDA IPC
handles with the other ranks (any callable that all-gathers a bytes object will do — bench.py
passes a torch.distributed one), maps the peers with ga_p2p_open and lays out the small device
tables the kernel needs.  No computation happens here.

Region layout (per rank):  inbox[parity 0][cap] f64 | inbox[parity 1][cap] f64 | flag[world] u64
where flag[s] is written by sender rank s.
"""
import ctypes as C

import numpy as np

from . import _lib as L


class Region:
    def __init__(self, cap, world):
        self.cap, self.world = max(int(cap), 1), world
        self.parity_stride = self.cap * 8
        self.flags_off = 2 * self.parity_stride
        self.bytes = self.flags_off + world * 8
        self.ptr = C.c_void_p()
        self.handle = (C.c_uint8 * 64)()
        L.check(L.load().ga_p2p_alloc(self.bytes, C.byref(self.ptr), self.handle))

    def free(self):
        if self.ptr:
            L.load().ga_p2p_free(self.ptr)
            self.ptr = C.c_void_p()


def open_peer(handle_bytes):
    buf = (C.c_uint8 * 64).from_buffer_copy(handle_bytes)
    p = C.c_void_p()
    L.check(L.load().ga_p2p_open(buf, C.byref(p)))
    return p.value


def build_tables(plan, bases, region):
    """bases[r] = address of rank r's region as seen from THIS process.  Returns numpy tables."""
    r, world = plan.rank, plan.world
    prev, nxt = (r - 1) % world, (r + 1) % world
    n_guest = plan.n_guest_streams
    guest_dst = (bases[prev] + np.arange(n_guest, dtype=np.uint64) * 8).astype(np.uint64)
    notify = np.array([bases[prev] + region.flags_off + r * 8], dtype=np.uint64) if n_guest else np.zeros(0, np.uint64)
    n_in = len(plan.route_dst)
    wait = np.array([bases[r] + region.flags_off + nxt * 8], dtype=np.uint64) if n_in else np.zeros(0, np.uint64)
    return guest_dst, notify, wait, plan.route_dst.astype(np.uint32)


class Exchange:
    """Device tables + argument block for one rank; .launch(stream) runs one step."""

    def __init__(self, plan, region, bases, dev_arrays, to_device, timeout_cycles=30_000_000_000):
        """dev_arrays: dict of device addresses {samples, wake_off, wake_n, status_in, seg, scores, winner,
        status_out}; to_device(np_array) -> object with .data_ptr() kept alive by the caller."""
        guest_dst, notify, wait, route_dst = build_tables(plan, bases, region)
        self._keep = [to_device(x.view(np.int64)) if x.dtype == np.uint64 else to_device(x.view(np.int32))
                      for x in (guest_dst, notify, wait, route_dst)]
        self._scratch = to_device(np.zeros(2, np.int32))
        a = L.P2PArgs()
        a.d_samples = dev_arrays["samples"]; a.d_wake_off = dev_arrays["wake_off"]; a.d_wake_n = dev_arrays["wake_n"]
        a.d_status_in = dev_arrays["status_in"]; a.d_seg_offsets = dev_arrays["seg"]; a.d_scores = dev_arrays["scores"]
        a.d_winner = dev_arrays["winner"]; a.d_status_out = dev_arrays["status_out"]
        a.d_guest_dst = self._keep[0].data_ptr() if guest_dst.size else None
        a.d_notify = self._keep[1].data_ptr() if notify.size else None
        a.d_wait = self._keep[2].data_ptr() if wait.size else None
        a.d_inbox = bases[plan.rank]
        a.d_route_dst = self._keep[3].data_ptr() if route_dst.size else None
        a.d_counter = self._scratch.data_ptr()
        a.d_error = self._scratch.data_ptr() + 4
        a.inbox_parity_stride = region.parity_stride
        a.epoch = 0
        a.timeout_cycles = timeout_cycles
        a.n_windows = plan.n_windows
        a.n_home_members = len(plan.member_stream)
        a.n_guests = plan.n_guest_streams
        a.n_notify, a.n_wait, a.n_in = notify.size, wait.size, route_dst.size
        self.args = a

    def launch(self, stream=0, epoch=None, wait=True):
        self.args.epoch = self.args.epoch + 1 if epoch is None else epoch
        saved = self.args.n_wait
        if not wait:
            self.args.n_wait = 0
        try:
            L.check(L.load().ga_k_arbitrate_p2p(C.byref(self.args), stream or None))
        finally:
            self.args.n_wait = saved

    def error_flag(self):
        return int(self._scratch.cpu().numpy()[1])

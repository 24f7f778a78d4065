"""Boundary behaviour of the context layer: ring wrap-around and back-pressure, ticket ordering,
in-flight limits and concurrent ga_ring_reserve from many threads (ctypes drops the GIL, so the
lock-free reserve path really runs concurrently)."""
import threading

import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


def make(**kw):
    import __graft_entry__ as entry
    entry.build()
    args = dict(device=0, max_streams=64, max_batch_msgs=1024, max_inflight=2, ring_bytes=1 << 16,
                arena_samples=1 << 20, wake_samples_max=64, mode=L.DO_SHA256 | L.DO_FEATURES)
    args.update(kw)
    return api.Context(**args)


def test_ring_wraps_and_frees_space_only_when_tickets_are_waited():
    with make() as c:
        rng = np.random.default_rng(4)
        for round_ in range(40):                       # 40 x 4 x 3,000 B through a 64 KiB ring: many wraps
            chunks = [rng.integers(0, 256, size=3000, dtype=np.uint8).tobytes() for _ in range(4)]
            res, _ = c.ingest(chunks, [0, 1, 2, 3], [0] * 4)
            for r, d in zip(res, chunks):
                assert r["digest"].tobytes() == ora.sha256(d)
        # fill the ring without waiting: the reservation that does not fit is refused, not overwritten
        offs = []
        with pytest.raises(L.GpuAudioError) as e:
            for _ in range(100):
                offs.append(c.push(b"\x05" * 4096))
        assert e.value.code == L.GA_ERR_RING_FULL and 10 <= len(offs) <= 16
        msgs = np.zeros(len(offs), dtype=L.MSG_DTYPE)
        msgs["ring_offset"] = offs; msgs["nbytes"] = 4096; msgs["stream_id"] = 7
        res, _ = c.wait(c.submit(msgs))
        assert all(r["digest"].tobytes() == ora.sha256(b"\x05" * 4096) for r in res)
        c.push(b"\x06" * 4096)                         # space is back


def test_tickets_in_flight_limit_and_submission_order():
    with make(ring_bytes=1 << 20) as c:
        def one(n):
            m = np.zeros(1, dtype=L.MSG_DTYPE)
            m["ring_offset"] = c.push(bytes([n]) * 100); m["nbytes"] = 100; m["stream_id"] = n
            return m
        t1, t2 = c.submit(one(1)), c.submit(one(2))
        with pytest.raises(L.GpuAudioError) as e:      # max_inflight = 2
            c.submit(one(3))
        assert e.value.code == L.GA_ERR_STATE
        with pytest.raises(L.GpuAudioError) as e:      # tickets are waited oldest first
            c.wait(t2)
        assert e.value.code == L.GA_ERR_STATE
        c._pending[t2] = 1
        assert c.wait(t1)[0]["digest"][0].tobytes() == ora.sha256(bytes([1]) * 100)
        assert c.wait(t2)[0]["digest"][0].tobytes() == ora.sha256(bytes([2]) * 100)
        with pytest.raises(L.GpuAudioError):           # a ticket can be waited once
            c._pending[t1] = 1
            c.wait(t1)


def test_concurrent_ring_reserve_from_many_threads():
    with make(ring_bytes=8 << 20, max_batch_msgs=8192) as c:
        n_threads, per = 8, 400
        got = [[] for _ in range(n_threads)]

        def worker(t):
            rng = np.random.default_rng(100 + t)
            for k in range(per):
                data = rng.integers(0, 256, size=int(rng.integers(1, 1500)), dtype=np.uint8).tobytes()
                got[t].append((c.push(data), data))

        threads = [threading.Thread(target=worker, args=(t,)) for t in range(n_threads)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        allm = sorted((off, d) for g in got for off, d in g)           # submit in ring-offset order
        for (o1, d1), (o2, _) in zip(allm, allm[1:]):
            assert o1 + ((len(d1) + 15) & ~15) <= o2                   # reservations never overlap
        msgs = np.zeros(len(allm), dtype=L.MSG_DTYPE)
        msgs["ring_offset"] = [o for o, _ in allm]
        msgs["nbytes"] = [len(d) for _, d in allm]
        msgs["stream_id"] = np.arange(len(allm)) % 64
        res, _ = c.wait(c.submit(msgs))
        for r, (_, d) in zip(res, allm):
            assert r["digest"].tobytes() == ora.sha256(d)


def test_arbitrate_reuses_its_scratch_and_handles_growing_window_sets():
    with make(mode=L.DO_ALL, arena_samples=4096, wake_samples_max=1024, ring_bytes=1 << 20) as c:
        rng = np.random.default_rng(6)
        sig = [rng.integers(-20000, 20000, size=200).astype("<i2") for _ in range(64)]
        c.ingest([s.tobytes() for s in sig], np.arange(64, dtype=np.uint32), np.full(64, L.MSG_WAKE | L.MSG_FIRST, np.uint32))
        want = np.array([ora.signal_strength(ora.bytes_to_float32(s.tobytes())[0])[0] for s in sig])
        for gs in (2, 4, 8, 64, 1, 16):                 # scratch grows and is reused
            seg = np.arange(0, 65, gs, dtype=np.uint32)
            winner, scores, status = c.arbitrate(seg, np.arange(64, dtype=np.uint32))
            assert np.array_equal(scores, want)
            for w in range(len(seg) - 1):
                a, b = seg[w], seg[w + 1]
                assert winner[w] == a + int(np.argmax(want[a:b]))
            assert (status == L.RELAY_WINNER).sum() == len(seg) - 1


def test_wait_view_records_survive_until_the_slot_comes_round_again():
    """ga_wait_view hands out the slot's pinned buffers; slots are used round-robin, so ticket t's
    records are overwritten only by the submit that returns ticket t + max_inflight."""
    with make(max_inflight=3, ring_bytes=1 << 20) as c:
        def one(n):
            m = np.zeros(2, dtype=L.MSG_DTYPE)
            m["ring_offset"] = [c.push(bytes([n]) * 200), c.push(bytes([n + 100]) * 64)]
            m["nbytes"] = [200, 64]; m["stream_id"] = [n % 64, (n + 1) % 64]
            return m
        want = lambda n: (ora.sha256(bytes([n]) * 200), ora.sha256(bytes([n + 100]) * 64))
        t1, t2 = c.submit(one(1)), c.submit(one(2))
        view1, _ = c.wait_view(t1)
        t3 = c.submit(one(3))                      # tickets 2 and 3 live in the other two slots
        c.wait(t2); c.wait(t3)
        assert (view1["digest"][0].tobytes(), view1["digest"][1].tobytes()) == want(1)
        view_ptr = view1.ctypes.data
        t4 = c.submit(one(4))                      # ticket 4 reuses ticket 1's slot ...
        view4, _ = c.wait_view(t4)
        assert view4.ctypes.data == view_ptr       # ... and therefore its buffers
        assert (view4["digest"][0].tobytes(), view4["digest"][1].tobytes()) == want(4)
        # in a FULL pipeline the very next submit is the one that comes round
        ts = [c.submit(one(n)) for n in (5, 6, 7)]
        v5, _ = c.wait_view(ts[0])
        assert v5["digest"][0].tobytes() == want(5)[0]
        t8 = c.submit(one(8))
        for t in ts[1:] + [t8]:
            last, _ = c.wait_view(t)
        assert last.ctypes.data == v5.ctypes.data and last["digest"][0].tobytes() == want(8)[0]


def test_a_failed_submit_keeps_the_fence_of_the_streams_it_touched():
    """A refused batch must restore, not clear, the ticket that last appended to each stream it had
    already walked over: a restart of such a stream in the next batch still has to wait for the
    batch in flight that writes the same arena space (tickets run on separate CUDA streams)."""
    rng = np.random.default_rng(21)
    S, chunk, T = 32, 6400, 10
    old = rng.integers(0, 256, size=(T, S, chunk), dtype=np.uint8)
    new = rng.integers(0, 256, size=(S, chunk), dtype=np.uint8)
    with make(max_streams=S + 1, max_batch_msgs=T * S, max_inflight=3, ring_bytes=16 << 20, arena_samples=40_000,
              wake_samples_max=400, mode=L.DO_ALL) as c:
        for _ in range(4):
            m1 = np.zeros(T * S, dtype=L.MSG_DTYPE)
            m1["ring_offset"] = [c.push(old[t, s]) for t in range(T) for s in range(S)]
            m1["nbytes"] = chunk; m1["stream_id"] = np.tile(np.arange(S), T)
            m1["flags"][:S] = L.MSG_FIRST
            t1 = c.submit(m1)                                   # in flight, not waited
            bad = np.zeros(S + 1, dtype=L.MSG_DTYPE)            # walks over every stream, then fails
            bad["ring_offset"][:S] = m1["ring_offset"][:S]; bad["nbytes"] = chunk
            bad["stream_id"] = np.arange(S + 1); bad["stream_id"][S] = 4000
            with pytest.raises(L.GpuAudioError):
                c.submit(bad)
            assert c.stream_info(3) == (T * chunk // 2, 0)
            m2 = np.zeros(S, dtype=L.MSG_DTYPE)
            m2["ring_offset"] = [c.push(new[s]) for s in range(S)]
            m2["nbytes"] = chunk; m2["stream_id"] = np.arange(S); m2["flags"] = L.MSG_FIRST
            t2 = c.submit(m2)
            c.wait(t1); c.wait(t2)
            for sid in (0, 13, S - 1):
                assert c.stream_info(sid) == (chunk // 2, 0)
                assert np.array_equal(c.fetch_samples(sid), ora.bytes_to_float32(new[sid].tobytes())[0])

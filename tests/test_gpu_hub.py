"""The C++ hub (csrc/hub.cu) against the reference's arbitration-window tests, restated one by
one from internal/grpc/collision_detection_test.go with injected time instead of sleeps.
Wake-word signals are fed as int16 PCM chunks (the hub's real input); scores and winners come
from the GPU arbitration."""
import struct

import numpy as np
import pytest

from loqa_hub_b200 import _lib as L
from loqa_hub_b200 import api, synth
from loqa_hub_b200.audio_service import (AudioService, RelayStatusCancelled, RelayStatusConnected,
                                         RelayStatusContending, RelayStatusWinner)
from oracle import oracle as ora

pytestmark = pytest.mark.gpu
MS = 1000          # microseconds per millisecond


@pytest.fixture()
def svc():
    import __graft_entry__ as entry
    entry.build()
    ctx = api.Context(device=0, max_streams=64, max_batch_msgs=4096, ring_bytes=8 << 20,
                      arena_samples=100_000, wake_samples_max=16_000, mode=L.DO_ALL)
    s = AudioService(ctx, max_streams=64, window_duration_us=100 * MS)
    yield s
    s.close()
    ctx.close()


def pcm(values):
    return np.asarray(values, dtype="<i2").tobytes()


def test_arbitration_window(svc):
    """TestArbitrationWindow :94-163."""
    svc.connect("relay1"); svc.connect("relay2")
    window = svc.start_arbitration_window("relay1", now_us=0)
    assert window is not None
    assert svc.window(window).is_active
    assert svc.window(window).n_relays == 1
    assert svc.join_arbitration_window("relay2", now_us=10 * MS)
    assert svc.window(window).n_relays == 2
    assert svc.poll(now_us=50 * MS) == 0                 # the 100 ms timer has not fired yet
    assert svc.poll(now_us=150 * MS) == 1                # "time.Sleep(150 ms)"
    w = svc.window(window)
    assert not w.is_active
    assert w.winner_id != ""
    cancelled = "relay2" if w.winner_id == "relay1" else "relay1"
    assert not svc.is_relay_active(cancelled)
    assert svc.relay(cancelled).cancel_closed            # close(relay.CancelChannel), audio_service.go:561
    assert svc.arbitration_window() is None              # as.arbitrationWindow = nil, :584-586


def test_join_arbitration_window_after_timeout(svc):
    """TestJoinArbitrationWindowAfterTimeout :165-183."""
    svc.connect("relay1"); svc.connect("relay2")
    svc.start_arbitration_window("relay1", now_us=0)
    assert svc.poll(now_us=100 * MS) == 1
    assert not svc.join_arbitration_window("relay2", now_us=100 * MS)
    # also without the timer having run: elapsed > WindowDuration is refused (:467-470)
    svc.start_arbitration_window("relay1", now_us=200 * MS)
    assert svc.join_arbitration_window("relay2", now_us=300 * MS)          # elapsed == duration still joins
    svc.connect("relay3")
    assert not svc.join_arbitration_window("relay3", now_us=300 * MS + 1)


def test_relay_status_checking_and_transitions(svc):
    """TestRelayStatusChecking :185-224 and TestRelayStatusTransitions :462-514."""
    assert not svc.is_relay_active("nonexistent")
    for rid, status, active in (("relay1", RelayStatusWinner, True), ("relay2", RelayStatusCancelled, False),
                                ("relay3", RelayStatusConnected, True), ("relay4", RelayStatusContending, True)):
        svc.connect(rid)
        assert not svc.is_relay_active(rid)              # connected but not yet in activeStreams
        svc.set_relay_status(rid, status)
        assert svc.is_relay_active(rid) == active


def test_signal_strength_arbitration(svc):
    """TestSignalStrengthArbitration :226-293: low / high / medium amplitude."""
    signals = {"relay1": [3277] * 4, "relay2": [32767] * 4, "relay3": [16384] * 4}
    for rid in signals:
        svc.connect(rid)
    for rid, v in signals.items():
        assert svc.on_chunk(rid, pcm(v), is_wake_word=True, is_end_of_speech=False, now_us=0)
    window = svc.arbitration_window()
    assert svc.window(window).n_relays == 3
    assert svc.flush() == 3
    svc.perform_arbitration(window)
    assert svc.window(window).winner_id == "relay2"
    r1, r2, r3 = (svc.relay(r) for r in ("relay1", "relay2", "relay3"))
    assert r1.signal_strength < r2.signal_strength and r3.signal_strength < r2.signal_strength
    assert (r2.status, r1.status, r3.status) == (RelayStatusWinner, RelayStatusCancelled, RelayStatusCancelled)
    for rid, v in signals.items():                       # SignalStrength is the reference's RMS, bit for bit
        want = ora.signal_strength(ora.bytes_to_float32(pcm(v))[0])[0]
        assert svc.relay(rid).signal_strength == want
    svc.perform_arbitration(window)                      # idempotent (:508-511)
    assert svc.window(window).winner_id == "relay2"


def test_cleanup_relay(svc):
    """TestCleanupRelay :295-318."""
    svc.connect("relay1")
    svc.set_relay_status("relay1", RelayStatusConnected)
    assert svc.active_streams() == 1
    svc.cleanup_relay("relay1")
    assert svc.active_streams() == 0
    svc.cleanup_relay("nonexistent")                     # must not fail


def test_request_scoped_lifecycle_and_consecutive_requests(svc):
    """TestRequestScopedArbitrationLifecycle :322-391, TestConsecutiveRequestsWithSingleRelay :395-459,
    TestArbitrationWindowClearance :517-582: a single relay wins every request, also with an EMPTY
    wake-word signal (first-member fallback, audio_service.go:542-547), and the window is cleared."""
    rid = "test-relay-001"
    svc.connect(rid)
    svc.set_relay_status(rid, RelayStatusConnected)
    window = svc.start_arbitration_window(rid, now_us=0)
    svc.perform_arbitration(window)                      # empty WakeWordSignal
    assert svc.window(window).winner_id == rid
    assert svc.relay(rid).status == RelayStatusWinner and svc.relay(rid).signal_strength == 0.0
    svc.complete_request(rid)
    assert svc.arbitration_window() is None
    assert svc.relay(rid).status == RelayStatusConnected
    assert svc.is_relay_active(rid)
    previous = window
    for i in range(3):
        now = (i + 1) * 1000 * MS
        assert svc.is_relay_active(rid)
        svc.connect(rid)                                 # a new gRPC stream = a new utterance
        assert svc.on_chunk(rid, pcm([16384, 16384, 16384]), True, False, now)
        window = svc.arbitration_window()
        assert window is not None and window != previous  # a different window instance
        svc.flush()
        svc.perform_arbitration(window)
        assert svc.window(window).winner_id == rid
        assert svc.relay(rid).status == RelayStatusWinner
        assert svc.relay(rid).signal_strength == ora.signal_strength(ora.bytes_to_float32(pcm([16384] * 3))[0])[0]
        svc.complete_request(rid)
        assert svc.arbitration_window() is None and svc.relay(rid).status == RelayStatusConnected
        previous = window


def test_stream_audio_flow_many_groups_and_wav_for_winners(svc):
    """StreamAudio end to end (:939-1041) for 4 collision groups x 4 relays in parallel windows:
    4 wake-word chunks (80 ms, inside the 100 ms window — as in the reference, where ~250 ms of
    wake word fit the 300 ms window) then speech, cancelled relays going quiet, late relays, and the
    winner's audio as the WAV the STT client posts (stt_client.go:365-398)."""
    groups, relays, n_msgs, wake, ms = 4, 4, 25, 4, 320
    ids = {(g, r): "g%d-relay%d" % (g, r) for g in range(groups) for r in range(relays)}
    for (g, r), rid in ids.items():
        svc.connect(rid, group=g)
    data = {(g, r): synth.relay_view(200 + g, r, n_msgs * ms) for g in range(groups) for r in range(relays)}
    for m in range(n_msgs):
        now = m * 20 * MS                                   # one 20 ms chunk per relay per tick
        if m == 5:                                          # the 100 ms windows opened at t=0 close now
            assert svc.poll(now) == groups
        for (g, r), rid in ids.items():
            if m >= 5 and not svc.is_relay_active(rid):
                continue                                    # cancelled relays stop streaming
            keep = svc.on_chunk(rid, data[g, r][m * ms:(m + 1) * ms].tobytes(), m < wake, m == n_msgs - 1, now)
            assert keep
        svc.flush()
    for g in range(groups):
        winner = ids[g, 0]                                  # relay 0 has the highest gain
        assert svc.relay(winner).status == RelayStatusWinner
        for r in range(1, relays):
            assert svc.relay(ids[g, r]).status == RelayStatusCancelled and not svc.is_relay_active(ids[g, r])
        # the score is the RMS of the wake-word buffer when the window closed: 4 chunks
        want = ora.signal_strength(ora.bytes_to_float32(data[g, 0][:wake * ms].tobytes())[0])[0]
        assert svc.relay(winner).signal_strength == want
        wav = svc.fetch_wav(winner)
        full, _ = ora.bytes_to_float32(data[g, 0].tobytes())               # WakeWordSignal ++ SpeechAudio (:592-594)
        assert wav == ora.float32_to_wav(full, 16000)
        assert struct.unpack("<I", wav[40:44])[0] == 4 * n_msgs * ms
    # a relay that shows up after its group's window was arbitrated starts a NEW window (the group's
    # pointer is nil again, :584-586); one that arrives while a window is open but expired is refused
    svc.connect("late", group=0)
    assert svc.on_chunk("late", pcm([100] * 320), True, False, 10_000 * MS)
    w = svc.arbitration_window(0)
    svc.connect("later", group=0)
    assert not svc.on_chunk("later", pcm([100] * 320), True, False, 10_000 * MS + 100 * MS + 1)   # :990-995
    svc.flush()
    assert svc.poll(10_200 * MS) == 1 and svc.window(w).winner_id == "late"


def test_concurrent_relay_threads_and_a_batching_thread(svc):
    """What grpc-go does to the hub: one goroutine per relay stream calling the per-chunk handler
    concurrently (StreamAudio :926) while ONE batching thread flushes ticks and runs the window
    timer.  8 threads x 4 relays, 4 collision groups of 8; ctypes releases the GIL during the C
    calls, so the hub's locking is really exercised."""
    import threading
    import time
    groups, per_group, n_msgs, wake, ms = 4, 8, 20, 4, 320
    ids = {(g, r): "t%d-%d" % (g, r) for g in range(groups) for r in range(per_group)}
    for (g, r), rid in ids.items():
        svc.connect(rid, group=g)
    data = {k: synth.relay_view(300 + k[0], k[1], n_msgs * ms) for k in ids}
    t0 = time.monotonic()
    now_us = lambda: int((time.monotonic() - t0) * 1e6)
    stop = threading.Event()
    errors = []

    def relay_thread(keys):
        try:
            for m in range(wake):                                  # wake word: all relays of a group within ~ms
                for k in keys:
                    assert svc.on_chunk(ids[k], data[k][m * ms:(m + 1) * ms].tobytes(), True, False, now_us())
            time.sleep(0.15)                                       # the 100 ms windows close meanwhile
            for m in range(wake, n_msgs):
                for k in keys:
                    if svc.is_relay_active(ids[k]):
                        svc.on_chunk(ids[k], data[k][m * ms:(m + 1) * ms].tobytes(), False, m == n_msgs - 1, now_us())
        except Exception as e:                                     # noqa: BLE001
            errors.append(e)

    def batcher():
        try:
            while not stop.is_set():
                svc.flush()
                svc.poll(now_us())
                time.sleep(0.002)
            svc.flush()
        except Exception as e:                                     # noqa: BLE001
            errors.append(e)

    keys = sorted(ids)
    threads = [threading.Thread(target=relay_thread, args=(keys[i::8],)) for i in range(8)]
    bt = threading.Thread(target=batcher)
    bt.start()
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    stop.set()
    bt.join()
    assert not errors, errors
    for g in range(groups):
        winner = ids[g, 0]
        assert svc.relay(winner).status == RelayStatusWinner
        assert sum(svc.relay(ids[g, r]).status == RelayStatusCancelled for r in range(per_group)) == per_group - 1
        want = ora.signal_strength(ora.bytes_to_float32(data[g, 0][:wake * ms].tobytes())[0])[0]
        assert svc.relay(winner).signal_strength == want
        full, _ = ora.bytes_to_float32(data[g, 0].tobytes())
        got = svc.ctx.fetch_samples(svc.relay(winner).slot)
        assert np.array_equal(got.view(np.uint32), full.view(np.uint32))


def test_two_utterances_of_one_relay_inside_one_tick(svc):
    """A relay whose next gRPC stream (= next utterance) starts before the batching tick has
    flushed the previous one: both sit in the queue, the flush submits them as two batches, and
    the stream ends up holding the second utterance only — as a fresh RelayStream would
    (audio_service.go:945-958)."""
    rid, other = "fast-relay", "steady-relay"
    svc.connect(rid); svc.connect(other, group=1)
    first = [pcm(np.full(320, 1000 + k)) for k in range(3)]
    second = [pcm(np.full(320, -7000 - k)) for k in range(2)]
    for k, c in enumerate(first):
        assert svc.on_chunk(rid, c, True, False, k * MS)
        assert svc.on_chunk(other, c, True, False, k * MS)
    svc.connect(rid)                                     # new stream from the same relay, same tick
    for k, c in enumerate(second):
        assert svc.on_chunk(rid, c, True, False, (10 + k) * MS)
    assert svc.flush() == 8
    want = np.concatenate([ora.bytes_to_float32(c)[0] for c in second])
    assert np.array_equal(svc.ctx.fetch_samples(svc.relay(rid).slot), want)
    want_other = np.concatenate([ora.bytes_to_float32(c)[0] for c in first])
    assert np.array_equal(svc.ctx.fetch_samples(svc.relay(other).slot), want_other)
    assert svc.flush() == 0


def test_relay_cleaned_up_while_its_window_is_open(svc):
    """cleanupRelay (:916-923) can run while the relay's window is still open (its gRPC stream
    ended).  Its audio is gone with its stream slot, so it no longer contends; the window closes
    for the others, and no ghost relay appears in the registry."""
    for r in ("a", "b", "c"):
        svc.connect(r)
    loud, quiet = pcm(np.full(320, 20000)), pcm(np.full(320, 3000))
    assert svc.on_chunk("a", loud, True, False, 0)
    assert svc.on_chunk("b", quiet, True, False, 1 * MS)
    assert svc.on_chunk("c", quiet, True, False, 2 * MS)
    svc.flush()
    window = svc.arbitration_window()
    svc.cleanup_relay("a")                               # the loudest relay hangs up
    assert svc.active_streams() == 2
    assert svc.poll(now_us=200 * MS) == 1
    assert svc.window(window).winner_id == "b"           # first of the two equals that are left
    assert svc.relay("b").status == RelayStatusWinner and svc.relay("c").status == RelayStatusCancelled
    assert svc.active_streams() == 2
    with pytest.raises(L.GpuAudioError):
        svc.relay("a")                                   # not resurrected by the arbitration


def test_cleanup_drops_the_relays_unflushed_chunks(svc):
    """A relay that disconnects before the batching tick: its queued chunks must not end up in the
    stream of the relay that inherits its slot."""
    svc.connect("old")
    slot = svc.relay("old").slot
    assert svc.on_chunk("old", pcm(np.full(320, 11111)), True, False, 0)
    svc.cleanup_relay("old")
    svc.connect("new", group=5)
    assert svc.relay("new").slot == slot                 # the freed slot is handed out again
    mine = pcm(np.full(320, -222))
    assert svc.on_chunk("new", mine, True, False, 1 * MS)
    assert svc.flush() == 1
    assert np.array_equal(svc.ctx.fetch_samples(slot), ora.bytes_to_float32(mine)[0])
    svc.ctx.push(b"\0" * 16)                             # the ring keeps working past the skipped reservation


def small_hub(**kw):
    import __graft_entry__ as entry
    entry.build()
    args = dict(device=0, max_streams=16, max_batch_msgs=8, max_inflight=2, ring_bytes=1 << 16,
                arena_samples=4_000, wake_samples_max=1_280, mode=L.DO_ALL)
    args.update(kw)
    ctx = api.Context(**args)
    return ctx, AudioService(ctx, max_streams=args["max_streams"], window_duration_us=100 * MS)


def test_flush_cuts_the_queue_at_the_contexts_batch_limit():
    """More chunks queued than one ga_submit takes (max_batch_msgs = 8): the flush goes through them in
    several batches instead of failing and re-queueing for ever."""
    ctx, svc = small_hub()
    try:
        rng = np.random.default_rng(1)
        data = {r: rng.integers(-20000, 20000, size=5 * 320).astype("<i2") for r in range(6)}
        for r in range(6):
            svc.connect("r%d" % r, group=r)
        for m in range(5):
            for r in range(6):
                assert svc.on_chunk("r%d" % r, data[r][m * 320:(m + 1) * 320].tobytes(), m < 2, m == 4, m * MS)
        assert svc.flush() == 30
        assert svc.flush() == 0
        for r in range(6):
            slot = svc.relay("r%d" % r).slot
            assert ctx.stream_info(slot) == (1600, 640)
            want, _ = ora.bytes_to_float32(data[r].tobytes())
            assert np.array_equal(ctx.fetch_samples(slot).view(np.uint32), want.view(np.uint32))
    finally:
        svc.close(); ctx.close()


def test_a_relay_that_overflows_its_arena_does_not_wedge_the_hub():
    """The reference appends speech without bound (audio_service.go:1002); a stream slot here has a
    fixed arena.  The relay that runs over loses the tick's chunks and leaves activeStreams; every
    other relay's chunks are ingested and later ticks work."""
    ctx, svc = small_hub(max_batch_msgs=64)
    try:
        svc.connect("talker", group=0)
        svc.connect("other", group=1)
        chunk = pcm(np.full(320, 1234))
        mine = [pcm(np.full(320, 100 + m)) for m in range(8)]
        assert svc.on_chunk("talker", chunk, True, False, 0)
        assert svc.on_chunk("other", mine[0], True, False, 0)
        assert svc.flush() == 2
        for m in range(9):                               # speech arena = 4,000 - 1,280 = 2,720 samples = 8.5 chunks
            assert svc.on_chunk("talker", chunk, False, False, (m + 1) * MS)
            if m < 5:
                assert svc.on_chunk("other", mine[m + 1], False, False, (m + 1) * MS)
        with pytest.raises(L.GpuAudioError) as e:
            svc.flush()
        assert e.value.code == L.GA_ERR_OVERFLOW and "dropped 9" in str(e.value)
        slot_o, slot_t = svc.relay("other").slot, svc.relay("talker").slot
        assert ctx.stream_info(slot_o) == (6 * 320, 320)            # everybody else's chunks went in
        assert ctx.stream_info(slot_t) == (320, 320)
        assert not svc.is_relay_active("talker")
        assert svc.flush() == 0                                      # nothing is stuck in the queue
        for m in range(6):
            svc.on_chunk("talker", chunk, False, False, (20 + m) * MS)   # not in activeStreams any more: ignored
            if m < 2:
                assert svc.on_chunk("other", mine[6 + m], False, False, (20 + m) * MS)
            assert svc.flush() == (1 if m < 2 else 0)
        want = np.concatenate([ora.bytes_to_float32(c)[0] for c in mine])
        assert np.array_equal(ctx.fetch_samples(slot_o).view(np.uint32), want.view(np.uint32))
    finally:
        svc.close(); ctx.close()


def test_ring_space_of_dropped_chunks_comes_back():
    """Chunks that never reach a ticket (their relay was cleaned up before the tick) must still give
    their ring space back, or a hub whose relays come and go fills its pinned ring."""
    ctx, svc = small_hub(ring_bytes=1 << 14, max_batch_msgs=64, arena_samples=16_000, wake_samples_max=8_000)
    try:
        chunk = pcm(np.arange(320))
        for round_ in range(6):                          # 6 x 20 x 640 B through a 16 KiB ring
            svc.connect("ghost", group=round_)
            for m in range(20):
                assert svc.on_chunk("ghost", chunk, True, False, m * MS)
            svc.cleanup_relay("ghost")
            assert svc.flush() == 0                      # all 20 belonged to a relay that is gone
        svc.connect("real", group=99)
        for m in range(20):
            assert svc.on_chunk("real", chunk, True, False, m * MS)
        assert svc.flush() == 20
        assert ctx.stream_info(svc.relay("real").slot) == (20 * 320, 20 * 320)
    finally:
        svc.close(); ctx.close()


def test_chunks_by_handle_and_stale_handles(svc):
    svc.connect("a", group=3)
    svc.connect("b", group=3)
    ha, hb = svc.relay_handle("a"), svc.relay_handle("b")
    loud, quiet = pcm(np.full(640, 9000)), pcm(np.full(640, 700))
    assert svc.on_chunk_h(hb, quiet, True, False, 0)
    assert svc.on_chunk_h(ha, loud, True, False, 1 * MS)
    svc.on_chunk_h(hb, quiet, False, True, 2 * MS)
    assert svc.flush() == 3
    assert svc.poll(200 * MS) == 1
    assert svc.relay("a").status == RelayStatusWinner and svc.relay("b").status == RelayStatusCancelled
    assert svc.relay("a").signal_strength == ora.signal_strength(ora.bytes_to_float32(loud)[0])[0]
    svc.cleanup_relay("b")
    with pytest.raises(L.GpuAudioError) as e:
        svc.on_chunk_h(hb, quiet, True, False, 300 * MS)
    assert e.value.code == L.GA_ERR_STATE
    svc.connect("c", group=9)                            # inherits b's slot; b's old handle stays dead
    assert svc.relay("c").slot == 1 and svc.relay_handle("c") != hb
    with pytest.raises(L.GpuAudioError):
        svc.on_chunk_h(hb, quiet, True, False, 300 * MS)
    assert svc.on_chunk_h(svc.relay_handle("c"), quiet, True, False, 300 * MS)
    assert svc.flush() == 1


def test_many_relay_threads_through_handles_while_a_batcher_flushes():
    """The per-chunk path under load: 8 threads push 20 ms chunks of 256 relays through
    ga_hub_on_chunk_h while one thread flushes; every stream must end up holding exactly its own
    utterance (order per relay kept, nothing lost, nothing duplicated)."""
    import threading
    import __graft_entry__ as entry
    entry.build()
    S, n_msgs, wake, ms, n_thr = 256, 30, 3, 320, 8
    ctx = api.Context(device=0, max_streams=S, max_batch_msgs=2048, max_inflight=2, ring_bytes=16 << 20,
                      arena_samples=n_msgs * ms, wake_samples_max=wake * ms, mode=L.DO_ALL)
    svc = AudioService(ctx, max_streams=S, window_duration_us=10_000 * MS)
    try:
        data = ora.synth_fill(40, S, 8, n_msgs, ms)                  # [S][n_msgs][ms]
        handles = []
        for s in range(S):
            svc.connect("r%d" % s, group=s // 8)
            handles.append(svc.relay_handle("r%d" % s))
        stop, errors, flushed = threading.Event(), [], [0]

        def relay_thread(mine):
            try:
                for m in range(n_msgs):
                    for s in mine:
                        while True:
                            try:
                                svc.on_chunk_h(handles[s], data[s, m].view(np.uint8), m < wake, m == n_msgs - 1, m * MS)
                                break
                            except L.GpuAudioError as e:             # back-pressure: wait for the batcher
                                if e.code != L.GA_ERR_RING_FULL:
                                    raise
            except Exception as e:                                   # noqa: BLE001
                errors.append(e)

        def batcher():
            try:
                while not stop.is_set():
                    flushed[0] += svc.flush()
                flushed[0] += svc.flush()
            except Exception as e:                                   # noqa: BLE001
                errors.append(e)

        threads = [threading.Thread(target=relay_thread, args=(range(i, S, n_thr),)) for i in range(n_thr)]
        bt = threading.Thread(target=batcher)
        bt.start()
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        stop.set()
        bt.join()
        assert not errors, errors
        assert flushed[0] == S * n_msgs
        for s in range(0, S, 7):
            want, _ = ora.bytes_to_float32(data[s].tobytes())
            assert ctx.stream_info(svc.relay("r%d" % s).slot) == (n_msgs * ms, wake * ms)
            assert np.array_equal(ctx.fetch_samples(svc.relay("r%d" % s).slot).view(np.uint32), want.view(np.uint32)), s
    finally:
        svc.close(); ctx.close()

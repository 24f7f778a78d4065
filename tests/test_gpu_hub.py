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

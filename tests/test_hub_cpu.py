"""The hub's relay registry and window bookkeeping (csrc/hub.cu) without a GPU: the parts of
internal/grpc/collision_detection_test.go that need no signal data.  (Scoring and selection need
samples on the device; they are covered by tests/test_gpu_hub.py.)"""
import pytest

import __graft_entry__ as entry
from loqa_hub_b200 import _lib as L
from loqa_hub_b200.audio_service import (AudioService, RelayStatusCancelled, RelayStatusConnected,
                                         RelayStatusContending, RelayStatusWinner)

MS = 1000


@pytest.fixture()
def svc():
    entry.build()
    s = AudioService(None, max_streams=8, window_duration_us=50 * MS)
    yield s
    s.close()


def test_window_membership_and_join_rules(svc):
    """TestArbitrationWindow :94-133 (up to the timer) and TestJoinArbitrationWindowAfterTimeout :165-183."""
    for r in ("relay1", "relay2", "relay3"):
        svc.connect(r)
    assert not svc.join_arbitration_window("relay2", 0)            # no window yet (audio_service.go:455-457)
    window = svc.start_arbitration_window("relay1", now_us=0)
    assert svc.window(window).is_active and svc.window(window).n_relays == 1
    assert svc.arbitration_window() == window
    assert svc.relay("relay1").status == RelayStatusContending
    assert svc.join_arbitration_window("relay2", 10 * MS)
    assert svc.join_arbitration_window("relay2", 20 * MS)          # joining twice keeps one entry (map insert)
    assert svc.window(window).n_relays == 2
    assert svc.join_arbitration_window("relay3", 50 * MS)          # elapsed == duration still joins
    svc.connect("relay4")
    assert not svc.join_arbitration_window("relay4", 50 * MS + 1)  # elapsed > WindowDuration (:467-470)
    assert svc.window(window).n_relays == 3
    svc.connect("other", group=7)                                  # another collision group has its own window
    assert not svc.join_arbitration_window("other", 0)
    w7 = svc.start_arbitration_window("other", 5 * MS)
    assert svc.arbitration_window(7) == w7 != window and svc.arbitration_window(0) == window


def test_relay_status_checking_transitions_and_cleanup(svc):
    """TestRelayStatusChecking :185-224, TestRelayStatusTransitions :462-514, TestCleanupRelay :295-318."""
    assert not svc.is_relay_active("nonexistent")
    svc.connect("relay1")
    assert not svc.is_relay_active("relay1")                       # not yet in activeStreams
    for status, active in ((RelayStatusConnected, True), (RelayStatusContending, True),
                           (RelayStatusWinner, True), (RelayStatusCancelled, False)):
        svc.set_relay_status("relay1", status)
        assert svc.is_relay_active("relay1") == active
    assert svc.active_streams() == 1
    svc.cleanup_relay("relay1")
    assert svc.active_streams() == 0
    svc.cleanup_relay("nonexistent")
    slots = [svc.connect("r%d" % i) for i in range(8)]
    assert sorted(slots) == list(range(8))
    with pytest.raises(L.GpuAudioError) as e:                      # all stream slots taken
        svc.connect("one-too-many")
    assert e.value.code == L.GA_ERR_OVERFLOW
    svc.cleanup_relay("r3")
    assert svc.connect("one-too-many") == slots[3]                 # the freed slot is reused


def test_anything_that_needs_samples_refuses_without_a_gpu_context(svc):
    svc.connect("relay1")
    w = svc.start_arbitration_window("relay1", 0)
    for call in (lambda: svc.on_chunk("relay1", b"\x00\x10", True, False, 0), lambda: svc.flush(),
                 lambda: svc.perform_arbitration(w), lambda: svc.poll(10_000 * MS)):
        with pytest.raises(L.GpuAudioError) as e:
            call()
        assert e.value.code == L.GA_ERR_STATE
    assert svc.poll(10 * MS) == 0                                  # nothing due yet: no GPU needed


def test_slot_bookkeeping_and_shard_placement_without_a_gpu(svc):
    """Stream slots are handed out densely, freed by cleanupRelay and reused; a hub over one context
    (or none) has a single shard, so ga_hub_connect_on accepts shard 0 only."""
    slots = [svc.connect("r%d" % i, group=i) for i in range(8)]
    assert slots == list(range(8))
    with pytest.raises(L.GpuAudioError) as e:
        svc.connect("one-too-many")
    assert e.value.code == L.GA_ERR_OVERFLOW
    svc.cleanup_relay("r3")
    assert svc.connect("again") == 3                                 # the freed slot
    assert svc.connect("r5", group=9) == 5                           # a known relay keeps its slot
    assert svc.relay_shard("r5") == 0
    svc.cleanup_relay("r0")
    assert svc.connect_on("pinned", 4, 0) == 0
    with pytest.raises(L.GpuAudioError) as e:
        svc.connect_on("elsewhere", 4, 1)
    assert e.value.code == L.GA_ERR_ARG
    with pytest.raises(L.GpuAudioError) as e:
        svc.relay_shard("nobody")
    assert e.value.code == L.GA_ERR_STATE
    with pytest.raises(L.GpuAudioError) as e:
        svc.fetch_wav("r5")
    assert e.value.code == L.GA_ERR_STATE                            # needs samples -> needs a GPU context

//go:build gpuaudio

package grpc

// audio_service_gpu.go — the bodies that change in internal/grpc/audio_service.go when the hub is
// built with `-tags gpuaudio` (CGO_ENABLED=1; Dockerfile:27 flips).  Everything not shown here —
// window start/join/timer (:408-502), status and cancel handling (:550-586), STT/LLM/TTS — stays as
// it is.  NOT compiled in the authoring image (no Go toolchain there); the same call sequence runs
// in C++ in csrc/hub.cu and is what tests/test_gpu_hub.py and tools/hub_realtime.cpp drive.
//
// Replaced reference lines:
//   :233-244  createAudioService      -> newGPU: context, batcher, slot free-list
//   :964/:999 bytesToFloat32Array + append under streamsMutex (:1000-1004) -> enqueueChunk
//   :517-547  scoring loop of performArbitration -> arbitrateOnGPU (segments in join order)
//   :592-594  fullAudio = WakeWordSignal ++ SpeechAudio -> winnerAudio (WAV + fingerprint)
//   :916-923  cleanupRelay -> also releaseSlot

import (
	"sync"
	"time"

	"github.com/loqalabs/loqa-hub/internal/gpuaudio"
)

type gpuState struct {
	ctx     *gpuaudio.Ctx
	batcher *gpuaudio.Batcher

	mu    sync.Mutex        // slot table only; never held across a cgo call that waits for the GPU
	slots map[string]uint32 // relay id -> stream slot
	free  []uint32
	next  uint32
	max   uint32

	joinOrder map[*ArbitrationWindow][]string // relay ids in the order they joined (Go maps have none)
}

// newGPU is called beside createAudioService (:233-244).  65,536 relays need 8 of these, one per
// GPU, relays placed by collision group (gpuaudio.NewCluster for windows that span GPUs).
func newGPU(device int, maxStreams uint32) (*gpuState, error) {
	const chunk, utterance, wake = 640, 80_000, 4_800 // 20 ms / 5 s / 300 ms at 16 kHz mono
	ctx, err := gpuaudio.New(gpuaudio.Config{Device: device, MaxStreams: maxStreams, MaxBatchMsgs: 4 * maxStreams,
		RingBytes: uint64(maxStreams) * chunk * 16, ArenaSamples: utterance, WakeSamples: wake})
	if err != nil {
		return nil, err // no CPU fallback: the hub refuses to start (or is built without the tag)
	}
	g := &gpuState{ctx: ctx, slots: map[string]uint32{}, max: maxStreams,
		joinOrder: map[*ArbitrationWindow][]string{}}
	g.batcher = gpuaudio.NewBatcher(ctx, int(4*maxStreams), 20*time.Millisecond, nil)
	return g, nil
}

// slotFor gives a relay its stream slot when its StreamAudio call begins (:952-954).
func (g *gpuState) slotFor(relayID string) (uint32, bool) {
	g.mu.Lock()
	defer g.mu.Unlock()
	if s, ok := g.slots[relayID]; ok {
		return s, true
	}
	var s uint32
	switch {
	case len(g.free) > 0:
		s, g.free = g.free[len(g.free)-1], g.free[:len(g.free)-1]
	case g.next < g.max:
		s, g.next = g.next, g.next+1
	default:
		return 0, false
	}
	g.slots[relayID] = s
	return s, true
}

// enqueueChunk stands where the handler converted and appended (:964-965 for wake-word chunks,
// :999-1004 for speech): the bytes go to the pinned ring, the float32 samples are produced on the
// GPU into the slot's WakeWordSignal / SpeechAudio regions.  No streamsMutex, no hub-wide lock at
// all: `first` is a local of the relay's own StreamAudio goroutine (true for the first chunk it
// receives, :952-954), the batcher's queue mutex covers a ring reservation and a counter.
func (as *AudioService) enqueueChunk(slot uint32, data []byte, isWakeWord, isEndOfSpeech, first bool) error {
	var flags uint32
	if isWakeWord {
		flags |= gpuaudio.FlagWake
	}
	if isEndOfSpeech {
		flags |= gpuaudio.FlagEOS
	}
	if first {
		flags |= gpuaudio.FlagFirst
	}
	return as.gpu.batcher.Enqueue(slot, data, flags)
}

// noteJoin is called from startArbitrationWindow (:432) and joinArbitrationWindow (:481) right where
// the relay is put into window.Relays: it records the join order the GPU arbitration uses in place
// of Go's random map order (:521).
func (g *gpuState) noteJoin(w *ArbitrationWindow, relayID string) {
	g.mu.Lock()
	defer g.mu.Unlock()
	for _, id := range g.joinOrder[w] {
		if id == relayID {
			return
		}
	}
	g.joinOrder[w] = append(g.joinOrder[w], relayID)
}

// arbitrateOnGPU replaces :517-547 of performArbitration for one or many windows: it fills
// relay.SignalStrength (:528) for contending relays and returns each window's winner id ("" for an
// empty window).  The caller holds window.mutex as before (:506); the status / cancel / winner
// code (:550-586) runs unchanged on the result.
func (as *AudioService) arbitrateOnGPU(windows []*ArbitrationWindow) ([]string, error) {
	g := as.gpu
	seg := []uint32{0}
	var ids []string
	var slots []uint32
	var status []uint8
	g.mu.Lock()
	for _, w := range windows {
		for _, id := range g.joinOrder[w] {
			relay, ok := w.Relays[id]
			if !ok {
				continue
			}
			ids = append(ids, id)
			slots = append(slots, g.slots[id])
			status = append(status, uint8(relay.Status)) // only RelayStatusContending is scored (:522-524)
		}
		seg = append(seg, uint32(len(slots)))
		delete(g.joinOrder, w)
	}
	g.mu.Unlock()
	winners := make([]string, len(windows))
	if len(slots) == 0 {
		return winners, nil
	}
	win, scores, _, err := g.ctx.Arbitrate(seg, slots, status)
	if err != nil {
		return nil, err
	}
	for k, w := range windows {
		for m := seg[k]; m < seg[k+1]; m++ {
			if status[m] == uint8(RelayStatusContending) {
				w.Relays[ids[m]].SignalStrength = scores[m] // :528
			}
		}
		if win[k] != 0xFFFFFFFF {
			winners[k] = ids[win[k]] // strict > from 0.0, first member if none exceeded it (:519, :535, :542-547)
		}
	}
	return winners, nil
}

// winnerAudio replaces :592-594 (and float32ToWAV, internal/llm/stt_client.go:365-398, plus the
// fingerprint of events/voice_event.go:84-88) for all windows that closed in this timer tick: one
// GPU gather + hash + copy instead of one pair of appends and one WAV build per winner.
func (as *AudioService) winnerAudio(relayIDs []string) (wavs [][]byte, audioHashes []string, err error) {
	g := as.gpu
	slots := make([]uint32, len(relayIDs))
	g.mu.Lock()
	for i, id := range relayIDs {
		slots[i] = g.slots[id]
	}
	g.mu.Unlock()
	return g.ctx.WinnerAudio(slots, 16000) // the literal 16000 of :609
}

// releaseSlot is the GPU half of cleanupRelay (:916-923).
func (g *gpuState) releaseSlot(relayID string) {
	g.mu.Lock()
	s, ok := g.slots[relayID]
	if ok {
		delete(g.slots, relayID)
		g.free = append(g.free, s)
	}
	g.mu.Unlock()
	if ok {
		_ = g.ctx.StreamReset(s)
	}
}

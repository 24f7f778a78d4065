package gpuaudio

// batcher.go — the batching goroutine between the per-relay StreamAudio goroutines
// (internal/grpc/audio_service.go:926) and the GPU.  NOT compiled in the authoring image (no Go
// toolchain there); csrc/hub.cu (ShardQueue, ga_hub_flush) is the same logic in C++ and is what the
// tests exercise (tests/test_gpu_hub.py::test_many_relay_threads_through_handles_while_a_batcher_flushes,
// tools/hub_realtime.cpp).
//
// Per chunk (any goroutine):  Enqueue  = reserve ring space + take a sequence number under a small
//                             mutex, copy the bytes into the pinned ring OUTSIDE it, publish.
// Per tick (one goroutine):   take the published prefix of the queue (complete both in arrival order
//                             and in ring offsets — the ring is released by offset when a ticket is
//                             waited, so no reservation may be left behind), cut it where a relay
//                             restarts an utterance, Submit, keep at most MaxInflight-1 tickets,
//                             Wait the oldest and fan its results out.

import (
	"sync"
	"sync/atomic"
	"time"
)

const (
	FlagWake  = 0x01 // chunk.IsWakeWord        (audio_service.go:962)
	FlagEOS   = 0x02 // chunk.IsEndOfSpeech     (audio_service.go:1008)
	FlagFirst = 0x04 // first chunk of the relay's stream: a new RelayStream (:945-958)
)

// ChunkResult is what the hub gets back per chunk: the per-message record (digest = the
// voice-event fingerprint of the received bytes, API.md:170) and the relay it belongs to.
type ChunkResult struct {
	Slot   uint32
	Flags  uint32
	Result Result
}

type entry struct {
	msg   Msg
	ready atomic.Uint64 // sequence number + 1 once msg is published
}

type Batcher struct {
	ctx  *Ctx
	tick time.Duration

	claim sync.Mutex // held for {ring reservation, sequence number} only
	tail  uint64     // next sequence number (under claim)
	head  atomic.Uint64
	q     []entry // power-of-two ring

	carry    []Msg // taken off the queue, not yet submitted (restart cut, pipeline full)
	inflight []inflightBatch
	maxBatch int
	OnResult func([]ChunkResult) // called from the batching goroutine
	stop     chan struct{}
	done     chan struct{}
}

type inflightBatch struct {
	ticket uint64
	msgs   []Msg
}

func NewBatcher(ctx *Ctx, maxBatchMsgs int, tick time.Duration, onResult func([]ChunkResult)) *Batcher {
	n := 1024
	for n < 4*maxBatchMsgs {
		n <<= 1
	}
	b := &Batcher{ctx: ctx, tick: tick, q: make([]entry, n), maxBatch: maxBatchMsgs, OnResult: onResult,
		stop: make(chan struct{}), done: make(chan struct{})}
	go b.run()
	return b
}

// Enqueue replaces `audioData := bytesToFloat32Array(chunk.AudioData)` + the append under
// streamsMutex (audio_service.go:964-965, :999-1004).  No hub-wide lock.
func (b *Batcher) Enqueue(slot uint32, audioData []byte, flags uint32) error {
	b.claim.Lock()
	if b.tail-b.head.Load() >= uint64(len(b.q)) {
		b.claim.Unlock()
		return ErrQueueFull
	}
	dst, off, err := b.ctx.Reserve(len(audioData))
	if err != nil {
		b.claim.Unlock()
		return err
	}
	seq := b.tail
	b.tail++
	b.claim.Unlock()

	copy(dst, audioData) // C-owned pinned memory; no Go pointer is kept by C
	e := &b.q[seq&uint64(len(b.q)-1)]
	e.msg = Msg{ring_offset: cu64(off), nbytes: cu32(uint32(len(audioData))), stream_id: cu32(slot), flags: cu32(flags)}
	e.ready.Store(seq + 1)
	return nil
}

func (b *Batcher) Close() {
	close(b.stop)
	<-b.done
}

func (b *Batcher) run() {
	defer close(b.done)
	t := time.NewTicker(b.tick)
	defer t.Stop()
	for {
		select {
		case <-b.stop:
			b.flush()
			b.drain(0)
			return
		case <-t.C:
			b.flush()
		}
	}
}

// take moves the published prefix of the queue behind whatever an earlier tick could not submit.
func (b *Batcher) take() []Msg {
	batch := b.carry
	b.carry = nil
	head := b.head.Load()
	for {
		e := &b.q[head&uint64(len(b.q)-1)]
		if e.ready.Load() != head+1 {
			break
		}
		batch = append(batch, e.msg)
		head++
	}
	b.head.Store(head)
	return batch
}

// cut returns the end of the sub-batch that starts at lo: at most maxBatch messages, and ending
// before a relay's restart (FlagFirst) that follows chunks of the same relay — ga_submit wants the
// new utterance in a later batch than the old one's chunks (they reuse the same sample arena).
func (b *Batcher) cut(batch []Msg, lo int) int {
	hi := lo + b.maxBatch
	if hi > len(batch) {
		hi = len(batch)
	}
	seen := make(map[uint32]struct{}, hi-lo)
	for k := lo; k < hi; k++ {
		s := uint32(batch[k].stream_id)
		if _, dup := seen[s]; dup && uint32(batch[k].flags)&FlagFirst != 0 {
			return k
		}
		seen[s] = struct{}{}
	}
	return hi
}

func (b *Batcher) flush() {
	batch := b.take()
	for lo := 0; lo < len(batch); {
		if len(b.inflight) >= MaxInflight-1 {
			b.drain(MaxInflight - 2) // the oldest ticket's records are read before its slot comes round again
		}
		hi := b.cut(batch, lo)
		ticket, err := b.ctx.Submit(batch[lo:hi])
		if err != nil {
			// arena overflow of one relay / CUDA failure: keep the rest for the next tick, in order
			// (the C++ hub additionally drops only the offending relay: ga_hub_flush)
			b.carry = append(b.carry, batch[lo:]...)
			break
		}
		b.inflight = append(b.inflight, inflightBatch{ticket, batch[lo:hi]})
		lo = hi
	}
	b.drain(MaxInflight - 2)
}

// drain waits tickets, oldest first, until at most keep are outstanding.
func (b *Batcher) drain(keep int) {
	for len(b.inflight) > keep {
		fb := b.inflight[0]
		b.inflight = b.inflight[1:]
		res, err := b.ctx.Wait(fb.ticket)
		if err != nil || b.OnResult == nil {
			continue
		}
		out := make([]ChunkResult, len(res))
		for i := range res {
			out[i] = ChunkResult{Slot: uint32(fb.msgs[i].stream_id), Flags: uint32(fb.msgs[i].flags), Result: res[i]}
		}
		b.OnResult(out) // copies: the pinned view is only valid until the slot is reused
	}
}

// Package gpuaudio binds libgpuaudio.so (include/gpuaudio.h) for loqa-hub.
// Written against the header; NOT compiled in the authoring image (no Go toolchain there).
// See INTEGRATION.md for the call-site changes in internal/grpc/audio_service.go.
package gpuaudio

/*
#cgo CFLAGS:  -I${SRCDIR}/../../third_party/gpuaudio/include
#cgo LDFLAGS: -L${SRCDIR}/../../third_party/gpuaudio/lib -lgpuaudio
#include <stdlib.h>
#include "gpuaudio.h"
*/
import "C"

import (
	"errors"
	"unsafe"
)

type Ctx struct{ c *C.ga_ctx }

type Config struct {
	Device                     int
	MaxStreams, MaxBatchMsgs   uint32
	RingBytes                  uint64
	ArenaSamples, WakeSamples  uint64
}

func lastErr(rc C.int) error { return errors.New(C.GoString(C.ga_last_error())) }

// MaxInflight is ga_config.max_inflight as New sets it: result slots are used round-robin, so the
// records of ticket t are overwritten by the Submit that returns ticket t+MaxInflight.
const MaxInflight = 4

var ErrQueueFull = errors.New("gpuaudio: chunk queue full until the next batch tick")

func cu64(v uint64) C.uint64_t { return C.uint64_t(v) }
func cu32(v uint32) C.uint32_t { return C.uint32_t(v) }

// New is called beside createAudioService (audio_service.go:233-244).
func New(cfg Config) (*Ctx, error) {
	cc := C.ga_config{device: C.int32_t(cfg.Device), max_streams: C.uint32_t(cfg.MaxStreams),
		max_batch_msgs: C.uint32_t(cfg.MaxBatchMsgs), max_inflight: 4,
		ring_bytes: C.uint64_t(cfg.RingBytes), arena_samples: C.uint64_t(cfg.ArenaSamples),
		wake_samples_max: C.uint64_t(cfg.WakeSamples), mode: C.GA_DO_ALL, vad_threshold: 0.01}
	var c *C.ga_ctx
	if rc := C.ga_init(&cc, &c); rc != 0 {
		return nil, lastErr(rc)
	}
	return &Ctx{c}, nil
}
func (x *Ctx) Close() { C.ga_shutdown(x.c) }

// Push copies one chunk.AudioData into the C-owned pinned ring and returns its offset.
// Called from the per-stream goroutines (StreamAudio, audio_service.go:964 and :999) instead of
// bytesToFloat32Array; lock-free on the C side.  C never keeps a Go pointer (cgo rule): the
// destination is C memory, wrapped with unsafe.Slice only for the copy.
func (x *Ctx) Push(audioData []byte) (uint64, error) {
	dst, off, err := x.Reserve(len(audioData))
	if err != nil {
		return 0, err
	}
	copy(dst, audioData)
	return off, nil
}

// Reserve is the first half of Push: n bytes of the pinned ring as a Go slice over C memory, and
// their ring offset.  The batcher reserves under its queue mutex and copies outside it.
func (x *Ctx) Reserve(n int) ([]byte, uint64, error) {
	var p unsafe.Pointer
	var off C.uint64_t
	if rc := C.ga_ring_reserve(x.c, C.uint64_t(n), &p, &off); rc != 0 {
		return nil, 0, lastErr(rc)
	}
	return unsafe.Slice((*byte)(p), n), uint64(off), nil
}

// StreamReset frees a stream slot's buffers for its next owner (cleanupRelay, audio_service.go:916-923).
func (x *Ctx) StreamReset(slot uint32) error {
	if rc := C.ga_stream_reset(x.c, C.uint32_t(slot)); rc != 0 {
		return lastErr(rc)
	}
	return nil
}

// WinnerAudio is processWinningRelayAudio's input for MANY winners at once (audio_service.go:590-609):
// per slot the WAV file float32ToWAV would build (internal/llm/stt_client.go:365-398) and the
// "sha256:<hex>" fingerprint of its samples (events/voice_event.go:84-88).  One launch + one copy per
// call; the WAV bytes are copied out of the C-owned staging buffer (valid only until the next call).
func (x *Ctx) WinnerAudio(slots []uint32, sampleRate int) (wavs [][]byte, hashes []string, err error) {
	if len(slots) == 0 {
		return nil, nil, nil
	}
	var base *C.uint8_t
	offs := make([]C.uint64_t, len(slots)+1)
	dig := make([]byte, 32*len(slots))
	if rc := C.ga_fetch_wavs(x.c, (*C.uint32_t)(&slots[0]), C.uint32_t(len(slots)), C.uint32_t(sampleRate),
		&base, &offs[0], (*C.uint8_t)(&dig[0])); rc != 0 {
		return nil, nil, lastErr(rc)
	}
	wavs, hashes = make([][]byte, len(slots)), make([]string, len(slots))
	for i := range slots {
		p := unsafe.Add(unsafe.Pointer(base), uintptr(offs[i]))
		hdr := unsafe.Slice((*byte)(p), 8)
		size := 8 + int(uint32(hdr[4])|uint32(hdr[5])<<8|uint32(hdr[6])<<16|uint32(hdr[7])<<24) // RIFF size + 8
		wavs[i] = append([]byte(nil), unsafe.Slice((*byte)(p), size)...)
		var h [C.GA_AUDIO_HASH_LEN]C.char
		C.ga_format_audio_hash((*C.uint8_t)(&dig[32*i]), &h[0])
		hashes[i] = C.GoString(&h[0])
	}
	return
}

type Msg = C.ga_msg       // {ring_offset, nbytes, stream_id, flags}
type Result = C.ga_result // {digest[32], sumsq, rms, n_samples, nonzero, peak, flags}

// Submit is called once per batch tick by the single batching goroutine (never per chunk:
// a cgo call is an OS-thread hop).  Messages in ring-offset order.
func (x *Ctx) Submit(msgs []Msg) (uint64, error) {
	var t C.uint64_t
	if rc := C.ga_submit(x.c, &msgs[0], C.uint32_t(len(msgs)), &t); rc != 0 {
		return 0, lastErr(rc)
	}
	return uint64(t), nil
}

// SubmitTicks is Submit for the steady state: ticks chunk periods of the same relays, every
// relay one chunk of chunkBytes per period, laid out period-major in ONE reservation at a stride
// of chunkBytes rounded up to 16.  One flag byte per chunk crosses PCIe instead of a 32-byte
// descriptor; the descriptors are built on the GPU.  flags may be nil.
func (x *Ctx) SubmitTicks(off uint64, ticks int, slots []uint32, chunkBytes int, flags []uint8) (uint64, error) {
	var t C.uint64_t
	var fp *C.uint8_t
	if flags != nil {
		fp = (*C.uint8_t)(&flags[0])
	}
	if rc := C.ga_submit_ticks(x.c, C.uint64_t(off), C.uint32_t(ticks), C.uint32_t(len(slots)),
		C.uint32_t(chunkBytes), (*C.uint32_t)(&slots[0]), fp, &t); rc != 0 {
		return 0, lastErr(rc)
	}
	return uint64(t), nil
}

// Wait returns a view of the C-owned pinned result records.  Ticket t's records are overwritten by the
// Submit that returns ticket t+4 (max_inflight): fan the results out before submitting again, or keep
// at most 3 tickets in flight.
func (x *Ctx) Wait(ticket uint64) ([]Result, error) {
	var r *C.ga_result
	var n C.uint32_t
	if rc := C.ga_wait_view(x.c, C.uint64_t(ticket), &r, nil, &n, nil); rc != 0 {
		return nil, lastErr(rc)
	}
	return unsafe.Slice((*Result)(r), int(n)), nil
}

// Arbitrate replaces the scoring loop of performArbitration (audio_service.go:517-547) for any
// number of windows.  members are stream slots in JOIN ORDER per window.
func (x *Ctx) Arbitrate(segOffsets, streamIDs []uint32, statusIn []uint8) (winner []uint32, scores []float64, statusOut []uint8, err error) {
	nw := len(segOffsets) - 1
	winner, scores, statusOut = make([]uint32, nw), make([]float64, len(streamIDs)), make([]uint8, len(streamIDs))
	rc := C.ga_arbitrate(x.c, (*C.uint32_t)(&segOffsets[0]), C.uint32_t(nw), (*C.uint32_t)(&streamIDs[0]),
		(*C.uint8_t)(&statusIn[0]), (*C.uint32_t)(&winner[0]), (*C.double)(&scores[0]), (*C.uint8_t)(&statusOut[0]))
	if rc != 0 {
		err = lastErr(rc)
	}
	return
}

// Samples fetches WakeWordSignal ++ SpeechAudio of one stream (audio_service.go:592-594) — only
// for the arbitration winner, right before TranscribeWithConfidence (:609).
func (x *Ctx) Samples(stream uint32) ([]float32, error) {
	var n, w C.uint64_t
	C.ga_stream_info(x.c, C.uint32_t(stream), &n, &w)
	out := make([]float32, int(n))
	if n > 0 {
		if rc := C.ga_fetch_samples(x.c, C.uint32_t(stream), 0, n, (*C.float)(&out[0])); rc != 0 {
			return nil, lastErr(rc)
		}
	}
	return out, nil
}

// Cluster arbitrates windows whose relays sit on different GPUs of this process (one Ctx per GPU).
type Cluster struct{ c *C.ga_cluster }

func NewCluster(ctxs []*Ctx, maxGuests int) (*Cluster, error) {
	hs := make([]*C.ga_ctx, len(ctxs))
	for i, x := range ctxs {
		hs[i] = x.c
	}
	var c *C.ga_cluster
	if rc := C.ga_cluster_create(&hs[0], C.uint32_t(len(hs)), C.uint32_t(maxGuests), &c); rc != 0 {
		return nil, lastErr(rc)
	}
	return &Cluster{c}, nil
}

func (k *Cluster) Close() { C.ga_cluster_destroy(k.c) }

// Arbitrate is performArbitration for windows across GPUs: member m is stream streamIDs[m] of
// context shardIDs[m], in join order per window.
func (k *Cluster) Arbitrate(segOffsets, shardIDs, streamIDs []uint32, statusIn []uint8) (winner []uint32, scores []float64, statusOut []uint8, err error) {
	nw := len(segOffsets) - 1
	winner, scores, statusOut = make([]uint32, nw), make([]float64, len(streamIDs)), make([]uint8, len(streamIDs))
	rc := C.ga_cluster_arbitrate(k.c, (*C.uint32_t)(&segOffsets[0]), C.uint32_t(nw), (*C.uint32_t)(&shardIDs[0]),
		(*C.uint32_t)(&streamIDs[0]), (*C.uint8_t)(&statusIn[0]), (*C.uint32_t)(&winner[0]), (*C.double)(&scores[0]),
		(*C.uint8_t)(&statusOut[0]))
	if rc != 0 {
		err = lastErr(rc)
	}
	return
}

/*
 * gpuaudio.h — C ABI of libgpuaudio.so: the B200 (sm_100a) relay-audio ingest path
 * that sits behind loqa-hub's Go audio handler (cgo package internal/gpuaudio).
 *
 * Plain C: pointers, sizes and PODs only — no CUDA, torch or C++ types.  Every
 * entry point names the reference code it replaces (paths relative to the
 * loqa-hub tree; see SURVEY.md §8a/§8b and INTEGRATION.md for the cgo binding).
 *
 * Conventions
 *   - int return: 0 = GA_OK, negative = ga_status error class.  Never aborts,
 *     never calls back into the caller.  ga_last_error() returns a thread-local,
 *     NUL-terminated description of the last failure on the calling thread.
 *   - There is NO CPU fallback: without a usable CUDA device every compute entry
 *     point fails with GA_ERR_CUDA.
 *   - "device pointer" arguments (ga_k_* layer) are CUDA device addresses owned by
 *     the caller; `stream` is a cudaStream_t passed as void* (NULL = default stream).
 *   - All structs are little-endian PODs with the exact layouts below (checked by
 *     ga_abi_sizes()).
 */
#ifndef GPUAUDIO_H
#define GPUAUDIO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GA_ABI_VERSION 1

typedef enum ga_status {
    GA_OK            =  0,
    GA_ERR_ARG       = -1,  /* bad argument (NULL, misaligned, out of range)     */
    GA_ERR_CUDA      = -2,  /* CUDA runtime / driver error, or no device          */
    GA_ERR_NOMEM     = -3,  /* host or device allocation failed                   */
    GA_ERR_RING_FULL = -4,  /* pinned ring has no room until tickets are waited   */
    GA_ERR_OVERFLOW  = -5,  /* a stream's sample arena / a batch limit overflowed */
    GA_ERR_STATE     = -6,  /* call not valid in the current state (bad ticket …) */
    GA_ERR_COMM      = -7   /* NCCL / multi-GPU exchange failure                  */
} ga_status;

/* RelayStatus — internal/grpc/audio_service.go:56-63 (same numeric values). */
enum {
    GA_RELAY_CONNECTED  = 0,
    GA_RELAY_CONTENDING = 1,
    GA_RELAY_WINNER     = 2,
    GA_RELAY_CANCELLED  = 3
};

/* ga_msg.flags / ga_desc.flags — AudioChunk.IsWakeWord / IsEndOfSpeech
 * (call sites audio_service.go:962, :1008) plus stream bookkeeping. */
#define GA_MSG_WAKE   0x01u   /* chunk.IsWakeWord                                    */
#define GA_MSG_EOS    0x02u   /* chunk.IsEndOfSpeech                                 */
#define GA_MSG_FIRST  0x04u   /* first chunk of an utterance: resets the stream state */

/* ga_k_ingest / ga_config mode bits: which parts of the path run. */
#define GA_DO_CONVERT   0x01u /* int16 -> f32 samples materialised (bytesToFloat32Array) */
#define GA_DO_FEATURES  0x02u /* per-message (+ optional per-frame) energy statistics    */
#define GA_DO_SHA256    0x04u /* per-message SHA-256 fingerprint of the received bytes   */
#define GA_DO_ALL       0x07u
#define GA_DO_UTTERANCE 0x08u /* context layer: also keep a running SHA-256 per stream over all
                                 received bytes of the utterance (ga_utterance_digest)       */

#define GA_FRAME_SAMPLES 320u /* 20 ms at 16 kHz                  */

/* Per-message result record, 64 bytes.  The feature half is what the reference
 * logs per conversion / per RMS calculation (audio_service.go:1093-1098, :843-849). */
typedef struct ga_result {
    uint8_t  digest[32];  /* SHA-256 of the message bytes exactly as received      */
    double   sumsq;       /* "sum":  Σ float64(float32(s*s)) in index order (:824-828) */
    double   rms;         /* "rms":  sqrt(sumsq / n_samples), 0 when n_samples == 0 (:841, :819-822) */
    uint32_t n_samples;   /* "output_samples" = nbytes / 2 (:1066)                  */
    uint32_t nonzero;     /* "non_zero_samples" (:1081-1083, :829-831)              */
    float    peak;        /* "max_abs_value" / "max_sample" (:1084-1090, :832-838)  */
    uint32_t flags;       /* bit 0: VAD (rms > vad_threshold); bits 8..: echo of ga_desc.flags */
} ga_result;

/* Per-frame record (20 ms framing, new spec — no reference counterpart), 32 bytes. */
typedef struct ga_frame {
    double   sumsq;
    double   rms;
    uint32_t n_samples;
    uint32_t nonzero;
    float    peak;
    uint32_t vad;
} ga_frame;

/* Device-side message descriptor, 32 bytes (ga_k_* layer). */
typedef struct ga_desc {
    uint64_t in_off;     /* byte offset of the message in the PCM slab; multiple of 16     */
    uint64_t out_off;    /* index of the message's first output sample in the f32 arena    */
    uint32_t nbytes;     /* message length in bytes, any value (odd trailing byte dropped
                            for conversion, kept for the digest)                           */
    uint32_t stream_id;  /* dense stream slot                                              */
    uint32_t flags;      /* GA_MSG_*                                                       */
    uint32_t frame_off;  /* index of the message's first ga_frame (when frames are wanted)  */
} ga_desc;

/* ===================================================================== kernels
 * Device-pointer layer: one call = one launch on `stream`, asynchronous.       */

/* Fused relay-message ingest.  For every message i < n_msgs:
 *   GA_DO_CONVERT : samples[desc[i].out_off + k] = float32(int16le(pcm, k)) / 32767.0f
 *                   — replaces bytesToFloat32Array, audio_service.go:1048-1101
 *   GA_DO_FEATURES: results[i].{sumsq,rms,n_samples,nonzero,peak,flags}; with
 *                   frames != NULL also one ga_frame per GA_FRAME_SAMPLES samples
 *                   — the statistics of calculateSignalStrength, audio_service.go:818-852,
 *                   applied per message / per 20 ms frame, in the reference's order
 *   GA_DO_SHA256  : results[i].digest = SHA-256(message bytes)
 *                   — the voice-event audio fingerprint (README.md:50, API.md:170; hook
 *                   events/voice_event.go:84-88)
 * d_pcm must be readable up to the next multiple of 16 past every message end.
 * d_samples may be NULL without GA_DO_CONVERT; d_frames may be NULL. */
int ga_k_ingest(const void *d_pcm, const ga_desc *d_desc, uint32_t n_msgs,
                float *d_samples, ga_result *d_results, ga_frame *d_frames,
                uint32_t mode, double vad_threshold, void *stream);

/* ga_k_ingest for a UNIFORM batch: message i is the msg_bytes bytes at
 * d_pcm + first_in_off + i * msg_stride (what a batch tick of equal-sized chunks copied into the
 * ring in order looks like); desc[i].in_off / nbytes are ignored, everything else is as in
 * ga_k_ingest and the results are identical.  The slab is described by a 2-D TMA tensor map, so one
 * hardware copy per 32-message slice replaces 32 per-message bulk copies.  first_in_off and
 * msg_stride must be multiples of 16. */
int ga_k_ingest_uniform(const void *d_pcm, const ga_desc *d_desc, uint32_t n_msgs,
                        uint64_t first_in_off, uint32_t msg_bytes, uint64_t msg_stride,
                        float *d_samples, ga_result *d_results, ga_frame *d_frames,
                        uint32_t mode, double vad_threshold, void *stream);

/* Segmented per-collision-window arbitration — replaces the scoring + selection of
 * performArbitration, audio_service.go:517-571, for n_windows windows at once.
 * Window w owns members [seg_offsets[w], seg_offsets[w+1]) in join order.  Member m's
 * wake-word buffer is d_samples[wake_off[m] .. +wake_n[m]) (RelayStream.WakeWordSignal).
 * scores[m]  = calculateSignalStrength(wake buffer) for contending members (exact
 *              reference order), untouched otherwise (:522-528);
 * winner[w]  = member index (global, into the member arrays) chosen by the strict-`>`
 *              running max from 0.0 (:519,:535), first member if none exceeded 0 (:542-547),
 *              UINT32_MAX for an empty window;
 * status_out[m] = GA_RELAY_WINNER / GA_RELAY_CANCELLED for every member (:553-571).
 * wake_off entries must be multiples of 4 samples.  All arrays are device pointers. */
int ga_k_arbitrate(const float *d_samples, const uint64_t *d_wake_off, const uint32_t *d_wake_n,
                   const uint8_t *d_status_in, const uint32_t *d_seg_offsets, uint32_t n_windows,
                   uint32_t n_members, double *d_scores, uint32_t *d_winner, uint8_t *d_status_out,
                   void *stream);

/* Scores only: calculateSignalStrength (audio_service.go:818-852) of n_members buffers.
 * d_stats (nullable) receives the logged side statistics {sumsq, rms, n_samples, nonzero,
 * peak} per member; its digest bytes are left untouched. */
int ga_k_signal_strength(const float *d_samples, const uint64_t *d_wake_off, const uint32_t *d_wake_n,
                         const uint8_t *d_status_in, uint32_t n_members, double *d_scores,
                         ga_result *d_stats, void *stream);

/* Selection only, from precomputed scores (cross-shard windows: the scores arrive as
 * exchanged records).  Same rules as above. */
int ga_k_select(const double *d_scores, const uint8_t *d_status_in, const uint32_t *d_seg_offsets,
                uint32_t n_windows, uint32_t *d_winner, uint8_t *d_status_out, void *stream);

/* Running SHA-256 of one stream, 128 bytes: midstate, bytes hashed so far, and the not yet
 * compressed tail.  A fresh state holds the FIPS 180-4 initial hash value (ga_k_sha256_reset). */
typedef struct ga_sha_state {
    uint32_t h[8];
    uint64_t total_len;
    uint32_t tail_len;    /* 0..63 */
    uint32_t reserved;
    uint8_t  tail[64];
    uint8_t  pad[16];
} ga_sha_state;

/* Per-utterance fingerprint (the "audio_hash" of a voice event: API.md:170, hook
 * events/voice_event.go:84-88; chunks arrive through StreamAudio, audio_service.go:939-1005).
 * Lane r walks the linked list of messages  i = run_first[r], next[i], next[next[i]] …
 * (UINT32_MAX ends it) — all of ONE stream, in arrival order — and continues that stream's
 * running SHA-256 in d_state[desc[i].stream_id]; a message flagged GA_MSG_FIRST restarts it.
 * Launches that touch the same stream must be ordered by the caller. */
int ga_k_sha256_chain(const void *d_pcm, const ga_desc *d_desc, const uint32_t *d_run_first,
                      const uint32_t *d_next, uint32_t n_runs, ga_sha_state *d_state, void *stream);
/* Digest of everything hashed so far for n streams (d_stream_ids NULL = streams 0..n-1), computed
 * on a copy of each state: d_digests[32*n], d_lengths[n] (nullable) = bytes hashed. */
int ga_k_sha256_final(const ga_sha_state *d_state, const uint32_t *d_stream_ids, uint32_t n,
                      uint8_t *d_digests, uint64_t *d_lengths, void *stream);
/* (Re)initialise n states starting at d_state to the empty-message state. */
int ga_k_sha256_reset(ga_sha_state *d_state, uint32_t n, void *stream);

/* Descriptor of one 48 kHz interleaved-stereo int16 message for ga_k_resample48, 64 bytes. */
#define GA_HIST_ZERO   0u   /* stream start: the 47 frames before the message are silence          */
#define GA_HIST_SLAB   1u   /* they are in the PCM slab at hist_off (tail of the previous message)  */
#define GA_HIST_STATE  2u   /* they are in d_hist_state[stream_id] (saved by an earlier launch)     */
#define GA_RS_SAVE_HIST 0x10000u  /* flags bit: save this message's last 47 frames to d_hist_state  */
#define GA_RS_MAX_MSG_BYTES 3840u /* one launch handles messages of up to 960 frames (20 ms)        */
typedef struct ga_rdesc {
    uint64_t in_off;     /* byte offset of the message in the PCM slab; multiple of 16            */
    uint64_t out_off;    /* index of its first OUTPUT sample in the f32 arena                     */
    uint64_t hist_off;   /* GA_HIST_SLAB: byte offset of the 47 preceding frames (multiple of 4)  */
    uint64_t n0;         /* absolute index (since stream start) of the message's first frame      */
    uint32_t nbytes;     /* <= GA_RS_MAX_MSG_BYTES; a trailing partial frame is dropped           */
    uint32_t stream_id;
    uint32_t flags;      /* GA_MSG_* | GA_RS_SAVE_HIST                                            */
    uint32_t frame_off;
    uint32_t hist_kind;  /* GA_HIST_*                                                             */
    uint32_t reserved[3];
} ga_rdesc;

/* 48 kHz stereo -> 16 kHz mono ÷3 decimator + statistics of the resampled samples (BASELINE
 * configs[3]).  New spec, no reference counterpart: the reference ignores AudioChunk.SampleRate
 * (audio_service.go:926-1043).  Per message: mono = (A1(L) + A1(R)) * 0.5f with A1 the
 * reference's conversion (:1078); y[M] = sum over the committed 48 taps (include/gpuaudio_taps.h)
 * in fixed order with fused multiply-adds, on the stream's absolute frame timeline; then
 * results[i].{sumsq,rms,n_samples,nonzero,peak,flags} (and frames) over y as in ga_k_ingest.
 * The digest field of results[i] is not written (run ga_k_ingest with GA_DO_SHA256 on the same
 * slab for the fingerprint of the received bytes).  d_hist_state: 192 bytes per stream (NULL: nothing
 * is carried).  At most ONE message per stream and call may carry GA_RS_SAVE_HIST — the stream's last;
 * the saves run as a second small launch behind the decimator on `stream`, so that a stream's first
 * message of the call (GA_HIST_STATE) reads the state the PREVIOUS call left, wherever in the batch
 * the two messages sit. */
int ga_k_resample48(const void *d_pcm, const ga_rdesc *d_desc, uint32_t n_msgs, void *d_hist_state,
                    float *d_samples, ga_result *d_results, ga_frame *d_frames,
                    double vad_threshold, void *stream);

/* ga_k_arbitrate for callers that cannot allocate GPU memory.  (No counterpart in the
 * reference: the Go hub has no device memory at all.  It exists for hosts of this library whose
 * arbitration call sits where cudaMalloc is off limits — a hub process that hands its whole HBM
 * budget to its own pool at start-up, or a window-close timer path that must not take the
 * device-wide synchronisation an allocation implies while other streams are ingesting.)
 *
 * Inputs, rules and outputs are those of ga_k_arbitrate, minus the score array: the caller gets
 * winner[] and status_out[] only.  The per-member float64 scores, which sit between the scoring
 * launch and the selection launch, live in a 512 KiB array built into the library's device image;
 * every GPU on which the CUDA runtime has loaded that image holds one copy until the process
 * exits.  Consequences:
 *   - n_members may be at most 65,536 per call; more is refused with GA_ERR_OVERFLOW (split the
 *     windows over several calls, or use ga_k_arbitrate with a score array of your own);
 *   - the call makes no cudaMalloc / cudaFree and uses no allocating library routine;
 *   - all calls on one GPU share the array and it keeps the last call's scores; each call first
 *     zeroes the slots it is going to use.  Two calls that overlap on the GPU (different streams,
 *     or different threads without ordering) race on it — serialise them on one stream or with
 *     events;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *     with that device current).  With CUDA's default lazy module loading that lookup can be the
 *     moment the driver places the library's image — this array included — on the device, which
 *     allocates and synchronises once.  So make one warm-up call per device at start-up, with
 *     that device current: before the host hands its HBM to its own pool and before any stream
 *     capture begins.  A call with n_windows == 0 does exactly the set-up and nothing else (all
 *     pointers may be NULL).  Later calls on that device only launch. */
int ga_k_arbitrate_noalloc(const float *d_samples, const uint64_t *d_wake_off, const uint32_t *d_wake_n,
                           const uint8_t *d_status_in, const uint32_t *d_seg_offsets, uint32_t n_windows,
                           uint32_t n_members, uint32_t *d_winner, uint8_t *d_status_out, void *stream);

/* Cross-shard windows: scatter scores that arrived from other shards
 * (d_gathered, the all-gathered guest scores) into this shard's member slots:
 * d_scores[dst_slot[i]] = d_gathered[src_idx[i]].  Runs between the scoring launch and
 * ga_k_select.  No reference counterpart: the reference is a single process. */
int ga_k_route_scores(const double *d_gathered, const uint32_t *d_src_idx,
                      const uint32_t *d_dst_slot, uint32_t n, double *d_scores, void *stream);

/* Fused cross-shard arbitration (SURVEY.md §8e end state): one launch per shard replaces
 * ga_k_signal_strength + all-gather + ga_k_route_scores + ga_k_select.  No reference counterpart
 * (the reference is one process).  Members [0, n_home_members) are the slots of this shard's home
 * windows (remote ones with wake_n = 0), members [n_home_members, +n_guests) are local relays of
 * windows homed elsewhere.  The kernel
 *   1. scores all of them (exact reference order, as ga_k_arbitrate);
 *   2. stores guest j's score to the 8-byte PEER address d_guest_dst[j] + (epoch & 1) *
 *      inbox_parity_stride — the home shard's inbox, mapped over NVLink with ga_p2p_open;
 *   3. its last block writes `epoch` to every address in d_notify (the home shards' flag words),
 *      waits until every LOCAL flag word in d_wait is >= epoch (bounded by timeout_cycles; on
 *      expiry d_error is set non-zero and the launch still completes), copies
 *      d_inbox[parity][k] into d_scores[d_route_dst[k]] and runs the selection for all windows.
 * epoch starts at 1 and grows by 1 per step on every shard; all shards must call once per step.
 * d_counter / d_error: one zero-initialised uint32 each, owned by the caller. */
typedef struct ga_p2p_args {
    const float    *d_samples;
    const uint64_t *d_wake_off;
    const uint32_t *d_wake_n;
    const uint8_t  *d_status_in;
    const uint32_t *d_seg_offsets;
    double         *d_scores;            /* [n_home_members + n_guests]                         */
    uint32_t       *d_winner;
    uint8_t        *d_status_out;        /* [n_home_members]                                    */
    const uint64_t *d_guest_dst;         /* [n_guests] peer addresses (parity 0)                */
    const uint64_t *d_notify;            /* [n_notify] peer flag addresses                      */
    const uint64_t *d_wait;              /* [n_wait]   local flag addresses                     */
    const double   *d_inbox;             /* local inbox, parity 0; parity 1 at +inbox_parity_stride */
    const uint32_t *d_route_dst;         /* [n_in] member slot of inbox entry k                 */
    uint32_t       *d_counter;
    uint32_t       *d_error;
    uint64_t        inbox_parity_stride; /* bytes                                               */
    uint64_t        epoch;
    long long       timeout_cycles;
    uint32_t        n_windows, n_home_members, n_guests, n_notify, n_wait, n_in;
} ga_p2p_args;
int ga_k_arbitrate_p2p(const ga_p2p_args *args, void *stream);
/* Device memory another process on the same node can map (cudaMalloc + CUDA IPC): zero-filled.
 * ga_p2p_open maps a peer's allocation from its 64-byte handle. */
int ga_p2p_alloc(uint64_t bytes, void **d_ptr, uint8_t handle[64]);
int ga_p2p_open(const uint8_t handle[64], void **d_ptr);
int ga_p2p_close(void *d_ptr);
int ga_p2p_free(void *d_ptr);

/* ===================================================================== context
 * Host-buffer layer: what internal/gpuaudio binds.  One context per GPU.        */

typedef struct ga_ctx ga_ctx;
typedef struct ga_cluster ga_cluster;   /* several contexts of one process: see "cluster" below */

typedef struct ga_config {
    int32_t  device;            /* CUDA ordinal                                        */
    uint32_t max_streams;       /* dense stream slots                                  */
    uint32_t max_batch_msgs;    /* largest n accepted by one ga_submit                 */
    uint32_t max_inflight;      /* tickets that may be outstanding (>= 1)              */
    uint64_t ring_bytes;        /* pinned host ring == device PCM slab size            */
    uint64_t arena_samples;     /* f32 samples kept per stream (wake word + speech);
                                   multiple of 4                                       */
    uint64_t wake_samples_max;  /* of which reserved for RelayStream.WakeWordSignal;
                                   multiple of 4, < arena_samples                      */
    uint32_t mode;              /* GA_DO_* bits                                        */
    uint32_t want_frames;       /* != 0: per-frame records are produced and returned   */
    double   vad_threshold;     /* VAD flag = rms > threshold (default 0.01)           */
    uint32_t input_format;      /* GA_FMT_*: what every message of this context holds  */
    uint32_t staging;           /* GA_STAGING_*: how message bytes reach shared memory  */
} ga_config;
#define GA_STAGING_AUTO  0u     /* uniform batches (equal chunk sizes, back to back in the ring) use
                                   one 2-D TMA tensor copy per 32 messages, others one bulk copy
                                   per message                                               */
#define GA_STAGING_BULK  1u     /* always one bulk copy per message                         */
#define GA_FMT_MONO16K   0u     /* 16 kHz mono int16 — what the reference assumes (:609) */
#define GA_FMT_STEREO48K 1u     /* 48 kHz interleaved stereo int16, resampled to 16 kHz mono;
                                   chunks of at most GA_RS_MAX_MSG_BYTES; a chunk that follows
                                   another chunk of its stream in the SAME ga_submit needs that
                                   predecessor to hold at least 47 frames                    */

/* audio_service.go:233-244 is where the hub would call this (createAudioService).  Everything the
 * real-time path needs is allocated here, and the window-close and winner-path kernels are launched
 * once on an empty stream, so that no allocation or module load lands in the first window close. */
int  ga_init(const ga_config *cfg, ga_ctx **out);
void ga_shutdown(ga_ctx *ctx);
const char *ga_last_error(void);

/* Reserve nbytes (rounded up to 16) in the C-owned pinned ring; the caller copies
 * chunk.AudioData there (audio_service.go:964, :999 hand the bytes over here instead of
 * converting them inline).  Callable concurrently from many threads (lock-free).
 * ring_offset is a monotonically increasing virtual offset (the library wraps it);
 * waiting a ticket releases the ring up to the highest offset that ticket contained, so
 * the batching goroutine submits reservations in offset order. */
int ga_ring_reserve(ga_ctx *ctx, uint64_t nbytes, void **host_ptr, uint64_t *ring_offset);

typedef struct ga_msg {
    uint64_t ring_offset;  /* virtual offset from ga_ring_reserve                       */
    uint32_t nbytes;
    uint32_t stream_id;
    uint32_t flags;        /* GA_MSG_*                                                  */
    uint32_t reserved;
} ga_msg;

/* One batch tick: H2D of the ring range + descriptors, the fused ingest launch, D2H of
 * the result records — all asynchronous.  Messages of one stream must appear in arrival
 * order.  Returns a ticket.  Called by the single batching goroutine.
 * A GA_MSG_FIRST message restarts its stream's sample buffers from empty (a new RelayStream,
 * audio_service.go:945-958).  Contexts that convert refuse (GA_ERR_ARG) a restart that follows
 * chunks of the same stream inside ONE batch — all messages of a batch are processed at once and
 * both utterances would be written to the same arena space; send it with the next batch
 * (ga_hub_flush splits its queue there).  Across batches the library orders the two itself. */
int ga_submit(ga_ctx *ctx, const ga_msg *msgs, uint32_t n, uint64_t *ticket);

/* The same batch in its regular steady-state shape, described by a handful of numbers instead
 * of one ga_msg per chunk: n_ticks chunk periods of n_streams relays, every relay sending one
 * chunk of chunk_bytes per period (the StreamAudio loop, audio_service.go:939-1005, seen across
 * all connected relays).  Message m = t * n_streams + j (t-th period, j-th relay) belongs to
 * stream stream_ids[j], lies at ring_offset + m * round_up(chunk_bytes, 16) (one contiguous
 * reservation) and carries flags[m] (GA_MSG_*, low 8 bits; NULL = none).  stream_ids must be
 * distinct.  The per-message descriptors are built on the GPU from the per-stream cursors, so
 * the host sends 1 byte per message and 12 per stream instead of 32 per message.  Results come
 * back exactly as from ga_submit of the same n_ticks * n_streams messages, the running utterance
 * digests (GA_DO_UTTERANCE) included.  48 kHz stereo contexts take chunks that are a multiple of
 * 12 bytes in [192, GA_RS_MAX_MSG_BYTES] (whole groups of three frames, at least the filter
 * history); other sizes go through ga_submit. */
int ga_submit_ticks(ga_ctx *ctx, uint64_t ring_offset, uint32_t n_ticks, uint32_t n_streams,
                    uint32_t chunk_bytes, const uint32_t *stream_ids, const uint8_t *flags,
                    uint64_t *ticket);

/* Block until the ticket's results are on the host, then copy them out (results[n],
 * frames[] if the context wants frames — sized by ga_ticket_frames) and release the
 * ticket's ring range.  Tickets must be waited in submission order. */
int ga_wait(ga_ctx *ctx, uint64_t ticket, ga_result *results, ga_frame *frames);
int ga_ticket_frames(ga_ctx *ctx, uint64_t ticket, uint32_t *n_frames);
/* Same as ga_wait without the copy: returns pointers into the C-owned pinned result buffers of
 * the ticket's slot (cgo wraps them with unsafe.Slice).  Slots are used round-robin, so the
 * records of ticket t stay untouched until the ga_submit / ga_submit_ticks call that returns
 * ticket t + max_inflight.  With max_inflight tickets outstanding when t is waited, that is the
 * very next submit: read (or copy) the records before submitting again, or keep at most
 * max_inflight - 1 tickets in flight.  Any out pointer may be NULL. */
int ga_wait_view(ga_ctx *ctx, uint64_t ticket, const ga_result **results, const ga_frame **frames,
                 uint32_t *n_results, uint32_t *n_frames);

/* Copy a stream's converted samples [first, first+n) to the host — what
 * RelayStream.WakeWordSignal ++ SpeechAudio holds (audio_service.go:592-594); used for
 * arbitration winners only. */
int ga_fetch_samples(ga_ctx *ctx, uint32_t stream_id, uint64_t first, uint64_t n, float *out);
int ga_stream_info(ga_ctx *ctx, uint32_t stream_id, uint64_t *n_samples, uint64_t *wake_samples);
int ga_stream_reset(ga_ctx *ctx, uint32_t stream_id);
/* Contexts with GA_DO_UTTERANCE: SHA-256 of all bytes the stream has received since its last
 * GA_MSG_FIRST (in arrival order), as of the last submitted batch; total_bytes may be NULL.
 * Typically read when the chunk flagged IsEndOfSpeech (audio_service.go:1008) has been waited. */
int ga_utterance_digest(ga_ctx *ctx, uint32_t stream_id, uint8_t digest[32], uint64_t *total_bytes);
/* Device address of the f32 sample arena (stream s starts at s * arena_samples; its wake-word
 * region first, its speech region at + wake_samples_max).  For harnesses that run the
 * ga_k_* arbitration sequence (e.g. with an exchange step between shards) on the context's
 * samples.  Valid until ga_shutdown; wait all tickets before reading. */
int ga_device_arena(ga_ctx *ctx, void **d_arena);
/* Fingerprint of a stream's CONVERTED audio: SHA-256 over the little-endian IEEE-754 bytes of
 * WakeWordSignal ++ SpeechAudio — the byte string the hub itself serialises for STT
 * (internal/llm/stt_client.go:390-395) and the natural input of the hook whose argument is
 * []float32 (VoiceEvent.SetAudioMetadata, events/voice_event.go:84-88).  Hashed on the device from
 * the sample arena; nothing is copied to the host but the digest. */
int ga_samples_digest(ga_ctx *ctx, uint32_t stream_id, uint8_t digest[32]);

/* performArbitration for many windows (audio_service.go:505-571).  Host arrays:
 * seg_offsets[n_windows+1], stream_ids / status_in / scores / status_out per member,
 * winner_idx[n_windows] (member index, UINT32_MAX for an empty window).  A member's wake
 * buffer is the samples its GA_MSG_WAKE messages produced. */
int ga_arbitrate(ga_ctx *ctx, const uint32_t *seg_offsets, uint32_t n_windows,
                 const uint32_t *stream_ids, const uint8_t *status_in,
                 uint32_t *winner_idx, double *scores, uint8_t *status_out);

/* Winner audio as a WAV file for the STT call (float32ToWAV, internal/llm/stt_client.go:365-398):
 * 44-byte RIFF header (format 3 = IEEE float, mono, 32 bit) + WakeWordSignal ++ SpeechAudio copied
 * device->host straight behind it.  out == NULL only reports the size in *n_bytes. */
int ga_fetch_wav(ga_ctx *ctx, uint32_t stream_id, uint32_t sample_rate, uint8_t *out, size_t cap,
                 size_t *n_bytes);

/* The winner path for MANY streams in one call — what runs once per closed window in the reference:
 * processWinningRelayAudio (audio_service.go:590-609, fullAudio = WakeWordSignal ++ SpeechAudio),
 * float32ToWAV (internal/llm/stt_client.go:365-398) and the fingerprint of those samples
 * (events/voice_event.go:84-88).  One gather launch packs the n WAV files, one chained SHA-256
 * launch hashes the samples where they lie, ONE device->host copy brings all of it into C-owned
 * pinned memory; no allocation per call (the staging grows on demand and is kept), and the
 * context is locked only while the work is queued, so ga_submit is not held up.
 *   *wavs + offsets[k]  = WAV file of stream_ids[k] (44 + 4 * n_samples bytes; offsets[n] ends the
 *                         last one; files are padded apart), valid until the next ga_fetch_wavs /
 *                         ga_samples_digest(s) call on this context;
 *   digests (nullable)  = 32 bytes per stream: SHA-256 of the samples' little-endian f32 bytes. */
int ga_fetch_wavs(ga_ctx *ctx, const uint32_t *stream_ids, uint32_t n, uint32_t sample_rate,
                  const uint8_t **wavs, uint64_t *offsets, uint8_t *digests);
/* ga_samples_digest for n streams in one launch. */
int ga_samples_digests(ga_ctx *ctx, const uint32_t *stream_ids, uint32_t n, uint8_t *digests);

/* ===================================================================== hub (window manager)
 * Host-side mirror of the reference's AudioService for this path (internal/grpc/audio_service.go):
 * relay registry, arbitration-window lifecycle, per-chunk routing.  It is the logic the Go code
 * around this library keeps, provided in C++ so it can run and be tested
 * without a Go toolchain, and generalised from the reference's single hub-wide window (:89) to one
 * window per collision group.  Time is injected in microseconds.  Thread-safe: one mutex for the
 * registry and the windows; the per-chunk path of speech chunks does not take it. */
typedef struct ga_hub ga_hub;
/* createAudioService :201-290; window_duration_us 0 = the reference's 300 ms (:243). */
int  ga_hub_create(ga_ctx *ctx, uint32_t max_streams, uint64_t window_duration_us, ga_hub **out);
void ga_hub_destroy(ga_hub *hub);
/* A relay opened its stream (StreamAudio :926); assigns it a stream slot of the context. */
int  ga_hub_connect(ga_hub *hub, const char *relay_id, uint32_t group, uint32_t *slot);
/* startArbitrationWindow :408-446 / joinArbitrationWindow :449-491 (joined = 0: no window, window
 * already arbitrated, or elapsed > duration). */
int  ga_hub_start_window(ga_hub *hub, const char *relay_id, uint64_t now_us, uint64_t *window_id);
int  ga_hub_join_window(ga_hub *hub, const char *relay_id, uint64_t now_us, int *joined);
/* The body of the StreamAudio loop for one AudioChunk (:939-1041): wake-word chunks start or join
 * the group's window and extend WakeWordSignal, other chunks extend SpeechAudio of relays that are
 * in activeStreams; the bytes go into the pinned ring and are queued for the next flush.
 * *keep_stream = 0 where the reference handler returns (late relay :990-995, cancelled at
 * end of speech :1018-1021). */
int  ga_hub_on_chunk(ga_hub *hub, const char *relay_id, const uint8_t *audio, uint32_t nbytes,
                     int is_wake_word, int is_end_of_speech, uint64_t now_us, int *keep_stream);
/* The same through a handle, for the per-chunk path of a hub with thousands of relays: the handle
 * (ga_hub_relay_handle, valid until the relay is cleaned up) stands for the relay's registry
 * entry, so no name lookup and — for speech chunks, all but the first ~15 of an utterance — no
 * hub-wide lock is taken: the reference's per-chunk streamsMutex (:1000-1004) is gone.  The bytes
 * are copied into the pinned ring outside any lock.  Callable from many threads at once; one
 * relay's chunks must come from one thread at a time, in order (its StreamAudio goroutine).
 * GA_ERR_RING_FULL: the ring or the chunk queue is full until the next ga_hub_flush. */
int  ga_hub_relay_handle(ga_hub *hub, const char *relay_id, uint64_t *handle);
int  ga_hub_on_chunk_h(ga_hub *hub, uint64_t handle, const uint8_t *audio, uint32_t nbytes,
                       int is_wake_word, int is_end_of_speech, uint64_t now_us, int *keep_stream);
/* Batch tick: ga_submit + wait for everything queued since the last flush — one batch per GPU,
 * more where a relay restarted an utterance inside the tick or the queue exceeds the context's
 * max_batch_msgs (*n_msgs = messages ingested).  A relay whose sample arena is full (the
 * reference appends without bound, :965 / :1002) loses its chunks of this tick and is taken out of
 * activeStreams; everybody else's chunks are ingested and the call returns GA_ERR_OVERFLOW.  On any
 * other failure the chunks not yet submitted stay queued for the next flush.  One flush at a time
 * (calls serialise); chunks keep arriving meanwhile. */
int  ga_hub_flush(ga_hub *hub, uint32_t *n_msgs);
/* performArbitration :505-587 for one window (idempotent, :508-511) … */
int  ga_hub_perform_arbitration(ga_hub *hub, uint64_t window_id);
/* … and runArbitrationTimer :494-502 for all groups: every active window whose duration has
 * elapsed at now_us is arbitrated in ONE ga_arbitrate call. */
int  ga_hub_poll(ga_hub *hub, uint64_t now_us, uint32_t *n_closed);
/* ga_hub_poll that also reports which windows it closed (at most cap ids; *n_closed = all). */
int  ga_hub_poll_windows(ga_hub *hub, uint64_t now_us, uint64_t *window_ids, uint32_t cap, uint32_t *n_closed);
/* processWinningRelayAudio (:590-609) for the winners of n closed windows at once: per GPU one
 * ga_fetch_wavs.  wav_ptrs[k] / wav_bytes[k] = the winner's WAV file for the STT request, inside
 * C-owned pinned memory (valid until the next call); audio_hashes (nullable, GA_AUDIO_HASH_LEN
 * bytes per window) = "sha256:<hex>" of its samples.  No winner: NULL / 0 / "". */
int  ga_hub_fetch_winner_wavs(ga_hub *hub, const uint64_t *window_ids, uint32_t n, uint32_t sample_rate,
                              const uint8_t **wav_ptrs, uint64_t *wav_bytes, char *audio_hashes);
int  ga_hub_window_info(ga_hub *hub, uint64_t window_id, int *is_active, char *winner_id,
                        size_t winner_cap, uint32_t *n_relays);
int  ga_hub_active_window(ga_hub *hub, uint32_t group, uint64_t *window_id);   /* 0 = nil */
int  ga_hub_relay_info(ga_hub *hub, const char *relay_id, int *status, double *signal_strength,
                       int *cancel_closed, uint32_t *slot);
int  ga_hub_is_relay_active(ga_hub *hub, const char *relay_id);                /* isRelayActive :880-913: 1 / 0, negative = error */
int  ga_hub_set_relay_status(ga_hub *hub, const char *relay_id, int status);   /* as.activeStreams[id].Status = … */
int  ga_hub_complete_request(ga_hub *hub, const char *relay_id);               /* collision_detection_test.go:361-369 */
int  ga_hub_cleanup_relay(ga_hub *hub, const char *relay_id);                  /* cleanupRelay :916-923 */
int  ga_hub_active_streams(ga_hub *hub, uint32_t *n);                          /* len(as.activeStreams) */
/* The same hub over every GPU of the box (one process, one context per GPU, held by `cluster`,
 * which must outlive the hub): ga_hub_connect places a relay on shard group % nGPU so that a
 * group's window normally stays on one GPU; ga_hub_connect_on forces a shard (the window then
 * spans GPUs and is arbitrated through ga_cluster_arbitrate).  Chunks go to the relay's own GPU;
 * ga_hub_flush submits one batch per GPU.  max_streams_per_shard <= the contexts' max_streams. */
int  ga_hub_create_cluster(ga_cluster *cluster, uint32_t max_streams_per_shard, uint64_t window_duration_us,
                           ga_hub **out);
int  ga_hub_connect_on(ga_hub *hub, const char *relay_id, uint32_t group, uint32_t shard, uint32_t *slot);
int  ga_hub_relay_shard(ga_hub *hub, const char *relay_id, uint32_t *shard);
/* ga_fetch_wav for a relay, from whichever GPU holds its audio. */
int  ga_hub_fetch_wav(ga_hub *hub, const char *relay_id, uint32_t sample_rate, uint8_t *out, size_t cap,
                      size_t *n_bytes);

/* --------------------------------------------------------------- one-shot mirrors
 * Same argument meaning and degenerate-input behaviour as the Go functions; used by
 * the parity tests and by callers that have a single buffer. */

/* bytesToFloat32Array(data) — audio_service.go:1048-1101.  Writes len/2 samples to out
 * (odd trailing byte dropped, empty input -> 0 samples, never an error); res may be NULL. */
int ga_bytes_to_float32(ga_ctx *ctx, const uint8_t *data, size_t len, float *out,
                        size_t *n_out, ga_result *res);
/* calculateSignalStrength(samples) — audio_service.go:818-852.  Empty -> rms 0.0. */
int ga_signal_strength(ga_ctx *ctx, const float *samples, size_t n, double *rms, ga_result *res);
/* SHA-256 of an arbitrary host byte string (the fingerprint hook, events/voice_event.go:84-88). */
int ga_sha256(ga_ctx *ctx, const uint8_t *data, size_t len, uint8_t digest[32]);
/* VoiceEvent.SetAudioMetadata(audioData []float32, sampleRate, isWakeWord) with the fingerprint
 * restored (events/voice_event.go:84-88; field removed by storage/migrations/
 * 001_remove_audio_hash.sql): *audio_duration = float64(n) / float64(sample_rate) as the hook
 * computes it, and audio_hash = "sha256:" + lowercase hex (the documented wire format,
 * API.md:170) of the SHA-256 of the samples' little-endian IEEE-754 bytes, hashed on the GPU.
 * zero_persistence != 0 (PrivacyConfig.ZeroPersistence, config/config.go:102-103) computes no
 * fingerprint and returns audio_hash = "".  audio_hash holds GA_AUDIO_HASH_LEN bytes. */
#define GA_AUDIO_HASH_LEN 72u   /* 7 + 64 + NUL */
int ga_set_audio_metadata(ga_ctx *ctx, const float *audio_data, size_t n, int sample_rate,
                          int zero_persistence, double *audio_duration, char *audio_hash);
/* The same for a stream whose samples are already in HBM (ga_samples_digest) — what the hub
 * calls for an arbitration winner after processWinningRelayAudio (audio_service.go:592-594). */
int ga_stream_audio_metadata(ga_ctx *ctx, uint32_t stream_id, int sample_rate, int zero_persistence,
                             double *audio_duration, char *audio_hash);
/* "sha256:<hex>" of a 32-byte digest; pure host code. */
void ga_format_audio_hash(const uint8_t digest[32], char *audio_hash);

/* ===================================================================== cluster
 * Several contexts of ONE process, one per GPU — how a single hub process drives an 8-GPU box.
 * A collision group is assigned to a shard when its window opens (startArbitrationWindow,
 * audio_service.go:408); relays that joined through another shard leave the window spanning GPUs.
 * ga_cluster_arbitrate is performArbitration (:505-571) for any mix of local and spanning windows:
 * member m is stream stream_ids[m] of context shard_ids[m] (join order per window, as in
 * ga_arbitrate); a window is homed on the shard of its FIRST member.  One fused launch per GPU
 * (ga_k_arbitrate_p2p): every shard scores its own relays, stores the scores of relays homed
 * elsewhere into the home shard's inbox over NVLink and raises a flag there; the home kernel
 * waits for its senders (bounded) and selects.  Outputs are indexed like the inputs: winner_idx[w]
 * is a member index or 0xFFFFFFFF for an empty window.  max_guests bounds the remote relays one
 * shard may receive per call.  The contexts must sit on distinct, peer-accessible devices and
 * outlive the cluster.  GA_ERR_COMM: no peer path, or a peer's scores did not arrive in time. */
int  ga_cluster_create(ga_ctx *const *ctxs, uint32_t n_ctx, uint32_t max_guests, ga_cluster **out);
void ga_cluster_destroy(ga_cluster *cl);
int  ga_cluster_arbitrate(ga_cluster *cl, const uint32_t *seg_offsets, uint32_t n_windows,
                          const uint32_t *shard_ids, const uint32_t *stream_ids, const uint8_t *status_in,
                          uint32_t *winner_idx, double *scores, uint8_t *status_out);
uint32_t ga_cluster_size(const ga_cluster *cl);
ga_ctx  *ga_cluster_context(const ga_cluster *cl, uint32_t shard);

/* ===================================================================== utilities */

/* Synthetic relay PCM: integer-only, bit-identical on host and device.
 * Fills n_samples int16 samples of relay view (group, relay) starting at sample index
 * `first`, for an utterance of utt_samples samples. */
int ga_synth_fill_host(int16_t *out, uint64_t seed, uint32_t group, uint32_t relay,
                       uint32_t first, uint32_t n_samples, uint32_t utt_samples);
/* Fill a whole time-major batch on the device: message m of stream s (= group*group_size
 * + relay, group offset by group_base) at d_pcm + (m*n_streams + s)*msg_samples*2. */
int ga_synth_fill_batch_device(void *d_pcm, uint64_t seed, uint32_t group_base,
                               uint32_t n_streams, uint32_t group_size, uint32_t n_msgs,
                               uint32_t msg_samples, void *stream);

/* As above with an explicit (group, relay) per stream slot (sharded layouts). */
int ga_synth_fill_batch_device_map(void *d_pcm, uint64_t seed, const uint32_t *d_group,
                                   const uint32_t *d_relay, uint32_t n_streams, uint32_t n_msgs,
                                   uint32_t msg_samples, void *stream);
/* 48 kHz interleaved-stereo version (configs[3]): 3x zero-order hold of the 16 kHz relay view,
 * R = L - (L >> 3); message m of stream s at d_pcm + (m*n_streams + s)*msg_frames*4. */
int ga_synth_fill_batch48_device_map(void *d_pcm, uint64_t seed, const uint32_t *d_group,
                                     const uint32_t *d_relay, uint32_t n_streams, uint32_t n_msgs,
                                     uint32_t msg_frames, void *stream);
/* Same batch layout, filled on the host with n_threads workers (CPU-baseline input). */
int ga_synth_fill_batch_host(void *pcm, uint64_t seed, uint32_t group_base, uint32_t n_streams,
                             uint32_t group_size, uint32_t n_msgs, uint32_t msg_samples,
                             uint32_t n_threads);

/* Introspection for harnesses. */
uint64_t ga_launch_count(void);            /* kernels launched by this library so far */
int      ga_abi_version(void);
/* sizes of {ga_result, ga_frame, ga_desc, ga_msg, ga_config, ga_rdesc, ga_sha_state} for binding
 * self-checks */
void     ga_abi_sizes(uint32_t out[7]);
int      ga_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* GPUAUDIO_H */

/*
 * gpuaudio_oracle.h — CPU restatement of the loqa-hub relay-audio ingest path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it, and there only as the checker or the timed CPU baseline.
 * The product (loqa_hub_b200/libgpuaudio.so) never links, loads or calls it.
 *
 * Parity status (see DESIGN.md "Oracle"):
 *   pinned   : ora_bytes_to_float32, ora_signal_strength, ora_arbitrate — against
 *              the reference's own test vectors (internal/grpc/audio_service_test.go:258-364,
 *              internal/grpc/collision_detection_test.go:48-92,226-293,395-459).
 *   pinned by standard : ora_sha256 — FIPS 180-4 / NIST known answers + OpenSSL
 *              (the reference has no SHA-256 call site left; it would be Go's
 *              crypto/sha256 from the Go 1.25.1 stdlib, go.mod:3).
 *   parity unpinned : ora_frame_features (20 ms framing), ora_resample_48k_stereo
 *              (÷3 polyphase), — new-spec pieces that have no reference code;
 *              this restatement IS their definition.
 *              The decimator is additionally held, bit for bit, to an independent
 *              exact-rational evaluation of the same definition
 *              (oracle/numpy_mirror.py Resampler, tests/golden/path_golden.json).
 *   pinned by format : ora_float32_to_wav — restates internal/llm/stt_client.go:365-398.
 *
 * All citations are relative to /root/reference.
 * Build: gcc -O2 -ffp-contract=off (never -ffast-math); see oracle/Makefile.
 */
#ifndef GPUAUDIO_ORACLE_H
#define GPUAUDIO_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* RelayStatus — internal/grpc/audio_service.go:56-63 */
enum {
    ORA_RELAY_CONNECTED  = 0,
    ORA_RELAY_CONTENDING = 1,
    ORA_RELAY_WINNER     = 2,
    ORA_RELAY_CANCELLED  = 3
};

/* side statistics the reference logs: audio_service.go:1093-1098 and :843-849 */
typedef struct {
    uint32_t n_samples;   /* output_samples / total_samples */
    uint32_t nonzero;     /* non_zero_samples               */
    float    peak;        /* max_abs_value / max_sample     */
    uint32_t _pad;
    double   sumsq;       /* "sum" (A2 only; 0 for A1)      */
    double   rms;         /* "rms" (A2 only; 0 for A1)      */
} ora_stats;

/* A1: bytesToFloat32Array — audio_service.go:1048-1101.
 * Returns the number of samples written to out (len/2; odd trailing byte dropped;
 * empty input gives 0).  out must hold len/2 floats.  st may be NULL. */
size_t ora_bytes_to_float32(const uint8_t *data, size_t len, float *out, ora_stats *st);

/* A2: calculateSignalStrength — audio_service.go:818-852.  Returns rms; st may be NULL. */
double ora_signal_strength(const float *samples, size_t n, ora_stats *st);

/* A3: the selection half of performArbitration — audio_service.go:517-571.
 * Members are visited in join order (our deterministic stand-in for Go's random
 * map order, SURVEY.md Appendix B rule 5).  wake[i]/wake_n[i] is member i's
 * WakeWordSignal.  status_in[i] is its RelayStatus before the call; scores[i]
 * receives SignalStrength for contending members and is left untouched
 * otherwise (the reference never writes it for skipped relays, :522-524).
 * status_out[i] = WINNER for the winner, CANCELLED for every other member (:553-571).
 * Returns the winner's index, or -1 for an empty window. */
int ora_arbitrate(uint32_t n_members, const float *const *wake, const size_t *wake_n,
                  const uint8_t *status_in, double *scores, uint8_t *status_out);

/* Same selection from precomputed scores (used for cross-shard windows where the
 * scores arrive as exchanged records). */
int ora_select_winner(uint32_t n_members, const double *scores, const uint8_t *status_in,
                      uint8_t *status_out);

/* N3: SHA-256 (FIPS 180-4), scalar restatement. */
typedef struct {
    uint32_t h[8];
    uint8_t  tail[64];
    uint32_t tail_len;
    uint32_t _pad;
    uint64_t total_len;
} ora_sha256_ctx;
void ora_sha256_init(ora_sha256_ctx *c);
void ora_sha256_update(ora_sha256_ctx *c, const uint8_t *data, size_t len);
void ora_sha256_final(const ora_sha256_ctx *c, uint8_t out[32]); /* does not modify c */
void ora_sha256(const uint8_t *data, size_t len, uint8_t out[32]);

/* N1: 20 ms framing.  Frame f covers samples [f*frame_len, min((f+1)*frame_len, n)).
 * Each frame record is A2 applied to that slice; vad = rms > vad_threshold. */
typedef struct {
    double   sumsq;
    double   rms;
    uint32_t n_samples;
    uint32_t nonzero;
    float    peak;
    uint32_t vad;
} ora_frame;
size_t ora_frame_features(const float *samples, size_t n, uint32_t frame_len,
                          double vad_threshold, ora_frame *out);

/* N2: 48 kHz interleaved-stereo int16 → 16 kHz mono f32.
 *   mono[j] = (A1(L[j]) + A1(R[j])) * 0.5f                        (f32, RN)
 *   y[M]    = fold_{k=0..T-1} acc = fmaf(taps[k], mono[3M+2-k], acc), acc0 = 0
 * on the stream's absolute input-frame timeline; frames before the stream start
 * are zero.  state carries the last T-1 mono frames and the absolute frame count.
 * Returns the number of output samples produced by this chunk.  A trailing
 * partial frame (len % 4 bytes) is dropped, mirroring A1's odd-byte rule. */
#define ORA_RESAMPLE_TAPS 48
typedef struct {
    float    hist[ORA_RESAMPLE_TAPS - 1]; /* mono frames n0-47 .. n0-1 */
    uint32_t _pad;
    uint64_t n_frames;                    /* absolute input frames consumed so far */
} ora_resample_state;
void   ora_resample_init(ora_resample_state *s);
size_t ora_resample_48k_stereo(ora_resample_state *s, const uint8_t *data, size_t len,
                               float *out);
const float *ora_resample_taps(void);

/* stt_client.go:365-398: 44-byte RIFF header (format 3 IEEE float, mono, 32-bit)
 * followed by little-endian f32 samples.  out must hold 44 + 4*n bytes. */
size_t ora_float32_to_wav(const float *samples, size_t n, uint32_t sample_rate, uint8_t *out);

/* ------------------------------------------------------------------ baseline */
/* Timed CPU baseline over one batch laid out like the GPU path's input:
 * n_streams streams, each n_msgs messages of msg_bytes bytes; message m of
 * stream s lives at pcm + ((uint64_t)m * n_streams + s) * msg_bytes (time-major,
 * the order chunks arrive in).  Per stream, one worker does what one reference
 * goroutine does per chunk (StreamAudio, audio_service.go:939-1005): fresh
 * allocation + convert (A1), append to the stream's buffer, SHA-256 of the
 * chunk; then per collision window (group_size consecutive streams) A2 over the
 * first wake_msgs messages' samples and the A3 selection.  No logging (the real
 * reference additionally writes 2-3 zap lines per chunk).
 * n_threads workers pull streams from a shared counter.
 * digests_out (n_streams*n_msgs*32, message-major as the input) and
 * winners_out (n_streams/group_size) may be NULL.
 * use_openssl != 0 selects OpenSSL's EVP SHA-256 (SHA-NI where the CPU has it —
 * the closest stand-in for Go's crypto/sha256) when the library was built with
 * it; ora_have_openssl() tells.  Returns elapsed seconds (CLOCK_MONOTONIC). */
int    ora_have_openssl(void);
double ora_cpu_baseline(const uint8_t *pcm, uint32_t n_streams, uint32_t n_msgs,
                        uint32_t msg_bytes, uint32_t wake_msgs, uint32_t group_size,
                        uint32_t n_threads, int use_openssl,
                        uint8_t *digests_out, int32_t *winners_out, double *scores_out);

/* The general form (ora_cpu_baseline above = mono, time-major, EVP).  See cpu_baseline.c for what
 * each choice stands for in the Go path. */
enum { ORA_FMT_MONO16K = 0, ORA_FMT_STEREO48K = 1 };         /* BASELINE configs[1,2,4] / configs[3] */
enum { ORA_LAYOUT_TIME_MAJOR = 0, ORA_LAYOUT_STREAM_MAJOR = 1 };
enum { ORA_SHA_PORT = 0, ORA_SHA_OPENSSL_EVP = 1, ORA_SHA_OPENSSL_LOW = 2 };
typedef struct {
    uint32_t n_streams, n_msgs, msg_bytes, wake_msgs, group_size;
    uint32_t format;      /* ORA_FMT_*: 48 kHz stereo chunks go through the decimator (N2) instead of A1 */
    uint32_t layout;      /* ORA_LAYOUT_*: chunk (s, m) at ((s*n_msgs + m) or (m*n_streams + s)) * msg_bytes */
    uint32_t sha_impl;    /* ORA_SHA_* (falls back to the scalar port without OpenSSL)                   */
    uint32_t n_threads;
} ora_baseline_cfg;
/* digests_out stays message-major [m][s][32] for either input layout. */
double ora_cpu_baseline_run(const ora_baseline_cfg *cfg, const uint8_t *pcm, uint8_t *digests_out,
                            int32_t *winners_out, double *scores_out);
/* Plain memcpy bandwidth of n_threads host threads (GB/s, read + write): the ceiling a
 * memory-bound baseline could not exceed, printed beside the thread-scaling table. */
double ora_memcpy_gbs(uint32_t n_threads, size_t bytes_per_thread, uint32_t reps);

/* ------------------------------------------------------------------ synthetic input */
/* The integer-only synthetic relay PCM of SURVEY.md §8d, restated for the CPU arms (the GPU
 * arm generates the same bytes on the device; tests compare the two).  Stream s is relay
 * s % group_size of collision group group_base + s / group_size; n_msgs messages of msg_samples
 * 16 kHz samples each.  stereo48 != 0: 3x zero-order hold to 48 kHz, R = L - (L >> 3), i.e.
 * msg_samples*3 interleaved stereo frames (12 bytes per 16 kHz sample) per message.
 * layout: ORA_LAYOUT_*.  n_threads workers. */
void ora_synth_fill(void *out, uint64_t seed, uint32_t group_base, uint32_t n_streams,
                    uint32_t group_size, uint32_t n_msgs, uint32_t msg_samples, int stereo48,
                    uint32_t layout, uint32_t n_threads);
int16_t ora_synth_sample(uint64_t seed, uint32_t group, uint32_t relay, uint32_t n, uint32_t utt_samples);

#ifdef __cplusplus
}
#endif
#endif

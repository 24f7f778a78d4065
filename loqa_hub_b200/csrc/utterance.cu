// utterance.cu — per-utterance running SHA-256 fingerprint, one lane per relay stream.
//
// The hub's docs describe ONE audio hash per voice event ("audio_hash": "sha256:…", API.md:170;
// hook VoiceEvent.SetAudioMetadata, internal/events/voice_event.go:84-88; the code was removed
// by internal/storage/migrations/001_remove_audio_hash.sql).  An utterance reaches the hub as
// many chunks (StreamAudio, internal/grpc/audio_service.go:939-1005), so the fingerprint of the
// whole utterance is a running hash over the received bytes in arrival order — wake-word chunks
// first, then speech chunks, the order processWinningRelayAudio concatenates them in (:592-594).
// The chain is inherently serial per stream; the parallelism is across streams: each lane owns
// one stream, walks that stream's messages of the batch in order (a linked list the host lays
// out), and carries {midstate, unprocessed tail <= 63 B, byte count} in a 128-byte per-stream
// state record between launches.  Finalisation (padding + length) works on a copy, so a digest
// can be taken at any time (at IsEndOfSpeech, :1008) without disturbing the running state.
#include "common.cuh"
#include "sha256.cuh"

namespace ga {

constexpr uint32_t kNone = 0xffffffffu;

// 64 bytes at p (any alignment) -> 16 big-endian words
__device__ __forceinline__ void load_block(const uint8_t *p, uint32_t (&w)[16])
{
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    if ((a & 15u) == 0) {
        const uint4 *q = reinterpret_cast<const uint4 *>(p);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint4 t = q[j];
            w[4 * j] = bswap32(t.x); w[4 * j + 1] = bswap32(t.y); w[4 * j + 2] = bswap32(t.z); w[4 * j + 3] = bswap32(t.w);
        }
    } else if ((a & 3u) == 0) {
        const uint32_t *q = reinterpret_cast<const uint32_t *>(p);
#pragma unroll
        for (int j = 0; j < 16; j++) w[j] = bswap32(q[j]);
    } else {
#pragma unroll
        for (int j = 0; j < 16; j++)
            w[j] = (static_cast<uint32_t>(p[4 * j]) << 24) | (static_cast<uint32_t>(p[4 * j + 1]) << 16) |
                   (static_cast<uint32_t>(p[4 * j + 2]) << 8) | static_cast<uint32_t>(p[4 * j + 3]);
    }
}

__global__ void __launch_bounds__(128)
sha256_chain_kernel(const uint8_t *__restrict__ pcm, const ga_desc *__restrict__ desc,
                    const uint32_t *__restrict__ run_first, const uint32_t *__restrict__ next,
                    uint32_t n_runs, ga_sha_state *__restrict__ state)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_runs) return;
    uint32_t i = run_first[r];
    if (i == kNone) return;
    ga_sha_state *st = state + desc[i].stream_id;
    Sha256State sha;
#pragma unroll
    for (int j = 0; j < 8; j++) sha.h[j] = st->h[j];
    uint64_t total = st->total_len;
    uint32_t tail_len = st->tail_len;

    while (i != kNone) {
        const ga_desc d = desc[i];
        if (d.flags & GA_MSG_FIRST) { sha.init(); total = 0; tail_len = 0; }
        const uint8_t *p = pcm + d.in_off;
        uint32_t n = d.nbytes;
        total += n;
        if (tail_len) {                                   // top up the carried partial block first
            const uint32_t take = min(64u - tail_len, n);
            for (uint32_t k = 0; k < take; k++) st->tail[tail_len + k] = p[k];
            tail_len += take; p += take; n -= take;
            if (tail_len == 64) {
                uint32_t w[16];
                load_block(st->tail, w);
                sha256_compress(sha, w);
                tail_len = 0;
            }
        }
        for (; n >= 64; n -= 64, p += 64) {               // whole blocks straight from the slab
            uint32_t w[16];
            load_block(p, w);
            sha256_compress(sha, w);
        }
        if (n) {                                          // keep the rest for the next message
            for (uint32_t k = 0; k < n; k++) st->tail[k] = p[k];
            tail_len = n;
        }
        i = next[i];
    }
#pragma unroll
    for (int j = 0; j < 8; j++) st->h[j] = sha.h[j];
    st->total_len = total;
    st->tail_len = tail_len;
}

__global__ void sha256_final_kernel(const ga_sha_state *__restrict__ state, const uint32_t *__restrict__ stream_ids,
                                    uint32_t n, uint8_t *__restrict__ digests, uint64_t *__restrict__ lengths)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const ga_sha_state *st = state + (stream_ids ? stream_ids[t] : t);
    Sha256State sha;
#pragma unroll
    for (int j = 0; j < 8; j++) sha.h[j] = st->h[j];
    const uint32_t rem = st->tail_len & 63u;
    const uint64_t bits = st->total_len * 8ull;
    uint32_t w[16];
#pragma unroll
    for (int j = 0; j < 16; j++) {
        uint32_t x = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const uint32_t pos = 4u * j + b;
            uint32_t byte = 0;
            if (pos < rem) byte = st->tail[pos];
            else if (pos == rem) byte = 0x80u;
            x |= byte << (24 - 8 * b);
        }
        w[j] = x;
    }
    if (rem >= 56) {
        sha256_compress(sha, w);
#pragma unroll
        for (int j = 0; j < 16; j++) w[j] = 0;
    }
    w[14] = static_cast<uint32_t>(bits >> 32);
    w[15] = static_cast<uint32_t>(bits);
    sha256_compress(sha, w);
    uint32_t *out = reinterpret_cast<uint32_t *>(digests + 32ull * t);
#pragma unroll
    for (int j = 0; j < 8; j++) out[j] = bswap32(sha.h[j]);
    if (lengths) lengths[t] = st->total_len;
}

__global__ void sha256_reset_kernel(ga_sha_state *state, uint32_t n)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    Sha256State iv;
    iv.init();
    ga_sha_state *st = state + t;
#pragma unroll
    for (int j = 0; j < 8; j++) st->h[j] = iv.h[j];
    st->total_len = 0; st->tail_len = 0; st->reserved = 0;
}

}  // namespace ga

extern "C" int ga_k_sha256_reset(ga_sha_state *d_state, uint32_t n, void *stream)
{
    using namespace ga;
    if (n == 0) return GA_OK;
    if (!d_state) { set_error("ga_k_sha256_reset: NULL argument"); return GA_ERR_ARG; }
    sha256_reset_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_state, n);
    GA_CHECK_LAUNCH("sha256_reset_kernel");
    return GA_OK;
}

extern "C" int ga_k_sha256_chain(const void *d_pcm, const ga_desc *d_desc, const uint32_t *d_run_first,
                                 const uint32_t *d_next, uint32_t n_runs, ga_sha_state *d_state, void *stream)
{
    using namespace ga;
    if (n_runs == 0) return GA_OK;
    if (!d_pcm || !d_desc || !d_run_first || !d_next || !d_state) { set_error("ga_k_sha256_chain: NULL argument"); return GA_ERR_ARG; }
    // a run is serial; few runs (the winners of a window close) are spread one warp per block over the SMs
    const uint32_t tpb = n_runs > 128u * 148u ? 128u : 32u;
    sha256_chain_kernel<<<(n_runs + tpb - 1) / tpb, tpb, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint8_t *>(d_pcm), d_desc, d_run_first, d_next, n_runs, d_state);
    GA_CHECK_LAUNCH("sha256_chain_kernel");
    return GA_OK;
}

extern "C" int ga_k_sha256_final(const ga_sha_state *d_state, const uint32_t *d_stream_ids, uint32_t n,
                                 uint8_t *d_digests, uint64_t *d_lengths, void *stream)
{
    using namespace ga;
    if (n == 0) return GA_OK;
    if (!d_state || !d_digests) { set_error("ga_k_sha256_final: NULL argument"); return GA_ERR_ARG; }
    if (reinterpret_cast<uintptr_t>(d_digests) & 3u) { set_error("ga_k_sha256_final: d_digests must be 4-byte aligned"); return GA_ERR_ARG; }
    sha256_final_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_state, d_stream_ids, n,
                                                                                      d_digests, d_lengths);
    GA_CHECK_LAUNCH("sha256_final_kernel");
    return GA_OK;
}

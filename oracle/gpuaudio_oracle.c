/*
 * gpuaudio_oracle.c — CPU restatement of the loqa-hub relay-audio ingest path.
 * TEST INFRASTRUCTURE ONLY (see gpuaudio_oracle.h for the rules and parity status).
 *
 * Every function cites the reference lines it follows (relative to /root/reference).
 * Numeric-semantics rules are SURVEY.md Appendix B; build with -ffp-contract=off so
 * the compiler cannot fuse the f32 multiply into the f64 add.
 */
#include "gpuaudio_oracle.h"
#include "../include/gpuaudio_taps.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ A1 */
/* internal/grpc/audio_service.go:1048-1101 */
size_t ora_bytes_to_float32(const uint8_t *data, size_t len, float *out, ora_stats *st)
{
    if (st) memset(st, 0, sizeof *st);
    if (len == 0) return 0;                      /* :1050-1053 empty -> []float32{} */
    size_t data_len = len;
    if (data_len % 2 != 0) data_len -= 1;        /* :1056-1063 drop the odd last byte */
    size_t n = data_len / 2;                     /* :1066 */
    uint32_t nonzero = 0;
    float peak = 0.0f;
    for (size_t i = 0; i < n; i++) {             /* :1070 */
        /* :1076 little-endian two's-complement int16 */
        int16_t val = (int16_t)((uint16_t)data[2 * i] | ((uint16_t)data[2 * i + 1] << 8));
        /* :1078 float32(val) / 32767.0 — an IEEE float32 division, RN-even */
        float q = (float)val / 32767.0f;   /* -ffp-contract=off, SSE: no excess precision */
        out[i] = q;
        if (q != 0.0f) nonzero++;                /* :1081-1083 */
        float a = q;                             /* :1084-1090 */
        if (a < 0) a = -a;
        if (a > peak) peak = a;
    }
    if (st) {
        st->n_samples = (uint32_t)n;
        st->nonzero = nonzero;
        st->peak = peak;
    }
    return n;
}

/* ------------------------------------------------------------------ A2 */
/* internal/grpc/audio_service.go:818-852 */
double ora_signal_strength(const float *samples, size_t n, ora_stats *st)
{
    if (st) memset(st, 0, sizeof *st);
    if (n == 0) return 0.0;                      /* :819-822 */
    double sum = 0.0;
    uint32_t nonzero = 0;
    float peak = 0.0f;
    for (size_t i = 0; i < n; i++) {             /* :827 */
        float s = samples[i];
        float sq = s * s;                        /* :828 f32 product rounded to f32 ... */
        sum += (double)sq;                       /*      ... then widened, f64 add in index order */
        if (s != 0.0f) nonzero++;                /* :829-831 */
        float a = s;                             /* :832-838 */
        if (a < 0) a = -a;
        if (a > peak) peak = a;
    }
    double rms = sqrt(sum / (double)n);          /* :841 */
    if (st) {
        st->n_samples = (uint32_t)n;
        st->nonzero = nonzero;
        st->peak = peak;
        st->sumsq = sum;
        st->rms = rms;
    }
    return rms;
}

/* ------------------------------------------------------------------ A3 */
/* internal/grpc/audio_service.go:517-571 (selection + status update) */
int ora_select_winner(uint32_t n_members, const double *scores, const uint8_t *status_in,
                      uint8_t *status_out)
{
    int winner = -1;
    double max_strength = 0.0;                   /* :519 zero value of float64 */
    for (uint32_t i = 0; i < n_members; i++) {   /* :521 (join order instead of map order) */
        if (status_in[i] != ORA_RELAY_CONTENDING) continue;   /* :522-524 */
        if (scores[i] > max_strength) {          /* :535 strict > */
            max_strength = scores[i];
            winner = (int)i;
        }
    }
    if (winner < 0 && n_members > 0) winner = 0; /* :542-547 first key of the range */
    if (status_out) {
        for (uint32_t i = 0; i < n_members; i++) /* :553-571 every relay in the window */
            status_out[i] = ((int)i == winner) ? ORA_RELAY_WINNER : ORA_RELAY_CANCELLED;
    }
    return winner;
}

int ora_arbitrate(uint32_t n_members, const float *const *wake, const size_t *wake_n,
                  const uint8_t *status_in, double *scores, uint8_t *status_out)
{
    for (uint32_t i = 0; i < n_members; i++) {
        if (status_in[i] != ORA_RELAY_CONTENDING) continue;   /* :522-524 */
        scores[i] = ora_signal_strength(wake[i], wake_n[i], NULL);  /* :527-528 */
    }
    return ora_select_winner(n_members, scores, status_in, status_out);
}

/* ------------------------------------------------------------------ N3: SHA-256 */
/* FIPS 180-4 §4.1.2, §4.2.2, §5.1.1, §5.3.3, §6.2.  Stands in for Go's
 * crypto/sha256 (no call site left in the reference). */
static const uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2
};

static inline uint32_t rotr32(uint32_t x, unsigned n) { return (x >> n) | (x << (32 - n)); }

static void sha256_block(uint32_t h[8], const uint8_t *p)
{
    uint32_t w[64];
    for (int t = 0; t < 16; t++)
        w[t] = ((uint32_t)p[4 * t] << 24) | ((uint32_t)p[4 * t + 1] << 16) |
               ((uint32_t)p[4 * t + 2] << 8) | (uint32_t)p[4 * t + 3];
    for (int t = 16; t < 64; t++) {
        uint32_t s0 = rotr32(w[t - 15], 7) ^ rotr32(w[t - 15], 18) ^ (w[t - 15] >> 3);
        uint32_t s1 = rotr32(w[t - 2], 17) ^ rotr32(w[t - 2], 19) ^ (w[t - 2] >> 10);
        w[t] = s1 + w[t - 7] + s0 + w[t - 16];
    }
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int t = 0; t < 64; t++) {
        uint32_t S1 = rotr32(e, 6) ^ rotr32(e, 11) ^ rotr32(e, 25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = hh + S1 + ch + K256[t] + w[t];
        uint32_t S0 = rotr32(a, 2) ^ rotr32(a, 13) ^ rotr32(a, 22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

void ora_sha256_init(ora_sha256_ctx *c)
{
    static const uint32_t iv[8] = { 0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a,
                                    0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19 };
    memset(c, 0, sizeof *c);
    memcpy(c->h, iv, sizeof iv);
}

void ora_sha256_update(ora_sha256_ctx *c, const uint8_t *data, size_t len)
{
    c->total_len += len;
    if (c->tail_len) {
        size_t take = 64 - c->tail_len;
        if (take > len) take = len;
        memcpy(c->tail + c->tail_len, data, take);
        c->tail_len += (uint32_t)take;
        data += take; len -= take;
        if (c->tail_len == 64) { sha256_block(c->h, c->tail); c->tail_len = 0; }
    }
    while (len >= 64) { sha256_block(c->h, data); data += 64; len -= 64; }
    if (len) { memcpy(c->tail, data, len); c->tail_len = (uint32_t)len; }
}

void ora_sha256_final(const ora_sha256_ctx *c, uint8_t out[32])
{
    uint32_t h[8];
    uint8_t blk[128];
    memcpy(h, c->h, sizeof h);
    memset(blk, 0, sizeof blk);
    memcpy(blk, c->tail, c->tail_len);
    blk[c->tail_len] = 0x80;
    size_t nblk = (c->tail_len + 1 + 8 <= 64) ? 1 : 2;
    uint64_t bits = c->total_len * 8;
    for (int i = 0; i < 8; i++) blk[nblk * 64 - 1 - i] = (uint8_t)(bits >> (8 * i));
    sha256_block(h, blk);
    if (nblk == 2) sha256_block(h, blk + 64);
    for (int i = 0; i < 8; i++) {
        out[4 * i] = (uint8_t)(h[i] >> 24); out[4 * i + 1] = (uint8_t)(h[i] >> 16);
        out[4 * i + 2] = (uint8_t)(h[i] >> 8); out[4 * i + 3] = (uint8_t)h[i];
    }
}

void ora_sha256(const uint8_t *data, size_t len, uint8_t out[32])
{
    ora_sha256_ctx c;
    ora_sha256_init(&c);
    ora_sha256_update(&c, data, len);
    ora_sha256_final(&c, out);
}

/* ------------------------------------------------------------------ N1: framing */
size_t ora_frame_features(const float *samples, size_t n, uint32_t frame_len,
                          double vad_threshold, ora_frame *out)
{
    size_t nf = 0;
    if (frame_len == 0) return 0;
    for (size_t off = 0; off < n; off += frame_len, nf++) {
        size_t m = n - off < frame_len ? n - off : frame_len;
        ora_stats st;
        ora_signal_strength(samples + off, m, &st);   /* A2 semantics per frame */
        out[nf].sumsq = st.sumsq;
        out[nf].rms = st.rms;
        out[nf].n_samples = st.n_samples;
        out[nf].nonzero = st.nonzero;
        out[nf].peak = st.peak;
        out[nf].vad = st.rms > vad_threshold ? 1u : 0u;
    }
    return nf;
}

/* ------------------------------------------------------------------ N2: ÷3 decimator */
static const uint32_t k_tap_bits[GA_RESAMPLE_TAPS] = GA_RESAMPLE_TAP_BITS;
static float k_taps[GA_RESAMPLE_TAPS];
static int k_taps_ready = 0;

const float *ora_resample_taps(void)
{
    if (!k_taps_ready) {
        memcpy(k_taps, k_tap_bits, sizeof k_taps);
        k_taps_ready = 1;
    }
    return k_taps;
}

void ora_resample_init(ora_resample_state *s) { memset(s, 0, sizeof *s); }

size_t ora_resample_48k_stereo(ora_resample_state *s, const uint8_t *data, size_t len,
                               float *out)
{
    enum { T = ORA_RESAMPLE_TAPS };
    const float *taps = ora_resample_taps();
    size_t nfr = len / 4;                         /* whole stereo frames only */
    float *mono = (float *)malloc((nfr + T - 1) * sizeof(float));
    memcpy(mono, s->hist, (T - 1) * sizeof(float));
    for (size_t j = 0; j < nfr; j++) {
        int16_t l = (int16_t)((uint16_t)data[4 * j] | ((uint16_t)data[4 * j + 1] << 8));
        int16_t r = (int16_t)((uint16_t)data[4 * j + 2] | ((uint16_t)data[4 * j + 3] << 8));
        volatile float fl = (float)l / 32767.0f;  /* A1's divide, per channel */
        volatile float fr = (float)r / 32767.0f;
        volatile float sum = fl + fr;
        mono[T - 1 + j] = sum * 0.5f;
    }
    uint64_t n0 = s->n_frames, n1 = n0 + nfr;
    size_t nout = 0;
    /* outputs M with n0 <= 3M+2 < n1 */
    uint64_t m0 = (n0 + 0) / 3;                   /* smallest M with 3M+2 >= n0 */
    while (3 * m0 + 2 < n0) m0++;
    for (uint64_t M = m0; 3 * M + 2 < n1; M++) {
        /* local index of absolute frame a is a - n0 + (T-1) */
        size_t top = (size_t)(3 * M + 2 - n0) + (T - 1);
        float acc = 0.0f;
        for (int k = 0; k < T; k++) acc = fmaf(taps[k], mono[top - k], acc);
        out[nout++] = acc;
    }
    /* carry the last T-1 mono frames (absolute n1-47 .. n1-1) */
    memmove(s->hist, mono + nfr, (T - 1) * sizeof(float));
    s->n_frames = n1;
    free(mono);
    return nout;
}

/* ------------------------------------------------------------------ WAV */
/* internal/llm/stt_client.go:365-398 */
static void put_le32(uint8_t *p, uint32_t v)
{
    p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

size_t ora_float32_to_wav(const float *samples, size_t n, uint32_t sample_rate, uint8_t *out)
{
    uint32_t data_size = (uint32_t)(n * 4);       /* :369 */
    uint32_t file_size = 36 + data_size;          /* :370 */
    memcpy(out, "RIFF", 4);  put_le32(out + 4, file_size);           /* :375-376 */
    memcpy(out + 8, "WAVE", 4);                                      /* :377 */
    memcpy(out + 12, "fmt ", 4);  put_le32(out + 16, 16);            /* :378-379 */
    out[20] = 3; out[21] = 0;                                        /* :380 IEEE float */
    out[22] = 1; out[23] = 0;                                        /* :381 mono */
    put_le32(out + 24, sample_rate);                                 /* :382 */
    put_le32(out + 28, sample_rate * 4);                             /* :383-384 byte rate */
    out[32] = 4; out[33] = 0;                                        /* :385 block align */
    out[34] = 32; out[35] = 0;                                       /* :386 bits/sample */
    memcpy(out + 36, "data", 4);  put_le32(out + 40, data_size);     /* :387-388 */
    for (size_t i = 0; i < n; i++) {                                 /* :391-396 */
        uint32_t bits;
        memcpy(&bits, &samples[i], 4);
        put_le32(out + 44 + 4 * i, bits);
    }
    return 44 + (size_t)data_size;
}

/*
 * cpu_baseline.c — the timed CPU baseline: the reference's per-chunk work, restated.
 * TEST / BENCH INFRASTRUCTURE ONLY (see gpuaudio_oracle.h).
 *
 * One worker thread plays the role of the goroutines grpc-go would run
 * (one per relay stream, internal/grpc/audio_service.go:926): it takes a whole
 * collision group and walks each relay's chunks in arrival order, doing per chunk what
 * StreamAudio does (:939-1005): bytesToFloat32Array into a FRESH allocation (:1066),
 * append to the stream's wake-word / speech buffer (:965, :1002, Go's amortised
 * append), plus the SHA-256 fingerprint of the chunk bytes (the voice-event
 * fingerprint the hub's docs describe, README.md:50 / API.md:170).  When all
 * streams of a collision window are in, the window is arbitrated as
 * performArbitration does (:505-551): calculateSignalStrength over each relay's
 * wake-word buffer, strict-> running max.  No logging — the real reference also
 * writes 2-3 zap lines per chunk (:956, :1093, :843), so this flatters it.
 *
 * What a goroutine is given, and what it is spared, so that the figure is the most
 * favourable honest reading of the Go path and scales with the host's cores:
 *   - input layout: a goroutine receives each chunk as its own contiguous []byte from
 *     gRPC.  ORA_LAYOUT_STREAM_MAJOR hands a worker its stream's chunks back to back;
 *     ORA_LAYOUT_TIME_MAJOR is the GPU slab's order (all streams' chunk t, then t+1),
 *     where consecutive chunks of one stream lie n_streams * msg_bytes apart — kept for
 *     the parity tests, not for timing.
 *   - heap: the growing wake-word / speech buffers come from a per-thread bump pool that
 *     is reset per stream — Go's allocator hands out spans without a system call, and
 *     glibc's mmap/munmap for every >128 KB growth step would serialise the threads in
 *     the kernel.  The re-copy on growth (what append costs) is still done.
 *   - SHA-256: OpenSSL with the digest implementation resolved ONCE per thread
 *     (ORA_SHA_OPENSSL_EVP: EVP_MD_fetch + EVP_DigestInit_ex2) or its low-level one-shot
 *     functions (ORA_SHA_OPENSSL_LOW: SHA256_Init/Update/Final, no provider lookup at
 *     all); both use SHA-NI where the CPU has it, like Go's crypto/sha256.  An implicit
 *     fetch per message (EVP_DigestInit_ex(ctx, EVP_sha256(), NULL)) takes shared
 *     reference counts and does not scale.
 *   - 48 kHz stereo input (BASELINE configs[3]): the decimator of gpuaudio_oracle.c
 *     restated without its per-call malloc, four outputs interleaved, compiled for the
 *     host's FMA unit — bit-identical to ora_resample_48k_stereo (tests/test_oracle_golden.py).
 */
#define _GNU_SOURCE
#include "gpuaudio_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#ifdef ORA_HAVE_OPENSSL
#define OPENSSL_SUPPRESS_DEPRECATED
#include <openssl/evp.h>
#include <openssl/sha.h>
#endif

int ora_have_openssl(void)
{
#ifdef ORA_HAVE_OPENSSL
    return 1;
#else
    return 0;
#endif
}

/* ------------------------------------------------------------------ per-thread heap */
typedef struct {
    uint8_t *base;
    size_t used, cap;
} pool;

static void *pool_alloc(pool *p, size_t bytes)
{
    bytes = (bytes + 63u) & ~(size_t)63u;
    if (p->used + bytes > p->cap) return NULL;   /* sized up front for the growth series; cannot happen */
    void *r = p->base + p->used;
    p->used += bytes;
    return r;
}

/* bytes a buffer that ends up holding `final` floats asks from the pool over its life: the growth
 * steps form a geometric series with ratio 1.25, so <= 5x the final capacity, plus the small steps */
static size_t pool_need(size_t final_floats)
{
    return (final_floats * 5 + 4096) * sizeof(float) * 5 / 4;
}

typedef struct {
    float *p;
    size_t len, cap;
} fbuf;

/* Go's append growth: double until 256 elements, then (cap + 3*256)/4 steps; the
 * exact policy does not matter for parity — only that growth re-copies. */
static void fbuf_append(fbuf *b, const float *src, size_t n, pool *pl)
{
    if (b->len + n > b->cap) {
        size_t nc = b->cap ? b->cap : 64;
        while (nc < b->len + n) nc = nc < 256 ? nc * 2 : nc + (nc + 3 * 256) / 4;
        float *np = (float *)pool_alloc(pl, nc * sizeof(float));
        if (!np) abort();
        if (b->len) memcpy(np, b->p, b->len * sizeof(float));
        b->p = np; b->cap = nc;           /* the old block is garbage (the GC's job in Go) */
    }
    memcpy(b->p + b->len, src, n * sizeof(float));
    b->len += n;
}

/* ------------------------------------------------------------------ 48 kHz leg */
/* ora_resample_48k_stereo without its allocation: mono[] is caller scratch of at least
 * (T-1) + nfr floats.  Same arithmetic in the same order. */
#if defined(__x86_64__)
__attribute__((target("fma")))
#endif
static size_t resample_chunk(ora_resample_state *s, const uint8_t *data, size_t len, float *out, float *mono)
{
    enum { T = ORA_RESAMPLE_TAPS };
    const float *taps = ora_resample_taps();
    const size_t nfr = len / 4;
    memcpy(mono, s->hist, (T - 1) * sizeof(float));
    for (size_t j = 0; j < nfr; j++) {
        int16_t l = (int16_t)((uint16_t)data[4 * j] | ((uint16_t)data[4 * j + 1] << 8));
        int16_t r = (int16_t)((uint16_t)data[4 * j + 2] | ((uint16_t)data[4 * j + 3] << 8));
        float fl = (float)l / 32767.0f, fr = (float)r / 32767.0f;
        float sum = fl + fr;
        mono[T - 1 + j] = sum * 0.5f;
    }
    const uint64_t n0 = s->n_frames, n1 = n0 + nfr;
    uint64_t m0 = n0 / 3;
    while (3 * m0 + 2 < n0) m0++;
    size_t nout = 0;
    uint64_t M = m0;
    for (; 3 * (M + 3) + 2 < n1; M += 4) {        /* four independent accumulation chains */
        const float *t0 = mono + (size_t)(3 * M + 2 - n0) + (T - 1);
        float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
        for (int k = 0; k < T; k++) {
            const float h = taps[k];
            a0 = __builtin_fmaf(h, t0[0 - k], a0);
            a1 = __builtin_fmaf(h, t0[3 - k], a1);
            a2 = __builtin_fmaf(h, t0[6 - k], a2);
            a3 = __builtin_fmaf(h, t0[9 - k], a3);
        }
        out[nout++] = a0; out[nout++] = a1; out[nout++] = a2; out[nout++] = a3;
    }
    for (; 3 * M + 2 < n1; M++) {
        const float *t0 = mono + (size_t)(3 * M + 2 - n0) + (T - 1);
        float acc = 0.0f;
        for (int k = 0; k < T; k++) acc = __builtin_fmaf(taps[k], t0[-k], acc);
        out[nout++] = acc;
    }
    memmove(s->hist, mono + nfr, (T - 1) * sizeof(float));
    s->n_frames = n1;
    return nout;
}

/* ------------------------------------------------------------------ the job */
typedef struct {
    ora_baseline_cfg c;
    const uint8_t *pcm;
    uint8_t *digests;
    int32_t *winners;
    double *scores;
    int fast_fma;
    atomic_uint next_group;
} job;

static void *worker(void *arg)
{
    job *j = (job *)arg;
    const ora_baseline_cfg *c = &j->c;
    const uint32_t gs = c->group_size, n_groups = c->n_streams / gs;
    const int fmt48 = c->format == ORA_FMT_STEREO48K;
    const size_t max_samples = fmt48 ? c->msg_bytes / 4 / 3 + 2 : c->msg_bytes / 2;
    /* wake-word buffers live until the group is arbitrated, speech buffers until the stream ends */
    pool wpool, spool;
    wpool.cap = gs * pool_need((size_t)c->wake_msgs * max_samples); wpool.used = 0;
    spool.cap = pool_need((size_t)c->n_msgs * max_samples); spool.used = 0;
    wpool.base = (uint8_t *)malloc(wpool.cap); spool.base = (uint8_t *)malloc(spool.cap);
    fbuf *wake = (fbuf *)calloc(gs, sizeof(fbuf));
    fbuf *speech = (fbuf *)calloc(gs, sizeof(fbuf));
    double *sc = (double *)calloc(gs, sizeof(double));
    uint8_t *st_in = (uint8_t *)malloc(gs);
    uint8_t *st_out = (uint8_t *)malloc(gs);
    const float **wp = (const float **)malloc(gs * sizeof(float *));
    size_t *wn = (size_t *)malloc(gs * sizeof(size_t));
    float *mono = fmt48 ? (float *)malloc((c->msg_bytes / 4 + ORA_RESAMPLE_TAPS) * sizeof(float)) : NULL;
#ifdef ORA_HAVE_OPENSSL
    EVP_MD_CTX *md = NULL;
    EVP_MD *alg = NULL;
    if (c->sha_impl == ORA_SHA_OPENSSL_EVP) {
        md = EVP_MD_CTX_new();
        alg = EVP_MD_fetch(NULL, "SHA256", NULL);      /* once per thread, not per message */
    }
#endif
    uint8_t dg[32];
    for (;;) {
        uint32_t g = atomic_fetch_add(&j->next_group, 1);
        if (g >= n_groups) break;
        wpool.used = 0;                                     /* the previous group's buffers are garbage now */
        for (uint32_t r = 0; r < gs; r++) {
            const uint32_t s = g * gs + r;
            memset(&wake[r], 0, sizeof(fbuf)); memset(&speech[r], 0, sizeof(fbuf));
            spool.used = 0;
            ora_resample_state rs;
            if (fmt48) ora_resample_init(&rs);
            for (uint32_t m = 0; m < c->n_msgs; m++) {
                const uint64_t idx = c->layout == ORA_LAYOUT_STREAM_MAJOR ? (uint64_t)s * c->n_msgs + m
                                                                          : (uint64_t)m * c->n_streams + s;
                const uint8_t *chunk = j->pcm + idx * c->msg_bytes;
                /* :964/:999 bytesToFloat32Array -> fresh slice */
                float *fresh = (float *)malloc((max_samples ? max_samples : 1) * sizeof(float));
                size_t n;
                if (!fmt48) n = ora_bytes_to_float32(chunk, c->msg_bytes, fresh, NULL);
                else if (j->fast_fma) n = resample_chunk(&rs, chunk, c->msg_bytes, fresh, mono);
                else n = ora_resample_48k_stereo(&rs, chunk, c->msg_bytes, fresh);
                /* :965 / :1002 append */
                if (m < c->wake_msgs) fbuf_append(&wake[r], fresh, n, &wpool);
                else fbuf_append(&speech[r], fresh, n, &spool);
                free(fresh);   /* the GC's job in Go */
                /* per-message fingerprint */
                uint8_t *dst = j->digests ? j->digests + ((uint64_t)m * c->n_streams + s) * 32 : dg;
#ifdef ORA_HAVE_OPENSSL
                if (md) {
                    unsigned int dl = 32;
                    EVP_DigestInit_ex2(md, alg, NULL);
                    EVP_DigestUpdate(md, chunk, c->msg_bytes);
                    EVP_DigestFinal_ex(md, dst, &dl);
                } else if (c->sha_impl == ORA_SHA_OPENSSL_LOW) {
                    SHA256_CTX sc256;
                    SHA256_Init(&sc256);
                    SHA256_Update(&sc256, chunk, c->msg_bytes);
                    SHA256_Final(dst, &sc256);
                } else
#endif
                    ora_sha256(chunk, c->msg_bytes, dst);
            }
            st_in[r] = ORA_RELAY_CONTENDING;
            wp[r] = wake[r].p; wn[r] = wake[r].len;
        }
        int w = ora_arbitrate(gs, wp, wn, st_in, sc, st_out);   /* :505-551 */
        if (j->winners) j->winners[g] = w;
        if (j->scores) memcpy(j->scores + (size_t)g * gs, sc, gs * sizeof(double));
    }
    free(wpool.base); free(spool.base);
    free(wake); free(speech); free(sc); free(st_in); free(st_out); free(wp); free(wn); free(mono);
#ifdef ORA_HAVE_OPENSSL
    if (md) EVP_MD_CTX_free(md);
    if (alg) EVP_MD_free(alg);
#endif
    return NULL;
}

double ora_cpu_baseline_run(const ora_baseline_cfg *cfg, const uint8_t *pcm, uint8_t *digests_out,
                            int32_t *winners_out, double *scores_out)
{
    job j;
    j.c = *cfg;
    if (j.c.group_size == 0) j.c.group_size = 1;
    if (j.c.n_threads == 0) j.c.n_threads = 1;
#ifndef ORA_HAVE_OPENSSL
    j.c.sha_impl = ORA_SHA_PORT;
#endif
    j.pcm = pcm; j.digests = digests_out; j.winners = winners_out; j.scores = scores_out;
#if defined(__x86_64__)
    j.fast_fma = __builtin_cpu_supports("fma") ? 1 : 0;
#else
    j.fast_fma = 0;
#endif
    atomic_init(&j.next_group, 0);
    pthread_t *th = (pthread_t *)malloc(j.c.n_threads * sizeof(pthread_t));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (uint32_t i = 1; i < j.c.n_threads; i++) pthread_create(&th[i], NULL, worker, &j);
    worker(&j);
    for (uint32_t i = 1; i < j.c.n_threads; i++) pthread_join(th[i], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

double ora_cpu_baseline(const uint8_t *pcm, uint32_t n_streams, uint32_t n_msgs,
                        uint32_t msg_bytes, uint32_t wake_msgs, uint32_t group_size,
                        uint32_t n_threads, int use_openssl,
                        uint8_t *digests_out, int32_t *winners_out, double *scores_out)
{
    ora_baseline_cfg c;
    memset(&c, 0, sizeof c);
    c.n_streams = n_streams; c.n_msgs = n_msgs; c.msg_bytes = msg_bytes; c.wake_msgs = wake_msgs;
    c.group_size = group_size; c.n_threads = n_threads;
    c.format = ORA_FMT_MONO16K; c.layout = ORA_LAYOUT_TIME_MAJOR;
    c.sha_impl = use_openssl ? ORA_SHA_OPENSSL_EVP : ORA_SHA_PORT;
    return ora_cpu_baseline_run(&c, pcm, digests_out, winners_out, scores_out);
}

/* ------------------------------------------------------------------ host memory ceiling */
/* n_threads workers each memcpy their own `bytes`-sized buffer `reps` times: the bandwidth the
 * host gives plain copies, printed beside the baseline so that a scaling shortfall can be told
 * from a memory-bound one.  Returns GB/s of bytes copied (read + write counted once each). */
typedef struct { uint8_t *a, *b; size_t bytes; uint32_t reps; } mcjob;
static void *mc_worker(void *arg)
{
    mcjob *m = (mcjob *)arg;
    for (uint32_t i = 0; i < m->reps; i++) {
        memcpy(m->b, m->a, m->bytes);
        __asm__ volatile("" ::"r"(m->b) : "memory");
    }
    return NULL;
}
double ora_memcpy_gbs(uint32_t n_threads, size_t bytes_per_thread, uint32_t reps)
{
    if (n_threads == 0) n_threads = 1;
    mcjob *m = (mcjob *)calloc(n_threads, sizeof(mcjob));
    pthread_t *th = (pthread_t *)malloc(n_threads * sizeof(pthread_t));
    for (uint32_t i = 0; i < n_threads; i++) {          /* pages touched outside the clock */
        m[i].a = (uint8_t *)malloc(bytes_per_thread); m[i].b = (uint8_t *)malloc(bytes_per_thread);
        m[i].bytes = bytes_per_thread; m[i].reps = reps;
        memset(m[i].a, 1, bytes_per_thread); memset(m[i].b, 2, bytes_per_thread);
    }
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (uint32_t i = 0; i < n_threads; i++) pthread_create(&th[i], NULL, mc_worker, &m[i]);
    for (uint32_t i = 0; i < n_threads; i++) pthread_join(th[i], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    for (uint32_t i = 0; i < n_threads; i++) { free(m[i].a); free(m[i].b); }
    free(th); free(m);
    double dt = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    return 2.0 * (double)bytes_per_thread * reps * n_threads / dt / 1e9;
}

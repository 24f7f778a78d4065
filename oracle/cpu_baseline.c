/*
 * cpu_baseline.c — the timed CPU baseline: the reference's per-chunk work, restated.
 * TEST / BENCH INFRASTRUCTURE ONLY (see gpuaudio_oracle.h).
 *
 * One worker thread plays the role of the goroutines grpc-go would run
 * (one per relay stream, internal/grpc/audio_service.go:926): it takes a whole
 * stream and walks its chunks in arrival order, doing per chunk what StreamAudio
 * does (:939-1005): bytesToFloat32Array into a FRESH allocation (:1066), append
 * to the stream's wake-word / speech buffer (:965, :1002, Go's amortised-doubling
 * append), plus the SHA-256 fingerprint of the chunk bytes (the voice-event
 * fingerprint the hub's docs describe, README.md:50 / API.md:170).  When all
 * streams of a collision window are in, the window is arbitrated as
 * performArbitration does (:505-551): calculateSignalStrength over each relay's
 * wake-word buffer, strict-> running max.  No logging — the real reference also
 * writes 2-3 zap lines per chunk (:956, :1093, :843), so this flatters it.
 */
#define _GNU_SOURCE
#include "gpuaudio_oracle.h"

#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#ifdef ORA_HAVE_OPENSSL
#include <openssl/evp.h>
#endif

int ora_have_openssl(void)
{
#ifdef ORA_HAVE_OPENSSL
    return 1;
#else
    return 0;
#endif
}

typedef struct {
    float *p;
    size_t len, cap;
} fbuf;

/* Go's append growth: double until 256 elements, then (cap + 3*256)/4 steps; the
 * exact policy does not matter for parity — only that growth re-copies. */
static void fbuf_append(fbuf *b, const float *src, size_t n)
{
    if (b->len + n > b->cap) {
        size_t nc = b->cap ? b->cap : 64;
        while (nc < b->len + n) nc = nc < 256 ? nc * 2 : nc + (nc + 3 * 256) / 4;
        float *np = (float *)malloc(nc * sizeof(float));
        if (b->len) memcpy(np, b->p, b->len * sizeof(float));
        free(b->p);
        b->p = np; b->cap = nc;
    }
    memcpy(b->p + b->len, src, n * sizeof(float));
    b->len += n;
}

typedef struct {
    const uint8_t *pcm;
    uint32_t n_streams, n_msgs, msg_bytes, wake_msgs, group_size;
    int use_openssl;
    uint8_t *digests;
    int32_t *winners;
    double *scores;
    atomic_uint next_group;
} job;

static void *worker(void *arg)
{
    job *j = (job *)arg;
    uint32_t n_groups = j->n_streams / j->group_size;
    size_t max_samples = j->msg_bytes / 2;
    fbuf *wake = (fbuf *)calloc(j->group_size, sizeof(fbuf));
    fbuf *speech = (fbuf *)calloc(j->group_size, sizeof(fbuf));
    double *sc = (double *)calloc(j->group_size, sizeof(double));
    uint8_t *st_in = (uint8_t *)malloc(j->group_size);
    uint8_t *st_out = (uint8_t *)malloc(j->group_size);
    const float **wp = (const float **)malloc(j->group_size * sizeof(float *));
    size_t *wn = (size_t *)malloc(j->group_size * sizeof(size_t));
#ifdef ORA_HAVE_OPENSSL
    EVP_MD_CTX *md = j->use_openssl ? EVP_MD_CTX_new() : NULL;
#endif
    uint8_t dg[32];
    for (;;) {
        uint32_t g = atomic_fetch_add(&j->next_group, 1);
        if (g >= n_groups) break;
        for (uint32_t r = 0; r < j->group_size; r++) {
            uint32_t s = g * j->group_size + r;
            wake[r].len = 0; speech[r].len = 0;
            for (uint32_t m = 0; m < j->n_msgs; m++) {
                const uint8_t *chunk = j->pcm + ((uint64_t)m * j->n_streams + s) * j->msg_bytes;
                /* :964/:999 bytesToFloat32Array -> fresh slice */
                float *fresh = (float *)malloc((max_samples ? max_samples : 1) * sizeof(float));
                size_t n = ora_bytes_to_float32(chunk, j->msg_bytes, fresh, NULL);
                /* :965 / :1002 append */
                fbuf_append(m < j->wake_msgs ? &wake[r] : &speech[r], fresh, n);
                free(fresh);   /* the GC's job in Go */
                /* per-message fingerprint */
                uint8_t *dst = j->digests ? j->digests + ((uint64_t)m * j->n_streams + s) * 32 : dg;
#ifdef ORA_HAVE_OPENSSL
                if (md) {
                    unsigned int dl = 32;
                    EVP_DigestInit_ex(md, EVP_sha256(), NULL);
                    EVP_DigestUpdate(md, chunk, j->msg_bytes);
                    EVP_DigestFinal_ex(md, dst, &dl);
                } else
#endif
                    ora_sha256(chunk, j->msg_bytes, dst);
            }
            st_in[r] = ORA_RELAY_CONTENDING;
            wp[r] = wake[r].p; wn[r] = wake[r].len;
        }
        int w = ora_arbitrate(j->group_size, wp, wn, st_in, sc, st_out);   /* :505-551 */
        if (j->winners) j->winners[g] = w;
        if (j->scores) memcpy(j->scores + (size_t)g * j->group_size, sc, j->group_size * sizeof(double));
    }
    for (uint32_t r = 0; r < j->group_size; r++) { free(wake[r].p); free(speech[r].p); }
    free(wake); free(speech); free(sc); free(st_in); free(st_out); free(wp); free(wn);
#ifdef ORA_HAVE_OPENSSL
    if (md) EVP_MD_CTX_free(md);
#endif
    return NULL;
}

double ora_cpu_baseline(const uint8_t *pcm, uint32_t n_streams, uint32_t n_msgs,
                        uint32_t msg_bytes, uint32_t wake_msgs, uint32_t group_size,
                        uint32_t n_threads, int use_openssl,
                        uint8_t *digests_out, int32_t *winners_out, double *scores_out)
{
    if (group_size == 0) group_size = 1;
    if (n_threads == 0) n_threads = 1;
    job j;
    j.pcm = pcm; j.n_streams = n_streams; j.n_msgs = n_msgs; j.msg_bytes = msg_bytes;
    j.wake_msgs = wake_msgs; j.group_size = group_size; j.use_openssl = use_openssl;
    j.digests = digests_out; j.winners = winners_out; j.scores = scores_out;
    atomic_init(&j.next_group, 0);
    pthread_t *th = (pthread_t *)malloc(n_threads * sizeof(pthread_t));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (uint32_t i = 1; i < n_threads; i++) pthread_create(&th[i], NULL, worker, &j);
    worker(&j);
    for (uint32_t i = 1; i < n_threads; i++) pthread_join(th[i], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

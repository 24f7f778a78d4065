/*
 * synth_oracle.c — the synthetic relay PCM of SURVEY.md §8d, for the CPU arms.
 * TEST / BENCH INFRASTRUCTURE ONLY (see gpuaudio_oracle.h).
 *
 * Integer-only, SplitMix64-keyed: per collision group three parabolic-sine tones at
 * group-dependent pitches under a trapezoid envelope (peak about +-12,000) plus +-200 noise;
 * relay r hears it at gain 0.9^r (Q15) plus +-64(r+1) noise of its own, clamped to int16.
 * The product has its own generator for the device (csrc/synth.h); tests hold the two and the
 * numpy statement (loqa_hub_b200/synth.py) byte-identical.  This copy exists so that
 * `bench.py --impl reference` never has to load the product's library.
 */
#define _GNU_SOURCE
#include "gpuaudio_oracle.h"

#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>

static uint64_t sm64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

/* parabolic sine: phase in [0, 65536), amplitude 4000 */
static int64_t tone(uint32_t phase)
{
    int64_t x = phase & 0x7FFFu;
    int64_t y = (x * (32768 - x)) >> 13;
    int64_t a = (y * 4000) >> 15;
    return (phase & 0x8000u) ? -a : a;
}

typedef struct {
    uint64_t gk, rk;
    uint32_t inc0, inc1, inc2;
    int64_t gain, amp;
} voice;

static void voice_init(voice *v, uint64_t seed, uint32_t group, uint32_t relay)
{
    static const int64_t g[8] = {32768, 29491, 26542, 23888, 21499, 19349, 17414, 15673};
    v->gk = sm64(seed ^ (0x10CA0B20ull + group));
    v->inc0 = 400u + (uint32_t)(v->gk % 1200u);
    v->inc1 = 1600u + (uint32_t)((v->gk >> 16) % 1600u);
    v->inc2 = 3200u + (uint32_t)((v->gk >> 32) % 2400u);
    v->rk = sm64(v->gk ^ (0x9E3779B97F4A7C15ull * (relay + 1)));
    v->gain = g[relay & 7u];
    v->amp = 64 * ((int64_t)relay + 1);
}

static int16_t voice_sample(const voice *v, uint32_t n, uint32_t utt)
{
    int64_t t = tone((v->inc0 * n) & 0xFFFFu) + tone((v->inc1 * n) & 0xFFFFu) + tone((v->inc2 * n) & 0xFFFFu);
    uint32_t edge = n < utt - 1 - n ? n : utt - 1 - n;
    if (n >= utt) edge = 0;
    int64_t env = (int64_t)(edge < 4000u ? edge : 4000u) * 8;
    int64_t u = ((t * env) >> 15) + ((int64_t)(sm64(v->gk + n) % 401u) - 200);
    int64_t noise = (int64_t)(sm64(v->rk + n) % (uint64_t)(2 * v->amp + 1)) - v->amp;
    int64_t x = ((u * v->gain) >> 15) + noise;
    if (x > 32767) x = 32767;
    if (x < -32768) x = -32768;
    return (int16_t)x;
}

int16_t ora_synth_sample(uint64_t seed, uint32_t group, uint32_t relay, uint32_t n, uint32_t utt_samples)
{
    voice v;
    voice_init(&v, seed, group, relay);
    return voice_sample(&v, n, utt_samples);
}

typedef struct {
    uint8_t *out;
    uint64_t seed;
    uint32_t group_base, n_streams, group_size, n_msgs, msg_samples, layout;
    int stereo48;
    atomic_uint next;
} sjob;

static void *synth_worker(void *arg)
{
    sjob *j = (sjob *)arg;
    const uint32_t utt = j->n_msgs * j->msg_samples;
    const size_t msg_bytes = (size_t)j->msg_samples * (j->stereo48 ? 12 : 2);
    for (;;) {
        uint32_t s = atomic_fetch_add(&j->next, 1);
        if (s >= j->n_streams) break;
        voice v;
        voice_init(&v, j->seed, j->group_base + s / j->group_size, s % j->group_size);
        for (uint32_t m = 0; m < j->n_msgs; m++) {
            const uint64_t idx = j->layout == ORA_LAYOUT_STREAM_MAJOR ? (uint64_t)s * j->n_msgs + m
                                                                      : (uint64_t)m * j->n_streams + s;
            uint8_t *dst = j->out + idx * msg_bytes;
            for (uint32_t k = 0; k < j->msg_samples; k++) {
                const int32_t l = voice_sample(&v, m * j->msg_samples + k, utt);
                if (!j->stereo48) {
                    dst[2 * k] = (uint8_t)l; dst[2 * k + 1] = (uint8_t)((uint32_t)l >> 8);
                } else {
                    const int32_t r = l - (l >> 3);
                    for (int h = 0; h < 3; h++) {        /* zero-order hold: three equal 48 kHz frames */
                        uint8_t *f = dst + (3 * (size_t)k + h) * 4;
                        f[0] = (uint8_t)l; f[1] = (uint8_t)((uint32_t)l >> 8);
                        f[2] = (uint8_t)r; f[3] = (uint8_t)((uint32_t)r >> 8);
                    }
                }
            }
        }
    }
    return NULL;
}

void ora_synth_fill(void *out, uint64_t seed, uint32_t group_base, uint32_t n_streams,
                    uint32_t group_size, uint32_t n_msgs, uint32_t msg_samples, int stereo48,
                    uint32_t layout, uint32_t n_threads)
{
    sjob j;
    j.out = (uint8_t *)out; j.seed = seed; j.group_base = group_base; j.n_streams = n_streams;
    j.group_size = group_size ? group_size : 1; j.n_msgs = n_msgs; j.msg_samples = msg_samples;
    j.layout = layout; j.stereo48 = stereo48;
    atomic_init(&j.next, 0);
    if (n_threads == 0) n_threads = 1;
    pthread_t *th = (pthread_t *)malloc(n_threads * sizeof(pthread_t));
    for (uint32_t i = 1; i < n_threads; i++) pthread_create(&th[i], NULL, synth_worker, &j);
    synth_worker(&j);
    for (uint32_t i = 1; i < n_threads; i++) pthread_join(th[i], NULL);
    free(th);
}

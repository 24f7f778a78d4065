// synth.h — integer-only synthetic relay PCM, host + device.
//
// Everything is 64-bit integer arithmetic keyed by SplitMix64, so the C++ host fill,
// the CUDA fill and the numpy mirror (loqa_hub_b200/synth.py) agree byte for byte.
//   base utterance u_g[n]: three parabolic-sine tones at group-dependent pitches under a
//       trapezoid envelope (peak ~ +-12,000) plus +-200 noise;
//   relay view x_{g,r}[n] = clamp_i16(((u_g[n] * G_r) >> 15) + noise_{g,r}[n]),
//       G_r = 0.9^r in Q15, noise amplitude 64*(r+1)  — same utterance, varied gain/noise,
//       so the relays of one collision group have well separated wake-word energies.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define GA_HD __host__ __device__ __forceinline__
#else
#define GA_HD inline
#endif

namespace ga {
namespace synth {

GA_HD uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// parabolic sine, phase in [0, 65536), amplitude 4000
GA_HD int64_t tone(uint32_t phase)
{
    const int64_t x = phase & 0x7FFFu;
    const int64_t y = (x * (32768 - x)) >> 13;          // 0 .. 32768
    const int64_t a = (y * 4000) >> 15;
    return (phase & 0x8000u) ? -a : a;
}

GA_HD int64_t gain_q15(uint32_t relay)
{
    // 0.9^r in Q15 for r = 0..7, repeating
    const int64_t g[8] = {32768, 29491, 26542, 23888, 21499, 19349, 17414, 15673};
    return g[relay & 7u];
}

GA_HD int16_t sample(uint64_t seed, uint32_t group, uint32_t relay, uint32_t n, uint32_t utt_samples)
{
    const uint64_t gk = splitmix64(seed ^ (0x10CA0B20ull + group));
    const uint32_t inc0 = 400u + static_cast<uint32_t>(gk % 1200u);             // ~100-390 Hz
    const uint32_t inc1 = 1600u + static_cast<uint32_t>((gk >> 16) % 1600u);    // ~390-780 Hz
    const uint32_t inc2 = 3200u + static_cast<uint32_t>((gk >> 32) % 2400u);    // ~780-1370 Hz
    const int64_t t = tone((inc0 * n) & 0xFFFFu) + tone((inc1 * n) & 0xFFFFu) + tone((inc2 * n) & 0xFFFFu);
    // trapezoid: 4000-sample ramps
    uint32_t edge = n < utt_samples - 1 - n ? n : utt_samples - 1 - n;
    if (n >= utt_samples) edge = 0;
    const int64_t env = static_cast<int64_t>(edge < 4000u ? edge : 4000u) * 8;  // 0 .. 32000 (Q15)
    const uint64_t rg = splitmix64(gk + n);
    const int64_t u = ((t * env) >> 15) + (static_cast<int64_t>(rg % 401u) - 200);
    const uint64_t rk = splitmix64(gk ^ (0x9E3779B97F4A7C15ull * (relay + 1)));
    const uint64_t rr = splitmix64(rk + n);
    const int64_t amp = 64 * (static_cast<int64_t>(relay) + 1);
    const int64_t noise = static_cast<int64_t>(rr % static_cast<uint64_t>(2 * amp + 1)) - amp;
    int64_t x = ((u * gain_q15(relay)) >> 15) + noise;
    if (x > 32767) x = 32767;
    if (x < -32768) x = -32768;
    return static_cast<int16_t>(x);
}

}  // namespace synth
}  // namespace ga

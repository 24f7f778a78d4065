// sha256.cuh — SHA-256 (FIPS 180-4) compression for one thread, all state in registers.
//
// Each lane hashes its own message; the 64 rounds are fully unrolled so the round
// constants become immediates and the 16-word message schedule lives in a rolling
// register window (no local memory).  Rotations are funnel shifts (SHF.R.W), Ch / Maj /
// the three-way XORs are single LOP3s — the kernel is bound by the integer ALU pipe.
#pragma once
#include <stdint.h>

namespace ga {

__device__ __forceinline__ uint32_t rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t bswap32(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

struct Sha256State {
    uint32_t h[8];
    __device__ __forceinline__ void init()
    {
        h[0] = 0x6a09e667u; h[1] = 0xbb67ae85u; h[2] = 0x3c6ef372u; h[3] = 0xa54ff53au;
        h[4] = 0x510e527fu; h[5] = 0x9b05688cu; h[6] = 0x1f83d9abu; h[7] = 0x5be0cd19u;
    }
};

// a + b issued as IMAD a*one+b: `one` is a kernel parameter equal to 1 that the assembler cannot
// fold, so the addition runs on the FMA pipe instead of the ALU pipe
__device__ __forceinline__ uint32_t add_fma_pipe(uint32_t a, uint32_t b, uint32_t one)
{
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(one), "r"(b));
    return r;
}

#define GA_SHA_ROUND(a, b, c, d, e, f, g, h, k, w)                                   \
    {                                                                                \
        uint32_t t1;                                                                 \
        if (kImadRound) {                                                            \
            /* h + k + w does not depend on this round's e: two FMA-pipe additions */ \
            const uint32_t hkw_ = add_fma_pipe(add_fma_pipe((w), (k), one), (h), one); \
            t1 = hkw_ + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) + (((e) & (f)) ^ (~(e) & (g))); \
        } else {                                                                     \
            t1 = (h) + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) +                    \
                 (((e) & (f)) ^ (~(e) & (g))) + (k) + (w);                           \
        }                                                                            \
        uint32_t t2 = (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) +                     \
                      (((a) & (b)) ^ ((a) & (c)) ^ ((b) & (c)));                     \
        (d) += t1;                                                                   \
        (h) = t1 + t2;                                                               \
    }

#define GA_SHA_SCHED(w, i)                                                           \
    {                                                                                \
        uint32_t x15 = w[((i) + 1) & 15], x2 = w[((i) + 14) & 15];                   \
        const uint32_t s0_ = rotr(x15, 7) ^ rotr(x15, 18) ^ (x15 >> 3);              \
        const uint32_t s1_ = rotr(x2, 17) ^ rotr(x2, 19) ^ (x2 >> 10);               \
        if (kImad) {                                                                 \
            uint32_t t_ = add_fma_pipe(s0_, w[(i) & 15], one);                       \
            t_ = add_fma_pipe(w[((i) + 9) & 15], t_, one);                           \
            w[(i) & 15] = add_fma_pipe(s1_, t_, one);                                \
        } else {                                                                     \
            w[(i) & 15] += s0_ + w[((i) + 9) & 15] + s1_;                            \
        }                                                                            \
    }

// w[16] holds the block as big-endian words on entry and is clobbered.
static __constant__ uint32_t ga_sha_k[64] = {
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u,
    0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u,
    0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau,
    0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u,
    0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
    0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
    0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u,
    0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};

template <bool kImad = false, bool kImadRound = false, int kRolled = 0>
__device__ __forceinline__ void sha256_compress(Sha256State &s, uint32_t (&w)[16], uint32_t one = 1u)
{
    uint32_t a = s.h[0], b = s.h[1], c = s.h[2], d = s.h[3];
    uint32_t e = s.h[4], f = s.h[5], g = s.h[6], h = s.h[7];

#define GA_R8(k0, k1, k2, k3, k4, k5, k6, k7, base)                                  \
    GA_SHA_ROUND(a, b, c, d, e, f, g, h, k0, w[(base + 0) & 15])                     \
    GA_SHA_ROUND(h, a, b, c, d, e, f, g, k1, w[(base + 1) & 15])                     \
    GA_SHA_ROUND(g, h, a, b, c, d, e, f, k2, w[(base + 2) & 15])                     \
    GA_SHA_ROUND(f, g, h, a, b, c, d, e, k3, w[(base + 3) & 15])                     \
    GA_SHA_ROUND(e, f, g, h, a, b, c, d, k4, w[(base + 4) & 15])                     \
    GA_SHA_ROUND(d, e, f, g, h, a, b, c, k5, w[(base + 5) & 15])                     \
    GA_SHA_ROUND(c, d, e, f, g, h, a, b, k6, w[(base + 6) & 15])                     \
    GA_SHA_ROUND(b, c, d, e, f, g, h, a, k7, w[(base + 7) & 15])
#define GA_S8(base)                                                                  \
    GA_SHA_SCHED(w, base + 0) GA_SHA_SCHED(w, base + 1) GA_SHA_SCHED(w, base + 2)    \
    GA_SHA_SCHED(w, base + 3) GA_SHA_SCHED(w, base + 4) GA_SHA_SCHED(w, base + 5)    \
    GA_SHA_SCHED(w, base + 6) GA_SHA_SCHED(w, base + 7)

    GA_R8(0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u, 0)
    GA_R8(0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u, 8)
    if (kRolled == 2) {
        // rounds 16..31 straight, 32..63 as two passes over the 16-round body: two thirds of the
        // unrolled form's code, one loop back-edge less than the three-pass form
        GA_S8(16)
        GA_R8(0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, 16)
        GA_S8(24)
        GA_R8(0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u, 24)
#pragma unroll 1
        for (int r = 32; r < 64; r += 16) {
            const uint32_t *kk = ga_sha_k + r;
            GA_S8(16)
            GA_R8(kk[0], kk[1], kk[2], kk[3], kk[4], kk[5], kk[6], kk[7], 16)
            GA_S8(24)
            GA_R8(kk[8], kk[9], kk[10], kk[11], kk[12], kk[13], kk[14], kk[15], 24)
        }
    } else if (kRolled) {
        // rounds 16..63 as three passes over one 16-round body (the register roles of a..h and
        // the 16-word window both return to their start after 16 rounds); the round constants
        // come from the constant bank.  A third of the code of the fully unrolled form — for
        // kernels whose loop body would otherwise overflow the 32 KB instruction cache.
#pragma unroll 1
        for (int r = 16; r < 64; r += 16) {
            const uint32_t *kk = ga_sha_k + r;
            GA_S8(16)
            GA_R8(kk[0], kk[1], kk[2], kk[3], kk[4], kk[5], kk[6], kk[7], 16)
            GA_S8(24)
            GA_R8(kk[8], kk[9], kk[10], kk[11], kk[12], kk[13], kk[14], kk[15], 24)
        }
    } else {
    GA_S8(16)
    GA_R8(0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, 16)
    GA_S8(24)
    GA_R8(0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u, 24)
    GA_S8(32)
    GA_R8(0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u, 32)
    GA_S8(40)
    GA_R8(0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u, 40)
    GA_S8(48)
    GA_R8(0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u, 48)
    GA_S8(56)
    GA_R8(0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u, 56)
    }

#undef GA_R8
#undef GA_S8
    s.h[0] += a; s.h[1] += b; s.h[2] += c; s.h[3] += d;
    s.h[4] += e; s.h[5] += f; s.h[6] += g; s.h[7] += h;
}

}  // namespace ga

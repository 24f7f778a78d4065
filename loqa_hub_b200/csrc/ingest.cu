// ingest.cu — fused relay-message ingest for B200 (sm_100a).
//
// Replaces, for a whole batch of relay messages per launch, the reference's per-chunk
// scalar loops (paths relative to the loqa-hub tree):
//   bytesToFloat32Array          internal/grpc/audio_service.go:1048-1101   (convert + side stats)
//   calculateSignalStrength      internal/grpc/audio_service.go:818-852     (sumsq / rms / nonzero / peak)
//   voice-event audio fingerprint (README.md:50, API.md:170; hook events/voice_event.go:84-88)
//
// Work decomposition.  A warp owns a tile of 32 consecutive messages; lane = message.
// Messages are streamed through shared memory in CH-byte slices ("rows"): every lane issues
// one TMA bulk copy (cp.async.bulk -> UBLKCP) for its own row into a STAGES-deep ring of
// stage buffers, completion is tracked by one mbarrier per stage.  The lane then walks ITS
// row's 64-byte blocks: each int16 pair is converted (exact division), fed to the reference's
// strictly sequential float64 sum of float32 squares (so sumsq/rms are bit-identical to the
// Go loop) and to the packed statistics, and the block is hashed (SHA-256 schedule in
// registers).  Rows are padded by 16 B so the lane-per-row LDS.128 pattern is bank-conflict
// free.  The f32 samples leave the SM in one of two ways (template parameter COOP):
//   COOP = 0  the lane stores its own samples, 16 bytes per instruction.  Uncoalesced across
//             lanes (L2 merges the sectors), but it needs no second pass, no shuffles and only
//             one division per sample — best when the kernel is bound by the integer ALU pipe,
//             i.e. whenever SHA-256 runs (measured 1.20 ms vs 1.36 ms per 20,480 audio-s).
//   COOP = 1  the lane writes its samples to a padded f32 tile in shared memory and a
//             warp-cooperative pass copies the tile out, 256 contiguous bytes per row per
//             instruction — best for the HBM-bound instantiations without SHA-256.
// A third variant staged the samples in shared memory and used TMA bulk stores; with 128-256 B
// per store and the proxy fence it was slower than both (tuning log).
// Staging has two forms (bit 2 of COOP).  Generic batches (ragged lengths, arbitrary offsets):
// one 1-D bulk copy per lane per slice — the instruction takes uniform registers, so the
// compiler serialises the 32 lanes (a 32-iteration R2UR/ELECT loop, ~17 % of the instructions of
// the HBM-bound instantiation).  Uniform batches (equal lengths at a constant stride, e.g. one
// 20 ms chunk per stream per tick laid out in arrival order): the slab is described by a 2-D TMA
// tensor map [message][byte] and ONE cp.async.bulk.tensor.2d per slice brings in the 32 x 128 B
// box, 128-byte swizzled so that lane-per-row 16-byte reads hit eight different bank groups.
// No tensor cores: nothing here is a contraction.
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "sha256.cuh"

namespace ga {

constexpr int kWarpsPerBlock = 4;

template <int CH, int STAGES, bool TMAP = false, bool OUT_TILE = false>
struct TileSmem {
    // generic: rows padded by 16 B ((CH/16 + 1) odd -> conflict-free LDS.128);
    // tensor-map: dense 128-byte rows, 128B-swizzled by the TMA unit, stages 1024-byte aligned
    static constexpr int kRowStride = TMAP ? CH : CH + 16;
    static constexpr int kStageBytes = 32 * kRowStride;
    // optional f32 output tile (one slice of every row, rows padded by 16 B: conflict-free
    // lane-per-row STS.128) followed by the tile's 32 output offsets
    static constexpr int kOutRowStride = 2 * CH + 16;
    static constexpr int kOutBase = STAGES * kStageBytes;
    static constexpr int kOutBytes = OUT_TILE ? 32 * kOutRowStride + 256 : 0;
    static constexpr int kBarBase = kOutBase + kOutBytes;
    static constexpr int kRawBytes = kBarBase + 128;
    static constexpr int kWarpBytes = TMAP ? ((kRawBytes + 1023) / 1024) * 1024 : kRawBytes;   // stages stay 1024-aligned
    static_assert(!TMAP || CH == 128, "the swizzled layout is written for 128-byte slices");
    static_assert(CH % 64 == 0, "slices must hold whole SHA-256 blocks");
    static_assert(((CH / 16) + 1) % 2 == 1, "row stride must be an odd number of 16-byte units");
};

struct IngestParams {
    const uint8_t *pcm;
    const ga_desc *desc;
    float *samples;
    ga_result *results;
    ga_frame *frames;
    double vad;
    uint32_t n_msgs;
    uint32_t n_tiles;
    uint32_t msg_bytes;   // uniform batches only
    uint32_t one;         // 1 (opaque to the assembler)
};
#ifndef GA_SCHED_IMAD
#define GA_SCHED_IMAD true
#endif
#ifndef GA_PACKED_CONVERT
#define GA_PACKED_CONVERT 1
#endif
#ifndef GA_SHA_ROLLED
#define GA_SHA_ROLLED 1
#endif
#ifndef GA_ROUND_IMAD
#define GA_ROUND_IMAD false
#endif

// One 2-D box [32 rows x 128 bytes] of the slab -> shared memory (SASS UTMALDG).
__device__ __forceinline__ void tma_load_box(void *smem_dst, const CUtensorMap *tmap, uint32_t col, uint32_t row,
                                             uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::
            "r"(smem_u32(smem_dst)),
        "l"(tmap), "r"(col), "r"(row), "r"(smem_u32(bar))
        : "memory");
}

// Address of byte `off` (16-byte granular swizzle) of row r in a stage.
template <bool TMAP, int ROW_STRIDE>
__device__ __forceinline__ const uint8_t *row_addr(const uint8_t *sbase, uint32_t r, uint32_t off)
{
    if (TMAP) return sbase + r * 128u + ((((off >> 4) ^ (r & 7u)) << 4) | (off & 15u));
    return sbase + r * ROW_STRIDE + off;
}

// Two int16 samples -> two exact floats without the conversion (XU) pipe: bias each half
// by 0x8000, splice it under the exponent of 2^23 and subtract 2^23 + 32768.
// kXu: use the conversion pipe after all (I2F.S16 with a half-word selector, no integer
// instruction) — right when the kernel is bound by the ALU pipe (SHA-256 modes).
template <bool kXu>
__device__ __forceinline__ void unpack2(uint32_t w, float &lo, float &hi)
{
    if (kXu) {
        lo = static_cast<float>(static_cast<short>(w & 0xffffu));
        hi = static_cast<float>(static_cast<short>(w >> 16));
        return;
    }
    const uint32_t x = w ^ 0x80008000u;
    lo = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7610)) - 8421376.0f;
    hi = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7632)) - 8421376.0f;
}
#ifndef GA_XU_UNPACK
#define GA_XU_UNPACK true
#endif

// Statistics of one span of samples in the reference's arithmetic.  `sum` is the strictly
// sequential float64 chain; the order-independent pieces are kept packed: max/min of both
// int16 halves in one VIMNMX.S16x2 each, non-zero counts as two 16-bit counters.
struct Span {
    double sum;
    uint32_t n;
    uint32_t mx, mn;     // packed running max / min of the raw int16 halves
    uint32_t nzp;        // packed non-zero counters (flush before 65,535 samples per half)
    float nzf;           // non-zero count of the whole-block path, kept on the FMA pipe (exact: < 2^24)
    __device__ __forceinline__ void reset() { sum = 0.0; n = 0; mx = 0x80008000u; mn = 0x7FFF7FFFu; nzp = 0; nzf = 0.0f; }
    __device__ __forceinline__ void minmax_word(uint32_t w)
    {
        mx = __vmaxs2(mx, w);
        mn = __vmins2(mn, w);
    }
    __device__ __forceinline__ void stats_word(uint32_t w)
    {
        minmax_word(w);
        const uint32_t t = (w & 0x7FFF7FFFu) + 0x7FFF7FFFu;      // bit 15/31 set iff the half is non-zero
        nzp += ((t | w) & 0x80008000u) >> 15;
    }
    // sample != 0 counted from its square: the smallest non-zero square is (1/32767)^2 ~ 9.3e-10, so
    // sq * 2^40 saturates to exactly 1.0f, and 0 stays 0 — FMUL.SAT + FADD on the FMA pipe instead of
    // five integer instructions per pair of samples on the ALU pipe SHA-256 is bound by
    __device__ __forceinline__ void count_nonzero(float sq0, float sq1, float sq2, float sq3)
    {
        nzf += (__saturatef(sq0 * 0x1p40f) + __saturatef(sq1 * 0x1p40f)) +
               (__saturatef(sq2 * 0x1p40f) + __saturatef(sq3 * 0x1p40f));
    }
    __device__ __forceinline__ uint32_t nonzero() const { return (nzp & 0xFFFFu) + (nzp >> 16) + static_cast<uint32_t>(nzf); }
    // max |v| = max(|max v|, |min v|) per half.  Written with abs(), not max(a, -b): nvcc 12.9
    // fuses the latter into VIMNMX3.RELU and drops the negation (observed on sm_100a; see
    // tools/dbg_vmin.cu and tests/test_gpu_ingest.py::test_peak_of_negative_samples_in_either_half).
    __device__ __forceinline__ int amax() const
    {
        const int a = abs(static_cast<int>(mx) >> 16), b = abs(static_cast<int>(mn) >> 16);
        const int c = abs(static_cast<int>(static_cast<int16_t>(mx & 0xFFFFu)));
        const int d = abs(static_cast<int>(static_cast<int16_t>(mn & 0xFFFFu)));
        return max(max(a, b), max(c, d));
    }
};

template <int CH, int STAGES, uint32_t MODE, int COOP>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, (MODE == 7u) ? 5 : 0)
ingest_kernel(const IngestParams p, const __grid_constant__ CUtensorMap tmap)
{
    constexpr bool kTmap = (COOP & 4) != 0;                                       // 2-D TMA staging (uniform batches)
    constexpr bool kNoFrames = (COOP & 8) != 0;                                   // launched without frame records: no per-frame energy chain
    constexpr bool kOutTile = (MODE & GA_DO_CONVERT) != 0 && (COOP & 1) == 1;
    using L = TileSmem<CH, STAGES, kTmap, kOutTile>;
    constexpr bool kCoop = (MODE & GA_DO_CONVERT) != 0 && (COOP & 1) == 1;      // cooperative store pass
    constexpr bool kLaneStore = (MODE & GA_DO_CONVERT) != 0 && (COOP & 1) == 0; // lane stores its own samples
    constexpr bool kFeat = (MODE & GA_DO_FEATURES) != 0;
    constexpr bool kSha = (MODE & GA_DO_SHA256) != 0;
    constexpr uint32_t FULL = 0xffffffffu;

    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *wbase = smem + warp * L::kWarpBytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(wbase + L::kBarBase);

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) mbar_init(&bars[s], 1);
        fence_mbar_init();
    }
    __syncwarp();

    const uint32_t gw = blockIdx.x * kWarpsPerBlock + warp;
    const uint32_t nw = gridDim.x * kWarpsPerBlock;

    // ------------------------------------------------------------ producer cursor
    uint32_t p_tile = gw, p_chunk = 0, p_nch = 0, p_nb = 0;
    uint64_t p_in = 0;
    auto p_load = [&]() {
        p_nb = 0; p_in = 0; p_nch = 0;
        if (kTmap) {
            p_nch = p.msg_bytes ? (p.msg_bytes - 1) / CH + 1 : 1;
        } else if (p_tile < p.n_tiles) {
            uint32_t i = p_tile * 32 + lane;
            uint32_t nch = 0;
            if (i < p.n_msgs) {
                p_in = p.desc[i].in_off;
                p_nb = p.desc[i].nbytes;
                nch = p_nb ? (p_nb - 1) / CH + 1 : 1;
            }
            p_nch = __reduce_max_sync(FULL, nch);
        }
    };
    auto produce = [&](int stage) {
        if (p_tile >= p.n_tiles) return;                       // warp-uniform
        const uint32_t lo = p_chunk * CH;
        if (kTmap) {
            // one box for the whole warp tile; rows / columns outside the tensor arrive as zeros
            if (lane == 0) {
                mbar_arrive_expect_tx(&bars[stage], L::kStageBytes);
                tma_load_box(wbase + stage * L::kStageBytes, &tmap, lo, p_tile * 32u, &bars[stage]);
            }
        } else {
            uint32_t bytes = p_nb > lo ? min(p_nb - lo, static_cast<uint32_t>(CH)) : 0u;
            bytes = (bytes + 15u) & ~15u;
            const uint32_t total = __reduce_add_sync(FULL, bytes);
            if (lane == 0) mbar_arrive_expect_tx(&bars[stage], total);
            __syncwarp();
            if (bytes)
                bulk_g2s(wbase + stage * L::kStageBytes + lane * L::kRowStride, p.pcm + p_in + lo,
                         bytes, &bars[stage]);
        }
        if (++p_chunk == p_nch) {
            p_tile += nw;
            p_chunk = 0;
            p_load();
        }
    };

    p_load();
#pragma unroll
    for (int s = 0; s < STAGES; s++) produce(s);

    // ------------------------------------------------------------ consumer
    uint32_t it = 0;   // stage jobs consumed so far by this warp
    for (uint32_t c_tile = gw; c_tile < p.n_tiles; c_tile += nw) {
        const uint32_t mi = c_tile * 32 + lane;
        const bool valid = mi < p.n_msgs;
        uint32_t nb = 0, flags = 0, frame_off = 0;
        uint64_t out_off = 0;
        if (valid) {
            const ga_desc d = p.desc[mi];
            nb = kTmap ? p.msg_bytes : d.nbytes; flags = d.flags; frame_off = d.frame_off; out_off = d.out_off;
        }
        const uint32_t my_nch = valid ? (nb ? (nb - 1) / CH + 1 : 1) : 0;
        const uint32_t c_nch = __reduce_max_sync(FULL, my_nch);
        const uint32_t nb_even = nb & ~1u;

        Sha256State sha;
        if (kSha) sha.init();
        // message-level: the f64 chain runs over the whole message; counts / extrema are
        // merged from the frames (they do not depend on order)
        double msg_sum = 0.0;
        uint32_t msg_nz = 0, fidx = 0;
        int msg_amax = 0;
        Span frm;
        frm.reset();

        auto emit_frame = [&]() {
            const uint32_t fnz = frm.nonzero();
            const int fa = frm.amax();
            if (!kNoFrames && p.frames) {
                ga_frame f;
                f.sumsq = frm.sum;
                f.rms = __dsqrt_rn(__ddiv_rn(frm.sum, static_cast<double>(frm.n)));
                f.n_samples = frm.n; f.nonzero = fnz;
                f.peak = pcm_to_f32(fa); f.vad = f.rms > p.vad ? 1u : 0u;
                p.frames[static_cast<uint64_t>(frame_off) + fidx] = f;
            }
            msg_nz += fnz;
            msg_amax = max(msg_amax, fa);
            fidx++;
            frm.reset();
        };
        // one converted sample in reference order (audio_service.go:827-828)
        auto chain = [&](float x) {
            const double q = sq_f64(x);
            msg_sum += q;
            if (!kNoFrames) frm.sum += q;
        };

        for (uint32_t c = 0; c < c_nch; c++, it++) {
            const int stage = it % STAGES;
            const uint32_t parity = (it / STAGES) & 1u;
            const uint8_t *sbase = wbase + stage * L::kStageBytes;
            const uint32_t lo = c * CH;
            mbar_wait(&bars[stage], parity);

            // ---------------- lane = message.  One loop over the 64-byte blocks this slice
            // contributes, including the padding block(s) when the message ends here, so the
            // unrolled compression exists once in the instruction stream.
            float *orow = reinterpret_cast<float *>(wbase + L::kOutBase + lane * L::kOutRowStride);
            if (kCoop) {
                const uint32_t hi_even = min(nb_even, lo + CH);
                const uint64_t ns = (c < my_nch && hi_even > lo) ? (hi_even - lo) >> 1 : 0;
                reinterpret_cast<uint64_t *>(wbase + L::kOutBase + 32 * L::kOutRowStride)[lane] = (out_off & 0xFFFFFFFFFFFFull) | (ns << 48);
            }
            if ((kFeat || kSha || kLaneStore || kCoop) && c < my_nch) {
                const uint32_t span = min(nb, lo + CH) - lo;
                const uint32_t nfull = span >> 6;
                const uint32_t rem = span & 63u;
                const bool last = (c == my_nch - 1);
                const uint32_t nblk = nfull + (last ? (kSha ? (rem >= 56 ? 2u : 1u) : (rem ? 1u : 0u)) : 0u);
                for (uint32_t b = 0; b < nblk; b++) {
                    const uint32_t vb = b < nfull ? 64u : (b == nfull ? rem : 0u);   // data bytes in this block
                    uint32_t w[16];
                    if (vb) {
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const uint4 t = *reinterpret_cast<const uint4 *>(row_addr<kTmap, L::kRowStride>(sbase, lane, b * 64 + 16 * j));
                            w[4 * j] = t.x; w[4 * j + 1] = t.y; w[4 * j + 2] = t.z; w[4 * j + 3] = t.w;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; j++) w[j] = 0;
                    }
                    if (kFeat || kLaneStore || kCoop) {
                        if (vb == 64) {
                            float *g = p.samples + out_off + (lo >> 1) + b * 32;
                            const bool galigned = (out_off & 3ull) == 0;
                            // stage 1: 32 independent conversions (and squares); stage 2: the two
                            // strictly ordered float64 chains — written apart so the scheduler can
                            // overlap the conversions of one word with the additions of another
                            float sqf[32];
#pragma unroll
                            for (int j = 0; j < 16; j += 2) {
                                float v0, v1, v2, v3;
                                unpack2<kSha && GA_XU_UNPACK>(w[j], v0, v1);
                                unpack2<kSha && GA_XU_UNPACK>(w[j + 1], v2, v3);
#if GA_PACKED_CONVERT
                                float x0, x1, x2, x3, q0_, q1_, q2_, q3_;
                                pcm_to_f32x2(v0, v1, x0, x1, q0_, q1_);
                                pcm_to_f32x2(v2, v3, x2, x3, q2_, q3_);
#else
                                const float x0 = pcm_to_f32(v0), x1 = pcm_to_f32(v1);
                                const float x2 = pcm_to_f32(v2), x3 = pcm_to_f32(v3);
#endif
                                if (kCoop) {
                                    *reinterpret_cast<float4 *>(orow + b * 32 + 2 * j) = make_float4(x0, x1, x2, x3);
                                } else if (kLaneStore) {
                                    if (galigned) st_global_v4(g + 2 * j, x0, x1, x2, x3);
                                    else { g[2 * j] = x0; g[2 * j + 1] = x1; g[2 * j + 2] = x2; g[2 * j + 3] = x3; }
                                }
                                if (kFeat) {
#if GA_PACKED_CONVERT
                                    sqf[2 * j] = q0_; sqf[2 * j + 1] = q1_; sqf[2 * j + 2] = q2_; sqf[2 * j + 3] = q3_;
#else
                                    sqf[2 * j] = __fmul_rn(x0, x0); sqf[2 * j + 1] = __fmul_rn(x1, x1);
                                    sqf[2 * j + 2] = __fmul_rn(x2, x2); sqf[2 * j + 3] = __fmul_rn(x3, x3);
#endif
                                    frm.minmax_word(w[j]);
                                    frm.minmax_word(w[j + 1]);
                                    frm.count_nonzero(sqf[2 * j], sqf[2 * j + 1], sqf[2 * j + 2], sqf[2 * j + 3]);
                                }
                            }
                            if (kFeat) {
#pragma unroll
                                for (int i = 0; i < 32; i++) {
                                    const double q = static_cast<double>(sqf[i]);
                                    msg_sum += q;
                                    if (!kNoFrames) frm.sum += q;   // the second strictly ordered chain: only when frames are wanted
                                }
                            }
                            if (kFeat) {
                                frm.n += 32;
                                if (frm.n == GA_FRAME_SAMPLES) emit_frame();
                            }
                        } else if (vb >= 2) {
                            // tail of the message: fewer than 32 samples, rolled loop from shared memory
                            const uint32_t ns = vb >> 1;           // an odd trailing byte is not a sample
                            for (uint32_t k = 0; k < ns; k++) {
                                const uint32_t h = *reinterpret_cast<const uint16_t *>(row_addr<kTmap, L::kRowStride>(sbase, lane, b * 64 + 2 * k));
                                float v0, v1;
                                unpack2<false>(h, v0, v1);
                                const float x0 = pcm_to_f32(v0);
                                if (kCoop) orow[b * 32 + k] = x0;
                                else if (kLaneStore) p.samples[out_off + (lo >> 1) + b * 32 + k] = x0;
                                if (kFeat) {
                                    chain(x0);
                                    frm.stats_word(h);              // upper half is a zero sample: neutral
                                }
                            }
                            if (kFeat) frm.n += ns;
                        }
                    }
                    if (kSha) {
                        if (vb < 64) {
                            // zero everything at and beyond byte vb; append 0x80 right after the data
#pragma unroll
                            for (int j = 0; j < 16; j++) {
                                const uint32_t b0 = 4u * j;
                                uint32_t m = 0;
                                if (b0 + 4 <= vb) m = 0xffffffffu;
                                else if (b0 < vb) m = (1u << (8 * (vb - b0))) - 1u;
                                uint32_t x = w[j] & m;
                                if (b == nfull && (vb >> 2) == static_cast<uint32_t>(j)) x |= 0x80u << (8 * (vb & 3u));
                                w[j] = x;
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 16; j++) w[j] = bswap32(w[j]);
                        if (last && b == nblk - 1) {
                            const uint64_t bits = static_cast<uint64_t>(nb) * 8ull;
                            w[14] = static_cast<uint32_t>(bits >> 32);
                            w[15] = static_cast<uint32_t>(bits);
                        }
                        // next to the conversion code the unrolled rounds overflow the 32 KB instruction
                        // cache: fused instantiations run rounds 16..63 as a 3-pass loop
                        sha256_compress<GA_SCHED_IMAD, GA_ROUND_IMAD, (kFeat || kLaneStore || kCoop) ? GA_SHA_ROLLED : 0>(sha, w, p.one);
                    }
                }
                if (kFeat && last && frm.n) emit_frame();

            }

            __syncwarp();          // every lane is done reading this stage
            produce(stage);        // refill it with the slice STAGES jobs ahead

            // ---------------- cooperative copy-out of the f32 tile (COOP bit 0): the lanes wrote their
            // rows above; now 16 lanes cover one row's 256 contiguous bytes per instruction
            if (kCoop) {
                constexpr int VPR = CH / 8;   // 16-byte vectors (4 samples) per row slice
                const uint64_t *otab = reinterpret_cast<const uint64_t *>(wbase + L::kOutBase + 32 * L::kOutRowStride);
#pragma unroll 4
                for (int k = 0; k < VPR; k++) {
                    const int v = k * 32 + lane;
                    const int row = v / VPR, col = v % VPR;
                    const uint64_t packed = otab[row];                 // out_off (48 bits) | valid samples of the row in this slice
                    const uint64_t r_out = packed & 0xFFFFFFFFFFFFull;
                    const uint32_t r_ns = static_cast<uint32_t>(packed >> 48);   // valid samples of the row in this slice (<= CH/2)
                    const uint32_t s0 = col * 4;
                    if (s0 < r_ns) {
                        const float4 t = *reinterpret_cast<const float4 *>(wbase + L::kOutBase + row * L::kOutRowStride + col * 16);
                        float *dst = p.samples + r_out + (lo >> 1) + s0;
                        const uint32_t nvalid = r_ns - s0;
                        if (nvalid >= 4 && (r_out & 3ull) == 0) {
                            st_global_v4(dst, t.x, t.y, t.z, t.w);
                        } else {
                            dst[0] = t.x;
                            if (nvalid > 1) dst[1] = t.y;
                            if (nvalid > 2) dst[2] = t.z;
                            if (nvalid > 3) dst[3] = t.w;
                        }
                    }
                }
                __syncwarp();      // the tile is free for the next slice
            }
        }

        // ---------------- result record
        if (valid) {
            uint4 *r = reinterpret_cast<uint4 *>(p.results + mi);
            if (kSha) {
                r[0] = make_uint4(bswap32(sha.h[0]), bswap32(sha.h[1]), bswap32(sha.h[2]), bswap32(sha.h[3]));
                r[1] = make_uint4(bswap32(sha.h[4]), bswap32(sha.h[5]), bswap32(sha.h[6]), bswap32(sha.h[7]));
            } else {
                r[0] = make_uint4(0, 0, 0, 0);
                r[1] = make_uint4(0, 0, 0, 0);
            }
            double rms = 0.0;
            uint32_t fl = flags << 8;
            float pk = 0.0f;
            if (kFeat) {
                const uint32_t n = nb >> 1;
                rms = n ? __dsqrt_rn(__ddiv_rn(msg_sum, static_cast<double>(n))) : 0.0;
                pk = pcm_to_f32(msg_amax);
                if (rms > p.vad) fl |= 1u;
            }
            r[2] = make_uint4(__double2loint(msg_sum), __double2hiint(msg_sum), __double2loint(rms), __double2hiint(rms));
            r[3] = make_uint4(nb >> 1, msg_nz, __float_as_uint(pk), fl);
        }
    }
}

// ---------------------------------------------------------------- host launcher
template <int CH, int STAGES, uint32_t MODE, int COOP>
static int launch_ingest(const IngestParams &p, const CUtensorMap &tmap, cudaStream_t st)
{
    using L = TileSmem<CH, STAGES, (COOP & 4) != 0, (MODE & GA_DO_CONVERT) != 0 && (COOP & 1) == 1>;
    constexpr int smem = kWarpsPerBlock * L::kWarpBytes;
    auto kern = ingest_kernel<CH, STAGES, MODE, COOP>;
    static thread_local int configured_dev = -1;
    static thread_local int blocks_per_sm = 0;
    int dev = 0;
    GA_CUDA_TRY(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        GA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        GA_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kern,
                                                                  kWarpsPerBlock * 32, smem));
        if (blocks_per_sm < 1) blocks_per_sm = 1;
        configured_dev = dev;
    }
    int sms = 0;
    int rc = num_sms(&sms);
    if (rc) return rc;
    uint32_t want = (p.n_tiles + kWarpsPerBlock - 1) / kWarpsPerBlock;
    uint32_t grid = static_cast<uint32_t>(sms * blocks_per_sm);
    if (grid > want) grid = want;
    if (grid == 0) return GA_OK;
    kern<<<grid, kWarpsPerBlock * 32, smem, st>>>(p, tmap);
    GA_CHECK_LAUNCH("ingest_kernel");
    return GA_OK;
}

static int check_ingest_args(const char *who, const void *d_pcm, const ga_desc *d_desc, float *d_samples,
                             ga_result *d_results, uint32_t &mode)
{
    if (!d_pcm || !d_desc || !d_results) { set_error("%s: NULL pcm/desc/results", who); return GA_ERR_ARG; }
    mode &= GA_DO_ALL;
    if (mode == 0) { set_error("%s: empty mode", who); return GA_ERR_ARG; }
    if ((mode & GA_DO_CONVERT) && !d_samples) { set_error("%s: GA_DO_CONVERT needs d_samples", who); return GA_ERR_ARG; }
    if ((reinterpret_cast<uintptr_t>(d_pcm) & 15u) || (reinterpret_cast<uintptr_t>(d_samples) & 15u) ||
        (reinterpret_cast<uintptr_t>(d_results) & 15u) || (reinterpret_cast<uintptr_t>(d_desc) & 15u)) {
        set_error("%s: device pointers must be 16-byte aligned", who);
        return GA_ERR_ARG;
    }
    return GA_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn get_encode_tiled()
{
    static encode_tiled_fn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<encode_tiled_fn>(p);
    }();
    return fn;
}

}  // namespace ga

extern "C" int ga_k_ingest(const void *d_pcm, const ga_desc *d_desc, uint32_t n_msgs,
                           float *d_samples, ga_result *d_results, ga_frame *d_frames,
                           uint32_t mode, double vad_threshold, void *stream)
{
    using namespace ga;
    if (n_msgs == 0) return GA_OK;
    int rc = check_ingest_args("ga_k_ingest", d_pcm, d_desc, d_samples, d_results, mode);
    if (rc) return rc;
    IngestParams p;
    p.pcm = static_cast<const uint8_t *>(d_pcm);
    p.desc = d_desc;
    p.samples = d_samples;
    p.results = d_results;
    p.frames = d_frames;
    p.vad = vad_threshold;
    p.n_msgs = n_msgs;
    p.n_tiles = (n_msgs + 31) / 32;
    p.msg_bytes = 0;
    p.one = 1;
    CUtensorMap none;
    memset(&none, 0, sizeof none);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // Defaults from the tuning sweep: lane stores when SHA-256 runs (ALU-bound),
    // the cooperative coalesced store pass otherwise (HBM-bound).
    switch (mode) {
        case 1: return launch_ingest<128, 2, 1, 1>(p, none, st);
        case 2: return p.frames ? launch_ingest<128, 2, 2, 1>(p, none, st) : launch_ingest<128, 2, 2, 1 | 8>(p, none, st);
        case 3: return p.frames ? launch_ingest<128, 2, 3, 1>(p, none, st) : launch_ingest<128, 2, 3, 1 | 8>(p, none, st);
        case 4: return launch_ingest<128, 3, 4, 0>(p, none, st);
        case 5: return launch_ingest<128, 2, 5, 0>(p, none, st);
        case 6: return p.frames ? launch_ingest<128, 2, 6, 0>(p, none, st) : launch_ingest<128, 2, 6, 0 | 8>(p, none, st);
        default: return p.frames ? launch_ingest<128, 2, 7, 0>(p, none, st) : launch_ingest<128, 2, 7, 0 | 8>(p, none, st);
    }
}

extern "C" int ga_k_ingest_uniform(const void *d_pcm, const ga_desc *d_desc, uint32_t n_msgs,
                                   uint64_t first_in_off, uint32_t msg_bytes, uint64_t msg_stride,
                                   float *d_samples, ga_result *d_results, ga_frame *d_frames,
                                   uint32_t mode, double vad_threshold, void *stream)
{
    using namespace ga;
    if (n_msgs == 0) return GA_OK;
    int rc = check_ingest_args("ga_k_ingest_uniform", d_pcm, d_desc, d_samples, d_results, mode);
    if (rc) return rc;
    if ((first_in_off & 15u) || (msg_stride & 15u) || msg_stride < msg_bytes || msg_stride >= (1ull << 40) || msg_bytes == 0) {
        set_error("ga_k_ingest_uniform: offset and stride must be multiples of 16, stride >= msg_bytes > 0");
        return GA_ERR_ARG;
    }
    encode_tiled_fn encode = get_encode_tiled();
    if (!encode) { set_error("ga_k_ingest_uniform: cuTensorMapEncodeTiled is not available from this driver"); return GA_ERR_CUDA; }
    CUtensorMap tmap;
    const cuuint64_t dims[2] = {msg_bytes, n_msgs};                 // [byte within message][message]
    const cuuint64_t strides[1] = {msg_stride};                     // bytes between consecutive messages
    const cuuint32_t box[2] = {128, 32};                            // one warp tile slice
    const cuuint32_t estr[2] = {1, 1};
    void *base = const_cast<uint8_t *>(static_cast<const uint8_t *>(d_pcm)) + first_in_off;
    CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { set_error("ga_k_ingest_uniform: cuTensorMapEncodeTiled failed (%d)", static_cast<int>(cr)); return GA_ERR_CUDA; }
    IngestParams p;
    p.pcm = static_cast<const uint8_t *>(d_pcm);
    p.desc = d_desc;
    p.samples = d_samples;
    p.results = d_results;
    p.frames = d_frames;
    p.vad = vad_threshold;
    p.n_msgs = n_msgs;
    p.n_tiles = (n_msgs + 31) / 32;
    p.msg_bytes = msg_bytes;
    p.one = 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (mode) {
        case 1: return launch_ingest<128, 2, 1, 5>(p, tmap, st);
        case 2: return p.frames ? launch_ingest<128, 2, 2, 5>(p, tmap, st) : launch_ingest<128, 2, 2, 5 | 8>(p, tmap, st);
        case 3: return p.frames ? launch_ingest<128, 2, 3, 5>(p, tmap, st) : launch_ingest<128, 2, 3, 5 | 8>(p, tmap, st);
        case 4: return launch_ingest<128, 3, 4, 4>(p, tmap, st);
        case 5: return launch_ingest<128, 2, 5, 4>(p, tmap, st);
        case 6: return p.frames ? launch_ingest<128, 2, 6, 4>(p, tmap, st) : launch_ingest<128, 2, 6, 4 | 8>(p, tmap, st);
        default: return p.frames ? launch_ingest<128, 2, 7, 4>(p, tmap, st) : launch_ingest<128, 2, 7, 4 | 8>(p, tmap, st);
    }
}

// resample.cu — 48 kHz interleaved-stereo int16 -> 16 kHz mono f32 (÷3 polyphase decimator)
// plus the per-message / per-frame energy statistics of the RESAMPLED samples, for B200.
//
// New-spec piece: the reference never reads AudioChunk.SampleRate
// (internal/grpc/audio_service.go:926-1043 treats every chunk as 16 kHz mono), so there is no
// reference code to restate; the definition (also in include/gpuaudio.h, ga_k_resample48) is:
//     mono[j] = (A1(L[j]) + A1(R[j])) * 0.5f          A1 = float32(int16)/32767.0 (:1078)
//     y[M]    = fold_{k=0..47} acc = fmaf(taps[k], mono[3M+2-k], acc),  acc0 = 0
// on the stream's absolute input-frame timeline, zero before the stream start.  The fixed tap
// order and explicit FMAs make the result bit-reproducible, so parity is exact, not 1e-5.
// The statistics are calculateSignalStrength's (:818-852) applied to y.
//
// A warp owns groups of 32 consecutive messages.  Pass 1 walks the group one message at a time
// with all 32 lanes: the message's bytes arrive in shared memory by ONE TMA bulk copy (3,840 B,
// issued one message ahead on a 2-stage mbarrier pipeline), 47 history frames + the message are
// converted to mono f32 in shared memory, each lane produces 10 CONSECUTIVE outputs from a
// 76-float register window (stride 30 floats between lanes = 15 x 8 B, odd, so the 64-bit loads
// are conflict-free; taps come from the constant bank; 480 FMAs per lane, all 32 lanes busy), and the outputs leave
// through shared memory as coalesced 16-byte stores.  Pass 2 is lane = message: each lane re-reads
// ITS message's outputs (still in L2) and runs the reference's strictly sequential float64 chain
// and the other statistics, so the serial chain costs 1/32 of a warp per message instead of a
// whole warp.  Only "computes every 3rd output of the 48-tap filter" is exploited (the polyphase
// saving); no tensor cores: a 48-tap FIR at 16 k outputs/s is 1.5 MFLOP per audio-second, four
// orders below the HBM cost.
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "../../include/gpuaudio_taps.h"

namespace ga {

constexpr int kTaps = GA_RESAMPLE_TAPS;          // 48
constexpr int kHist = kTaps - 1;                 // 47 frames carried between messages
constexpr int kRsWarps = 4;
constexpr int kRsMaxFrames = 960;                // frames per message (20 ms at 48 kHz)
constexpr int kOutPerLane = 10;                  // 32 lanes x 10 outputs = one 20 ms message
constexpr int kWin = 3 * (kOutPerLane - 1) + kTaps + 1;   // 76 floats: a lane window of 75, +1 when it starts at an odd float
constexpr int kRsMaxOut = kRsMaxFrames / 3 + 1;
constexpr int kPad = 48;                         // message frames start at mono[48]: 16-byte aligned rows
// mono[] holds kPad + kRsMaxFrames floats, padded so the 88-float register windows never run off
constexpr int kMonoFloats = ((kPad + kRsMaxFrames + 36 + 3) / 4) * 4 + 96;
constexpr int kYFloats = ((kRsMaxOut + 3) / 4) * 4 + 4;
constexpr int kPcmStage = kRsMaxFrames * 4 + 16;
constexpr int kRsWarpBytes = 2 * kPcmStage + kMonoFloats * 4 + kYFloats * 4 + 64;
// pass 2 staging: slices of 32 outputs of the group's 32 messages, rows padded by 16 B (conflict-free
// lane-per-row LDS.128), two buffers laid over the PCM stages and mono[] that pass 1 is done with
constexpr int kP2Slice = 32;
constexpr int kP2RowStride = kP2Slice * 4 + 16;
constexpr int kP2BufBytes = 32 * kP2RowStride;
static_assert(2 * kP2BufBytes <= 2 * kPcmStage + kMonoFloats * 4, "pass-2 buffers must fit the pass-1 buffers");

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__constant__ uint32_t c_tap_bits[kTaps] = GA_RESAMPLE_TAP_BITS;

#ifndef GA_RS_ROLLCONV
#define GA_RS_ROLLCONV 1
#endif
#ifndef GA_RS_EVICT_FIRST
#define GA_RS_EVICT_FIRST 1
#endif
#ifndef GA_RS_FMA1
#define GA_RS_FMA1 1
#endif
// mono = (A1(L) + A1(R)) * 0.5f for one interleaved stereo frame (L in the low half, R in the high half).
//
// A1(v) = float32(v) / 32767.0f is an IEEE division (audio_service.go:1078).  For the 65,536 possible
// inputs it can be had from ONE fused multiply-add:  with a = v * 2^-15 (exact),
//     RN(v / 32767) == fma(a, c, a),   c = 2^-15 + 2^-30   (exactly representable; 32768/32767 = 1 + c + 2^-45 + …)
// — a * c is exact inside the FMA and the one rounding sees a * (1 + c), whose relative distance from
// the true quotient (2^-45) never reaches a rounding boundary of the result; verified for all 65,536
// values against the division (exact rational arithmetic, and tests/test_gpu_resample.py on the GPU).
// Scaling by a power of two commutes with rounding, so the halves are taken first:
//     (A1(L) + A1(R)) * 0.5f == fma(l, c, l) + fma(r, c, r)   with l = L * 2^-16, r = R * 2^-16,
// and v * 2^-16 costs nothing: the int16, biased by 0x8000, is spliced under the exponent of 2^7
// (ulp 2^-16) and 128.5 is subtracted.  Per frame: LOP3, 2 PRMT, FADD2, FFMA2, FADD.
__device__ __forceinline__ float mono_of(uint32_t lr)
{
    const uint32_t x = lr ^ 0x80008000u;
#if GA_RS_FMA1
    unsigned long long v, q, kb, kc;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "r"(__byte_perm(x, 0x43000000u, 0x7610)), "r"(__byte_perm(x, 0x43000000u, 0x7632)));
    asm("mov.b64 %0, {%1, %1};" : "=l"(kb) : "f"(-128.5f));
    asm("mov.b64 %0, {%1, %1};" : "=l"(kc) : "f"(0x1.0002p-15f));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(v) : "l"(v), "l"(kb));
    asm("fma.rn.f32x2 %0, %1, %2, %1;" : "=l"(q) : "l"(v), "l"(kc));
    float l, r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(r) : "l"(q));
    return __fadd_rn(l, r);
#else
    // the division as multiply + residual correction (common.cuh pcm_to_f32), both channels per instruction
    const float rcp = 1.0f / 32767.0f;
    unsigned long long v, q0, e, q, kb, kr, kn;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "r"(__byte_perm(x, 0x4B000000u, 0x7610)), "r"(__byte_perm(x, 0x4B000000u, 0x7632)));
    asm("mov.b64 %0, {%1, %1};" : "=l"(kb) : "f"(-8421376.0f));
    asm("mov.b64 %0, {%1, %1};" : "=l"(kr) : "f"(rcp));
    asm("mov.b64 %0, {%1, %1};" : "=l"(kn) : "f"(-32767.0f));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(v) : "l"(v), "l"(kb));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(q0) : "l"(v), "l"(kr));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(e) : "l"(kn), "l"(q0), "l"(v));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(q) : "l"(e), "l"(kr), "l"(q0));
    float l, r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(r) : "l"(q));
    return __fmul_rn(__fadd_rn(l, r), 0.5f);
#endif
}

// Outputs are the M with n0 <= 3M+2 < n0 + nfr.  With r0 = n0 mod 3 and M = n0/3 + j that is
// 0 <= 3j + 2 - r0 < nfr: the message yields floor((nfr + r0) / 3) outputs and the newest tap of
// its first output is frame 2 - r0 of the message — only n0 mod 3 matters, which every lane works
// out once for its own message of the group (one 64-bit modulo instead of two per message and lane).
__device__ __forceinline__ uint32_t out_count(uint32_t r0, uint32_t nfr, uint32_t *top0)
{
    if (top0) *top0 = 2u - r0 + kPad;
    return (nfr + r0) / 3u;
}

struct ResampleParams {
    const uint8_t *pcm;
    const ga_rdesc *desc;
    uint8_t *hist_state;     // [max_streams][192] raw stereo frames: the last 47 of each stream
    float *samples;
    ga_result *results;
    ga_frame *frames;
    double vad;
    uint32_t n_msgs;
};

__global__ void __launch_bounds__(kRsWarps * 32)
resample48_kernel(const ResampleParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *wbase = smem + warp * kRsWarpBytes;
    float *mono = reinterpret_cast<float *>(wbase + 2 * kPcmStage);
    float *ybuf = mono + kMonoFloats;
    uint64_t *bars = reinterpret_cast<uint64_t *>(ybuf + kYFloats);
    if (lane == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_mbar_init();
    }
    __syncwarp();

    // the PCM is read once: evict-first in L2, so that the outputs pass 2 comes back for stay there
    const uint64_t pcm_policy = l2_policy_evict_first();
    const uint32_t n_groups = (p.n_msgs + 31) / 32;
    uint32_t job = 0;                                           // messages staged so far by this warp
    for (uint32_t grp = blockIdx.x * kRsWarps + warp; grp < n_groups; grp += gridDim.x * kRsWarps) {
        const uint32_t base = grp * 32;
        const uint32_t cnt = min(32u, p.n_msgs - base);

        // every lane holds the descriptor of "its" message of the group; pass 1 fetches message i's
        // fields from lane i by shuffle, so no global-load latency sits at the head of a message
        ga_rdesc dl;
        memset(&dl, 0, sizeof dl);
        if (static_cast<uint32_t>(lane) < cnt) dl = p.desc[base + lane];
        const uint32_t r0_l = static_cast<uint32_t>(dl.n0 % 3ull);
        constexpr uint32_t FULL = 0xffffffffu;

        // one bulk copy per message, issued one message ahead
        auto issue = [&](uint32_t i, uint32_t j) {
            const uint64_t in_off = __shfl_sync(FULL, dl.in_off, i);
            const uint32_t nb = __shfl_sync(FULL, dl.nbytes, i);
            if (lane == 0) {
                const uint32_t bytes = (min(nb, static_cast<uint32_t>(kRsMaxFrames * 4)) + 15u) & ~15u;
                mbar_arrive_expect_tx(&bars[j & 1], bytes);
#if GA_RS_EVICT_FIRST
                if (bytes) bulk_g2s_hint(wbase + (j & 1) * kPcmStage, p.pcm + in_off, bytes, &bars[j & 1], pcm_policy);
#else
                if (bytes) bulk_g2s(wbase + (j & 1) * kPcmStage, p.pcm + in_off, bytes, &bars[j & 1]);
#endif
            }
        };
        issue(0, job);
        const uint32_t *hsrc_next = nullptr;
        uint32_t h0_next = 0, h1_next = 0;
        auto load_hist = [&](uint32_t i) {
            const uint32_t kind = __shfl_sync(FULL, dl.hist_kind, i);
            const uint64_t hoff = __shfl_sync(FULL, dl.hist_off, i);
            const uint32_t sid = __shfl_sync(FULL, dl.stream_id, i);
            hsrc_next = nullptr;
            if (kind == GA_HIST_SLAB) hsrc_next = reinterpret_cast<const uint32_t *>(p.pcm + hoff);
            else if (kind == GA_HIST_STATE) hsrc_next = reinterpret_cast<const uint32_t *>(p.hist_state + static_cast<uint64_t>(sid) * 192);
            h0_next = 0; h1_next = 0;
            if (hsrc_next) {
                h0_next = hsrc_next[lane];
                if (lane + 32 < kHist) h1_next = hsrc_next[lane + 32];
            }
        };
        load_hist(0);

        // ---------------- pass 1: the whole warp on one message at a time
        for (uint32_t i = 0; i < cnt; i++, job++) {
            if (i + 1 < cnt) issue(i + 1, job + 1);
            ga_rdesc d;
            d.in_off = 0;
            d.out_off = __shfl_sync(FULL, dl.out_off, i);
            d.nbytes = __shfl_sync(FULL, dl.nbytes, i);
            const uint32_t nfr = min(d.nbytes >> 2, static_cast<uint32_t>(kRsMaxFrames));   // a partial frame is dropped
            uint32_t top0;
            const uint32_t n_out = out_count(__shfl_sync(FULL, r0_l, i), nfr, &top0);
            // the usual message: 20 ms = 960 frames -> 320 outputs at a 16-byte aligned place in the arena
            const bool std_msg = nfr == kRsMaxFrames && n_out == kRsMaxFrames / 3 && (d.out_off & 3ull) == 0;

            // history: the 47 frames before the message.  The loads for message i+1 are issued now
            // and consumed one iteration later, so their latency hides behind this message's work.
            const uint32_t *hsrc = hsrc_next;
            const uint32_t h0 = h0_next, h1 = h1_next;
            if (i + 1 < cnt) load_hist(i + 1);
            mbar_wait(&bars[job & 1], (job >> 1) & 1u);
            const uint8_t *msg = wbase + (job & 1) * kPcmStage;

            mono[1 + lane] = hsrc ? mono_of(h0) : 0.0f;
            if (lane + 32 < kHist) mono[33 + lane] = hsrc ? mono_of(h1) : 0.0f;
            if (nfr == kRsMaxFrames) {                          // 240 vectors of 4 frames: 7.5 per lane, no loop
#if GA_RS_ROLLCONV
#pragma unroll 1
#else
#pragma unroll
#endif
                for (int it = 0; it < kRsMaxFrames / 128 + 1; it++) {
                    const uint32_t v = lane + 32 * it;
                    if (it < kRsMaxFrames / 128 || v < kRsMaxFrames / 4) {
                        const uint4 q = reinterpret_cast<const uint4 *>(msg)[v];
                        reinterpret_cast<float4 *>(mono + kPad)[v] = make_float4(mono_of(q.x), mono_of(q.y), mono_of(q.z), mono_of(q.w));
                    }
                }
            } else {
                const uint32_t nvec = (nfr + 3) >> 2;
#pragma unroll 2
                for (uint32_t v = lane; v < nvec; v += 32) {    // frames past nfr in the last vector are never read
                    const uint4 q = reinterpret_cast<const uint4 *>(msg)[v];
                    reinterpret_cast<float4 *>(mono + kPad)[v] = make_float4(mono_of(q.x), mono_of(q.y), mono_of(q.z), mono_of(q.w));
                }
            }
            __syncwarp();

            // FIR: lane produces outputs [10*lane, 10*lane + 10) — 32 lanes x 10 = the 320 outputs of a
            // 20 ms message, every lane busy.  Its window (75 floats) starts at an even float for every
            // lane (lane stride 30), so it is read with 64-bit loads: 15 eight-byte words between lanes,
            // odd, hence conflict-free.
            const uint32_t o0 = kOutPerLane * lane;
            if (o0 < n_out) {
                // newest tap of output o is mono[top0 + 3*o]; oldest is 47 below.  Window base (aligned down to 2):
                const uint32_t lo_idx = top0 + 3 * o0 - kHist;              // oldest float needed
                const uint32_t wbase2 = lo_idx & ~1u;
                const uint32_t sh = lo_idx - wbase2;                        // 0 or 1, the same for every lane
                float win[kWin];
                const float2 *w2 = reinterpret_cast<const float2 *>(mono + wbase2);
#pragma unroll
                for (int k = 0; k < kWin / 2; k++) {
                    const float2 t = w2[k];
                    win[2 * k] = t.x; win[2 * k + 1] = t.y;
                }
                float acc[kOutPerLane];
#pragma unroll
                for (int o = 0; o < kOutPerLane; o++) acc[o] = 0.0f;
                // the window offset `sh` is warp-uniform and takes two values: a branch makes every
                // win[] index a compile-time constant (registers, not local memory)
#define GA_FIR(SH)                                                                              \
    _Pragma("unroll") for (int k = 0; k < kTaps; k++) {                                         \
        const float h = __uint_as_float(c_tap_bits[k]);                                         \
        _Pragma("unroll") for (int o = 0; o < kOutPerLane; o++)                                 \
            acc[o] = __fmaf_rn(h, win[(SH) + 3 * o + kHist - k], acc[o]);                       \
    }
                if (sh == 0) { GA_FIR(0) } else { GA_FIR(1) }
#undef GA_FIR
                // outputs leave through shared memory: pairs as 64-bit stores (5 words between lanes: conflict-free)
                if (o0 + kOutPerLane <= n_out) {
#pragma unroll
                    for (int o = 0; o < kOutPerLane; o += 2)
                        *reinterpret_cast<float2 *>(ybuf + o0 + o) = make_float2(acc[o], acc[o + 1]);
                } else {
#pragma unroll
                    for (int o = 0; o < kOutPerLane; o++)
                        if (o0 + o < n_out) ybuf[o0 + o] = acc[o];
                }
            }
            __syncwarp();

            // coalesced store of the resampled samples
            float *out = p.samples + d.out_off;
            if (std_msg) {                                      // 80 vectors: 2.5 per lane, no loop
#pragma unroll
                for (int it = 0; it < 3; it++) {
                    const uint32_t v = lane + 32 * it;
                    if (it < 2 || v < kRsMaxFrames / 12)
                        *reinterpret_cast<float4 *>(out + 4 * v) = reinterpret_cast<const float4 *>(ybuf)[v];
                }
            } else if ((d.out_off & 3ull) == 0) {
                const uint32_t n4 = n_out >> 2;
                for (uint32_t v = lane; v < n4; v += 32) {
                    const float4 t = reinterpret_cast<const float4 *>(ybuf)[v];
                    *reinterpret_cast<float4 *>(out + 4 * v) = t;
                }
                for (uint32_t k = (n4 << 2) + lane; k < n_out; k += 32) out[k] = ybuf[k];
            } else {
                for (uint32_t k = lane; k < n_out; k += 32) out[k] = ybuf[k];
            }

            // (the carry of the stream's last 47 raw frames into hist_state is NOT done here: another
            // warp of this launch may still have to read the stream's OLD state for an earlier chunk of
            // the batch — resample48_save_hist_kernel runs behind this launch on the same stream)
            __syncwarp();              // mono / ybuf / this PCM stage are free again
        }

        // ---------------- pass 2: lane = message.  calculateSignalStrength (audio_service.go:818-852)
        // over the message's outputs: the float64 sum strictly in index order, non-zero count, peak.
        {
            const bool mine = static_cast<uint32_t>(lane) < cnt;
            const uint32_t mi = base + lane;
            const ga_rdesc &d = dl;
            const uint32_t nfr = min(d.nbytes >> 2, static_cast<uint32_t>(kRsMaxFrames));
            const uint32_t n_out = mine ? out_count(r0_l, nfr, nullptr) : 0u;
            const float *y = p.samples + d.out_off;
            double msg_sum = 0.0, frm_sum = 0.0;
            uint32_t nz = 0, fnz = 0, fn = 0, fidx = 0;
            float pk = 0.0f, fpk = 0.0f;
            // wf: per-frame records wanted (compile-time, so the common no-frames case carries no test per sample)
            auto step = [&](float v, auto wf) {
                const double q = sq_f64(v);
                msg_sum += q;
                const uint32_t z = (v != 0.0f);
                nz += z;
                const float a = fabsf(v);
                pk = fmaxf(pk, a);
                if constexpr (decltype(wf)::value) {
                    frm_sum += q; fnz += z; fpk = fmaxf(fpk, a);
                    if (++fn == GA_FRAME_SAMPLES) {
                        ga_frame f;
                        f.sumsq = frm_sum; f.rms = __dsqrt_rn(__ddiv_rn(frm_sum, static_cast<double>(fn)));
                        f.n_samples = fn; f.nonzero = fnz; f.peak = fpk; f.vad = f.rms > p.vad ? 1u : 0u;
                        p.frames[static_cast<uint64_t>(d.frame_off) + fidx++] = f;
                        frm_sum = 0.0; fnz = 0; fn = 0; fpk = 0.0f;
                    }
                }
            };
            // c samples of this lane's row of a staged slice
            auto consume = [&](const float4 *row, uint32_t c, auto wf) {
                if (c == kP2Slice) {
#pragma unroll
                    for (int k = 0; k < kP2Slice / 4; k++) {
                        const float4 t = row[k];
                        step(t.x, wf); step(t.y, wf); step(t.z, wf); step(t.w, wf);
                    }
                    return;
                }
                uint32_t k = 0;
                for (; k + 4 <= c; k += 4) {
                    const float4 t = row[k >> 2];
                    step(t.x, wf); step(t.y, wf); step(t.z, wf); step(t.w, wf);
                }
                if (k < c) {
                    const float4 t = row[k >> 2];
                    step(t.x, wf);
                    if (k + 1 < c) step(t.y, wf);
                    if (k + 2 < c) step(t.z, wf);
                }
            };
            const bool with_frames = p.frames != nullptr;
            // The outputs were written cooperatively in pass 1 and are read back per lane here.  Each
            // lane reading its own row from L2 exposes the full memory latency once per 16 bytes, so
            // the rows come back through shared memory instead: slices of 32 samples of all 32 messages,
            // fetched by asynchronous copies (cp.async: coalesced, no registers, two slices in flight)
            // into the buffers pass 1 no longer needs.
            const uint32_t max_out = __reduce_max_sync(FULL, n_out);
            const bool staged = __all_sync(FULL, !mine || (d.out_off & 3ull) == 0) && max_out >= kP2Slice;
            if (staged) {
                const uint32_t n_slices = (max_out + kP2Slice - 1) / kP2Slice;
                auto issue_slice = [&](uint32_t sl) {
                    uint8_t *buf = wbase + (sl & 1u) * kP2BufBytes;
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        const int msg_i = i * 4 + (lane >> 3), piece = lane & 7;
                        const uint64_t off_m = __shfl_sync(FULL, d.out_off, msg_i);
                        const uint32_t n_m = __shfl_sync(FULL, n_out, msg_i);
                        const uint32_t s0 = sl * kP2Slice + piece * 4;
                        if (s0 < n_m)
                            cp_async16(buf + msg_i * kP2RowStride + piece * 16, p.samples + off_m + s0);
                    }
                    cp_async_commit();
                };
                issue_slice(0);
                if (n_slices > 1) issue_slice(1);
                for (uint32_t sl = 0; sl < n_slices; sl++) {
                    if (sl + 1 < n_slices) cp_async_wait<1>(); else cp_async_wait<0>();
                    __syncwarp();
                    const uint32_t lo = sl * kP2Slice;
                    if (lo < n_out) {
                        const uint32_t c = min(n_out - lo, static_cast<uint32_t>(kP2Slice));
                        const float4 *row = reinterpret_cast<const float4 *>(wbase + (sl & 1u) * kP2BufBytes + lane * kP2RowStride);
                        if (with_frames) consume(row, c, std::true_type{});
                        else consume(row, c, std::false_type{});
                    }
                    __syncwarp();
                    if (sl + 2 < n_slices) issue_slice(sl + 2);
                }
                // the next group's PCM arrives in these buffers through the TMA (async proxy)
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            } else if (mine) {
                // rare: unaligned outputs or short messages — straight from L2, always with the frame test
                uint32_t k = 0;
                if ((d.out_off & 3ull) == 0) {
                    for (; k + 4 <= n_out; k += 4) {
                        const float4 t = __ldcg(reinterpret_cast<const float4 *>(y + k));
                        if (with_frames) { step(t.x, std::true_type{}); step(t.y, std::true_type{}); step(t.z, std::true_type{}); step(t.w, std::true_type{}); }
                        else { step(t.x, std::false_type{}); step(t.y, std::false_type{}); step(t.z, std::false_type{}); step(t.w, std::false_type{}); }
                    }
                }
                for (; k < n_out; k++) {
                    if (with_frames) step(__ldcg(y + k), std::true_type{});
                    else step(__ldcg(y + k), std::false_type{});
                }
            }
            if (mine) {
                if (p.frames && fn) {
                    ga_frame f;
                    f.sumsq = frm_sum; f.rms = __dsqrt_rn(__ddiv_rn(frm_sum, static_cast<double>(fn)));
                    f.n_samples = fn; f.nonzero = fnz; f.peak = fpk; f.vad = f.rms > p.vad ? 1u : 0u;
                    p.frames[static_cast<uint64_t>(d.frame_off) + fidx] = f;
                }
                ga_result *r = p.results + mi;
                const double rms = n_out ? __dsqrt_rn(__ddiv_rn(msg_sum, static_cast<double>(n_out))) : 0.0;
                r->sumsq = msg_sum; r->rms = rms; r->n_samples = n_out; r->nonzero = nz; r->peak = pk;
                r->flags = (d.flags << 8) | (rms > p.vad ? 1u : 0u);
            }
        }
        __syncwarp();
    }
}


// Carry of the per-stream filter history to the next batch: for every message flagged
// GA_RS_SAVE_HIST (the context flags the stream's LAST chunk of the batch), the 47 stereo frames
// that precede the end of the message go to hist_state[stream_id].  This runs as its own launch
// BEHIND resample48_kernel on the same CUDA stream: inside the main kernel the save of a stream's
// last chunk raced with the GA_HIST_STATE read of the same stream's first chunk of the batch when
// the two sat in different 32-message groups (different warps, no ordering).  Here a stream's old
// state is read (short last chunk: its history is part of the new state) and overwritten by ONE
// warp with a __syncwarp between, and nothing else reads hist_state during this launch.
// A warp scans 32 descriptors, then handles its flagged messages one at a time.
constexpr int kSaveWarps = 8;
__global__ void __launch_bounds__(kSaveWarps * 32)
resample48_save_hist_kernel(const uint8_t *__restrict__ pcm, const ga_rdesc *__restrict__ desc, uint8_t *hist_state,
                            uint32_t n_msgs)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t base = (blockIdx.x * kSaveWarps + (threadIdx.x >> 5)) * 32u;
    if (base >= n_msgs) return;
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t mi = base + lane;
    uint4 tail = make_uint4(0u, 0u, 0u, 0u);               // {nbytes, stream_id, flags, frame_off}
    if (mi < n_msgs) tail = __ldg(reinterpret_cast<const uint4 *>(desc + mi) + 2);
    uint32_t todo = __ballot_sync(FULL, (tail.z & GA_RS_SAVE_HIST) != 0u);
    uint64_t in_off = 0, hist_off = 0;
    uint32_t hist_kind = GA_HIST_ZERO;
    if (todo >> lane & 1u) {
        in_off = desc[mi].in_off;
        hist_off = desc[mi].hist_off;
        hist_kind = desc[mi].hist_kind;
    }
    while (todo) {
        const int i = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint64_t m_in = __shfl_sync(FULL, in_off, i), m_hist = __shfl_sync(FULL, hist_off, i);
        const uint32_t kind = __shfl_sync(FULL, hist_kind, i), nb = __shfl_sync(FULL, tail.x, i), sid = __shfl_sync(FULL, tail.y, i);
        const uint32_t nfr = min(nb >> 2, static_cast<uint32_t>(kRsMaxFrames));
        uint32_t *hs = reinterpret_cast<uint32_t *>(hist_state + static_cast<uint64_t>(sid) * 192);
        const uint32_t *mw = reinterpret_cast<const uint32_t *>(pcm + m_in);
        const uint32_t *hsrc = kind == GA_HIST_SLAB ? reinterpret_cast<const uint32_t *>(pcm + m_hist)
                             : kind == GA_HIST_STATE ? hs : nullptr;
        uint32_t v[2] = {0u, 0u};
#pragma unroll
        for (int t = 0; t < 2; t++) {
            const int j = static_cast<int>(lane) + 32 * t;
            // frame (nfr - 47 + j) of this message, or — for a short message — of the history before it
            const int a = static_cast<int>(nfr) - kHist + j;
            if (j < kHist) v[t] = a >= 0 ? mw[a] : (hsrc ? hsrc[kHist + a] : 0u);
        }
        __syncwarp();          // all reads of the old state are done before it is overwritten
#pragma unroll
        for (int t = 0; t < 2; t++) {
            const int j = static_cast<int>(lane) + 32 * t;
            if (j < kHist) hs[j] = v[t];
        }
        __syncwarp();
    }
}

}  // namespace ga

extern "C" int ga_k_resample48(const void *d_pcm, const ga_rdesc *d_desc, uint32_t n_msgs, void *d_hist_state,
                               float *d_samples, ga_result *d_results, ga_frame *d_frames,
                               double vad_threshold, void *stream)
{
    using namespace ga;
    if (n_msgs == 0) return GA_OK;
    if (!d_pcm || !d_desc || !d_samples || !d_results) { set_error("ga_k_resample48: NULL argument"); return GA_ERR_ARG; }
    if ((reinterpret_cast<uintptr_t>(d_pcm) & 15u) || (reinterpret_cast<uintptr_t>(d_samples) & 15u) ||
        (reinterpret_cast<uintptr_t>(d_desc) & 15u) || (reinterpret_cast<uintptr_t>(d_hist_state) & 15u)) {
        set_error("ga_k_resample48: device pointers must be 16-byte aligned");
        return GA_ERR_ARG;
    }
    ResampleParams p;
    p.pcm = static_cast<const uint8_t *>(d_pcm);
    p.desc = d_desc;
    p.hist_state = static_cast<uint8_t *>(d_hist_state);
    p.samples = d_samples;
    p.results = d_results;
    p.frames = d_frames;
    p.vad = vad_threshold;
    p.n_msgs = n_msgs;
    static thread_local int configured_dev = -1;
    static thread_local int blocks_per_sm = 1;
    constexpr int smem = kRsWarps * kRsWarpBytes;
    int dev = 0;
    GA_CUDA_TRY(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        GA_CUDA_TRY(cudaFuncSetAttribute(resample48_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        GA_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, resample48_kernel, kRsWarps * 32, smem));
        if (blocks_per_sm < 1) blocks_per_sm = 1;
        configured_dev = dev;
    }
    int sms = 0;
    int rc = num_sms(&sms);
    if (rc) return rc;
    uint32_t grid = static_cast<uint32_t>(sms * blocks_per_sm);
    const uint32_t want = ((n_msgs + 31) / 32 + kRsWarps - 1) / kRsWarps;
    if (grid > want) grid = want;
    resample48_kernel<<<grid, kRsWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(p);
    GA_CHECK_LAUNCH("resample48_kernel");
    if (d_hist_state) {
        const uint32_t sgrid = ((n_msgs + 31) / 32 + kSaveWarps - 1) / kSaveWarps;
        resample48_save_hist_kernel<<<sgrid, kSaveWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            p.pcm, d_desc, p.hist_state, n_msgs);
        GA_CHECK_LAUNCH("resample48_save_hist_kernel");
    }
    return GA_OK;
}

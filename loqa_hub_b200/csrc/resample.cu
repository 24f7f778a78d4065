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
// One warp per message.  The 47 history frames + the message's frames are converted to mono
// f32 in shared memory with coalesced 16-byte loads; each lane then produces 12 CONSECUTIVE
// outputs from an 84-float register window (stride 36 floats between lanes = 9 x 16 B, odd, so
// the LDS.128 loads are conflict-free; taps come from the constant bank), 576 FMAs per lane;
// outputs go back through shared memory so the global stores are coalesced 16-byte bursts and
// lane 0 can run the reference's strictly sequential float64 chain.  Only "computes every 3rd
// output of the 48-tap filter" is exploited (the polyphase saving); no tensor cores: a 48-tap
// FIR at 16 k outputs/s is 1.5 MFLOP per audio-second, four orders below the HBM cost.
#include "common.cuh"
#include "../../include/gpuaudio_taps.h"

namespace ga {

constexpr int kTaps = GA_RESAMPLE_TAPS;          // 48
constexpr int kHist = kTaps - 1;                 // 47 frames carried between messages
constexpr int kRsWarps = 4;
constexpr int kRsMaxFrames = 960;                // frames per message handled per pass (20 ms at 48 kHz)
constexpr int kOutPerLane = 12;
constexpr int kRsMaxOut = kRsMaxFrames / 3 + 1;
// mono[] holds kHist + kRsMaxFrames floats, padded so the 84-float register windows never run off
constexpr int kPad = 48;                         // message frames start at mono[48]: 16-byte aligned rows
constexpr int kMonoFloats = ((kPad + kRsMaxFrames + 36 + 3) / 4) * 4 + 96;
constexpr int kYFloats = ((kRsMaxOut + 3) / 4) * 4 + 4;
constexpr int kRsWarpBytes = kMonoFloats * 4 + kYFloats * 4 + 64;

__constant__ uint32_t c_tap_bits[kTaps] = GA_RESAMPLE_TAP_BITS;

__device__ __forceinline__ float mono_of(uint32_t lr)
{
    // lr = one interleaved stereo frame: L in the low half, R in the high half
    const uint32_t x = lr ^ 0x80008000u;
    const float l = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7610)) - 8421376.0f;
    const float r = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7632)) - 8421376.0f;
    return __fmul_rn(__fadd_rn(pcm_to_f32(l), pcm_to_f32(r)), 0.5f);
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
    constexpr uint32_t FULL = 0xffffffffu;
    extern __shared__ __align__(16) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *mono = reinterpret_cast<float *>(smem + warp * kRsWarpBytes);
    float *ybuf = mono + kMonoFloats;

    for (uint32_t mi = blockIdx.x * kRsWarps + warp; mi < p.n_msgs; mi += gridDim.x * kRsWarps) {
        const ga_rdesc d = p.desc[mi];
        const uint32_t nfr = d.nbytes >> 2;                 // whole stereo frames; a partial one is dropped
        const uint64_t n0 = d.n0;                           // absolute index of the first frame
        // outputs M with n0 <= 3M+2 < n0 + nfr
        const uint64_t m_first = (n0 + 0) / 3 + ((3 * (n0 / 3) + 2 < n0) ? 1 : 0);
        const uint64_t n1 = n0 + nfr;
        const uint32_t n_out = (n1 >= 3 * m_first + 3) ? static_cast<uint32_t>((n1 - 3 * m_first - 3) / 3 + 1) : 0u;
        // local index of absolute frame a in mono[] is a - n0 + kPad (history in mono[1..47])
        const uint32_t top0 = static_cast<uint32_t>(3 * m_first + 2 - n0) + kPad;    // mono index of output 0's newest tap

        // ---- history: 47 frames before the message
        const uint32_t *hsrc = nullptr;
        if (d.hist_kind == GA_HIST_SLAB) hsrc = reinterpret_cast<const uint32_t *>(p.pcm + d.hist_off);
        else if (d.hist_kind == GA_HIST_STATE) hsrc = reinterpret_cast<const uint32_t *>(p.hist_state + static_cast<uint64_t>(d.stream_id) * 192);
        // all global loads of the message are issued before any is used (8 x 16 B per lane in flight)
        const uint4 *src = reinterpret_cast<const uint4 *>(p.pcm + d.in_off);
        const uint32_t nvec = (nfr + 3) >> 2;
        uint4 q[kRsMaxFrames / 4 / 32 + 1];
#pragma unroll
        for (int i = 0; i < kRsMaxFrames / 4 / 32 + 1; i++) {
            const uint32_t v = lane + 32 * i;
            q[i] = v < nvec ? src[v] : make_uint4(0, 0, 0, 0);
        }
        uint32_t h0 = 0, h1 = 0;
        if (hsrc) {
            h0 = hsrc[lane];
            if (lane + 32 < kHist) h1 = hsrc[lane + 32];
        }
        mono[1 + lane] = hsrc ? mono_of(h0) : 0.0f;
        if (lane + 32 < kHist) mono[33 + lane] = hsrc ? mono_of(h1) : 0.0f;
#pragma unroll
        for (int i = 0; i < kRsMaxFrames / 4 / 32 + 1; i++) {
            const uint32_t v = lane + 32 * i;
            if (v < nvec)   // frames past nfr in the last vector are never read by a valid output
                reinterpret_cast<float4 *>(mono + kPad)[v] =
                    make_float4(mono_of(q[i].x), mono_of(q[i].y), mono_of(q[i].z), mono_of(q[i].w));
        }
        __syncwarp();

        // ---- FIR: lane produces outputs [12*lane, 12*lane + 12)
        const uint32_t o0 = kOutPerLane * lane;
        if (o0 < n_out) {
            // newest tap of output o is mono[top0 + 3*o]; oldest is 47 below.  Window base (aligned down to 4):
            const uint32_t lo_idx = top0 + 3 * o0 - kHist;              // oldest float needed
            const uint32_t base = lo_idx & ~3u;
            const uint32_t sh = lo_idx - base;                          // 0..3
            float win[88];
            const float4 *w4 = reinterpret_cast<const float4 *>(mono + base);
#pragma unroll
            for (int i = 0; i < 22; i++) {
                const float4 t = w4[i];
                win[4 * i] = t.x; win[4 * i + 1] = t.y; win[4 * i + 2] = t.z; win[4 * i + 3] = t.w;
            }
            float acc[kOutPerLane];
#pragma unroll
            for (int o = 0; o < kOutPerLane; o++) acc[o] = 0.0f;
            // the window offset `sh` is warp-uniform (36*lane is a multiple of 4) and takes only 4
            // values: a switch makes every win[] index a compile-time constant (registers, not local memory)
#define GA_FIR(SH)                                                                              \
    _Pragma("unroll") for (int k = 0; k < kTaps; k++) {                                         \
        const float h = __uint_as_float(c_tap_bits[k]);                                         \
        _Pragma("unroll") for (int o = 0; o < kOutPerLane; o++)                                 \
            acc[o] = __fmaf_rn(h, win[(SH) + 3 * o + kHist - k], acc[o]);                       \
    }
            switch (sh) {
                case 0: GA_FIR(0) break;
                case 1: GA_FIR(1) break;
                case 2: GA_FIR(2) break;
                default: GA_FIR(3) break;
            }
#undef GA_FIR
#pragma unroll
            for (int o = 0; o < kOutPerLane; o++)
                if (o0 + o < n_out) ybuf[o0 + o] = acc[o];
        }
        __syncwarp();

        // ---- coalesced store of the resampled samples
        float *out = p.samples + d.out_off;
        if ((d.out_off & 3ull) == 0) {
            const uint32_t n4 = n_out >> 2;
            for (uint32_t v = lane; v < n4; v += 32) {
                const float4 t = reinterpret_cast<const float4 *>(ybuf)[v];
                st_global_v4(out + 4 * v, t.x, t.y, t.z, t.w);
            }
            for (uint32_t k = (n4 << 2) + lane; k < n_out; k += 32) out[k] = ybuf[k];
        } else {
            for (uint32_t k = lane; k < n_out; k += 32) out[k] = ybuf[k];
        }

        // ---- statistics: order-independent ones and the squares in parallel; only the strictly
        // sequential float64 additions (audio_service.go:827-828) stay on lane 0.  The squares
        // reuse mono[] (the FIR is done with it) as a double array.
        double *sq = reinterpret_cast<double *>(mono);
        uint32_t nz = 0;
        float pk = 0.0f;
        for (uint32_t k = lane; k < n_out; k += 32) {
            const float y = ybuf[k];
            nz += (y != 0.0f);
            pk = fmaxf(pk, fabsf(y));
            sq[k] = sq_f64(y);
        }
        nz = __reduce_add_sync(FULL, nz);
        pk = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(pk)));   // non-negative floats order as uints
        __syncwarp();
        if (lane == 0) {
            double msg_sum = 0.0;
            uint32_t k = 0;
            for (; k + 8 <= n_out; k += 8) {
                double q[8];
#pragma unroll
                for (int j = 0; j < 8; j++) q[j] = sq[k + j];
#pragma unroll
                for (int j = 0; j < 8; j++) msg_sum += q[j];
            }
            for (; k < n_out; k++) msg_sum += sq[k];
            if (p.frames) {
                // 20 ms frames of the output: the same chain restarted every 320 samples
                uint32_t fidx = 0;
                for (uint32_t f0 = 0; f0 < n_out; f0 += GA_FRAME_SAMPLES) {
                    const uint32_t f1 = min(f0 + GA_FRAME_SAMPLES, n_out);
                    double fs = 0.0;
                    uint32_t fnz = 0;
                    float fpk = 0.0f;
                    for (uint32_t j = f0; j < f1; j++) {
                        const float y = ybuf[j];
                        fs += sq[j];
                        fnz += (y != 0.0f);
                        fpk = fmaxf(fpk, fabsf(y));
                    }
                    ga_frame f;
                    f.sumsq = fs; f.rms = __dsqrt_rn(__ddiv_rn(fs, static_cast<double>(f1 - f0)));
                    f.n_samples = f1 - f0; f.nonzero = fnz; f.peak = fpk; f.vad = f.rms > p.vad ? 1u : 0u;
                    p.frames[static_cast<uint64_t>(d.frame_off) + fidx++] = f;
                }
            }
            ga_result *r = p.results + mi;
            const double rms = n_out ? __dsqrt_rn(__ddiv_rn(msg_sum, static_cast<double>(n_out))) : 0.0;
            r->sumsq = msg_sum; r->rms = rms; r->n_samples = n_out; r->nonzero = nz; r->peak = pk;
            r->flags = (d.flags << 8) | (rms > p.vad ? 1u : 0u);
        }

        // ---- carry: the stream's last 47 raw frames, for the first message of the next batch
        if ((d.flags & GA_RS_SAVE_HIST) && p.hist_state) {
            uint32_t *hs = reinterpret_cast<uint32_t *>(p.hist_state + static_cast<uint64_t>(d.stream_id) * 192);
            const uint32_t *msg = reinterpret_cast<const uint32_t *>(p.pcm + d.in_off);
            uint32_t v[2] = {0u, 0u};
#pragma unroll
            for (int t = 0; t < 2; t++) {
                const int j = lane + 32 * t;
                // frame (nfr - 47 + j) of this message, or — for a short message — of the history before it
                const int64_t a = static_cast<int64_t>(nfr) - kHist + j;
                if (j < kHist) v[t] = a >= 0 ? msg[a] : (hsrc ? hsrc[kHist + a] : 0u);
            }
            __syncwarp();          // all reads of the old state are done before it is overwritten
#pragma unroll
            for (int t = 0; t < 2; t++) {
                const int j = lane + 32 * t;
                if (j < kHist) hs[j] = v[t];
            }
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
    const uint32_t want = (n_msgs + kRsWarps - 1) / kRsWarps;
    if (grid > want) grid = want;
    resample48_kernel<<<grid, kRsWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(p);
    GA_CHECK_LAUNCH("resample48_kernel");
    return GA_OK;
}

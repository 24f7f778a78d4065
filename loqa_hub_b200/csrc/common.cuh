// common.cuh — shared device/host helpers for libgpuaudio (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <new>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/gpuaudio.h"

namespace ga {

// ---------------------------------------------------------------- errors (host)
void set_error(const char *fmt, ...);
extern std::atomic<uint64_t> g_launches;

#define GA_CUDA_TRY(expr)                                                             \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            ga::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),     \
                          __FILE__, __LINE__);                                        \
            return GA_ERR_CUDA;                                                       \
        }                                                                             \
    } while (0)

// The C ABI never lets a C++ exception out (it would terminate the hub): every entry point that
// can allocate on the host wraps its body in these.
#define GA_GUARD_BEGIN try {
#define GA_GUARD_END                                                                   \
    } catch (const std::bad_alloc &) {                                                 \
        ga::set_error("%s: out of host memory", __func__);                             \
        return GA_ERR_NOMEM;                                                           \
    } catch (...) {                                                                    \
        ga::set_error("%s: unexpected C++ exception", __func__);                       \
        return GA_ERR_STATE;                                                           \
    }

#define GA_CHECK_LAUNCH(name)                                                         \
    do {                                                                              \
        cudaError_t _e = cudaGetLastError();                                          \
        if (_e != cudaSuccess) {                                                      \
            ga::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));   \
            return GA_ERR_CUDA;                                                       \
        }                                                                             \
        ga::g_launches.fetch_add(1, std::memory_order_relaxed);                       \
    } while (0)

int num_sms(int *out);   // cached per current device

// Library-internal entry points of the context layer used by the hub (csrc/hub.cu).
// ga_submit with the index of the message a refusal is about (UINT32_MAX: not about one message).
int submit_impl(ga_ctx *ctx, const ga_msg *msgs, uint32_t n, uint64_t *ticket, uint32_t *bad_index);
void ctx_limits(const ga_ctx *ctx, uint32_t *max_batch_msgs, uint64_t *frames_cap);
// With no ticket in flight: the ring is free up to `off` (reservations that were dropped, not submitted).
void ring_release_to(ga_ctx *ctx, uint64_t off);

// ---------------------------------------------------------------- PTX: mbarrier + bulk copy
// Shared-memory staging of message bytes uses the TMA engine's 1-D bulk copy
// (cp.async.bulk -> SASS UBLKCP) completing on an mbarrier; one barrier per pipeline
// stage, armed by one lane with the stage's byte total.
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void fence_mbar_init()
{
    // make the barrier initialisation visible to the async proxy
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes,
                                         uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// The same with an L2 eviction policy for the lines it reads: evict_first for data that is streamed
// through once, so that it does not push out lines another part of the kernel will come back for.
__device__ __forceinline__ uint64_t l2_policy_evict_first()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar, uint64_t policy)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}

__device__ __forceinline__ void st_global_v4(float *p, float a, float b, float c, float d)
{
    // streaming store: the f32 samples are written once and not re-read by this kernel
    asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
                 "f"(d)
                 : "memory");
}

// ---------------------------------------------------------------- reference arithmetic
// float32(val) / 32767.0 (audio_service.go:1078) must be an IEEE RN division.  With a
// constant divisor the quotient is reached by one multiply and one residual correction
// (Markstein): q0 = v*r, e = v - 32767*q0 (exact in one fma), q = q0 + e*r.  Verified
// bit-exact against the division for all 65,536 int16 inputs (tests/test_gpu_ingest.py)
// — multiply-by-reciprocal alone is wrong for 1,536 of them.
__device__ __forceinline__ float pcm_to_f32(float v)
{
    const float r = 1.0f / 32767.0f;   // correctly rounded reciprocal (compile-time)
    float q0 = __fmul_rn(v, r);
    float e = __fmaf_rn(-32767.0f, q0, v);
    return __fmaf_rn(e, r, q0);
}

__device__ __forceinline__ float pcm_to_f32(int v) { return pcm_to_f32(static_cast<float>(v)); }

// Two samples per instruction: Blackwell's packed FP32 forms (FMUL2 / FFMA2) perform two
// independent IEEE operations, so the result is bit-identical to two scalar calls; it halves the
// issue slots of the conversion.  sq0/sq1 receive the squares (f32-rounded, audio_service.go:828).
__device__ __forceinline__ void pcm_to_f32x2(float v0, float v1, float &x0, float &x1, float &sq0, float &sq1)
{
    const float r = 1.0f / 32767.0f;
    unsigned long long v, rr, nn, q0, e, q, sq;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(v0), "f"(v1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(rr) : "f"(r));
    asm("mov.b64 %0, {%1, %1};" : "=l"(nn) : "f"(-32767.0f));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(q0) : "l"(v), "l"(rr));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(e) : "l"(nn), "l"(q0), "l"(v));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(q) : "l"(e), "l"(rr), "l"(q0));
    asm("mul.rn.f32x2 %0, %1, %1;" : "=l"(sq) : "l"(q));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x0), "=f"(x1) : "l"(q));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(sq0), "=f"(sq1) : "l"(sq));
}

// float64(sample*sample): product rounded to f32 first, then widened (audio_service.go:828)
__device__ __forceinline__ double sq_f64(float s) { return static_cast<double>(__fmul_rn(s, s)); }

#endif  // __CUDACC__

}  // namespace ga

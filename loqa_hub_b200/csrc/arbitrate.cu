// arbitrate.cu — segmented per-collision-window arbitration for B200 (sm_100a).
//
// Replaces the scoring and selection of performArbitration
// (internal/grpc/audio_service.go:505-571) for many windows per launch:
//   score_kernel : calculateSignalStrength (:818-852) of every contending member's
//                  wake-word buffer.  The reference sums float64(float32(s*s)) strictly in
//                  index order and the winner must never differ, so each LANE owns one
//                  member and adds its samples sequentially; the 32 members of a warp tile
//                  are streamed through shared memory with one TMA bulk copy per member
//                  per slice so global reads stay coalesced 512-byte bursts.
//   select_kernel: per window, the reference's loop — skip non-contending (:522-524),
//                  strict `>` against a running max that starts at 0.0 (:519, :535), first
//                  member when nothing exceeded 0.0 (:542-547) — then Winner / Cancelled
//                  for every member (:553-571).  Join order stands in for Go's random map
//                  iteration order.
#include <cstring>
#include <mutex>

#include "common.cuh"

namespace ga {

constexpr int kArbWarps = 1;                    // 4,096 members are only 128 tiles: one warp per block spreads them over the SMs
constexpr int kArbSlice = 2048;                 // bytes of f32 per member per stage: per-lane bulk copies are issued
                                                // one at a time (~60 cycles each), so few and large
constexpr int kArbStages = 3;
constexpr int kArbRowStride = kArbSlice + 16;   // 129 x 16 B (odd): conflict-free lane-per-row LDS.128
constexpr int kArbStageBytes = 32 * kArbRowStride;
constexpr int kArbWarpBytes = kArbStages * kArbStageBytes + 128;

// Scores of members [0, n_members).  Members >= mirror_first additionally store their score to
// mirror[m - mirror_first] + mirror_byte_off — an address in ANOTHER GPU's memory (NVLink peer
// mapping) when a collision window spans shards; nullptr switches that off.
__device__ __forceinline__ void
score_tiles(const float *__restrict__ samples, const uint64_t *__restrict__ wake_off,
            const uint32_t *__restrict__ wake_n, const uint8_t *__restrict__ status_in,
            uint32_t n_members, double *__restrict__ scores, ga_result *__restrict__ stats,
            uint32_t mirror_first, const uint64_t *__restrict__ mirror, uint64_t mirror_byte_off)
{
    constexpr uint32_t FULL = 0xffffffffu;
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *wbase = smem + warp * kArbWarpBytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(wbase + kArbStages * kArbStageBytes);
    if (lane == 0) {
        for (int s = 0; s < kArbStages; s++) mbar_init(&bars[s], 1);
        fence_mbar_init();
    }
    __syncwarp();

    const uint32_t n_tiles = (n_members + 31) / 32;
    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x * kArbWarps + warp; tile < n_tiles; tile += gridDim.x * kArbWarps) {
        const uint32_t m = tile * 32 + lane;
        const bool active = m < n_members && (!status_in || status_in[m] == GA_RELAY_CONTENDING);
        const uint32_t n = active ? wake_n[m] : 0;
        const uint8_t *src = reinterpret_cast<const uint8_t *>(samples + (active ? wake_off[m] : 0));
        const uint32_t nbytes = n * 4u;
        const uint32_t my_nch = (nbytes + kArbSlice - 1) / kArbSlice;
        const uint32_t nch = __reduce_max_sync(FULL, my_nch);

        auto issue = [&](uint32_t c, uint32_t job) {
            const int stage = job % kArbStages;
            const uint32_t lo = c * kArbSlice;
            uint32_t bytes = nbytes > lo ? min(nbytes - lo, static_cast<uint32_t>(kArbSlice)) : 0u;
            bytes = (bytes + 15u) & ~15u;
            const uint32_t total = __reduce_add_sync(FULL, bytes);
            if (lane == 0) mbar_arrive_expect_tx(&bars[stage], total);
            __syncwarp();
            if (bytes) bulk_g2s(wbase + stage * kArbStageBytes + lane * kArbRowStride, src + lo, bytes, &bars[stage]);
        };

        for (uint32_t c = 0; c < nch && c < kArbStages; c++) issue(c, it + c);
        double sum = 0.0;
        uint32_t nz = 0;       // side statistics the reference logs (:829-838)
        float peak = 0.0f;
        auto add = [&](float x) {
            sum += sq_f64(x);
            nz += (x != 0.0f);
            peak = fmaxf(peak, fabsf(x));
        };
        for (uint32_t c = 0; c < nch; c++) {
            const uint32_t job = it + c;
            const int stage = job % kArbStages;
            mbar_wait(&bars[stage], (job / kArbStages) & 1u);
            const uint32_t lo = c * (kArbSlice / 4);
            if (lo < n) {
                const uint32_t cnt = min(n - lo, static_cast<uint32_t>(kArbSlice / 4));
                const float4 *row = reinterpret_cast<const float4 *>(wbase + stage * kArbStageBytes + lane * kArbRowStride);
                uint32_t k = 0;
                if (!stats) {
                    // arbitration needs only the sum: multiply, widen, add — nothing else per sample
                    for (; k + 4 <= cnt; k += 4) {
                        const float4 v = row[k >> 2];
                        sum += sq_f64(v.x); sum += sq_f64(v.y); sum += sq_f64(v.z); sum += sq_f64(v.w);
                    }
                } else {
                    for (; k + 4 <= cnt; k += 4) {
                        const float4 v = row[k >> 2];
                        add(v.x); add(v.y); add(v.z); add(v.w);
                    }
                }
                if (k < cnt) {
                    const float4 v = row[k >> 2];
                    add(v.x);
                    if (k + 1 < cnt) add(v.y);
                    if (k + 2 < cnt) add(v.z);
                }
            }
            __syncwarp();
            if (c + kArbStages < nch) issue(c + kArbStages, job + kArbStages);
        }
        it += nch;
        if (active) {   // :841; empty buffer -> 0.0 (:819-822)
            const double rms = n ? __dsqrt_rn(__ddiv_rn(sum, static_cast<double>(n))) : 0.0;
            scores[m] = rms;
            if (stats) {
                ga_result *r = stats + m;
                r->sumsq = sum; r->rms = rms; r->n_samples = n; r->nonzero = nz; r->peak = peak; r->flags = 0;
            }
            if (mirror && m >= mirror_first)
                *reinterpret_cast<volatile double *>(mirror[m - mirror_first] + mirror_byte_off) = rms;
        }
    }
}

__global__ void __launch_bounds__(kArbWarps * 32)
score_kernel(const float *__restrict__ samples, const uint64_t *__restrict__ wake_off,
             const uint32_t *__restrict__ wake_n, const uint8_t *__restrict__ status_in,
             uint32_t n_members, double *__restrict__ scores, ga_result *__restrict__ stats)
{
    score_tiles(samples, wake_off, wake_n, status_in, n_members, scores, stats, 0, nullptr, 0);
}

// One window: the reference's selection loop (:519-547) and status update (:553-571).
__device__ __forceinline__ void select_window(const double *__restrict__ scores, const uint8_t *__restrict__ status_in,
                                              const uint32_t *__restrict__ seg, uint32_t w,
                                              uint32_t *__restrict__ winner, uint8_t *__restrict__ status_out)
{
    const uint32_t a = seg[w], b = seg[w + 1];
    uint32_t win = 0xffffffffu;
    double best = 0.0;                                            // :519
    for (uint32_t m = a; m < b; m++) {                            // :521 (join order)
        if (status_in[m] != GA_RELAY_CONTENDING) continue;        // :522-524
        const double s = scores[m];
        if (s > best) { best = s; win = m; }                      // :535 strict >
    }
    if (win == 0xffffffffu && b > a) win = a;                     // :542-547
    winner[w] = win;
    for (uint32_t m = a; m < b; m++)                              // :553-571
        status_out[m] = (m == win) ? GA_RELAY_WINNER : GA_RELAY_CANCELLED;
}

// ---------------------------------------------------------------- fused cross-shard arbitration
// score -> store guest scores into their home shard's inbox over NVLink -> raise a flag there ->
// wait for the shards that owe us scores -> scatter the inbox into member slots -> select:
// ONE launch per shard instead of score + all-gather + route + select.  Blocks score tiles; the
// block that finishes last (atomic ticket) does the exchange and the selection.  The inbox is
// double-buffered by step parity: a neighbour can be at most one step ahead of our selection,
// because its next step cannot finish before it has seen our flag for that step.
__device__ __forceinline__ void st_release_sys(uint64_t addr, uint64_t v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(uint64_t addr)
{
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(addr) : "memory");
    return v;
}

__global__ void __launch_bounds__(kArbWarps * 32)
arbitrate_p2p_kernel(const ga_p2p_args a)
{
    __shared__ bool s_last;
    const uint32_t n_all = a.n_home_members + a.n_guests;
    const uint64_t parity_off = (a.epoch & 1ull) * a.inbox_parity_stride;
    score_tiles(a.d_samples, a.d_wake_off, a.d_wake_n, a.d_status_in, n_all, a.d_scores, nullptr,
                a.n_home_members, a.d_guest_dst, parity_off);
    __threadfence_system();                     // my peer stores are visible before my block counts itself done
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(a.d_counter, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!s_last) return;

    // ---- last block: every score of this shard is written (locally and to the peers)
    __threadfence_system();
    if (threadIdx.x == 0) *a.d_counter = 0;     // ready for the next launch
    for (uint32_t i = threadIdx.x; i < a.n_notify; i += blockDim.x) st_release_sys(a.d_notify[i], a.epoch);
    for (uint32_t i = threadIdx.x; i < a.n_wait; i += blockDim.x) {
        const long long t0 = clock64();
        while (ld_acquire_sys(a.d_wait[i]) < a.epoch) {
            if (clock64() - t0 > a.timeout_cycles) { atomicExch(a.d_error, 1u + i); break; }   // never hang the GPU
            __nanosleep(200);
        }
    }
    __syncthreads();
    const volatile double *inbox = reinterpret_cast<const volatile double *>(
        reinterpret_cast<const uint8_t *>(a.d_inbox) + parity_off);
    for (uint32_t k = threadIdx.x; k < a.n_in; k += blockDim.x) a.d_scores[a.d_route_dst[k]] = inbox[k];
    __syncthreads();
    for (uint32_t w = threadIdx.x; w < a.n_windows; w += blockDim.x)
        select_window(a.d_scores, a.d_status_in, a.d_seg_offsets, w, a.d_winner, a.d_status_out);
}

__global__ void select_kernel(const double *__restrict__ scores, const uint8_t *__restrict__ status_in,
                              const uint32_t *__restrict__ seg, uint32_t n_windows,
                              uint32_t *__restrict__ winner, uint8_t *__restrict__ status_out)
{
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < n_windows) select_window(scores, status_in, seg, w, winner, status_out);
}

__global__ void route_kernel(const double *__restrict__ gathered, const uint32_t *__restrict__ src,
                             const uint32_t *__restrict__ dst, uint32_t n, double *__restrict__ scores)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) scores[dst[i]] = gathered[src[i]];
}

static int launch_select(const double *d_scores, const uint8_t *d_status_in, const uint32_t *d_seg,
                         uint32_t n_windows, uint32_t *d_winner, uint8_t *d_status_out, cudaStream_t st)
{
    if (n_windows == 0) return GA_OK;
    const int threads = 128;
    select_kernel<<<(n_windows + threads - 1) / threads, threads, 0, st>>>(d_scores, d_status_in, d_seg,
                                                                          n_windows, d_winner, d_status_out);
    GA_CHECK_LAUNCH("select_kernel");
    return GA_OK;
}

}  // namespace ga

extern "C" int ga_k_select(const double *d_scores, const uint8_t *d_status_in, const uint32_t *d_seg_offsets,
                           uint32_t n_windows, uint32_t *d_winner, uint8_t *d_status_out, void *stream)
{
    using namespace ga;
    if (n_windows && (!d_scores || !d_status_in || !d_seg_offsets || !d_winner || !d_status_out)) {
        set_error("ga_k_select: NULL argument");
        return GA_ERR_ARG;
    }
    return launch_select(d_scores, d_status_in, d_seg_offsets, n_windows, d_winner, d_status_out,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int ga_k_route_scores(const double *d_gathered, const uint32_t *d_src_idx,
                                 const uint32_t *d_dst_slot, uint32_t n, double *d_scores, void *stream)
{
    using namespace ga;
    if (n == 0) return GA_OK;
    if (!d_gathered || !d_src_idx || !d_dst_slot || !d_scores) { set_error("ga_k_route_scores: NULL argument"); return GA_ERR_ARG; }
    route_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(d_gathered, d_src_idx, d_dst_slot, n, d_scores);
    GA_CHECK_LAUNCH("route_kernel");
    return GA_OK;
}

extern "C" int ga_k_arbitrate_p2p(const ga_p2p_args *args, void *stream)
{
    using namespace ga;
    if (!args) { set_error("ga_k_arbitrate_p2p: NULL argument"); return GA_ERR_ARG; }
    const ga_p2p_args &a = *args;
    // a shard without windows of its own still has to score and send its guests
    if (a.n_windows == 0 && a.n_guests == 0 && a.n_notify == 0) return GA_OK;
    const uint32_t n_all = a.n_home_members + a.n_guests;
    if ((a.n_windows && (!a.d_seg_offsets || !a.d_winner)) || !a.d_counter || !a.d_error ||
        (n_all && (!a.d_samples || !a.d_wake_off || !a.d_wake_n || !a.d_status_in || !a.d_scores)) ||
        (a.n_home_members && !a.d_status_out) ||
        (a.n_guests && !a.d_guest_dst) || (a.n_notify && !a.d_notify) || (a.n_wait && !a.d_wait) ||
        (a.n_in && (!a.d_inbox || !a.d_route_dst))) {
        set_error("ga_k_arbitrate_p2p: NULL array in arguments");
        return GA_ERR_ARG;
    }
    if (a.epoch == 0) { set_error("ga_k_arbitrate_p2p: epoch must start at 1 and increase by 1 per step"); return GA_ERR_ARG; }
    static thread_local int configured_dev = -1;
    int dev = 0;
    GA_CUDA_TRY(cudaGetDevice(&dev));
    constexpr int smem = kArbWarps * kArbWarpBytes;
    if (configured_dev != dev) {
        GA_CUDA_TRY(cudaFuncSetAttribute(arbitrate_p2p_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured_dev = dev;
    }
    const uint32_t n_tiles = (n_all + 31) / 32;
    uint32_t grid = (n_tiles + kArbWarps - 1) / kArbWarps;
    if (grid == 0) grid = 1;
    arbitrate_p2p_kernel<<<grid, kArbWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(a);
    GA_CHECK_LAUNCH("arbitrate_p2p_kernel");
    return GA_OK;
}

// Peer-visible device memory for the inbox + flags (cudaMalloc + CUDA IPC handle).
extern "C" int ga_p2p_alloc(uint64_t bytes, void **d_ptr, uint8_t handle[64])
{
    using namespace ga;
    if (!d_ptr || !handle || bytes == 0) { set_error("ga_p2p_alloc: bad argument"); return GA_ERR_ARG; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); set_error("ga_p2p_alloc: %s", cudaGetErrorString(e)); return GA_ERR_NOMEM; }
    GA_CUDA_TRY(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); cudaGetLastError(); set_error("ga_p2p_alloc: cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); return GA_ERR_CUDA; }
    memcpy(handle, &h, 64);
    *d_ptr = p;
    return GA_OK;
}

extern "C" int ga_p2p_open(const uint8_t handle[64], void **d_ptr)
{
    using namespace ga;
    if (!d_ptr || !handle) { set_error("ga_p2p_open: bad argument"); return GA_ERR_ARG; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    GA_CUDA_TRY(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return GA_OK;
}

extern "C" int ga_p2p_close(void *d_ptr)
{
    using namespace ga;
    if (d_ptr) GA_CUDA_TRY(cudaIpcCloseMemHandle(d_ptr));
    return GA_OK;
}

extern "C" int ga_p2p_free(void *d_ptr)
{
    using namespace ga;
    if (d_ptr) GA_CUDA_TRY(cudaFree(d_ptr));
    return GA_OK;
}

static int launch_scores(const float *d_samples, const uint64_t *d_wake_off, const uint32_t *d_wake_n,
                         const uint8_t *d_status_in, uint32_t n_members, double *d_scores,
                         ga_result *d_stats, cudaStream_t st)
{
    using namespace ga;
    if (n_members == 0) return GA_OK;
    static thread_local int configured_dev = -1;
    int dev = 0;
    GA_CUDA_TRY(cudaGetDevice(&dev));
    constexpr int smem = kArbWarps * kArbWarpBytes;
    if (configured_dev != dev) {
        GA_CUDA_TRY(cudaFuncSetAttribute(score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured_dev = dev;
    }
    const uint32_t n_tiles = (n_members + 31) / 32;
    const uint32_t grid = (n_tiles + kArbWarps - 1) / kArbWarps;
    score_kernel<<<grid, kArbWarps * 32, smem, st>>>(d_samples, d_wake_off, d_wake_n, d_status_in,
                                                     n_members, d_scores, d_stats);
    GA_CHECK_LAUNCH("score_kernel");
    return GA_OK;
}

extern "C" int ga_k_signal_strength(const float *d_samples, const uint64_t *d_wake_off,
                                    const uint32_t *d_wake_n, const uint8_t *d_status_in,
                                    uint32_t n_members, double *d_scores, ga_result *d_stats, void *stream)
{
    using namespace ga;
    if (n_members == 0) return GA_OK;
    if (!d_samples || !d_wake_off || !d_wake_n || !d_scores) {
        set_error("ga_k_signal_strength: NULL argument");
        return GA_ERR_ARG;
    }
    if (reinterpret_cast<uintptr_t>(d_samples) & 15u) {
        set_error("ga_k_signal_strength: d_samples must be 16-byte aligned");
        return GA_ERR_ARG;
    }
    return launch_scores(d_samples, d_wake_off, d_wake_n, d_status_in, n_members, d_scores, d_stats,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int ga_k_arbitrate(const float *d_samples, const uint64_t *d_wake_off, const uint32_t *d_wake_n,
                              const uint8_t *d_status_in, const uint32_t *d_seg_offsets, uint32_t n_windows,
                              uint32_t n_members, double *d_scores, uint32_t *d_winner,
                              uint8_t *d_status_out, void *stream)
{
    using namespace ga;
    if (n_windows == 0) return GA_OK;
    if (!d_seg_offsets || !d_winner) { set_error("ga_k_arbitrate: NULL argument"); return GA_ERR_ARG; }
    if (n_members && (!d_samples || !d_wake_off || !d_wake_n || !d_status_in || !d_scores || !d_status_out)) {
        set_error("ga_k_arbitrate: NULL argument");
        return GA_ERR_ARG;
    }
    if (reinterpret_cast<uintptr_t>(d_samples) & 15u) {
        set_error("ga_k_arbitrate: d_samples must be 16-byte aligned");
        return GA_ERR_ARG;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = launch_scores(d_samples, d_wake_off, d_wake_n, d_status_in, n_members, d_scores, nullptr, st);
    if (rc) return rc;
    return launch_select(d_scores, d_status_in, d_seg_offsets, n_windows, d_winner, d_status_out, st);
}

// ---------------------------------------------------------------- arbitration without allocation
// Scratch for ga_k_arbitrate_noalloc: the per-member float64 scores that sit between the
// scoring launch and the selection launch (512 KiB).  It is part of the library's device image:
// every GPU on which the CUDA runtime has loaded that image holds one copy until the process
// exits.  With CUDA's default lazy module loading that happens at the first use of the library
// on a device, which can be this entry point's own first call there — the set-up below.
#define GA_NOALLOC_MAX_MEMBERS 65536u
static __device__ __align__(256) double g_noalloc_scores[GA_NOALLOC_MAX_MEMBERS];

namespace {
constexpr int kMaxDevices = 64;
double *g_noalloc_ptr[kMaxDevices];     // device address of g_noalloc_scores, per device ordinal
bool g_noalloc_ready[kMaxDevices];
std::mutex g_noalloc_mu;
}  // namespace

// Winners and statuses of many collision windows for callers that must not allocate.
// Same inputs, rules and outputs as ga_k_arbitrate, except that the scores are not returned:
// they are kept in g_noalloc_scores.  See include/gpuaudio.h for the contract.
extern "C" int ga_k_arbitrate_noalloc(const float *d_samples, const uint64_t *d_wake_off,
                                      const uint32_t *d_wake_n, const uint8_t *d_status_in,
                                      const uint32_t *d_seg_offsets, uint32_t n_windows, uint32_t n_members,
                                      uint32_t *d_winner, uint8_t *d_status_out, void *stream)
{
    using namespace ga;
    int dev = 0;
    GA_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { set_error("ga_k_arbitrate_noalloc: device ordinal %d not supported", dev); return GA_ERR_ARG; }
    double *scores = nullptr;
    {
        // Set-up, once per device, with that device current: the name g_noalloc_scores is only a
        // host-side handle, a kernel needs the address the symbol has on THIS device.  Done before
        // anything else so that a call with n_windows == 0 is a pure warm-up: under lazy module
        // loading this lookup is where the driver may place the library's image (this array
        // included) on the device, which allocates and synchronises — the one thing a caller that
        // cannot allocate has to get out of the way at start-up.
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ready[dev]) {
            double *p = nullptr;
            GA_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_scores));
            g_noalloc_ptr[dev] = p;
            g_noalloc_ready[dev] = true;
        }
        scores = g_noalloc_ptr[dev];
    }
    if (n_windows == 0) return GA_OK;
    if (!d_seg_offsets || !d_winner) { set_error("ga_k_arbitrate_noalloc: NULL argument"); return GA_ERR_ARG; }
    if (n_members && (!d_samples || !d_wake_off || !d_wake_n || !d_status_in || !d_status_out)) {
        set_error("ga_k_arbitrate_noalloc: NULL argument");
        return GA_ERR_ARG;
    }
    if (n_members > GA_NOALLOC_MAX_MEMBERS) {
        set_error("ga_k_arbitrate_noalloc: %u window members exceed the built-in scratch of %u; "
                  "split the windows over several calls or use ga_k_arbitrate with your own score array",
                  n_members, GA_NOALLOC_MAX_MEMBERS);
        return GA_ERR_OVERFLOW;
    }
    if (reinterpret_cast<uintptr_t>(d_samples) & 15u) {
        set_error("ga_k_arbitrate_noalloc: d_samples must be 16-byte aligned");
        return GA_ERR_ARG;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // The array still holds the previous call's scores: start this call's slots at 0.0 so a member
    // the scorer skips (not contending) reads as "no signal", never as someone else's old score.
    if (n_members) GA_CUDA_TRY(cudaMemsetAsync(scores, 0, sizeof(double) * n_members, st));
    int rc = launch_scores(d_samples, d_wake_off, d_wake_n, d_status_in, n_members, scores, nullptr, st);
    if (rc) return rc;
    return launch_select(scores, d_status_in, d_seg_offsets, n_windows, d_winner, d_status_out, st);
}

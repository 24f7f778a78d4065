// context.cu — host side of libgpuaudio: the C-owned pinned ring, batch submit/wait,
// the per-stream sample arena, and the one-shot mirrors of the Go functions.
//
// This is the layer the cgo package internal/gpuaudio binds.  It keeps the
// reference's observable behaviour for the path: degenerate inputs (empty / odd-length
// chunks, empty wake-word buffers) produce empty / 0.0 results and never an error
// (internal/grpc/audio_service.go:1050-1063, :819-822); errors are reserved for misuse of
// the boundary and CUDA failures.  There is no CPU fallback anywhere in this file.
#include <algorithm>
#include <atomic>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <thread>
#include <utility>
#include <vector>

#include "common.cuh"
#include "synth.h"

namespace ga {

// ---------------------------------------------------------------- errors / globals
static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof t_err, fmt, ap);
    va_end(ap);
}

int num_sms(int *out)
{
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    GA_CUDA_TRY(cudaGetDevice(&dev));
    if (dev != cached_dev) {
        GA_CUDA_TRY(cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev));
        cached_dev = dev;
    }
    *out = cached;
    return GA_OK;
}

// ---------------------------------------------------------------- context
struct StreamState {
    uint64_t wake_n = 0;     // samples in the wake-word region   (RelayStream.WakeWordSignal)
    uint64_t speech_n = 0;   // samples in the speech region      (RelayStream.SpeechAudio)
    uint64_t n_frames = 0;   // 48 kHz input only: stereo frames consumed since the stream start
    uint64_t stamp = 0;      // ticket of the batch that last appended to the stream
};

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    cudaEvent_t kdone = nullptr;             // kernels of this ticket finished (48 kHz input: orders the state carry)
    ga_desc *h_desc = nullptr, *d_desc = nullptr;
    ga_rdesc *h_rdesc = nullptr, *d_rdesc = nullptr;   // 48 kHz input only
    uint32_t *h_link = nullptr, *d_link = nullptr;     // GA_DO_UTTERANCE: next[n] then run_first[n]
    ga_result *h_res = nullptr, *d_res = nullptr;
    ga_frame *h_frames = nullptr, *d_frames = nullptr;
    uint8_t *h_tick = nullptr, *d_tick = nullptr;      // ga_submit_ticks: ids | wake0 | speech0 | flags
    uint64_t ticket = 0;
    uint64_t release_to = 0;
    uint32_t n_msgs = 0, n_frames = 0;
    bool busy = false;
};

}  // namespace ga

struct ga_ctx {
    ga_config cfg;
    uint8_t *h_ring = nullptr, *d_ring = nullptr;
    float *d_arena = nullptr;
    uint8_t *d_hist = nullptr;                 // 48 kHz input: last 47 raw frames per stream (192 B each)
    ga_sha_state *d_sha = nullptr;             // GA_DO_UTTERANCE: running digest per stream
    std::vector<uint64_t> batch_stamp;         // 48 kHz input: ticket that last touched a stream ...
    std::vector<uint32_t> batch_last;          // ... and the index of that stream's latest message in it
    std::vector<uint64_t> tick_stamp;          // ga_submit_ticks: duplicate-stream check
    std::vector<std::pair<uint32_t, uint64_t>> touched;   // ga_submit: (stream, stamp before this batch), for undo
    uint64_t frames_cap = 0;
    std::atomic<uint64_t> ring_head{0}, ring_tail{0};
    std::vector<ga::StreamState> streams;
    std::vector<ga::Slot> slots;
    uint64_t next_ticket = 1, oldest_ticket = 1;
    cudaEvent_t prev_kdone = nullptr;
    uint64_t fence_ticket = 0;                 // tickets up to this one must finish before the next batch's kernels
                                               // start: a restarted stream reuses arena space they may still write
    uint8_t *d_arb = nullptr;                  // grow-only device scratch of ga_arbitrate (no per-call cudaMalloc)
    uint8_t *h_arb = nullptr;                  // its pinned host mirror: one copy up, one copy down per call
    size_t arb_bytes = 0;
    std::mutex mu;   // guards submit/wait/arbitrate bookkeeping (not the ring reserve path)
    // scratch for ga_arbitrate
    cudaStream_t aux = nullptr;
    // Readers of the sample arena on `aux` (arbitration, the winner path) queue their work under `mu`
    // and wait for it WITHOUT `mu`, so the batching thread's ga_submit is never held up by a window
    // close.  A later submit that restarts a stream must not overwrite samples such a reader is still
    // reading: its CUDA stream waits for aux_ev (recorded behind the reader's work).
    cudaEvent_t aux_ev = nullptr;
    bool aux_used = false;
    std::mutex arb_mu;      // one ga_arbitrate at a time (they share d_arb / h_arb)
    std::mutex win_mu;      // one winner-path call at a time (they share d_win / h_win)
    uint8_t *d_win = nullptr, *h_win = nullptr;   // grow-only staging of ga_fetch_wavs / ga_samples_digests
    size_t win_bytes = 0;
};

namespace ga {

static int set_device(const ga_ctx *ctx)
{
    GA_CUDA_TRY(cudaSetDevice(ctx->cfg.device));
    return GA_OK;
}

static void destroy(ga_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->cfg.device);
    for (auto &s : c->slots) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        if (s.done) cudaEventDestroy(s.done);
        if (s.kdone) cudaEventDestroy(s.kdone);
        if (s.h_desc) cudaFreeHost(s.h_desc);
        if (s.d_desc) cudaFree(s.d_desc);
        if (s.h_link) cudaFreeHost(s.h_link);
        if (s.d_link) cudaFree(s.d_link);
        if (s.h_rdesc) cudaFreeHost(s.h_rdesc);
        if (s.d_rdesc) cudaFree(s.d_rdesc);
        if (s.h_res) cudaFreeHost(s.h_res);
        if (s.d_res) cudaFree(s.d_res);
        if (s.h_frames) cudaFreeHost(s.h_frames);
        if (s.d_frames) cudaFree(s.d_frames);
        if (s.h_tick) cudaFreeHost(s.h_tick);
        if (s.d_tick) cudaFree(s.d_tick);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    if (c->aux) cudaStreamDestroy(c->aux);
    if (c->aux_ev) cudaEventDestroy(c->aux_ev);
    if (c->d_win) cudaFree(c->d_win);
    if (c->h_win) cudaFreeHost(c->h_win);
    if (c->h_ring) cudaFreeHost(c->h_ring);
    if (c->d_ring) cudaFree(c->d_ring);
    if (c->d_arena) cudaFree(c->d_arena);
    if (c->d_hist) cudaFree(c->d_hist);
    if (c->d_sha) cudaFree(c->d_sha);
    if (c->d_arb) cudaFree(c->d_arb);
    if (c->h_arb) cudaFreeHost(c->h_arb);
    delete c;
}

// ga_submit_ticks side buffer: five u32 arrays of up to cap streams (ids, wake0, speech0 and, for
// 48 kHz input, the two halves of the stream's frame count), then cap flag bytes
static size_t tick_bytes(uint32_t cap) { return (static_cast<size_t>(cap) * 21u + 63u) & ~static_cast<size_t>(63); }

#define GA_ALLOC_TRY(expr)                                                            \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            ga::set_error("%s failed: %s", #expr, cudaGetErrorString(_e));            \
            cudaGetLastError();                                                       \
            ga::destroy(c);                                                           \
            return _e == cudaErrorMemoryAllocation ? GA_ERR_NOMEM : GA_ERR_CUDA;      \
        }                                                                             \
    } while (0)

}  // namespace ga

using namespace ga;

extern "C" const char *ga_last_error(void) { return t_err; }
extern "C" uint64_t ga_launch_count(void) { return g_launches.load(); }
extern "C" int ga_abi_version(void) { return GA_ABI_VERSION; }
extern "C" void ga_abi_sizes(uint32_t out[7])
{
    out[0] = sizeof(ga_result); out[1] = sizeof(ga_frame); out[2] = sizeof(ga_desc);
    out[3] = sizeof(ga_msg); out[4] = sizeof(ga_config); out[5] = sizeof(ga_rdesc);
    out[6] = sizeof(ga_sha_state);
}
extern "C" int ga_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

static_assert(sizeof(ga_result) == 64, "ga_result layout");
static_assert(sizeof(ga_frame) == 32, "ga_frame layout");
static_assert(sizeof(ga_desc) == 32, "ga_desc layout");
static_assert(sizeof(ga_msg) == 24, "ga_msg layout");
static_assert(sizeof(ga_rdesc) == 64, "ga_rdesc layout");
static_assert(sizeof(ga_config) == 64, "ga_config layout");
static_assert(sizeof(ga_sha_state) == 128, "ga_sha_state layout");

static int arb_reserve(ga_ctx *ctx, size_t need);
namespace ga {
static int win_reserve(ga_ctx *ctx, size_t need);
static int winners_impl(ga_ctx *ctx, const uint32_t *stream_ids, uint32_t n, bool want_wav, uint32_t sample_rate,
                        bool want_digest, const uint8_t **base, uint64_t *wav_off, const uint8_t **digests);
}

extern "C" int ga_init(const ga_config *cfg, ga_ctx **out)
{
    GA_GUARD_BEGIN
    if (!cfg || !out) { set_error("ga_init: NULL argument"); return GA_ERR_ARG; }
    *out = nullptr;
    if (cfg->max_streams == 0 || cfg->max_batch_msgs == 0 || cfg->ring_bytes < 4096 ||
        (cfg->ring_bytes & 15u) || (cfg->mode & (GA_DO_ALL | GA_DO_UTTERANCE)) == 0) {
        set_error("ga_init: bad config (streams/batch/ring/mode)");
        return GA_ERR_ARG;
    }
    if ((cfg->mode & GA_DO_CONVERT) &&
        (cfg->arena_samples == 0 || (cfg->arena_samples & 3u) || (cfg->wake_samples_max & 3u) ||
         cfg->wake_samples_max >= cfg->arena_samples)) {
        set_error("ga_init: arena_samples / wake_samples_max must be multiples of 4 with wake < arena");
        return GA_ERR_ARG;
    }
    if (cfg->input_format > GA_FMT_STEREO48K ||
        (cfg->input_format == GA_FMT_STEREO48K && !(cfg->mode & GA_DO_CONVERT))) {
        set_error("ga_init: bad input_format (48 kHz stereo input needs GA_DO_CONVERT)");
        return GA_ERR_ARG;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("ga_init: no usable CUDA device (%s); this library has no CPU fallback",
                  e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        return GA_ERR_CUDA;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { set_error("ga_init: device %d out of range", cfg->device); return GA_ERR_ARG; }
    ga_ctx *c = new (std::nothrow) ga_ctx;
    if (!c) { set_error("ga_init: out of host memory"); return GA_ERR_NOMEM; }
    c->cfg = *cfg;
    if (c->cfg.max_inflight == 0) c->cfg.max_inflight = 1;
    if (c->cfg.vad_threshold == 0.0) c->cfg.vad_threshold = 0.01;
    GA_ALLOC_TRY(cudaSetDevice(cfg->device));
    GA_ALLOC_TRY(cudaHostAlloc(reinterpret_cast<void **>(&c->h_ring), cfg->ring_bytes, cudaHostAllocDefault));
    GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&c->d_ring), cfg->ring_bytes + 16));
    if (cfg->mode & GA_DO_CONVERT) {
        const uint64_t arena = static_cast<uint64_t>(cfg->max_streams) * cfg->arena_samples;
        GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&c->d_arena), (arena + 4) * sizeof(float)));
    }
    const bool fmt48 = cfg->input_format == GA_FMT_STEREO48K;
    if (fmt48) {
        GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&c->d_hist), static_cast<size_t>(cfg->max_streams) * 192));
        GA_ALLOC_TRY(cudaMemset(c->d_hist, 0, static_cast<size_t>(cfg->max_streams) * 192));
    }
    const bool utter = (cfg->mode & GA_DO_UTTERANCE) != 0;
    if (utter) {
        GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&c->d_sha), sizeof(ga_sha_state) * cfg->max_streams));
        GA_ALLOC_TRY(cudaMemset(c->d_sha, 0, sizeof(ga_sha_state) * cfg->max_streams));
        if (ga_k_sha256_reset(c->d_sha, cfg->max_streams, nullptr) != GA_OK) { destroy(c); return GA_ERR_CUDA; }
        GA_ALLOC_TRY(cudaDeviceSynchronize());
    }
    try {
        c->streams.resize(cfg->max_streams);
        c->slots.resize(c->cfg.max_inflight);
        if (fmt48 || utter) { c->batch_stamp.assign(cfg->max_streams, 0); c->batch_last.assign(cfg->max_streams, 0); }
    } catch (...) {
        set_error("ga_init: out of host memory");
        destroy(c);
        return GA_ERR_NOMEM;
    }
    c->frames_cap = cfg->want_frames ? cfg->max_batch_msgs + cfg->ring_bytes / (2 * GA_FRAME_SAMPLES) : 0;
    for (auto &s : c->slots) {
        GA_ALLOC_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        GA_ALLOC_TRY(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        GA_ALLOC_TRY(cudaEventCreateWithFlags(&s.kdone, cudaEventDisableTiming));
        GA_ALLOC_TRY(cudaHostAlloc(reinterpret_cast<void **>(&s.h_desc), sizeof(ga_desc) * cfg->max_batch_msgs, cudaHostAllocDefault));
        GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&s.d_desc), sizeof(ga_desc) * cfg->max_batch_msgs));
        if (utter) {
            GA_ALLOC_TRY(cudaHostAlloc(reinterpret_cast<void **>(&s.h_link), sizeof(uint32_t) * 2 * cfg->max_batch_msgs, cudaHostAllocDefault));
            GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&s.d_link), sizeof(uint32_t) * 2 * cfg->max_batch_msgs));
        }
        if (fmt48) {
            GA_ALLOC_TRY(cudaHostAlloc(reinterpret_cast<void **>(&s.h_rdesc), sizeof(ga_rdesc) * cfg->max_batch_msgs, cudaHostAllocDefault));
            GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&s.d_rdesc), sizeof(ga_rdesc) * cfg->max_batch_msgs));
        }
        GA_ALLOC_TRY(cudaHostAlloc(reinterpret_cast<void **>(&s.h_res), sizeof(ga_result) * cfg->max_batch_msgs, cudaHostAllocDefault));
        GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&s.d_res), sizeof(ga_result) * cfg->max_batch_msgs));
        if (c->frames_cap) {
            GA_ALLOC_TRY(cudaHostAlloc(reinterpret_cast<void **>(&s.h_frames), sizeof(ga_frame) * c->frames_cap, cudaHostAllocDefault));
            GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&s.d_frames), sizeof(ga_frame) * c->frames_cap));
        }
        {
            GA_ALLOC_TRY(cudaHostAlloc(reinterpret_cast<void **>(&s.h_tick), tick_bytes(cfg->max_batch_msgs), cudaHostAllocDefault));
            GA_ALLOC_TRY(cudaMalloc(reinterpret_cast<void **>(&s.d_tick), tick_bytes(cfg->max_batch_msgs)));
        }
    }
    {
        try { c->tick_stamp.assign(cfg->max_streams, 0); }
        catch (...) { set_error("ga_init: out of host memory"); destroy(c); return GA_ERR_NOMEM; }
    }
    GA_ALLOC_TRY(cudaStreamCreateWithFlags(&c->aux, cudaStreamNonBlocking));
    GA_ALLOC_TRY(cudaEventCreateWithFlags(&c->aux_ev, cudaEventDisableTiming));
    if ((cfg->mode & GA_DO_CONVERT) && arb_reserve(c, 26ull * cfg->max_streams + 4096) != GA_OK) { destroy(c); return GA_ERR_NOMEM; }
    if (cfg->mode & GA_DO_CONVERT) {
        // The winner path (ga_fetch_wavs / ga_samples_digests) runs at window close, in real time: its
        // staging is sized here for a window close of every eighth stream with its wake-word buffer and
        // as much speech again (it still grows on demand), and its kernels are launched once on an
        // empty stream so that neither an allocation nor a module load lands in the first window close.
        const size_t per = 44 + 8 * static_cast<size_t>(cfg->wake_samples_max) + 32;
        const size_t want = std::min<size_t>(512ull * cfg->max_streams + per * (cfg->max_streams / 8 + 1), 96ull << 20);
        if (win_reserve(c, want) != GA_OK) { destroy(c); return GA_ERR_NOMEM; }
        const uint32_t sid = 0;
        const uint8_t *b = nullptr, *dg = nullptr;
        uint64_t off[2];
        if (winners_impl(c, &sid, 1, true, 16000, true, &b, off, &dg) != GA_OK) { destroy(c); return GA_ERR_CUDA; }
        // the same for the window-close kernels (ga_arbitrate): one window of one stream without audio
        const uint32_t seg[2] = {0, 1};
        const uint8_t st_in = GA_RELAY_CONTENDING;
        uint8_t st_out = 0;
        uint32_t win = 0;
        double sc = 0.0;
        if (ga_arbitrate(c, seg, 1, &sid, &st_in, &win, &sc, &st_out) != GA_OK) { destroy(c); return GA_ERR_CUDA; }
    }
    *out = c;
    return GA_OK;
    GA_GUARD_END
}

extern "C" void ga_shutdown(ga_ctx *ctx) { destroy(ctx); }

// ---------------------------------------------------------------- pinned ring (lock-free MPSC bump)
extern "C" int ga_ring_reserve(ga_ctx *ctx, uint64_t nbytes, void **host_ptr, uint64_t *ring_offset)
{
    if (!ctx || !host_ptr || !ring_offset) { set_error("ga_ring_reserve: NULL argument"); return GA_ERR_ARG; }
    const uint64_t size = ctx->cfg.ring_bytes;
    uint64_t need = (nbytes + 15u) & ~15ull;
    if (need == 0) need = 16;
    if (need > size) { set_error("ga_ring_reserve: %llu bytes exceed the ring", (unsigned long long)nbytes); return GA_ERR_ARG; }
    uint64_t head = ctx->ring_head.load(std::memory_order_relaxed);
    for (;;) {
        uint64_t start = head;
        const uint64_t phys = start % size;
        if (phys + need > size) start += size - phys;          // never split a message across the wrap
        const uint64_t end = start + need;
        if (end - ctx->ring_tail.load(std::memory_order_acquire) > size) {
            set_error("ga_ring_reserve: ring full (wait for outstanding tickets)");
            return GA_ERR_RING_FULL;
        }
        if (ctx->ring_head.compare_exchange_weak(head, end, std::memory_order_acq_rel)) {
            *ring_offset = start;
            *host_ptr = ctx->h_ring + start % size;
            return GA_OK;
        }
    }
}

// ---------------------------------------------------------------- submit / wait
// Tickets run on their own CUDA streams.  When a batch restarts a stream (GA_MSG_FIRST, or after
// ga_stream_reset) whose previous chunks belong to a ticket still in flight, that ticket's kernel
// may still be writing the arena space the new utterance starts over in: order the two.
static int fence_wait(ga_ctx *ctx, cudaStream_t s, uint64_t fence)
{
    // arena readers on the aux stream that were queued before this batch (see ga_ctx::aux_ev)
    if (ctx->aux_used) GA_CUDA_TRY(cudaStreamWaitEvent(s, ctx->aux_ev, 0));
    if (fence < ctx->oldest_ticket) return GA_OK;
    for (auto &o : ctx->slots)
        if (o.busy && o.ticket <= fence) GA_CUDA_TRY(cudaStreamWaitEvent(s, o.done, 0));
    return GA_OK;
}

extern "C" int ga_submit(ga_ctx *ctx, const ga_msg *msgs, uint32_t n, uint64_t *ticket)
{
    return ga::submit_impl(ctx, msgs, n, ticket, nullptr);
}

void ga::ctx_limits(const ga_ctx *ctx, uint32_t *max_batch_msgs, uint64_t *frames_cap)
{
    if (max_batch_msgs) *max_batch_msgs = ctx->cfg.max_batch_msgs;
    if (frames_cap) *frames_cap = ctx->frames_cap;
}

void ga::ring_release_to(ga_ctx *ctx, uint64_t off)
{
    std::lock_guard<std::mutex> lk(ctx->mu);
    for (auto &s : ctx->slots) if (s.busy) return;
    const uint64_t head = ctx->ring_head.load(std::memory_order_acquire);
    if (off > head) off = head;
    if (off > ctx->ring_tail.load(std::memory_order_relaxed)) ctx->ring_tail.store(off, std::memory_order_release);
}

int ga::submit_impl(ga_ctx *ctx, const ga_msg *msgs, uint32_t n, uint64_t *ticket, uint32_t *bad_index)
{
    GA_GUARD_BEGIN
    if (bad_index) *bad_index = 0xffffffffu;
    if (!ctx || !ticket || (n && !msgs)) { set_error("ga_submit: NULL argument"); return GA_ERR_ARG; }
    if (n > ctx->cfg.max_batch_msgs) { set_error("ga_submit: %u messages exceed max_batch_msgs %u", n, ctx->cfg.max_batch_msgs); return GA_ERR_OVERFLOW; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = set_device(ctx);
    if (rc) return rc;
    // slots are used round-robin (ticket t lives in slot (t-1) % max_inflight), so the pinned result
    // buffers a ga_wait_view handed out are not written again before ticket t + max_inflight is submitted
    Slot *slot = &ctx->slots[(ctx->next_ticket - 1) % ctx->slots.size()];
    if (slot->busy) { set_error("ga_submit: %u tickets already in flight", ctx->cfg.max_inflight); return GA_ERR_STATE; }

    const uint64_t size = ctx->cfg.ring_bytes;
    const uint64_t arena = ctx->cfg.arena_samples, wake_cap = ctx->cfg.wake_samples_max;
    const bool convert = (ctx->cfg.mode & GA_DO_CONVERT) != 0;
    const uint64_t head = ctx->ring_head.load(std::memory_order_acquire);
    const uint64_t tail = ctx->ring_tail.load(std::memory_order_acquire);

    // One lean pass: validate, advance the stream cursors, write the descriptor, grow the H2D
    // runs.  A bad message rolls the cursors of the messages before it back, so a failed
    // submit leaves the context exactly as it was.
    struct Saved { StreamState before; };
    std::vector<Saved> resets;                 // only GA_MSG_FIRST messages need a saved state
    uint64_t release_to = tail, frames = 0, fence = ctx->fence_ticket;
    struct Run { uint64_t lo, hi; };
    Run runs[16];
    int n_runs = 0;
    bool scattered = false;
    int err = GA_OK;
    uint32_t i = 0;
    ga_desc *descs = slot->h_desc;
    StreamState *streams = ctx->streams.data();
    const uint32_t max_streams = ctx->cfg.max_streams;
    const bool fmt48 = ctx->cfg.input_format == GA_FMT_STEREO48K;
    const bool utter = (ctx->cfg.mode & GA_DO_UTTERANCE) != 0;
    uint32_t *link_next = slot->h_link, *link_first = slot->h_link ? slot->h_link + ctx->cfg.max_batch_msgs : nullptr;
    uint32_t n_chains = 0;
    ga_rdesc *rdescs = slot->h_rdesc;
    const uint64_t stamp = ctx->next_ticket;      // marks streams already seen in THIS batch
    ctx->touched.clear();
    // equal-sized chunks copied into the ring back to back (a batch tick) can be staged by one
    // 2-D TMA copy per 32 messages instead of one bulk copy per message
    bool uniform = n > 0;
    uint64_t phys0 = 0, ustride = 0;
    for (; i < n; i++) {
        const ga_msg &m = msgs[i];
        const uint64_t need = (static_cast<uint64_t>(m.nbytes) + 15u) & ~15ull;
        const uint64_t phys = m.ring_offset % size;
        if (m.stream_id >= max_streams) { set_error("ga_submit: msg %u: stream_id %u out of range", i, m.stream_id); err = GA_ERR_ARG; break; }
        if ((m.ring_offset & 15u) || m.ring_offset < tail || m.ring_offset + need > head || phys + need > size) {
            set_error("ga_submit: msg %u: ring_offset %llu not a live reservation", i, (unsigned long long)m.ring_offset);
            err = GA_ERR_ARG;
            break;
        }
        if (i == 0) { phys0 = phys; ustride = need; }
        else if (uniform) uniform = (m.nbytes == msgs[0].nbytes) && (phys == phys0 + i * ustride);
        StreamState &st = streams[m.stream_id];
        if (st.stamp != stamp) ctx->touched.emplace_back(m.stream_id, st.stamp);   // first message of the stream in this batch
        if (m.flags & GA_MSG_FIRST) {
            // a new utterance reuses the stream's arena from its start while the batch's kernel may
            // still be writing the old utterance's chunks there: it has to wait for the next batch
            if (convert && st.stamp == stamp) {
                set_error("ga_submit: msg %u starts a new utterance on stream %u after chunks of the previous one "
                          "in the same batch; submit it with the next batch", i, m.stream_id);
                err = GA_ERR_ARG;
                break;
            }
            if (st.stamp > fence) fence = st.stamp;
            resets.push_back({st});
            st = StreamState();
        }
        st.stamp = stamp;
        uint64_t ns = m.nbytes >> 1;
        if (fmt48) {
            // 48 kHz stereo: the message yields one output per input frame index 3M+2 it contains
            if (m.nbytes > GA_RS_MAX_MSG_BYTES) {
                if (m.flags & GA_MSG_FIRST) { st = resets.back().before; resets.pop_back(); }
                set_error("ga_submit: msg %u: 48 kHz chunks are limited to %u bytes", i, GA_RS_MAX_MSG_BYTES);
                err = GA_ERR_ARG;
                break;
            }
            const uint64_t nfr = m.nbytes >> 2, n0 = st.n_frames, n1 = n0 + nfr, mf = n0 / 3;
            ns = (n1 >= 3 * mf + 3) ? (n1 - 3 * mf - 3) / 3 + 1 : 0;
            ga_rdesc &r = rdescs[i];
            r.in_off = phys; r.nbytes = m.nbytes; r.stream_id = m.stream_id; r.n0 = n0;
            r.flags = m.flags | GA_RS_SAVE_HIST; r.hist_off = 0; r.hist_kind = GA_HIST_ZERO;
            r.reserved[0] = r.reserved[1] = r.reserved[2] = 0;
            if (ctx->batch_stamp[m.stream_id] == stamp) {
                // the stream's previous chunk is in this very batch: take the history from the slab
                ga_rdesc &prev = rdescs[ctx->batch_last[m.stream_id]];
                const uint64_t pfr = prev.nbytes >> 2;
                if (pfr < 47) {
                    if (m.flags & GA_MSG_FIRST) { st = resets.back().before; resets.pop_back(); }
                    set_error("ga_submit: msg %u: its stream's previous chunk in this batch has %llu < 47 frames; "
                              "submit short 48 kHz chunks in separate batches", i, (unsigned long long)pfr);
                    err = GA_ERR_ARG;
                    break;
                }
                if (!(m.flags & GA_MSG_FIRST)) { r.hist_kind = GA_HIST_SLAB; r.hist_off = prev.in_off + (pfr - 47) * 4; }
                prev.flags &= ~GA_RS_SAVE_HIST;             // only the stream's last chunk of the batch saves
            } else if (n0 != 0) {
                r.hist_kind = GA_HIST_STATE;
            }
        }
        const bool is_wake = (m.flags & GA_MSG_WAKE) != 0;
        uint64_t &cursor = is_wake ? st.wake_n : st.speech_n;   // audio_service.go:965 / :1002 append
        ga_desc &d = descs[i];
        d.in_off = phys;
        d.nbytes = m.nbytes;
        d.stream_id = m.stream_id;
        d.flags = m.flags;
        d.frame_off = static_cast<uint32_t>(frames);
        d.out_off = 0;
        if (convert) {
            const uint64_t region = is_wake ? 0 : wake_cap, cap = is_wake ? wake_cap : arena - wake_cap;
            if (cursor + ns > cap) {
                if (m.flags & GA_MSG_FIRST) { st = resets.back().before; resets.pop_back(); }
                set_error("ga_submit: msg %u overflows stream %u's %s sample arena", i, m.stream_id, is_wake ? "wake-word" : "speech");
                err = GA_ERR_OVERFLOW;
                break;
            }
            d.out_off = static_cast<uint64_t>(m.stream_id) * arena + region + cursor;
        }
        cursor += ns;
        if (utter) {
            // thread this message onto its stream's list for the running-digest kernel
            link_next[i] = 0xffffffffu;
            if (ctx->batch_stamp[m.stream_id] == stamp) link_next[ctx->batch_last[m.stream_id]] = i;
            else link_first[n_chains++] = i;
        }
        if (fmt48) {
            rdescs[i].out_off = d.out_off;
            rdescs[i].frame_off = d.frame_off;
            st.n_frames += m.nbytes >> 2;
        }
        if (fmt48 || utter) {
            ctx->batch_stamp[m.stream_id] = stamp;
            ctx->batch_last[m.stream_id] = i;
        }
        frames += (ns + GA_FRAME_SAMPLES - 1) / GA_FRAME_SAMPLES;
        const uint64_t end = m.ring_offset + need;
        if (end > release_to) release_to = end;
        if (need) {
            const uint64_t hi = phys + need;
            if (n_runs && phys >= runs[n_runs - 1].lo && phys <= runs[n_runs - 1].hi + 4096) {
                if (hi > runs[n_runs - 1].hi) runs[n_runs - 1].hi = hi;
            } else if (n_runs < 16) {
                runs[n_runs++] = {phys, hi};
            } else {
                scattered = true;
                if (phys < runs[0].lo) runs[0].lo = phys;
                if (hi > runs[0].hi) runs[0].hi = hi;
            }
        }
    }
    if (!err && ctx->frames_cap && frames > ctx->frames_cap) {
        set_error("ga_submit: batch needs %llu frame records, capacity %llu", (unsigned long long)frames, (unsigned long long)ctx->frames_cap);
        err = GA_ERR_OVERFLOW;
    }
    // undo messages upto-1 .. 0: a submit that fails — here or while queueing the GPU work
    // below — leaves every stream cursor as it found it
    auto undo = [&](uint32_t upto) {
        while (upto-- > 0) {
            const ga_msg &m = msgs[upto];
            StreamState &st = streams[m.stream_id];
            uint64_t ns = m.nbytes >> 1;
            if (fmt48) {
                const uint64_t nfr = m.nbytes >> 2;
                st.n_frames -= nfr;
                const uint64_t n0 = st.n_frames, n1 = n0 + nfr, mf = n0 / 3;
                ns = (n1 >= 3 * mf + 3) ? (n1 - 3 * mf - 3) / 3 + 1 : 0;
            }
            if (fmt48 || utter) ctx->batch_stamp[m.stream_id] = 0;
            ((m.flags & GA_MSG_WAKE) ? st.wake_n : st.speech_n) -= ns;
            if (m.flags & GA_MSG_FIRST) { st = resets.back().before; resets.pop_back(); }
        }
        // every stream this batch touched (the refused message's included) gets back the ticket that
        // last appended to it: a later GA_MSG_FIRST / ga_stream_reset must still fence against that
        // ticket if it is in flight, and the failed batch's ticket number will be used again
        for (auto it = ctx->touched.rbegin(); it != ctx->touched.rend(); ++it) streams[it->first].stamp = it->second;
        ctx->touched.clear();
    };
    if (err) {
        if (bad_index && i < n) *bad_index = i;     // the refusal is about message i (and so about its stream)
        undo(i);
        return err;
    }
    if (scattered) {   // more than 16 disjoint ranges: one covering copy
        for (int r = 1; r < n_runs; r++) {
            if (runs[r].lo < runs[0].lo) runs[0].lo = runs[r].lo;
            if (runs[r].hi > runs[0].hi) runs[0].hi = runs[r].hi;
        }
        n_runs = 1;
    }

    cudaStream_t s = slot->stream;
    auto enqueue = [&]() -> int {
    rc = fence_wait(ctx, s, fence);
    if (rc) return rc;
    for (int r = 0; r < n_runs; r++)
        GA_CUDA_TRY(cudaMemcpyAsync(ctx->d_ring + runs[r].lo, ctx->h_ring + runs[r].lo, runs[r].hi - runs[r].lo, cudaMemcpyHostToDevice, s));
    if (n) {
        GA_CUDA_TRY(cudaMemcpyAsync(slot->d_desc, slot->h_desc, sizeof(ga_desc) * n, cudaMemcpyHostToDevice, s));
        if (!fmt48) {
            if (ctx->cfg.mode & GA_DO_ALL) {
                if (uniform && msgs[0].nbytes && ctx->cfg.staging != GA_STAGING_BULK)
                    rc = ga_k_ingest_uniform(ctx->d_ring, slot->d_desc, n, phys0, msgs[0].nbytes, ustride, ctx->d_arena, slot->d_res,
                                             ctx->frames_cap ? slot->d_frames : nullptr, ctx->cfg.mode & GA_DO_ALL, ctx->cfg.vad_threshold, s);
                else
                    rc = ga_k_ingest(ctx->d_ring, slot->d_desc, n, ctx->d_arena, slot->d_res,
                                     ctx->frames_cap ? slot->d_frames : nullptr, ctx->cfg.mode & GA_DO_ALL, ctx->cfg.vad_threshold, s);
                if (rc) return rc;
            } else {
                GA_CUDA_TRY(cudaMemsetAsync(slot->d_res, 0, sizeof(ga_result) * n, s));
            }
        } else {
            // fingerprint of the received bytes first (writes whole records), then the decimator
            // fills the statistics of the resampled samples into the same records
            GA_CUDA_TRY(cudaMemcpyAsync(slot->d_rdesc, slot->h_rdesc, sizeof(ga_rdesc) * n, cudaMemcpyHostToDevice, s));
            if (ctx->cfg.mode & GA_DO_SHA256) {
                // equal-sized chunks back to back (a tick of 20 ms chunks): tensor-map staging here too
                if (uniform && msgs[0].nbytes && ctx->cfg.staging != GA_STAGING_BULK)
                    rc = ga_k_ingest_uniform(ctx->d_ring, slot->d_desc, n, phys0, msgs[0].nbytes, ustride, nullptr, slot->d_res,
                                             nullptr, GA_DO_SHA256, ctx->cfg.vad_threshold, s);
                else
                    rc = ga_k_ingest(ctx->d_ring, slot->d_desc, n, nullptr, slot->d_res, nullptr, GA_DO_SHA256, ctx->cfg.vad_threshold, s);
                if (rc) return rc;
            } else {
                GA_CUDA_TRY(cudaMemsetAsync(slot->d_res, 0, sizeof(ga_result) * n, s));
            }
            // tickets run on their own CUDA streams; the per-stream history carried in d_hist makes
            // the decimator launches of consecutive tickets dependent, so chain them
            if (ctx->prev_kdone) GA_CUDA_TRY(cudaStreamWaitEvent(s, ctx->prev_kdone, 0));
            rc = ga_k_resample48(ctx->d_ring, slot->d_rdesc, n, ctx->d_hist, ctx->d_arena, slot->d_res,
                                 ctx->frames_cap ? slot->d_frames : nullptr, ctx->cfg.vad_threshold, s);
            if (rc) return rc;
        }
        if (utter) {
            // the running digests make consecutive tickets dependent: chain them like the 48 kHz carry
            GA_CUDA_TRY(cudaMemcpyAsync(slot->d_link, slot->h_link, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, s));
            GA_CUDA_TRY(cudaMemcpyAsync(slot->d_link + ctx->cfg.max_batch_msgs, slot->h_link + ctx->cfg.max_batch_msgs,
                                        sizeof(uint32_t) * n_chains, cudaMemcpyHostToDevice, s));
            if (ctx->prev_kdone && !fmt48) GA_CUDA_TRY(cudaStreamWaitEvent(s, ctx->prev_kdone, 0));
            rc = ga_k_sha256_chain(ctx->d_ring, slot->d_desc, slot->d_link + ctx->cfg.max_batch_msgs, slot->d_link, n_chains, ctx->d_sha, s);
            if (rc) return rc;
        }
        if (fmt48 || utter) {
            GA_CUDA_TRY(cudaEventRecord(slot->kdone, s));
            ctx->prev_kdone = slot->kdone;
        }
        GA_CUDA_TRY(cudaMemcpyAsync(slot->h_res, slot->d_res, sizeof(ga_result) * n, cudaMemcpyDeviceToHost, s));
        if (ctx->frames_cap && frames)
            GA_CUDA_TRY(cudaMemcpyAsync(slot->h_frames, slot->d_frames, sizeof(ga_frame) * frames, cudaMemcpyDeviceToHost, s));
    }
    GA_CUDA_TRY(cudaEventRecord(slot->done, s));
    return GA_OK;
    };
    rc = enqueue();
    if (rc) {
        undo(n);
        return rc;
    }
    ctx->fence_ticket = 0;
    slot->busy = true;
    slot->ticket = ctx->next_ticket++;
    slot->n_msgs = n;
    slot->n_frames = static_cast<uint32_t>(frames);
    slot->release_to = release_to;
    *ticket = slot->ticket;
    return GA_OK;
    GA_GUARD_END
}

// ---------------------------------------------------------------- regular tick batches
namespace ga {

// Descriptor of message m = t * n_streams + j from the stream's cursors at the start of the batch
// and the flags of its earlier periods in the batch — the bookkeeping ga_submit does on the host
// (append to WakeWordSignal or SpeechAudio, audio_service.go:965 / :1002), replayed per message.
__global__ void __launch_bounds__(256)
tick_desc_kernel(const uint32_t *__restrict__ ids, const uint32_t *__restrict__ wake0,
                 const uint32_t *__restrict__ speech0, const uint8_t *__restrict__ flags, uint32_t n,
                 uint32_t n_streams, uint32_t chunk_bytes, uint32_t stride, uint64_t phys0, uint64_t arena,
                 uint64_t wake_cap, uint32_t frames_per_msg, int convert, ga_desc *__restrict__ desc,
                 uint32_t *__restrict__ link_next, uint32_t *__restrict__ run_first,
                 ga_rdesc *__restrict__ rdesc, const uint32_t *__restrict__ n0_lo, const uint32_t *__restrict__ n0_hi,
                 uint32_t n_ticks)
{
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n) return;
    if (link_next) {                  // running digests: a stream's chunks of the batch, in order, start at period 0
        link_next[m] = m + n_streams < n ? m + n_streams : 0xffffffffu;
        if (m < n_streams) run_first[m] = m;
    }
    const uint32_t t = m / n_streams, j = m - t * n_streams;
    // 48 kHz stereo chunks of 3k frames give k outputs each, wherever the stream stands
    const uint32_t ns = rdesc ? chunk_bytes / 12u : chunk_bytes >> 1;
    uint32_t w = wake0[j], s = speech0[j];
    if (flags) {
        for (uint32_t k = 0; k < t; k++) {
            const uint32_t f = flags[static_cast<size_t>(k) * n_streams + j];
            if (f & GA_MSG_FIRST) { w = 0; s = 0; }
            if (f & GA_MSG_WAKE) w += ns; else s += ns;
        }
    } else {
        s += t * ns;
    }
    const uint32_t f = flags ? flags[m] : 0u;
    if (f & GA_MSG_FIRST) { w = 0; s = 0; }
    const uint32_t id = ids[j];
    ga_desc d;
    d.in_off = phys0 + static_cast<uint64_t>(m) * stride;
    d.out_off = convert ? static_cast<uint64_t>(id) * arena + ((f & GA_MSG_WAKE) ? w : wake_cap + s) : 0;
    d.nbytes = chunk_bytes;
    d.stream_id = id;
    d.flags = f;
    d.frame_off = m * frames_per_msg;
    uint4 *o = reinterpret_cast<uint4 *>(desc + m);
    const uint4 *v = reinterpret_cast<const uint4 *>(&d);
    o[0] = v[0];
    o[1] = v[1];
    if (rdesc) {
        // decimator descriptor: position on the stream's frame timeline and where the 47 frames
        // before the chunk come from.  A restart is only possible in period 0 here (conversion is on).
        const uint32_t frames = chunk_bytes >> 2;
        const bool first = flags && (flags[j] & GA_MSG_FIRST);
        const uint64_t base = first ? 0ull : (static_cast<uint64_t>(n0_hi[j]) << 32) | n0_lo[j];
        ga_rdesc r;
        r.in_off = d.in_off;
        r.out_off = d.out_off;
        r.n0 = base + static_cast<uint64_t>(t) * frames;
        r.nbytes = chunk_bytes;
        r.stream_id = id;
        r.flags = f | (t + 1 == n_ticks ? GA_RS_SAVE_HIST : 0u);    // only the stream's last chunk of the batch saves
        r.frame_off = d.frame_off;
        r.reserved[0] = r.reserved[1] = r.reserved[2] = 0;
        if (t > 0) {
            r.hist_kind = GA_HIST_SLAB;                                // tail of the stream's previous chunk, one period back
            r.hist_off = d.in_off - static_cast<uint64_t>(n_streams) * stride + static_cast<uint64_t>(frames - 47u) * 4u;
        } else {
            r.hist_kind = (first || base == 0) ? GA_HIST_ZERO : GA_HIST_STATE;
            r.hist_off = 0;
        }
        uint4 *ro = reinterpret_cast<uint4 *>(rdesc + m);
        const uint4 *rv = reinterpret_cast<const uint4 *>(&r);
        ro[0] = rv[0]; ro[1] = rv[1]; ro[2] = rv[2]; ro[3] = rv[3];
    }
}

}  // namespace ga

extern "C" int ga_submit_ticks(ga_ctx *ctx, uint64_t ring_offset, uint32_t n_ticks, uint32_t n_streams,
                               uint32_t chunk_bytes, const uint32_t *stream_ids, const uint8_t *flags,
                               uint64_t *ticket)
{
    GA_GUARD_BEGIN
    if (!ctx || !ticket || !stream_ids) { set_error("ga_submit_ticks: NULL argument"); return GA_ERR_ARG; }
    if (n_ticks == 0 || n_streams == 0) { set_error("ga_submit_ticks: empty batch"); return GA_ERR_ARG; }
    const uint64_t n64 = static_cast<uint64_t>(n_ticks) * n_streams;
    if (n64 > ctx->cfg.max_batch_msgs) {
        set_error("ga_submit_ticks: %llu messages exceed max_batch_msgs %u", (unsigned long long)n64, ctx->cfg.max_batch_msgs);
        return GA_ERR_OVERFLOW;
    }
    const uint32_t n = static_cast<uint32_t>(n64);
    std::lock_guard<std::mutex> lk(ctx->mu);
    const bool fmt48 = ctx->cfg.input_format == GA_FMT_STEREO48K;
    if (fmt48 && (chunk_bytes % 12u != 0 || chunk_bytes < 47u * 4u || chunk_bytes > GA_RS_MAX_MSG_BYTES)) {
        // whole groups of 3 stereo frames (so every chunk yields the same number of outputs) and at
        // least the 47 frames of filter history the next period takes from the slab
        set_error("ga_submit_ticks: 48 kHz chunks must be a multiple of 12 bytes in [192, %u]; use ga_submit", GA_RS_MAX_MSG_BYTES);
        return GA_ERR_ARG;
    }
    const uint64_t arena = ctx->cfg.arena_samples, wake_cap = ctx->cfg.wake_samples_max;
    const bool convert = (ctx->cfg.mode & GA_DO_CONVERT) != 0;
    const bool utter = (ctx->cfg.mode & GA_DO_UTTERANCE) != 0;
    if (convert && arena > 0xffffffffull) { set_error("ga_submit_ticks: arena_samples beyond 2^32; use ga_submit"); return GA_ERR_STATE; }
    int rc = set_device(ctx);
    if (rc) return rc;
    Slot *slot = &ctx->slots[(ctx->next_ticket - 1) % ctx->slots.size()];      // round-robin, as in ga_submit
    if (slot->busy) { set_error("ga_submit_ticks: %u tickets already in flight", ctx->cfg.max_inflight); return GA_ERR_STATE; }

    const uint64_t size = ctx->cfg.ring_bytes;
    const uint64_t stride = (static_cast<uint64_t>(chunk_bytes) + 15u) & ~15ull;
    const uint64_t span = stride * n, phys0 = ring_offset % size;
    const uint64_t head = ctx->ring_head.load(std::memory_order_acquire);
    const uint64_t tail = ctx->ring_tail.load(std::memory_order_acquire);
    if ((ring_offset & 15u) || ring_offset < tail || ring_offset + span > head || phys0 + span > size) {
        set_error("ga_submit_ticks: [%llu, +%llu) is not one live reservation", (unsigned long long)ring_offset, (unsigned long long)span);
        return GA_ERR_ARG;
    }
    // side buffer, packed so that one copy moves it: ids | wake0 | speech0 | n0 lo | n0 hi (n_streams each) | flags (n)
    uint32_t *h_ids = reinterpret_cast<uint32_t *>(slot->h_tick), *h_wake0 = h_ids + n_streams, *h_speech0 = h_wake0 + n_streams;
    uint32_t *h_n0lo = h_speech0 + n_streams, *h_n0hi = h_n0lo + n_streams;
    uint8_t *h_flags = reinterpret_cast<uint8_t *>(h_n0hi + n_streams);
    const uint64_t ns = fmt48 ? chunk_bytes / 12u : chunk_bytes >> 1, frames_per_msg = (ns + GA_FRAME_SAMPLES - 1) / GA_FRAME_SAMPLES;
    const uint64_t frames = frames_per_msg * n;
    if (ctx->frames_cap && frames > ctx->frames_cap) {
        set_error("ga_submit_ticks: batch needs %llu frame records, capacity %llu", (unsigned long long)frames, (unsigned long long)ctx->frames_cap);
        return GA_ERR_OVERFLOW;
    }
    // per stream: remember where its cursors start, then replay its n_ticks chunks
    std::vector<StreamState> before;
    try { before.resize(n_streams); } catch (...) { set_error("ga_submit_ticks: out of host memory"); return GA_ERR_NOMEM; }
    const uint64_t stamp = ctx->next_ticket;
    uint64_t fence = ctx->fence_ticket;
    StreamState *streams = ctx->streams.data();
    uint32_t j = 0;
    int err = GA_OK;
    for (; j < n_streams && !err; j++) {
        const uint32_t id = stream_ids[j];
        if (id >= ctx->cfg.max_streams) { set_error("ga_submit_ticks: stream_ids[%u] = %u out of range", j, id); err = GA_ERR_ARG; break; }
        if (ctx->tick_stamp[id] == stamp) { set_error("ga_submit_ticks: stream %u listed twice", id); err = GA_ERR_ARG; break; }
        ctx->tick_stamp[id] = stamp;
        StreamState &st = streams[id];
        before[j] = st;
        h_ids[j] = id;
        h_wake0[j] = static_cast<uint32_t>(st.wake_n);
        h_speech0[j] = static_cast<uint32_t>(st.speech_n);
        h_n0lo[j] = static_cast<uint32_t>(st.n_frames);
        h_n0hi[j] = static_cast<uint32_t>(st.n_frames >> 32);
        for (uint32_t t = 0; t < n_ticks; t++) {
            const uint32_t f = flags ? flags[static_cast<size_t>(t) * n_streams + j] : 0u;
            if (f & GA_MSG_FIRST) {
                if (convert && t > 0) {       // same rule as ga_submit: the old utterance's chunks are still in flight
                    set_error("ga_submit_ticks: period %u starts a new utterance on stream %u after chunks of the "
                              "previous one in the same batch; submit it with the next batch", t, id);
                    err = GA_ERR_ARG;
                    break;
                }
                if (st.stamp > fence) fence = st.stamp;
                st = StreamState();
            }
            const bool is_wake = (f & GA_MSG_WAKE) != 0;
            uint64_t &cursor = is_wake ? st.wake_n : st.speech_n;
            if (convert && cursor + ns > (is_wake ? wake_cap : arena - wake_cap)) {
                set_error("ga_submit_ticks: period %u overflows stream %u's %s sample arena", t, id, is_wake ? "wake-word" : "speech");
                err = GA_ERR_OVERFLOW;
                break;
            }
            cursor += ns;
            if (fmt48) st.n_frames += chunk_bytes >> 2;
        }
        if (err) { st = before[j]; ctx->tick_stamp[id] = 0; }
        else st.stamp = stamp;
    }
    auto undo = [&](uint32_t upto) {
        while (upto-- > 0) { streams[stream_ids[upto]] = before[upto]; ctx->tick_stamp[stream_ids[upto]] = 0; }
    };
    if (err) { undo(j); return err; }
    if (flags) memcpy(h_flags, flags, n);

    cudaStream_t s = slot->stream;
    auto enqueue = [&]() -> int {
        int r = fence_wait(ctx, s, fence);
        if (r) return r;
        if (span) GA_CUDA_TRY(cudaMemcpyAsync(ctx->d_ring + phys0, ctx->h_ring + phys0, span, cudaMemcpyHostToDevice, s));
        uint32_t *d_ids = reinterpret_cast<uint32_t *>(slot->d_tick), *d_wake0 = d_ids + n_streams, *d_speech0 = d_wake0 + n_streams;
        uint32_t *d_n0lo = d_speech0 + n_streams, *d_n0hi = d_n0lo + n_streams;
        uint8_t *d_flags = reinterpret_cast<uint8_t *>(d_n0hi + n_streams);
        GA_CUDA_TRY(cudaMemcpyAsync(slot->d_tick, slot->h_tick, 20ull * n_streams + (flags ? n : 0), cudaMemcpyHostToDevice, s));
        tick_desc_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_ids, d_wake0, d_speech0, flags ? d_flags : nullptr, n, n_streams, chunk_bytes,
                                                        static_cast<uint32_t>(stride), phys0, arena, wake_cap,
                                                        static_cast<uint32_t>(frames_per_msg), convert ? 1 : 0, slot->d_desc,
                                                        utter ? slot->d_link : nullptr,
                                                        utter ? slot->d_link + ctx->cfg.max_batch_msgs : nullptr,
                                                        fmt48 ? slot->d_rdesc : nullptr, d_n0lo, d_n0hi, n_ticks);
        GA_CHECK_LAUNCH("tick_desc_kernel");
        if (fmt48) {
            // as in ga_submit: fingerprint of the received bytes (whole records), then the decimator
            // fills in the statistics; its per-stream history chains consecutive tickets
            if (ctx->cfg.mode & GA_DO_SHA256) {
                if (ctx->cfg.staging != GA_STAGING_BULK)
                    r = ga_k_ingest_uniform(ctx->d_ring, slot->d_desc, n, phys0, chunk_bytes, stride, nullptr, slot->d_res, nullptr,
                                            GA_DO_SHA256, ctx->cfg.vad_threshold, s);
                else
                    r = ga_k_ingest(ctx->d_ring, slot->d_desc, n, nullptr, slot->d_res, nullptr, GA_DO_SHA256, ctx->cfg.vad_threshold, s);
                if (r) return r;
            } else {
                GA_CUDA_TRY(cudaMemsetAsync(slot->d_res, 0, sizeof(ga_result) * n, s));
            }
            if (ctx->prev_kdone) GA_CUDA_TRY(cudaStreamWaitEvent(s, ctx->prev_kdone, 0));
            r = ga_k_resample48(ctx->d_ring, slot->d_rdesc, n, ctx->d_hist, ctx->d_arena, slot->d_res,
                                ctx->frames_cap ? slot->d_frames : nullptr, ctx->cfg.vad_threshold, s);
        } else if (!(ctx->cfg.mode & GA_DO_ALL))
            GA_CUDA_TRY(cudaMemsetAsync(slot->d_res, 0, sizeof(ga_result) * n, s));
        else if (chunk_bytes && ctx->cfg.staging != GA_STAGING_BULK)
            r = ga_k_ingest_uniform(ctx->d_ring, slot->d_desc, n, phys0, chunk_bytes, stride, ctx->d_arena, slot->d_res,
                                    ctx->frames_cap ? slot->d_frames : nullptr, ctx->cfg.mode & GA_DO_ALL, ctx->cfg.vad_threshold, s);
        else
            r = ga_k_ingest(ctx->d_ring, slot->d_desc, n, ctx->d_arena, slot->d_res,
                            ctx->frames_cap ? slot->d_frames : nullptr, ctx->cfg.mode & GA_DO_ALL, ctx->cfg.vad_threshold, s);
        if (r) return r;
        if (utter) {
            // the running digests make consecutive tickets dependent, as in ga_submit
            if (ctx->prev_kdone && !fmt48) GA_CUDA_TRY(cudaStreamWaitEvent(s, ctx->prev_kdone, 0));
            r = ga_k_sha256_chain(ctx->d_ring, slot->d_desc, slot->d_link + ctx->cfg.max_batch_msgs, slot->d_link, n_streams, ctx->d_sha, s);
            if (r) return r;
        }
        if (fmt48 || utter) {
            GA_CUDA_TRY(cudaEventRecord(slot->kdone, s));
            ctx->prev_kdone = slot->kdone;
        }
        GA_CUDA_TRY(cudaMemcpyAsync(slot->h_res, slot->d_res, sizeof(ga_result) * n, cudaMemcpyDeviceToHost, s));
        if (ctx->frames_cap && frames)
            GA_CUDA_TRY(cudaMemcpyAsync(slot->h_frames, slot->d_frames, sizeof(ga_frame) * frames, cudaMemcpyDeviceToHost, s));
        GA_CUDA_TRY(cudaEventRecord(slot->done, s));
        return GA_OK;
    };
    rc = enqueue();
    if (rc) { undo(n_streams); return rc; }
    ctx->fence_ticket = 0;
    slot->busy = true;
    slot->ticket = ctx->next_ticket++;
    slot->n_msgs = n;
    slot->n_frames = static_cast<uint32_t>(frames);
    slot->release_to = ring_offset + span > tail ? ring_offset + span : tail;
    *ticket = slot->ticket;
    return GA_OK;
    GA_GUARD_END
}

static Slot *find_ticket(ga_ctx *ctx, uint64_t ticket)
{
    for (auto &s : ctx->slots) if (s.busy && s.ticket == ticket) return &s;
    return nullptr;
}

extern "C" int ga_ticket_frames(ga_ctx *ctx, uint64_t ticket, uint32_t *n_frames)
{
    if (!ctx || !n_frames) { set_error("ga_ticket_frames: NULL argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    Slot *s = find_ticket(ctx, ticket);
    if (!s) { set_error("ga_ticket_frames: unknown ticket %llu", (unsigned long long)ticket); return GA_ERR_STATE; }
    *n_frames = ctx->frames_cap ? s->n_frames : 0;
    return GA_OK;
}

extern "C" int ga_wait(ga_ctx *ctx, uint64_t ticket, ga_result *results, ga_frame *frames)
{
    if (!ctx) { set_error("ga_wait: NULL context"); return GA_ERR_ARG; }
    Slot *s;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        s = find_ticket(ctx, ticket);
        if (!s) { set_error("ga_wait: unknown ticket %llu", (unsigned long long)ticket); return GA_ERR_STATE; }
        if (ticket != ctx->oldest_ticket) {
            set_error("ga_wait: tickets must be waited in submission order (oldest is %llu)", (unsigned long long)ctx->oldest_ticket);
            return GA_ERR_STATE;
        }
        if (s->n_msgs && !results) { set_error("ga_wait: NULL results"); return GA_ERR_ARG; }
    }
    int rc = set_device(ctx);
    if (rc) return rc;
    GA_CUDA_TRY(cudaEventSynchronize(s->done));     // outside the lock: submit may proceed meanwhile
    if (s->n_msgs) std::memcpy(results, s->h_res, sizeof(ga_result) * s->n_msgs);
    if (frames && ctx->frames_cap && s->n_frames) std::memcpy(frames, s->h_frames, sizeof(ga_frame) * s->n_frames);
    std::lock_guard<std::mutex> lk(ctx->mu);
    uint64_t t = ctx->ring_tail.load(std::memory_order_relaxed);
    if (s->release_to > t) ctx->ring_tail.store(s->release_to, std::memory_order_release);
    s->busy = false;
    ctx->oldest_ticket = ticket + 1;
    return GA_OK;
}

extern "C" int ga_wait_view(ga_ctx *ctx, uint64_t ticket, const ga_result **results, const ga_frame **frames,
                            uint32_t *n_results, uint32_t *n_frames)
{
    if (!ctx) { set_error("ga_wait_view: NULL context"); return GA_ERR_ARG; }
    Slot *s;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        s = find_ticket(ctx, ticket);
        if (!s) { set_error("ga_wait_view: unknown ticket %llu", (unsigned long long)ticket); return GA_ERR_STATE; }
        if (ticket != ctx->oldest_ticket) {
            set_error("ga_wait_view: tickets must be waited in submission order (oldest is %llu)", (unsigned long long)ctx->oldest_ticket);
            return GA_ERR_STATE;
        }
    }
    int rc = set_device(ctx);
    if (rc) return rc;
    GA_CUDA_TRY(cudaEventSynchronize(s->done));
    if (results) *results = s->h_res;
    if (frames) *frames = ctx->frames_cap ? s->h_frames : nullptr;
    if (n_results) *n_results = s->n_msgs;
    if (n_frames) *n_frames = ctx->frames_cap ? s->n_frames : 0;
    std::lock_guard<std::mutex> lk(ctx->mu);
    uint64_t t = ctx->ring_tail.load(std::memory_order_relaxed);
    if (s->release_to > t) ctx->ring_tail.store(s->release_to, std::memory_order_release);
    s->busy = false;
    ctx->oldest_ticket = ticket + 1;
    return GA_OK;
}

// every in-flight ingest must be visible to work queued on `st`
static int join_inflight(ga_ctx *ctx, cudaStream_t st)
{
    for (auto &s : ctx->slots)
        if (s.busy) GA_CUDA_TRY(cudaStreamWaitEvent(st, s.done, 0));
    return GA_OK;
}

extern "C" int ga_stream_info(ga_ctx *ctx, uint32_t stream_id, uint64_t *n_samples, uint64_t *wake_samples)
{
    if (!ctx || stream_id >= ctx->cfg.max_streams) { set_error("ga_stream_info: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    const StreamState &st = ctx->streams[stream_id];
    if (n_samples) *n_samples = st.wake_n + st.speech_n;
    if (wake_samples) *wake_samples = st.wake_n;
    return GA_OK;
}

extern "C" int ga_stream_reset(ga_ctx *ctx, uint32_t stream_id)
{
    if (!ctx || stream_id >= ctx->cfg.max_streams) { set_error("ga_stream_reset: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (ctx->streams[stream_id].stamp > ctx->fence_ticket) ctx->fence_ticket = ctx->streams[stream_id].stamp;
    ctx->streams[stream_id] = StreamState();   // cleanupRelay, audio_service.go:916-923
    if (ctx->d_sha) {
        int rc = set_device(ctx);
        if (rc) return rc;
        rc = join_inflight(ctx, ctx->aux);
        if (rc) return rc;
        rc = ga_k_sha256_reset(ctx->d_sha + stream_id, 1, ctx->aux);
        if (rc) return rc;
        GA_CUDA_TRY(cudaStreamSynchronize(ctx->aux));
    }
    return GA_OK;
}

extern "C" int ga_utterance_digest(ga_ctx *ctx, uint32_t stream_id, uint8_t digest[32], uint64_t *total_bytes)
{
    if (!ctx || !digest || stream_id >= ctx->cfg.max_streams) { set_error("ga_utterance_digest: bad argument"); return GA_ERR_ARG; }
    if (!ctx->d_sha) { set_error("ga_utterance_digest: context was not created with GA_DO_UTTERANCE"); return GA_ERR_STATE; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = set_device(ctx);
    if (rc) return rc;
    rc = join_inflight(ctx, ctx->aux);
    if (rc) return rc;
    // 32-byte digest + 8-byte length land in the first 40 bytes of a scratch state record (the last one is spare)
    struct Out { uint8_t d[32]; uint64_t len; } out;
    uint8_t *d_out = nullptr;
    GA_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&d_out), 64 + sizeof(uint32_t)));
    uint32_t sid = stream_id;
    cudaError_t e = cudaMemcpyAsync(d_out + 64, &sid, sizeof sid, cudaMemcpyHostToDevice, ctx->aux);
    if (e == cudaSuccess) {
        rc = ga_k_sha256_final(ctx->d_sha, reinterpret_cast<uint32_t *>(d_out + 64), 1, d_out,
                               reinterpret_cast<uint64_t *>(d_out + 32), ctx->aux);
        if (rc == GA_OK) e = cudaMemcpyAsync(&out, d_out, 40, cudaMemcpyDeviceToHost, ctx->aux);
        if (rc == GA_OK && e == cudaSuccess) e = cudaStreamSynchronize(ctx->aux);
    }
    cudaFree(d_out);
    if (rc) return rc;
    if (e != cudaSuccess) { set_error("ga_utterance_digest: %s", cudaGetErrorString(e)); return GA_ERR_CUDA; }
    std::memcpy(digest, out.d, 32);
    if (total_bytes) *total_bytes = out.len;
    return GA_OK;
}

extern "C" int ga_device_arena(ga_ctx *ctx, void **d_arena)
{
    if (!ctx || !d_arena) { set_error("ga_device_arena: NULL argument"); return GA_ERR_ARG; }
    if (!ctx->d_arena) { set_error("ga_device_arena: context does not materialise samples"); return GA_ERR_STATE; }
    *d_arena = ctx->d_arena;
    return GA_OK;
}

extern "C" int ga_samples_digests(ga_ctx *ctx, const uint32_t *stream_ids, uint32_t n, uint8_t *digests);
extern "C" int ga_samples_digest(ga_ctx *ctx, uint32_t stream_id, uint8_t digest[32])
{
    if (!ctx || !digest || stream_id >= ctx->cfg.max_streams) { set_error("ga_samples_digest: bad argument"); return GA_ERR_ARG; }
    if (!ctx->d_arena) { set_error("ga_samples_digest: context does not materialise samples"); return GA_ERR_STATE; }
    return ga_samples_digests(ctx, &stream_id, 1, digest);      // the batched path: no allocation, lock dropped before the wait
}

extern "C" int ga_fetch_samples(ga_ctx *ctx, uint32_t stream_id, uint64_t first, uint64_t n, float *out)
{
    if (!ctx || stream_id >= ctx->cfg.max_streams || (n && !out)) { set_error("ga_fetch_samples: bad argument"); return GA_ERR_ARG; }
    if (!(ctx->cfg.mode & GA_DO_CONVERT)) { set_error("ga_fetch_samples: context does not materialise samples"); return GA_ERR_STATE; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = set_device(ctx);
    if (rc) return rc;
    const StreamState st = ctx->streams[stream_id];
    if (first + n > st.wake_n + st.speech_n) { set_error("ga_fetch_samples: range beyond the stream's %llu samples", (unsigned long long)(st.wake_n + st.speech_n)); return GA_ERR_ARG; }
    rc = join_inflight(ctx, ctx->aux);
    if (rc) return rc;
    // logical order is WakeWordSignal ++ SpeechAudio (audio_service.go:592-594)
    const float *base = ctx->d_arena + static_cast<uint64_t>(stream_id) * ctx->cfg.arena_samples;
    uint64_t done = 0;
    if (first < st.wake_n && n) {
        const uint64_t k = std::min(n, st.wake_n - first);
        GA_CUDA_TRY(cudaMemcpyAsync(out, base + first, k * sizeof(float), cudaMemcpyDeviceToHost, ctx->aux));
        done = k;
    }
    if (done < n) {
        const uint64_t sp_first = first + done - st.wake_n;
        GA_CUDA_TRY(cudaMemcpyAsync(out + done, base + ctx->cfg.wake_samples_max + sp_first, (n - done) * sizeof(float), cudaMemcpyDeviceToHost, ctx->aux));
    }
    GA_CUDA_TRY(cudaStreamSynchronize(ctx->aux));
    return GA_OK;
}


// ---------------------------------------------------------------- the winner path, batched
// What runs once per closed window in the reference — processWinningRelayAudio
// (audio_service.go:590-609: fullAudio = WakeWordSignal ++ SpeechAudio), float32ToWAV
// (internal/llm/stt_client.go:365-398) and the voice-event fingerprint of those samples
// (events/voice_event.go:84-88) — for MANY winners in one call: one gather launch packs every
// winner's WAV file into a staging buffer, one chained SHA-256 launch hashes their f32 bytes in
// place in the arena, ONE device->host copy brings everything into C-owned pinned memory.  The
// context lock is held only while the work is queued.
namespace ga {

struct WavItem {            // one winner, 48 bytes
    uint64_t wake_src;      // index of the stream's first wake-word sample in the arena
    uint64_t speech_src;    // index of its first speech sample
    uint64_t dst;           // byte offset of the WAV file in the staging buffer (== 4 mod 16: data 16-byte aligned)
    uint32_t wake_n, speech_n;
    uint32_t sample_rate, pad;
};

constexpr int kWavBlocksPerItem = 8;

__global__ void __launch_bounds__(256)
wav_pack_kernel(const float *__restrict__ arena, const WavItem *items, uint8_t *out)   // items and out share the staging buffer
{
    const WavItem it = items[blockIdx.x / kWavBlocksPerItem];
    const uint32_t part = blockIdx.x % kWavBlocksPerItem;
    const uint32_t n = it.wake_n + it.speech_n, data_size = n * 4u;
    uint32_t *hdr = reinterpret_cast<uint32_t *>(out + it.dst);
    if (part == 0 && threadIdx.x < 11) {
        // stt_client.go:375-388, as eleven little-endian words
        const uint32_t w[11] = {0x46464952u /* RIFF */, 36u + data_size, 0x45564157u /* WAVE */, 0x20746d66u /* "fmt " */,
                                16u, 3u | (1u << 16) /* IEEE float, mono */, it.sample_rate, it.sample_rate * 4u,
                                4u | (32u << 16) /* block align, bits */, 0x61746164u /* data */, data_size};
        hdr[threadIdx.x] = w[threadIdx.x];
    }
    float *dst = reinterpret_cast<float *>(out + it.dst + 44);
    const float *wk = arena + it.wake_src, *sp = arena + it.speech_src;
    for (uint32_t i = part * blockDim.x + threadIdx.x; i < n; i += kWavBlocksPerItem * blockDim.x)
        dst[i] = i < it.wake_n ? wk[i] : sp[i - it.wake_n];       // :592-594 order, :391-396 payload
}

static int win_reserve(ga_ctx *ctx, size_t need)
{
    if (need <= ctx->win_bytes) return GA_OK;
    GA_CUDA_TRY(cudaStreamSynchronize(ctx->aux));
    if (ctx->d_win) cudaFree(ctx->d_win);
    if (ctx->h_win) cudaFreeHost(ctx->h_win);
    ctx->d_win = nullptr; ctx->h_win = nullptr; ctx->win_bytes = 0;
    const size_t want = need + need / 2;
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&ctx->d_win), want);
    if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void **>(&ctx->h_win), want, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        if (ctx->d_win) { cudaFree(ctx->d_win); ctx->d_win = nullptr; }
        set_error("winner staging: allocating %zu bytes failed: %s", want, cudaGetErrorString(e));
        return GA_ERR_NOMEM;
    }
    ctx->win_bytes = want;
    return GA_OK;
}

// wavs / digests for n streams.  Outputs point into the pinned staging buffer (valid until the next
// winner-path call on this context).
static int winners_impl(ga_ctx *ctx, const uint32_t *stream_ids, uint32_t n, bool want_wav, uint32_t sample_rate,
                        bool want_digest, const uint8_t **base, uint64_t *wav_off, const uint8_t **digests)
{
    if (!(ctx->cfg.mode & GA_DO_CONVERT)) { set_error("winner path: context does not materialise samples"); return GA_ERR_STATE; }
    std::lock_guard<std::mutex> wlk(ctx->win_mu);
    std::unique_lock<std::mutex> lk(ctx->mu);
    int rc = set_device(ctx);
    if (rc) return rc;
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    // staging layout: items | descs (2 per winner) | next (2) | first (1) || states | digests | wav files
    const size_t o_item = 0, o_desc = o_item + up(sizeof(WavItem) * n), o_next = o_desc + up(sizeof(ga_desc) * 2 * n),
                 o_first = o_next + up(8ull * n), o_state = o_first + up(4ull * n), o_dig = o_state + up(sizeof(ga_sha_state) * n),
                 o_wav = o_dig + up(32ull * n);
    size_t wav_total = 0;
    for (uint32_t k = 0; k < n; k++) {
        if (stream_ids[k] >= ctx->cfg.max_streams) { set_error("winner path: stream_id %u out of range", stream_ids[k]); return GA_ERR_ARG; }
        const StreamState &st = ctx->streams[stream_ids[k]];
        if ((st.wake_n + st.speech_n) >> 30) { set_error("winner path: stream %u is too long", stream_ids[k]); return GA_ERR_ARG; }
        if (want_wav) wav_total += ((44 + 4 * (st.wake_n + st.speech_n) + 15) & ~size_t(15)) + 16;
    }
    rc = win_reserve(ctx, o_wav + wav_total + 256);
    if (rc) return rc;
    uint8_t *hb = ctx->h_win, *db = ctx->d_win;
    WavItem *items = reinterpret_cast<WavItem *>(hb + o_item);
    ga_desc *descs = reinterpret_cast<ga_desc *>(hb + o_desc);
    uint32_t *next = reinterpret_cast<uint32_t *>(hb + o_next), *first = reinterpret_cast<uint32_t *>(hb + o_first);
    size_t cur = o_wav;
    for (uint32_t k = 0; k < n; k++) {
        const StreamState &st = ctx->streams[stream_ids[k]];
        const uint64_t base_idx = static_cast<uint64_t>(stream_ids[k]) * ctx->cfg.arena_samples;
        WavItem &it = items[k];
        it.wake_src = base_idx; it.speech_src = base_idx + ctx->cfg.wake_samples_max;
        it.wake_n = static_cast<uint32_t>(st.wake_n); it.speech_n = static_cast<uint32_t>(st.speech_n);
        it.sample_rate = sample_rate; it.pad = 0;
        it.dst = cur + 4;                                     // header at == 4 (mod 16): the samples start 16-byte aligned
        if (want_wav) {
            if (wav_off) wav_off[k] = it.dst;
            cur += ((44 + 4 * (st.wake_n + st.speech_n) + 15) & ~size_t(15)) + 16;
        }
        // the two arena regions as two chained "messages" of one running SHA-256 (stt_client.go:390-395 byte order)
        ga_desc &d0 = descs[2 * k], &d1 = descs[2 * k + 1];
        std::memset(&d0, 0, 2 * sizeof(ga_desc));
        d0.in_off = base_idx * 4; d0.nbytes = static_cast<uint32_t>(st.wake_n * 4); d0.stream_id = k; d0.flags = GA_MSG_FIRST;
        d1.in_off = (base_idx + ctx->cfg.wake_samples_max) * 4; d1.nbytes = static_cast<uint32_t>(st.speech_n * 4); d1.stream_id = k;
        next[2 * k] = 2 * k + 1; next[2 * k + 1] = 0xffffffffu; first[k] = 2 * k;
    }
    if (want_wav && wav_off) wav_off[n] = cur + 4;
    cudaStream_t s = ctx->aux;
    rc = join_inflight(ctx, s);
    if (rc) return rc;
    GA_CUDA_TRY(cudaMemcpyAsync(db, hb, o_state, cudaMemcpyHostToDevice, s));
    if (want_digest) {
        rc = ga_k_sha256_chain(ctx->d_arena, reinterpret_cast<ga_desc *>(db + o_desc), reinterpret_cast<uint32_t *>(db + o_first),
                               reinterpret_cast<uint32_t *>(db + o_next), n, reinterpret_cast<ga_sha_state *>(db + o_state), s);
        if (rc) return rc;
        rc = ga_k_sha256_final(reinterpret_cast<ga_sha_state *>(db + o_state), nullptr, n, db + o_dig, nullptr, s);
        if (rc) return rc;
    }
    if (want_wav) {
        wav_pack_kernel<<<n * kWavBlocksPerItem, 256, 0, s>>>(ctx->d_arena, reinterpret_cast<WavItem *>(db + o_item), db);
        GA_CHECK_LAUNCH("wav_pack_kernel");
    }
    const size_t lo = want_digest ? o_dig : o_wav, hi = want_wav ? cur : o_wav;
    if (hi > lo) GA_CUDA_TRY(cudaMemcpyAsync(hb + lo, db + lo, hi - lo, cudaMemcpyDeviceToHost, s));
    GA_CUDA_TRY(cudaEventRecord(ctx->aux_ev, s));
    ctx->aux_used = true;
    lk.unlock();                                              // ga_submit goes on meanwhile
    GA_CUDA_TRY(cudaEventSynchronize(ctx->aux_ev));
    if (base) *base = hb;
    if (digests) *digests = hb + o_dig;
    return GA_OK;
}

}  // namespace ga

extern "C" int ga_fetch_wavs(ga_ctx *ctx, const uint32_t *stream_ids, uint32_t n, uint32_t sample_rate,
                             const uint8_t **wavs, uint64_t *offsets, uint8_t *digests)
{
    GA_GUARD_BEGIN
    if (!ctx || !wavs || !offsets || (n && !stream_ids)) { set_error("ga_fetch_wavs: NULL argument"); return GA_ERR_ARG; }
    *wavs = nullptr;
    offsets[0] = 0;
    if (n == 0) return GA_OK;
    const uint8_t *dg = nullptr;
    int rc = ga::winners_impl(ctx, stream_ids, n, true, sample_rate, digests != nullptr, wavs, offsets, &dg);
    if (rc) return rc;
    if (digests) std::memcpy(digests, dg, 32ull * n);
    return GA_OK;
    GA_GUARD_END
}

extern "C" int ga_samples_digests(ga_ctx *ctx, const uint32_t *stream_ids, uint32_t n, uint8_t *digests)
{
    GA_GUARD_BEGIN
    if (!ctx || (n && (!stream_ids || !digests))) { set_error("ga_samples_digests: NULL argument"); return GA_ERR_ARG; }
    if (n == 0) return GA_OK;
    const uint8_t *dg = nullptr;
    int rc = ga::winners_impl(ctx, stream_ids, n, false, 0, true, nullptr, nullptr, &dg);
    if (rc) return rc;
    std::memcpy(digests, dg, 32ull * n);
    return GA_OK;
    GA_GUARD_END
}

// ---------------------------------------------------------------- arbitration (host arrays)
namespace {
struct DevBuf {   // RAII for the transient device arrays of the host-array entry points
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t bytes)
    {
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
        if (e != cudaSuccess) { cudaGetLastError(); set_error("cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e)); return GA_ERR_NOMEM; }
        return GA_OK;
    }
    template <class T> T *as() { return static_cast<T *>(p); }
};
}  // namespace

// Tables of ga_arbitrate: grow-only, device + pinned mirror.  ga_init sizes them for every stream
// contending in windows of one, so a hub never allocates at window close.
static int arb_reserve(ga_ctx *ctx, size_t need)
{
    if (need <= ctx->arb_bytes) return GA_OK;
    if (ctx->aux) GA_CUDA_TRY(cudaStreamSynchronize(ctx->aux));
    if (ctx->d_arb) cudaFree(ctx->d_arb);
    if (ctx->h_arb) cudaFreeHost(ctx->h_arb);
    ctx->d_arb = nullptr; ctx->h_arb = nullptr; ctx->arb_bytes = 0;
    const size_t want = need + need / 2;
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&ctx->d_arb), want);
    if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void **>(&ctx->h_arb), want, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        if (ctx->d_arb) { cudaFree(ctx->d_arb); ctx->d_arb = nullptr; }
        set_error("ga_arbitrate: allocating %zu bytes of tables failed: %s", want, cudaGetErrorString(e));
        return GA_ERR_NOMEM;
    }
    ctx->arb_bytes = want;
    return GA_OK;
}

extern "C" int ga_arbitrate(ga_ctx *ctx, const uint32_t *seg_offsets, uint32_t n_windows,
                            const uint32_t *stream_ids, const uint8_t *status_in,
                            uint32_t *winner_idx, double *scores, uint8_t *status_out)
{
    GA_GUARD_BEGIN
    if (!ctx) { set_error("ga_arbitrate: NULL context"); return GA_ERR_ARG; }
    if (n_windows == 0) return GA_OK;
    if (!seg_offsets || !winner_idx) { set_error("ga_arbitrate: NULL argument"); return GA_ERR_ARG; }
    const uint32_t n_members = seg_offsets[n_windows];
    if (n_members && (!stream_ids || !status_in || !scores || !status_out)) { set_error("ga_arbitrate: NULL member array"); return GA_ERR_ARG; }
    if (!(ctx->cfg.mode & GA_DO_CONVERT)) { set_error("ga_arbitrate: context does not materialise samples"); return GA_ERR_STATE; }
    for (uint32_t w = 0; w < n_windows; w++)
        if (seg_offsets[w] > seg_offsets[w + 1]) { set_error("ga_arbitrate: seg_offsets not monotone at %u", w); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> alk(ctx->arb_mu);      // the tables are shared; held to the end
    std::unique_lock<std::mutex> lk(ctx->mu);          // stream cursors + queueing; dropped before the wait
    int rc = set_device(ctx);
    if (rc) return rc;
    // device arrays carved from one grow-only scratch allocation owned by the context, filled
    // through a pinned host mirror: inputs | scores (in and out) | winners, statuses (out)
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    const size_t o_off = 0, o_cnt = o_off + up(8ull * n_members), o_st = o_cnt + up(4ull * n_members),
                 o_seg = o_st + up(n_members), o_sc = o_seg + up(4ull * (n_windows + 1)),
                 o_win = o_sc + up(8ull * n_members), o_so = o_win + up(4ull * n_windows),
                 end = o_so + up(n_members), need = end + 256;
    cudaStream_t s = ctx->aux;
    rc = arb_reserve(ctx, need);
    if (rc) return rc;
    uint8_t *b = ctx->d_arb, *hb = ctx->h_arb;
    uint64_t *h_off = reinterpret_cast<uint64_t *>(hb + o_off);
    uint32_t *h_cnt = reinterpret_cast<uint32_t *>(hb + o_cnt);
    for (uint32_t m = 0; m < n_members; m++) {
        if (stream_ids[m] >= ctx->cfg.max_streams) { set_error("ga_arbitrate: member %u: stream_id out of range", m); return GA_ERR_ARG; }
        h_off[m] = static_cast<uint64_t>(stream_ids[m]) * ctx->cfg.arena_samples;
        h_cnt[m] = static_cast<uint32_t>(ctx->streams[stream_ids[m]].wake_n);
    }
    if (n_members) {
        std::memcpy(hb + o_st, status_in, n_members);
        // scores of skipped relays stay what the caller passed in (audio_service.go:522-524)
        std::memcpy(hb + o_sc, scores, 8ull * n_members);
    }
    std::memcpy(hb + o_seg, seg_offsets, 4ull * (n_windows + 1));
    uint64_t *d_off = reinterpret_cast<uint64_t *>(b + o_off);
    uint32_t *d_cnt = reinterpret_cast<uint32_t *>(b + o_cnt), *d_seg = reinterpret_cast<uint32_t *>(b + o_seg),
             *d_win = reinterpret_cast<uint32_t *>(b + o_win);
    uint8_t *d_st = b + o_st, *d_so = b + o_so;
    double *d_sc = reinterpret_cast<double *>(b + o_sc);
    rc = join_inflight(ctx, s);
    if (rc) return rc;
    GA_CUDA_TRY(cudaMemcpyAsync(b, hb, o_win, cudaMemcpyHostToDevice, s));
    rc = ga_k_arbitrate(ctx->d_arena, d_off, d_cnt, d_st, d_seg, n_windows, n_members, d_sc, d_win, d_so, s);
    if (rc) return rc;
    GA_CUDA_TRY(cudaMemcpyAsync(hb + o_sc, b + o_sc, end - o_sc, cudaMemcpyDeviceToHost, s));
    GA_CUDA_TRY(cudaEventRecord(ctx->aux_ev, s));
    ctx->aux_used = true;
    lk.unlock();                                       // ga_submit may go on while the GPU arbitrates
    GA_CUDA_TRY(cudaEventSynchronize(ctx->aux_ev));
    if (n_members) {
        std::memcpy(scores, hb + o_sc, 8ull * n_members);
        std::memcpy(status_out, hb + o_so, n_members);
    }
    std::memcpy(winner_idx, hb + o_win, 4ull * n_windows);
    return GA_OK;
    GA_GUARD_END
}

// ---------------------------------------------------------------- one-shot mirrors
static int one_shot_ingest(ga_ctx *ctx, const uint8_t *data, size_t len, float *out, ga_result *res, uint32_t mode)
{
    if (len > 0xfffffff0ull) { set_error("one-shot message of %zu bytes exceeds the 4 GiB message limit", len); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = set_device(ctx);
    if (rc) return rc;
    DevBuf d_pcm, d_out, d_desc, d_res;
    const size_t padded = (len + 15u) & ~size_t(15);
    const size_t ns = len / 2;
    if ((rc = d_pcm.alloc(padded + 16)) || (rc = d_out.alloc((ns + 4) * sizeof(float))) ||
        (rc = d_desc.alloc(sizeof(ga_desc))) || (rc = d_res.alloc(sizeof(ga_result))))
        return rc;
    ga_desc d{};
    d.nbytes = static_cast<uint32_t>(len);
    cudaStream_t s = ctx->aux;
    if (len) GA_CUDA_TRY(cudaMemcpyAsync(d_pcm.p, data, len, cudaMemcpyHostToDevice, s));
    GA_CUDA_TRY(cudaMemcpyAsync(d_desc.p, &d, sizeof d, cudaMemcpyHostToDevice, s));
    rc = ga_k_ingest(d_pcm.p, d_desc.as<ga_desc>(), 1, d_out.as<float>(), d_res.as<ga_result>(), nullptr, mode,
                     ctx->cfg.vad_threshold, s);
    if (rc) return rc;
    ga_result r;
    GA_CUDA_TRY(cudaMemcpyAsync(&r, d_res.p, sizeof r, cudaMemcpyDeviceToHost, s));
    if (out && ns) GA_CUDA_TRY(cudaMemcpyAsync(out, d_out.p, ns * sizeof(float), cudaMemcpyDeviceToHost, s));
    GA_CUDA_TRY(cudaStreamSynchronize(s));
    if (res) *res = r;
    return GA_OK;
}

extern "C" int ga_bytes_to_float32(ga_ctx *ctx, const uint8_t *data, size_t len, float *out,
                                   size_t *n_out, ga_result *res)
{
    if (!ctx || (len && !data) || (len >= 2 && !out)) { set_error("ga_bytes_to_float32: NULL argument"); return GA_ERR_ARG; }
    if (n_out) *n_out = len / 2;
    if (len == 0) {                       // audio_service.go:1050-1053: empty in, empty out, no error
        if (res) std::memset(res, 0, sizeof *res);
        return GA_OK;
    }
    return one_shot_ingest(ctx, data, len, out, res, GA_DO_CONVERT | GA_DO_FEATURES);
}

extern "C" int ga_sha256(ga_ctx *ctx, const uint8_t *data, size_t len, uint8_t digest[32])
{
    if (!ctx || (len && !data) || !digest) { set_error("ga_sha256: NULL argument"); return GA_ERR_ARG; }
    ga_result r;
    int rc = one_shot_ingest(ctx, data, len, nullptr, &r, GA_DO_SHA256);
    if (rc) return rc;
    std::memcpy(digest, r.digest, 32);
    return GA_OK;
}

// ---------------------------------------------------------------- voice-event hook
extern "C" void ga_format_audio_hash(const uint8_t digest[32], char *audio_hash)
{
    static const char hex[] = "0123456789abcdef";
    std::memcpy(audio_hash, "sha256:", 7);
    for (int i = 0; i < 32; i++) {
        audio_hash[7 + 2 * i] = hex[digest[i] >> 4];
        audio_hash[8 + 2 * i] = hex[digest[i] & 15];
    }
    audio_hash[71] = 0;
}

// events/voice_event.go:84-88: float64(len(audioData)) / float64(sampleRate), Go semantics (a zero
// rate gives +Inf / NaN, never an error); the fingerprint is over the samples as the hub
// serialises them (stt_client.go:390-395).
extern "C" int ga_set_audio_metadata(ga_ctx *ctx, const float *audio_data, size_t n, int sample_rate,
                                     int zero_persistence, double *audio_duration, char *audio_hash)
{
    if (!ctx || (n && !audio_data) || !audio_duration || !audio_hash) { set_error("ga_set_audio_metadata: NULL argument"); return GA_ERR_ARG; }
    *audio_duration = static_cast<double>(n) / static_cast<double>(sample_rate);
    audio_hash[0] = 0;
    if (zero_persistence) return GA_OK;
    uint8_t digest[32];
    int rc = ga_sha256(ctx, reinterpret_cast<const uint8_t *>(audio_data), n * sizeof(float), digest);
    if (rc) return rc;
    ga_format_audio_hash(digest, audio_hash);
    return GA_OK;
}

extern "C" int ga_stream_audio_metadata(ga_ctx *ctx, uint32_t stream_id, int sample_rate, int zero_persistence,
                                        double *audio_duration, char *audio_hash)
{
    if (!ctx || !audio_duration || !audio_hash) { set_error("ga_stream_audio_metadata: NULL argument"); return GA_ERR_ARG; }
    uint64_t n = 0;
    int rc = ga_stream_info(ctx, stream_id, &n, nullptr);
    if (rc) return rc;
    *audio_duration = static_cast<double>(n) / static_cast<double>(sample_rate);
    audio_hash[0] = 0;
    if (zero_persistence) return GA_OK;
    uint8_t digest[32];
    rc = ga_samples_digest(ctx, stream_id, digest);
    if (rc) return rc;
    ga_format_audio_hash(digest, audio_hash);
    return GA_OK;
}

extern "C" int ga_signal_strength(ga_ctx *ctx, const float *samples, size_t n, double *rms, ga_result *res)
{
    if (!ctx || (n && !samples) || !rms) { set_error("ga_signal_strength: NULL argument"); return GA_ERR_ARG; }
    if (n == 0) {                         // audio_service.go:819-822
        *rms = 0.0;
        if (res) std::memset(res, 0, sizeof *res);
        return GA_OK;
    }
    if (n > 0x3fffffffull) { set_error("ga_signal_strength: %zu samples exceed the 2^30 limit", n); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = set_device(ctx);
    if (rc) return rc;
    DevBuf d_s, d_off, d_cnt, d_sc, d_stats;
    if ((rc = d_s.alloc((n + 4) * sizeof(float))) || (rc = d_off.alloc(8)) || (rc = d_cnt.alloc(4)) ||
        (rc = d_sc.alloc(8)) || (rc = d_stats.alloc(sizeof(ga_result))))
        return rc;
    const uint64_t off = 0;
    const uint32_t cnt = static_cast<uint32_t>(n);
    cudaStream_t s = ctx->aux;
    GA_CUDA_TRY(cudaMemcpyAsync(d_s.p, samples, n * sizeof(float), cudaMemcpyHostToDevice, s));
    GA_CUDA_TRY(cudaMemcpyAsync(d_off.p, &off, 8, cudaMemcpyHostToDevice, s));
    GA_CUDA_TRY(cudaMemcpyAsync(d_cnt.p, &cnt, 4, cudaMemcpyHostToDevice, s));
    GA_CUDA_TRY(cudaMemsetAsync(d_stats.p, 0, sizeof(ga_result), s));
    rc = ga_k_signal_strength(d_s.as<float>(), d_off.as<uint64_t>(), d_cnt.as<uint32_t>(), nullptr, 1,
                              d_sc.as<double>(), d_stats.as<ga_result>(), s);
    if (rc) return rc;
    ga_result r;
    GA_CUDA_TRY(cudaMemcpyAsync(&r, d_stats.p, sizeof r, cudaMemcpyDeviceToHost, s));
    GA_CUDA_TRY(cudaMemcpyAsync(rms, d_sc.p, 8, cudaMemcpyDeviceToHost, s));
    GA_CUDA_TRY(cudaStreamSynchronize(s));
    if (res) *res = r;
    return GA_OK;
}

// ---------------------------------------------------------------- synthetic PCM
extern "C" int ga_synth_fill_host(int16_t *out, uint64_t seed, uint32_t group, uint32_t relay,
                                  uint32_t first, uint32_t n_samples, uint32_t utt_samples)
{
    if (!out && n_samples) { set_error("ga_synth_fill_host: NULL output"); return GA_ERR_ARG; }
    for (uint32_t i = 0; i < n_samples; i++) out[i] = synth::sample(seed, group, relay, first + i, utt_samples);
    return GA_OK;
}

extern "C" int ga_synth_fill_batch_host(void *pcm, uint64_t seed, uint32_t group_base, uint32_t n_streams,
                                        uint32_t group_size, uint32_t n_msgs, uint32_t msg_samples,
                                        uint32_t n_threads)
{
    if (!pcm || group_size == 0) { set_error("ga_synth_fill_batch_host: bad argument"); return GA_ERR_ARG; }
    if (n_threads == 0) n_threads = 1;
    int16_t *out = static_cast<int16_t *>(pcm);
    const uint32_t utt = n_msgs * msg_samples;
    std::atomic<uint32_t> next{0};
    auto work = [&]() {
        for (;;) {
            const uint32_t s = next.fetch_add(1);
            if (s >= n_streams) break;
            for (uint32_t m = 0; m < n_msgs; m++) {
                int16_t *dst = out + (static_cast<uint64_t>(m) * n_streams + s) * msg_samples;
                for (uint32_t k = 0; k < msg_samples; k++)
                    dst[k] = synth::sample(seed, group_base + s / group_size, s % group_size, m * msg_samples + k, utt);
            }
        }
    };
    std::vector<std::thread> th;
    for (uint32_t i = 1; i < n_threads; i++) th.emplace_back(work);
    work();
    for (auto &t : th) t.join();
    return GA_OK;
}

__global__ void synth_batch_kernel(int16_t *pcm, uint64_t seed, uint32_t group_base, uint32_t n_streams,
                                   uint32_t group_size, uint32_t n_msgs, uint32_t msg_samples)
{
    const uint64_t total = static_cast<uint64_t>(n_msgs) * n_streams * msg_samples;
    const uint32_t utt = n_msgs * msg_samples;
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint32_t k = static_cast<uint32_t>(i % msg_samples);
        const uint64_t ms = i / msg_samples;
        const uint32_t s = static_cast<uint32_t>(ms % n_streams);
        const uint32_t m = static_cast<uint32_t>(ms / n_streams);
        pcm[i] = synth::sample(seed, group_base + s / group_size, s % group_size, m * msg_samples + k, utt);
    }
}

__global__ void synth_batch_map_kernel(int16_t *pcm, uint64_t seed, const uint32_t *__restrict__ group,
                                       const uint32_t *__restrict__ relay, uint32_t n_streams,
                                       uint32_t n_msgs, uint32_t msg_samples)
{
    const uint64_t total = static_cast<uint64_t>(n_msgs) * n_streams * msg_samples;
    const uint32_t utt = n_msgs * msg_samples;
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint32_t k = static_cast<uint32_t>(i % msg_samples);
        const uint64_t ms = i / msg_samples;
        const uint32_t s = static_cast<uint32_t>(ms % n_streams);
        const uint32_t m = static_cast<uint32_t>(ms / n_streams);
        pcm[i] = synth::sample(seed, group[s], relay[s], m * msg_samples + k, utt);
    }
}

extern "C" int ga_synth_fill_batch_device_map(void *d_pcm, uint64_t seed, const uint32_t *d_group,
                                              const uint32_t *d_relay, uint32_t n_streams,
                                              uint32_t n_msgs, uint32_t msg_samples, void *stream)
{
    if (!d_pcm || !d_group || !d_relay) { set_error("ga_synth_fill_batch_device_map: bad argument"); return GA_ERR_ARG; }
    const uint64_t total = static_cast<uint64_t>(n_msgs) * n_streams * msg_samples;
    if (total == 0) return GA_OK;
    int sms = 0;
    int rc = num_sms(&sms);
    if (rc) return rc;
    synth_batch_map_kernel<<<sms * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<int16_t *>(d_pcm), seed, d_group, d_relay, n_streams, n_msgs, msg_samples);
    GA_CHECK_LAUNCH("synth_batch_map_kernel");
    return GA_OK;
}

// 48 kHz stereo view of the same relay PCM: 3x zero-order hold, R = L - (L >> 3)
__global__ void synth_batch48_map_kernel(uint32_t *pcm, uint64_t seed, const uint32_t *__restrict__ group,
                                         const uint32_t *__restrict__ relay, uint32_t n_streams,
                                         uint32_t n_msgs, uint32_t msg_frames)
{
    const uint64_t total = static_cast<uint64_t>(n_msgs) * n_streams * msg_frames;
    const uint32_t utt = (n_msgs * msg_frames + 2) / 3;
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint32_t k = static_cast<uint32_t>(i % msg_frames);
        const uint64_t ms = i / msg_frames;
        const uint32_t s = static_cast<uint32_t>(ms % n_streams);
        const uint32_t m = static_cast<uint32_t>(ms / n_streams);
        const int32_t l = synth::sample(seed, group[s], relay[s], (m * msg_frames + k) / 3, utt);
        const int32_t r = l - (l >> 3);
        pcm[i] = (static_cast<uint32_t>(l) & 0xFFFFu) | (static_cast<uint32_t>(r) << 16);
    }
}

extern "C" int ga_synth_fill_batch48_device_map(void *d_pcm, uint64_t seed, const uint32_t *d_group,
                                                const uint32_t *d_relay, uint32_t n_streams,
                                                uint32_t n_msgs, uint32_t msg_frames, void *stream)
{
    if (!d_pcm || !d_group || !d_relay) { set_error("ga_synth_fill_batch48_device_map: bad argument"); return GA_ERR_ARG; }
    const uint64_t total = static_cast<uint64_t>(n_msgs) * n_streams * msg_frames;
    if (total == 0) return GA_OK;
    int sms = 0;
    int rc = num_sms(&sms);
    if (rc) return rc;
    synth_batch48_map_kernel<<<sms * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<uint32_t *>(d_pcm), seed, d_group, d_relay, n_streams, n_msgs, msg_frames);
    GA_CHECK_LAUNCH("synth_batch48_map_kernel");
    return GA_OK;
}

extern "C" int ga_synth_fill_batch_device(void *d_pcm, uint64_t seed, uint32_t group_base,
                                          uint32_t n_streams, uint32_t group_size, uint32_t n_msgs,
                                          uint32_t msg_samples, void *stream)
{
    if (!d_pcm || group_size == 0) { set_error("ga_synth_fill_batch_device: bad argument"); return GA_ERR_ARG; }
    const uint64_t total = static_cast<uint64_t>(n_msgs) * n_streams * msg_samples;
    if (total == 0) return GA_OK;
    int sms = 0;
    int rc = num_sms(&sms);
    if (rc) return rc;
    synth_batch_kernel<<<sms * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<int16_t *>(d_pcm), seed, group_base, n_streams, group_size, n_msgs, msg_samples);
    GA_CHECK_LAUNCH("synth_batch_kernel");
    return GA_OK;
}

// ---------------------------------------------------------------- cluster: one process, several GPUs
// The hub is ONE process; on an 8-GPU box it holds one context per GPU and assigns a collision group
// to a shard when its window opens.  When the relays of a window ended up on different shards,
// performArbitration (audio_service.go:505-571) needs the scores of all of them on one GPU: every
// shard scores its own relays, writes the scores of relays whose window is homed elsewhere straight
// into that shard's inbox over NVLink (peer stores from inside the kernel) and raises a flag there;
// the home shard's kernel waits for its senders' flags and runs the selection — ONE launch per GPU
// (ga_k_arbitrate_p2p), no host round trip between scoring and selection.
namespace ga {

struct ClusterShard {
    ga_ctx *ctx = nullptr;
    int dev = 0;
    cudaStream_t stream = nullptr;
    uint8_t *region = nullptr;            // inbox[2][cap] f64 | flags[n_shards] u64 | counter, error u32
    uint8_t *scratch = nullptr;           // grow-only device tables ...
    uint8_t *h_stage = nullptr;           // ... and their pinned host mirror: one copy in, one copy out per call
    size_t scratch_bytes = 0;
    // per-call host tables
    std::vector<uint64_t> off, guest_dst, notify, wait;
    std::vector<uint32_t> cnt, seg, route_dst, member_of, windows;
    std::vector<uint8_t> st;
    std::vector<double> sc;
    uint32_t n_home = 0, n_guests = 0;
    struct Incoming { uint32_t from, member, pos; };
    std::vector<Incoming> incoming;
    std::vector<uint32_t> win_out;
    std::vector<uint8_t> st_out;
    uint32_t err = 0;
    bool launched = false;
    size_t o_win = 0, o_sc = 0, o_so = 0, o_err = 0;
};

}  // namespace ga

struct ga_cluster {
    std::vector<ga::ClusterShard> shards;
    uint32_t cap = 0;                      // inbox slots per shard and parity
    uint64_t parity_stride = 0, flags_off = 0, ctl_off = 0, region_bytes = 0;
    uint64_t epoch = 0;
    std::mutex mu;
};

extern "C" void ga_cluster_destroy(ga_cluster *cl)
{
    if (!cl) return;
    for (auto &s : cl->shards) {
        cudaSetDevice(s.dev);
        if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); }
        if (s.region) cudaFree(s.region);
        if (s.scratch) cudaFree(s.scratch);
        if (s.h_stage) cudaFreeHost(s.h_stage);
    }
    delete cl;
}

extern "C" int ga_cluster_create(ga_ctx *const *ctxs, uint32_t n_ctx, uint32_t max_guests, ga_cluster **out)
{
    GA_GUARD_BEGIN
    if (!ctxs || !out || n_ctx == 0 || n_ctx > 64) { set_error("ga_cluster_create: bad argument"); return GA_ERR_ARG; }
    *out = nullptr;
    for (uint32_t i = 0; i < n_ctx; i++) {
        if (!ctxs[i]) { set_error("ga_cluster_create: NULL context %u", i); return GA_ERR_ARG; }
        if (!(ctxs[i]->cfg.mode & GA_DO_CONVERT)) { set_error("ga_cluster_create: context %u does not materialise samples", i); return GA_ERR_STATE; }
        for (uint32_t j = 0; j < i; j++)
            if (ctxs[j]->cfg.device == ctxs[i]->cfg.device) {
                // two waiting kernels on one GPU could starve each other: one shard per device
                set_error("ga_cluster_create: contexts %u and %u share device %d", j, i, ctxs[i]->cfg.device);
                return GA_ERR_ARG;
            }
    }
    ga_cluster *cl = new (std::nothrow) ga_cluster;
    if (!cl) { set_error("ga_cluster_create: out of host memory"); return GA_ERR_NOMEM; }
    cl->cap = max_guests ? max_guests : 1;
    cl->parity_stride = 8ull * cl->cap;
    cl->flags_off = 2 * cl->parity_stride;
    cl->ctl_off = cl->flags_off + 8ull * n_ctx;
    cl->region_bytes = cl->ctl_off + 16;
    try { cl->shards.resize(n_ctx); } catch (...) { delete cl; set_error("ga_cluster_create: out of host memory"); return GA_ERR_NOMEM; }
    auto fail = [&](int code) { ga_cluster_destroy(cl); return code; };
    for (uint32_t i = 0; i < n_ctx; i++) {
        ClusterShard &s = cl->shards[i];
        s.ctx = ctxs[i];
        s.dev = ctxs[i]->cfg.device;
        cudaError_t e = cudaSetDevice(s.dev);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&s.region), cl->region_bytes);
        if (e == cudaSuccess) e = cudaMemset(s.region, 0, cl->region_bytes);
        if (e != cudaSuccess) { cudaGetLastError(); set_error("ga_cluster_create: device %d: %s", s.dev, cudaGetErrorString(e)); return fail(GA_ERR_CUDA); }
        for (uint32_t j = 0; j < n_ctx; j++) {
            if (j == i) continue;
            int can = 0;
            e = cudaDeviceCanAccessPeer(&can, s.dev, ctxs[j]->cfg.device);
            if (e != cudaSuccess || !can) {
                cudaGetLastError();
                set_error("ga_cluster_create: device %d cannot address device %d (no NVLink / PCIe peer path)", s.dev, ctxs[j]->cfg.device);
                return fail(GA_ERR_COMM);
            }
            e = cudaDeviceEnablePeerAccess(ctxs[j]->cfg.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                set_error("ga_cluster_create: cudaDeviceEnablePeerAccess(%d -> %d): %s", s.dev, ctxs[j]->cfg.device, cudaGetErrorString(e));
                return fail(GA_ERR_COMM);
            }
            cudaGetLastError();
        }
    }
    for (auto &s : cl->shards) { cudaSetDevice(s.dev); cudaDeviceSynchronize(); }
    *out = cl;
    return GA_OK;
    GA_GUARD_END
}

extern "C" uint32_t ga_cluster_size(const ga_cluster *cl) { return cl ? static_cast<uint32_t>(cl->shards.size()) : 0u; }
extern "C" ga_ctx *ga_cluster_context(const ga_cluster *cl, uint32_t shard)
{
    return (cl && shard < cl->shards.size()) ? cl->shards[shard].ctx : nullptr;
}

extern "C" int ga_cluster_arbitrate(ga_cluster *cl, const uint32_t *seg_offsets, uint32_t n_windows,
                                    const uint32_t *shard_ids, const uint32_t *stream_ids, const uint8_t *status_in,
                                    uint32_t *winner_idx, double *scores, uint8_t *status_out)
{
    GA_GUARD_BEGIN
    if (!cl) { set_error("ga_cluster_arbitrate: NULL cluster"); return GA_ERR_ARG; }
    if (n_windows == 0) return GA_OK;
    if (!seg_offsets || !winner_idx) { set_error("ga_cluster_arbitrate: NULL argument"); return GA_ERR_ARG; }
    const uint32_t n_members = seg_offsets[n_windows];
    if (n_members && (!shard_ids || !stream_ids || !status_in || !scores || !status_out)) { set_error("ga_cluster_arbitrate: NULL member array"); return GA_ERR_ARG; }
    const uint32_t n_sh = static_cast<uint32_t>(cl->shards.size());
    for (uint32_t w = 0; w < n_windows; w++)
        if (seg_offsets[w] > seg_offsets[w + 1]) { set_error("ga_cluster_arbitrate: seg_offsets not monotone at %u", w); return GA_ERR_ARG; }
    for (uint32_t m = 0; m < n_members; m++)
        if (shard_ids[m] >= n_sh || stream_ids[m] >= cl->shards[shard_ids[m]].ctx->cfg.max_streams) {
            set_error("ga_cluster_arbitrate: member %u: shard %u / stream %u out of range", m, shard_ids[m], stream_ids[m]);
            return GA_ERR_ARG;
        }
    std::lock_guard<std::mutex> lk(cl->mu);

    // ---- plan: a window is homed on the shard of its first member (join order)
    for (auto &s : cl->shards) {
        s.off.clear(); s.cnt.clear(); s.st.clear(); s.sc.clear(); s.seg.assign(1, 0); s.member_of.clear(); s.windows.clear();
        s.incoming.clear(); s.guest_dst.clear(); s.notify.clear(); s.wait.clear(); s.route_dst.clear();
        s.n_home = s.n_guests = 0; s.err = 0;
    }
    // the streams' wake-word cursors are read under the contexts' locks, taken once each, in shard order
    std::vector<std::unique_lock<std::mutex>> ctx_locks;
    ctx_locks.reserve(n_sh);
    for (auto &s : cl->shards) ctx_locks.emplace_back(s.ctx->mu);
    auto wake_of = [&](uint32_t shard, uint32_t sid, uint64_t &off, uint32_t &cnt) {
        ga_ctx *c = cl->shards[shard].ctx;
        off = static_cast<uint64_t>(sid) * c->cfg.arena_samples;
        cnt = static_cast<uint32_t>(c->streams[sid].wake_n);
    };
    for (uint32_t w = 0; w < n_windows; w++) {
        const uint32_t lo = seg_offsets[w], hi = seg_offsets[w + 1];
        if (lo == hi) { winner_idx[w] = 0xffffffffu; continue; }
        const uint32_t h = shard_ids[lo];
        ClusterShard &s = cl->shards[h];
        s.windows.push_back(w);
        for (uint32_t m = lo; m < hi; m++) {
            uint64_t off = 0; uint32_t cnt = 0;
            const uint32_t pos = static_cast<uint32_t>(s.off.size());
            if (shard_ids[m] == h) wake_of(h, stream_ids[m], off, cnt);
            else if (status_in[m] == GA_RELAY_CONTENDING) s.incoming.push_back({shard_ids[m], m, pos});
            s.off.push_back(off); s.cnt.push_back(cnt); s.st.push_back(status_in[m]); s.sc.push_back(scores[m]);
            s.member_of.push_back(m);
        }
        s.seg.push_back(static_cast<uint32_t>(s.off.size()));
    }
    for (uint32_t h = 0; h < n_sh; h++) {
        ClusterShard &s = cl->shards[h];
        s.n_home = static_cast<uint32_t>(s.off.size());
        if (s.incoming.size() > cl->cap) {
            set_error("ga_cluster_arbitrate: shard %u receives %zu remote relays, cluster was created for %u", h, s.incoming.size(), cl->cap);
            return GA_ERR_OVERFLOW;
        }
    }
    // guests: the k-th remote relay of home shard h lands in inbox slot k there
    std::vector<uint8_t> sends(static_cast<size_t>(n_sh) * n_sh, 0);     // sends[from * n + to]
    for (uint32_t h = 0; h < n_sh; h++) {
        ClusterShard &home = cl->shards[h];
        for (uint32_t k = 0; k < home.incoming.size(); k++) {
            const auto &in = home.incoming[k];
            ClusterShard &g = cl->shards[in.from];
            uint64_t off; uint32_t cnt;
            wake_of(in.from, stream_ids[in.member], off, cnt);
            g.off.push_back(off); g.cnt.push_back(cnt); g.st.push_back(GA_RELAY_CONTENDING); g.sc.push_back(0.0);
            g.guest_dst.push_back(reinterpret_cast<uint64_t>(home.region) + 8ull * k);
            g.n_guests++;
            home.route_dst.push_back(in.pos);
            sends[static_cast<size_t>(in.from) * n_sh + h] = 1;
        }
    }
    for (uint32_t a = 0; a < n_sh; a++)
        for (uint32_t b = 0; b < n_sh; b++)
            if (sends[static_cast<size_t>(a) * n_sh + b]) {
                cl->shards[a].notify.push_back(reinterpret_cast<uint64_t>(cl->shards[b].region) + cl->flags_off + 8ull * a);
                cl->shards[b].wait.push_back(reinterpret_cast<uint64_t>(cl->shards[b].region) + cl->flags_off + 8ull * a);
            }

    ctx_locks.clear();

    // ---- one launch per GPU; nothing is waited for until all of them are queued
    const uint64_t epoch = ++cl->epoch;
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    int rc = GA_OK;
    for (uint32_t i = 0; i < n_sh && rc == GA_OK; i++) {
        ClusterShard &s = cl->shards[i];
        const size_t n_all = s.off.size(), nw = s.windows.size();
        if (n_all == 0 && s.notify.empty()) continue;
        auto step = [&]() -> int {
            GA_CUDA_TRY(cudaSetDevice(s.dev));
            // inputs | error word, scores (in and out) | winners, statuses (out): the tables go up in ONE
            // copy from the pinned mirror and the results come back in ONE
            const size_t o_off = 0, o_cnt = o_off + up(8 * n_all), o_st = o_cnt + up(4 * n_all), o_seg = o_st + up(n_all),
                         o_gd = o_seg + up(4 * (nw + 1)), o_no = o_gd + up(8 * s.guest_dst.size()),
                         o_wa = o_no + up(8 * s.notify.size()), o_rd = o_wa + up(8 * s.wait.size()),
                         o_err = o_rd + up(4 * s.route_dst.size()), o_sc = o_err + 256, o_win = o_sc + up(8 * n_all),
                         o_so = o_win + up(4 * nw), need = o_so + up(s.n_home) + 256;
            if (need > s.scratch_bytes) {
                GA_CUDA_TRY(cudaStreamSynchronize(s.stream));
                if (s.scratch) cudaFree(s.scratch);
                if (s.h_stage) cudaFreeHost(s.h_stage);
                s.scratch = nullptr; s.h_stage = nullptr; s.scratch_bytes = 0;
                const size_t want = need + need / 2;
                cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&s.scratch), want);
                if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void **>(&s.h_stage), want, cudaHostAllocDefault);
                if (e != cudaSuccess) {
                    cudaGetLastError();
                    if (s.scratch) { cudaFree(s.scratch); s.scratch = nullptr; }
                    set_error("ga_cluster_arbitrate: allocating %zu bytes of tables failed: %s", want, cudaGetErrorString(e));
                    return GA_ERR_NOMEM;
                }
                s.scratch_bytes = want;
            }
            uint8_t *b = s.scratch;
            auto put = [&](size_t o, const void *src, size_t bytes) -> int {
                if (bytes) std::memcpy(s.h_stage + o, src, bytes);
                return GA_OK;
            };
            std::memset(s.h_stage + o_err, 0, 4);
            {   // the wake-word buffers must be complete: wait for every ticket of this context in flight
                std::lock_guard<std::mutex> g(s.ctx->mu);
                int r = join_inflight(s.ctx, s.stream);
                if (r) return r;
            }
            int r;
            if ((r = put(o_off, s.off.data(), 8 * n_all)) || (r = put(o_cnt, s.cnt.data(), 4 * n_all)) ||
                (r = put(o_st, s.st.data(), n_all)) || (r = put(o_sc, s.sc.data(), 8 * n_all)) ||
                (r = put(o_seg, s.seg.data(), 4 * (nw + 1))) || (r = put(o_gd, s.guest_dst.data(), 8 * s.guest_dst.size())) ||
                (r = put(o_no, s.notify.data(), 8 * s.notify.size())) || (r = put(o_wa, s.wait.data(), 8 * s.wait.size())) ||
                (r = put(o_rd, s.route_dst.data(), 4 * s.route_dst.size())))
                return r;
            GA_CUDA_TRY(cudaMemcpyAsync(b, s.h_stage, o_win, cudaMemcpyHostToDevice, s.stream));
            ga_p2p_args a;
            std::memset(&a, 0, sizeof a);
            a.d_samples = s.ctx->d_arena;
            a.d_wake_off = reinterpret_cast<uint64_t *>(b + o_off); a.d_wake_n = reinterpret_cast<uint32_t *>(b + o_cnt);
            a.d_status_in = b + o_st; a.d_seg_offsets = reinterpret_cast<uint32_t *>(b + o_seg);
            a.d_scores = reinterpret_cast<double *>(b + o_sc); a.d_winner = reinterpret_cast<uint32_t *>(b + o_win);
            a.d_status_out = b + o_so;
            a.d_guest_dst = reinterpret_cast<uint64_t *>(b + o_gd); a.d_notify = reinterpret_cast<uint64_t *>(b + o_no);
            a.d_wait = reinterpret_cast<uint64_t *>(b + o_wa); a.d_inbox = reinterpret_cast<double *>(s.region);
            a.d_route_dst = reinterpret_cast<uint32_t *>(b + o_rd);
            a.d_counter = reinterpret_cast<uint32_t *>(s.region + cl->ctl_off);
            a.d_error = reinterpret_cast<uint32_t *>(b + o_err);
            a.inbox_parity_stride = cl->parity_stride;
            a.epoch = epoch;
            a.timeout_cycles = 4'000'000'000ll;          // ~2 s: a peer that never launched must not hang this GPU
            a.n_windows = static_cast<uint32_t>(nw); a.n_home_members = s.n_home; a.n_guests = s.n_guests;
            a.n_notify = static_cast<uint32_t>(s.notify.size()); a.n_wait = static_cast<uint32_t>(s.wait.size());
            a.n_in = static_cast<uint32_t>(s.route_dst.size());
            r = ga_k_arbitrate_p2p(&a, s.stream);
            if (r) return r;
            // into PINNED memory, so this does not block the host: the kernel may be waiting for a
            // shard that is only launched later in this loop
            GA_CUDA_TRY(cudaMemcpyAsync(s.h_stage + o_err, b + o_err, need - 256 - o_err, cudaMemcpyDeviceToHost, s.stream));
            s.launched = true; s.o_win = o_win; s.o_sc = o_sc; s.o_so = o_so; s.o_err = o_err;
            return GA_OK;
        };
        rc = step();
    }
    // collect (also after a failure: the shards already launched finish on their own, by data or by timeout)
    for (uint32_t i = 0; i < n_sh; i++) {
        ClusterShard &s = cl->shards[i];
        if (!s.launched) continue;
        s.launched = false;
        cudaSetDevice(s.dev);
        const size_t nw = s.windows.size();
        cudaError_t e = cudaStreamSynchronize(s.stream);
        if (e != cudaSuccess) {
            cudaGetLastError();
            if (rc == GA_OK) { set_error("ga_cluster_arbitrate: device %d: %s", s.dev, cudaGetErrorString(e)); rc = GA_ERR_CUDA; }
            continue;
        }
        s.win_out.resize(nw); s.st_out.resize(s.n_home);
        std::memcpy(&s.err, s.h_stage + s.o_err, 4);
        if (nw) std::memcpy(s.win_out.data(), s.h_stage + s.o_win, 4 * nw);
        if (s.n_home) {
            std::memcpy(s.sc.data(), s.h_stage + s.o_sc, 8ull * s.n_home);
            std::memcpy(s.st_out.data(), s.h_stage + s.o_so, s.n_home);
        }
        if (s.err && rc == GA_OK) {
            set_error("ga_cluster_arbitrate: shard %u timed out waiting for a peer's scores", i);
            rc = GA_ERR_COMM;
        }
        s.err = 0;
    }
    if (rc) return rc;
    for (auto &s : cl->shards) {
        for (size_t k = 0; k < s.windows.size(); k++) {
            const uint32_t w = s.windows[k], p = s.win_out[k];
            winner_idx[w] = p == 0xffffffffu ? 0xffffffffu : s.member_of[p];
        }
        for (uint32_t p = 0; p < s.n_home; p++) {
            const uint32_t m = s.member_of[p];
            if (status_in[m] == GA_RELAY_CONTENDING) scores[m] = s.sc[p];   // :528; skipped relays keep theirs (:522-524)
            status_out[m] = s.st_out[p];
        }
    }
    return GA_OK;
    GA_GUARD_END
}

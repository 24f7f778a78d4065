// hub.cu — host-side mirror of the reference's AudioService for the ingest path: the relay
// registry, the arbitration-window lifecycle and the per-chunk routing of StreamAudio, on top
// of a ga_ctx.  Pure host C++ (no kernels here); it is what the Go code around the C ABI does,
// written out so that the window rules can be tested here where no Go toolchain exists and so
// that a maintainer can see exactly which reference line each rule comes from.
//
// Reference (internal/grpc/audio_service.go):
//   RelayStream :44-53, RelayStatus :56-63, ArbitrationWindow :66-73
//   startArbitrationWindow :408-446, joinArbitrationWindow :449-491, runArbitrationTimer :494-502
//   performArbitration :505-587, isRelayActive :880-913, cleanupRelay :916-923
//   StreamAudio per-chunk body :939-1041
// Difference by design: the reference holds ONE hub-wide window
// (AudioService.arbitrationWindow, :89); here there is one per collision group, so thousands of
// rooms arbitrate concurrently.  With a single group the behaviour is the reference's.
// Time is injected (now_us) instead of read from a clock, and the timer goroutine (:494-502) is
// ga_hub_poll: every window whose duration has elapsed is arbitrated in ONE ga_arbitrate call.
//
// Locking.  The reference takes its global streamsMutex for EVERY chunk (:1000-1004).  Here the hub
// mutex guards the registry and the windows only: a wake-word chunk takes it to start / join its
// group's window (the first ~15 of an utterance's 250 chunks); a speech chunk takes no hub-wide
// lock at all.  The chunk's bytes are copied into the pinned ring outside any lock, and the message
// is published on a per-GPU multi-producer queue whose sequence numbers are handed out together
// with the ring reservation (a mutex per GPU held for those two operations only), so that the
// flush always takes a prefix of the queue that is complete both in arrival order and in ring
// offsets — the ring is released by offset when a ticket is waited.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "common.cuh"

namespace {

struct Relay {                       // RelayStream :44-53; lives in a fixed table, index = shard * max_slots + slot
    std::string id;
    uint32_t group = 0;
    uint32_t shard = 0;              // which GPU's context holds the relay's audio
    uint32_t slot = 0;               // ga_ctx stream slot there: WakeWordSignal / SpeechAudio live on the GPU
    std::atomic<int> status{GA_RELAY_CONNECTED};   // :56-63 (read by the relay's own chunk path without the hub lock)
    double signal_strength = 0.0;    // :50
    bool cancel_closed = false;      // close(relay.CancelChannel) :561
    std::atomic<bool> in_active_streams{false};    // as.activeStreams[relayID] exists (:91)
    std::atomic<bool> first_chunk{true};
    uint64_t window = 0;             // window it last joined (0 = none)
    std::atomic<uint32_t> gen{0};    // bumped when the slot changes hands: stale handles and queued chunks are refused
    bool alive = false;
    uint64_t dropped_chunks = 0;     // chunks the flush had to drop (sample arena full)
};

struct Window {                      // ArbitrationWindow :66-73
    uint64_t id = 0;
    uint32_t group = 0;
    uint64_t start_us = 0;           // StartTime
    uint64_t duration_us = 0;        // WindowDuration
    std::vector<std::string> relays; // Relays, in JOIN ORDER (the deterministic stand-in for map order)
    bool is_active = true;           // IsActive
    std::string winner;              // WinnerID
};

struct QMsg {
    ga_msg m;
    uint32_t gen;                    // the slot's generation when the chunk arrived
};

struct alignas(64) QEntry {          // one cache line per entry: neighbouring entries are written by different threads
    QMsg v;
    std::atomic<uint64_t> ready{0};  // sequence number + 1 once v is published
};

// Chunks pushed since the last flush, per GPU: bounded multi-producer / single-consumer queue.
// Held for a ring reservation plus one increment (tens of nanoseconds) by every relay thread of a
// GPU: a spin lock — a futex-based mutex puts contending threads to sleep for longer than that.
struct SpinLock {
    std::atomic<bool> held{false};
    void lock()
    {
        for (int spins = 0;; spins++) {
            if (!held.load(std::memory_order_relaxed) && !held.exchange(true, std::memory_order_acquire)) return;
            if (spins < 256) {
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            } else {
                std::this_thread::yield();
            }
        }
    }
    void unlock() { held.store(false, std::memory_order_release); }
};

struct ShardQueue {
    alignas(64) SpinLock claim;      // held for {ring reservation, sequence number} only
    uint64_t tail = 0;               // next sequence number (under claim; same cache line as the lock)
    alignas(64) std::atomic<uint64_t> head{0};   // first sequence number the flush has not taken yet (written by the flush only)
    std::unique_ptr<QEntry[]> q;
    uint64_t cap = 0;                // power of two
    std::vector<QMsg> carry;         // taken off the queue but not submitted (flush only)
    uint64_t taken_end = 0;          // ring offset up to which the flush has taken chunks
};

}  // namespace

struct ga_hub {
    // no context: bookkeeping only; one: single GPU; several (with their cluster): one hub process
    // driving a multi-GPU box, relays placed by collision group, windows may span GPUs
    std::vector<ga_ctx *> ctxs;
    ga_cluster *cluster = nullptr;
    uint64_t window_duration_us = 300000;                 // arbitrationWindowDuration, 300 ms (:243)
    std::unique_ptr<Relay[]> table;                       // n_shards * max_slots entries, never moved
    std::unordered_map<std::string, uint32_t> by_id;      // relay id -> table index
    std::map<uint64_t, Window> windows;                   // every window ever opened and not yet dropped
    std::unordered_map<uint32_t, uint64_t> group_window;  // as.arbitrationWindow per group (absent = nil)
    std::vector<std::vector<uint32_t>> free_slots;        // per shard
    std::vector<uint32_t> next_slot;
    uint32_t max_slots = 0;                               // stream slots per shard
    uint32_t n_shards = 1;
    uint64_t next_window = 1;
    std::unique_ptr<ShardQueue[]> queues;                 // per shard
    std::mutex mu;                                        // registry + windows
    std::mutex flush_mu;                                  // one flush at a time
};

using namespace ga;

static Relay *find_relay(ga_hub *h, const char *id)
{
    auto it = h->by_id.find(id);
    return it == h->by_id.end() ? nullptr : &h->table[it->second];
}

static inline uint64_t handle_of(const ga_hub *h, const Relay *r)
{
    return (static_cast<uint64_t>(r->gen.load(std::memory_order_relaxed)) << 32) | (r->shard * h->max_slots + r->slot);
}

static Relay *relay_of_handle(ga_hub *h, uint64_t handle)
{
    const uint32_t idx = static_cast<uint32_t>(handle), gen = static_cast<uint32_t>(handle >> 32);
    if (idx >= h->n_shards * h->max_slots) return nullptr;
    Relay *r = &h->table[idx];
    return r->gen.load(std::memory_order_acquire) == gen && (gen & 1u) ? r : nullptr;     // odd generation = slot in use
}

static int hub_alloc(ga_hub *h, uint32_t n_shards, uint32_t max_slots, uint64_t window_duration_us)
{
    h->max_slots = max_slots;
    h->n_shards = n_shards;
    h->free_slots.resize(n_shards);
    h->next_slot.assign(n_shards, 0);
    h->table.reset(new Relay[static_cast<size_t>(n_shards) * max_slots]);
    h->queues.reset(new ShardQueue[n_shards]);
    for (uint32_t sh = 0; sh < h->ctxs.size(); sh++) {
        uint32_t max_batch = 0;
        ga::ctx_limits(h->ctxs[sh], &max_batch, nullptr);
        uint64_t cap = 1024;
        while (cap < 4ull * max_batch) cap <<= 1;           // four full batches may wait for a flush
        h->queues[sh].cap = cap;
        h->queues[sh].q.reset(new QEntry[cap]);
    }
    if (window_duration_us) h->window_duration_us = window_duration_us;
    return GA_OK;
}

extern "C" int ga_hub_create(ga_ctx *ctx, uint32_t max_streams, uint64_t window_duration_us, ga_hub **out)
{
    GA_GUARD_BEGIN
    // ctx may be NULL: the registry and the window bookkeeping then work on their own (that is how
    // the CPU-only tests run them); everything that needs samples reports GA_ERR_STATE
    if (!out || max_streams == 0) { set_error("ga_hub_create: bad argument"); return GA_ERR_ARG; }
    std::unique_ptr<ga_hub> h(new ga_hub);
    if (ctx) h->ctxs.push_back(ctx);
    hub_alloc(h.get(), 1, max_streams, window_duration_us);
    *out = h.release();
    return GA_OK;
    GA_GUARD_END
}

// The hub over every GPU of the box: one context per GPU held by `cluster`.
extern "C" int ga_hub_create_cluster(ga_cluster *cluster, uint32_t max_streams_per_shard, uint64_t window_duration_us,
                                     ga_hub **out)
{
    GA_GUARD_BEGIN
    if (!out || !cluster || max_streams_per_shard == 0) { set_error("ga_hub_create_cluster: bad argument"); return GA_ERR_ARG; }
    const uint32_t n = ga_cluster_size(cluster);
    std::unique_ptr<ga_hub> h(new ga_hub);
    for (uint32_t i = 0; i < n; i++) h->ctxs.push_back(ga_cluster_context(cluster, i));
    h->cluster = cluster;
    hub_alloc(h.get(), n, max_streams_per_shard, window_duration_us);
    *out = h.release();
    return GA_OK;
    GA_GUARD_END
}

extern "C" void ga_hub_destroy(ga_hub *h) { delete h; }

// A relay device opened its gRPC stream (StreamAudio :926).  It is known to the hub but not yet
// in activeStreams — the reference only adds it there when it starts or joins a window.
static int connect_relay(ga_hub *h, const char *relay_id, uint32_t group, uint32_t shard, uint32_t *slot, uint64_t *handle)
{
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) {
        uint32_t s;
        auto &fs = h->free_slots[shard];
        if (!fs.empty()) { s = fs.back(); fs.pop_back(); }
        else if (h->next_slot[shard] < h->max_slots) s = h->next_slot[shard]++;
        else { set_error("ga_hub_connect: all %u stream slots of shard %u are in use", h->max_slots, shard); return GA_ERR_OVERFLOW; }
        const uint32_t idx = shard * h->max_slots + s;
        r = &h->table[idx];
        r->id = relay_id; r->slot = s; r->shard = shard;
        r->status.store(GA_RELAY_CONNECTED); r->signal_strength = 0.0; r->cancel_closed = false;
        r->in_active_streams.store(false); r->window = 0; r->alive = true; r->dropped_chunks = 0;
        r->gen.store(r->gen.load() + 1, std::memory_order_release);       // even -> odd: in use
        h->by_id.emplace(relay_id, idx);
    }
    r->group = group;                 // a known relay keeps its shard and slot: its audio lives there
    r->first_chunk.store(true);
    if (slot) *slot = r->slot;
    if (handle) *handle = handle_of(h, r);
    return GA_OK;
}

extern "C" int ga_hub_connect(ga_hub *h, const char *relay_id, uint32_t group, uint32_t *slot)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_connect: bad argument"); return GA_ERR_ARG; }
    // a collision group lives on shard group % nGPU, so its windows normally stay on one GPU
    return connect_relay(h, relay_id, group, group % h->n_shards, slot, nullptr);
    GA_GUARD_END
}

// A relay whose audio has to live on a given GPU (it connected through another front end, or the
// group's shard is full): its group's windows then span GPUs.
extern "C" int ga_hub_connect_on(ga_hub *h, const char *relay_id, uint32_t group, uint32_t shard, uint32_t *slot)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_connect_on: bad argument"); return GA_ERR_ARG; }
    if (shard >= h->n_shards) { set_error("ga_hub_connect_on: shard %u of %u", shard, h->n_shards); return GA_ERR_ARG; }
    return connect_relay(h, relay_id, group, shard, slot, nullptr);
    GA_GUARD_END
}

extern "C" int ga_hub_relay_handle(ga_hub *h, const char *relay_id, uint64_t *handle)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id || !handle) { set_error("ga_hub_relay_handle: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_relay_handle: unknown relay %s", relay_id); return GA_ERR_STATE; }
    *handle = handle_of(h, r);
    return GA_OK;
    GA_GUARD_END
}

extern "C" int ga_hub_relay_shard(ga_hub *h, const char *relay_id, uint32_t *shard)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id || !shard) { set_error("ga_hub_relay_shard: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_relay_shard: unknown relay %s", relay_id); return GA_ERR_STATE; }
    *shard = r->shard;
    return GA_OK;
    GA_GUARD_END
}

// startArbitrationWindow :408-446.  Replaces the group's window pointer, adds the relay as the
// first contender (fresh RelayStream: status Contending, open cancel channel).
static uint64_t start_window(ga_hub *h, Relay *r, uint64_t now_us)
{
    Window w;
    w.id = h->next_window++;
    w.group = r->group;
    w.start_us = now_us;                        // :415
    w.duration_us = h->window_duration_us;      // :416
    w.relays.push_back(r->id);                  // :432
    r->status = GA_RELAY_CONTENDING;            // :427
    r->cancel_closed = false;                   // :428
    r->signal_strength = 0.0;
    r->in_active_streams = true;                // :435
    r->window = w.id;
    h->windows[w.id] = w;
    h->group_window[r->group] = w.id;           // :436 as.arbitrationWindow = window
    return w.id;
}

// joinArbitrationWindow :449-491
static bool join_window(ga_hub *h, Relay *r, uint64_t now_us)
{
    auto gw = h->group_window.find(r->group);
    if (gw == h->group_window.end()) return false;            // :455-457 window == nil
    Window &w = h->windows[gw->second];
    if (!w.is_active) return false;                           // :462-464
    const uint64_t elapsed = now_us > w.start_us ? now_us - w.start_us : 0;   // time.Since; a clock behind the window's start counts as 0
    if (elapsed > w.duration_us) return false;                // :467-470 elapsed > WindowDuration
    bool present = false;
    for (auto &id : w.relays) present |= (id == r->id);
    if (!present) w.relays.push_back(r->id);                  // :481 (map insert: re-joining replaces the entry)
    r->status = GA_RELAY_CONTENDING;                          // :477
    r->cancel_closed = false;
    r->in_active_streams = true;                              // :482
    r->window = w.id;
    return true;
}

extern "C" int ga_hub_start_window(ga_hub *h, const char *relay_id, uint64_t now_us, uint64_t *window_id)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_start_window: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_start_window: unknown relay %s", relay_id); return GA_ERR_STATE; }
    const uint64_t id = start_window(h, r, now_us);
    if (window_id) *window_id = id;
    return GA_OK;
    GA_GUARD_END
}

extern "C" int ga_hub_join_window(ga_hub *h, const char *relay_id, uint64_t now_us, int *joined)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id || !joined) { set_error("ga_hub_join_window: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_join_window: unknown relay %s", relay_id); return GA_ERR_STATE; }
    *joined = join_window(h, r, now_us) ? 1 : 0;
    return GA_OK;
    GA_GUARD_END
}

// The chunk's bytes go into the pinned ring and the message onto the GPU's queue.  Only the ring
// reservation and the sequence number are taken under the queue's mutex; the copy is not.
static int enqueue_chunk(ga_hub *h, Relay *r, const uint8_t *audio, uint32_t nbytes, uint32_t flags)
{
    ShardQueue &q = h->queues[r->shard];
    void *dst = nullptr;
    uint64_t off = 0, seq = 0;
    {
        std::lock_guard<SpinLock> lk(q.claim);
        if (q.tail - q.head.load(std::memory_order_acquire) >= q.cap) {
            set_error("ga_hub_on_chunk: %llu chunks are waiting for a flush on GPU %u", (unsigned long long)q.cap, r->shard);
            return GA_ERR_RING_FULL;
        }
        int rc = ga_ring_reserve(h->ctxs[r->shard], nbytes, &dst, &off);
        if (rc) return rc;
        seq = q.tail++;
    }
    if (nbytes) std::memcpy(dst, audio, nbytes);
    QEntry &e = q.q[seq & (q.cap - 1)];
    e.v.m = ga_msg{};
    e.v.m.ring_offset = off; e.v.m.nbytes = nbytes; e.v.m.stream_id = r->slot; e.v.m.flags = flags;
    e.v.gen = r->gen.load(std::memory_order_relaxed);
    e.ready.store(seq + 1, std::memory_order_release);
    return GA_OK;
}

// One received AudioChunk: the body of the StreamAudio loop, :939-1041.  keep_stream = 0 is the
// reference's `return nil` (the handler ends and the relay's stream is closed).
static int on_chunk(ga_hub *h, Relay *r, const uint8_t *audio, uint32_t nbytes, int is_wake_word, int is_end_of_speech,
                    uint64_t now_us, int *keep_stream)
{
    if (keep_stream) *keep_stream = 1;
    bool ingest = false;
    if (is_wake_word) {                                           // :962
        ingest = true;                                            // :964-965 convert + append to wakeWordBuffer
        std::lock_guard<std::mutex> lk(h->mu);                    // the window registry: wake-word chunks only
        if (h->group_window.find(r->group) == h->group_window.end()) {
            start_window(h, r, now_us);                           // :972-980
        } else if (!join_window(h, r, now_us)) {                  // :983-995 too late
            if (keep_stream) *keep_stream = 0;
            return GA_OK;
        }
    } else {
        ingest = r->in_active_streams.load(std::memory_order_acquire);   // :1000-1004: only relays in activeStreams keep speech
    }
    if (ingest) {
        const bool first = r->first_chunk.exchange(false);
        int rc = enqueue_chunk(h, r, audio, nbytes,
                               (is_wake_word ? GA_MSG_WAKE : 0u) | (is_end_of_speech ? GA_MSG_EOS : 0u) | (first ? GA_MSG_FIRST : 0u));
        if (rc) { if (first) r->first_chunk.store(true); return rc; }
    }
    if (is_end_of_speech) {                                       // :1008-1041
        const int st = r->status.load(std::memory_order_acquire);
        const bool active = r->in_active_streams.load(std::memory_order_acquire) &&
                            (st == GA_RELAY_WINNER || st == GA_RELAY_CONNECTED || st == GA_RELAY_CONTENDING);
        if (!active && keep_stream) *keep_stream = 0;             // :1018-1021
    }
    return GA_OK;
}

extern "C" int ga_hub_on_chunk(ga_hub *h, const char *relay_id, const uint8_t *audio, uint32_t nbytes,
                               int is_wake_word, int is_end_of_speech, uint64_t now_us, int *keep_stream)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id || (nbytes && !audio)) { set_error("ga_hub_on_chunk: bad argument"); return GA_ERR_ARG; }
    Relay *r;
    {
        std::lock_guard<std::mutex> lk(h->mu);      // name lookup only; ga_hub_on_chunk_h skips it
        r = find_relay(h, relay_id);
    }
    if (!r) { set_error("ga_hub_on_chunk: relay %s is not connected", relay_id); return GA_ERR_STATE; }
    if (h->ctxs.empty()) { set_error("ga_hub_on_chunk: this hub has no GPU context"); return GA_ERR_STATE; }
    return on_chunk(h, r, audio, nbytes, is_wake_word, is_end_of_speech, now_us, keep_stream);
    GA_GUARD_END
}

extern "C" int ga_hub_on_chunk_h(ga_hub *h, uint64_t handle, const uint8_t *audio, uint32_t nbytes,
                                 int is_wake_word, int is_end_of_speech, uint64_t now_us, int *keep_stream)
{
    GA_GUARD_BEGIN
    if (!h || (nbytes && !audio)) { set_error("ga_hub_on_chunk_h: bad argument"); return GA_ERR_ARG; }
    Relay *r = relay_of_handle(h, handle);
    if (!r) { set_error("ga_hub_on_chunk_h: stale or unknown relay handle"); return GA_ERR_STATE; }
    if (h->ctxs.empty()) { set_error("ga_hub_on_chunk_h: this hub has no GPU context"); return GA_ERR_STATE; }
    return on_chunk(h, r, audio, nbytes, is_wake_word, is_end_of_speech, now_us, keep_stream);
    GA_GUARD_END
}

// The batch tick: everything pushed since the last flush goes through ga_submit, one batch per GPU
// at a time, all GPUs working at the same time.  A batch is cut where a relay restarts an utterance
// after chunks of its previous one, and at the context's max_batch_msgs.  A chunk the context
// refuses for its stream's sake (the relay's sample arena is full — the reference appends without
// bound, :965 / :1002) takes that relay's chunks out of the batch and the rest goes ahead; the call
// then returns GA_ERR_OVERFLOW after everything else has been ingested.
extern "C" int ga_hub_flush(ga_hub *h, uint32_t *n_msgs)
{
    GA_GUARD_BEGIN
    if (!h) { set_error("ga_hub_flush: NULL hub"); return GA_ERR_ARG; }
    if (h->ctxs.empty()) { set_error("ga_hub_flush: this hub has no GPU context"); return GA_ERR_STATE; }
    std::lock_guard<std::mutex> flk(h->flush_mu);
    struct Work {
        std::vector<ga_msg> batch;
        size_t lo = 0;                // first message not yet submitted
        uint32_t max_batch = 0;
        size_t limit = 0;             // current sub-batch size limit (halved when the context refuses a batch as a whole)
        uint64_t ticket = 0;
        uint64_t end_off = 0;         // ring offset behind the last chunk taken
        bool inflight = false, done = false, failed = false;
    };
    const uint32_t n_sh = h->n_shards;
    std::vector<Work> work(n_sh);
    uint32_t total = 0, dropped = 0, kept_back = 0;
    for (uint32_t sh = 0; sh < n_sh; sh++) {
        ShardQueue &q = h->queues[sh];
        Work &w = work[sh];
        ga::ctx_limits(h->ctxs[sh], &w.max_batch, nullptr);
        w.limit = w.max_batch ? w.max_batch : 1;
        std::vector<QMsg> taken;
        taken.swap(q.carry);
        uint64_t head = q.head.load(std::memory_order_relaxed);
        for (;; head++) {               // the published prefix: complete in arrival order and in ring offsets
            QEntry &e = q.q[head & (q.cap - 1)];
            if (e.ready.load(std::memory_order_acquire) != head + 1) break;
            taken.push_back(e.v);
        }
        q.head.store(head, std::memory_order_release);
        for (const QMsg &t : taken) {
            const uint64_t end = t.m.ring_offset + ((static_cast<uint64_t>(t.m.nbytes) + 15u) & ~15ull);
            if (end > q.taken_end) q.taken_end = end;
            // chunks of a relay that was cleaned up before the flush go with it: the slot may have a new owner
            if (h->table[sh * h->max_slots + t.m.stream_id].gen.load(std::memory_order_acquire) != t.gen) continue;
            w.batch.push_back(t.m);
        }
        w.end_off = q.taken_end;
        total += static_cast<uint32_t>(w.batch.size());
        w.done = w.batch.empty();
    }
    // end of the sub-batch that starts at w.lo
    auto next_cut = [](const Work &w) {
        size_t hi = std::min(w.batch.size(), w.lo + w.limit);
        bool any_first = false;
        for (size_t k = w.lo + 1; k < hi && !any_first; k++) any_first = (w.batch[k].flags & GA_MSG_FIRST) != 0;
        if (any_first) {
            std::unordered_set<uint32_t> seen;
            for (size_t k = w.lo; k < hi; k++) {
                if ((w.batch[k].flags & GA_MSG_FIRST) && seen.count(w.batch[k].stream_id)) { hi = k; break; }
                seen.insert(w.batch[k].stream_id);
            }
        }
        return hi;
    };
    int first_err = GA_OK;
    for (bool busy = true; busy;) {
        busy = false;
        for (uint32_t sh = 0; sh < n_sh; sh++) {            // round k: the next sub-batch of every GPU
            Work &w = work[sh];
            if (w.done) continue;
            for (;;) {
                const size_t hi = next_cut(w);
                uint32_t bad = 0xffffffffu;
                const int rc = ga::submit_impl(h->ctxs[sh], w.batch.data() + w.lo, static_cast<uint32_t>(hi - w.lo), &w.ticket, &bad);
                if (rc == GA_OK) { w.inflight = true; w.lo = hi; w.limit = w.max_batch ? w.max_batch : 1; break; }
                if (rc == GA_ERR_OVERFLOW && bad == 0xffffffffu && hi - w.lo > 1) {
                    w.limit = (hi - w.lo) / 2;                  // the batch as a whole is too big (frame records): halve it
                    continue;
                }
                if ((rc == GA_ERR_OVERFLOW || rc == GA_ERR_ARG) && bad != 0xffffffffu) {
                    // one relay's chunk cannot be taken: drop that relay's chunks, keep everybody else's
                    const uint32_t slot = w.batch[w.lo + bad].stream_id;
                    const size_t before = w.batch.size();
                    w.batch.erase(std::remove_if(w.batch.begin() + w.lo, w.batch.end(),
                                                 [slot](const ga_msg &m) { return m.stream_id == slot; }), w.batch.end());
                    const uint32_t k = static_cast<uint32_t>(before - w.batch.size());
                    dropped += k;
                    {
                        std::lock_guard<std::mutex> lk(h->mu);
                        Relay &r = h->table[sh * h->max_slots + slot];
                        r.dropped_chunks += k;
                        r.in_active_streams.store(false);       // no more speech is kept for it (as for a relay outside activeStreams)
                    }
                    if (!first_err) first_err = GA_ERR_OVERFLOW;
                    if (w.lo == w.batch.size()) { w.done = !w.inflight; break; }
                    continue;                                   // the same sub-batch without that relay
                }
                // not about one stream (CUDA failure, pipeline full …): keep the rest for the next flush, in order
                ShardQueue &q = h->queues[sh];
                for (size_t k = w.lo; k < w.batch.size(); k++)
                    q.carry.push_back(QMsg{w.batch[k], h->table[sh * h->max_slots + w.batch[k].stream_id].gen.load()});
                kept_back += static_cast<uint32_t>(w.batch.size() - w.lo);
                w.failed = true;
                w.done = true;
                if (!first_err || first_err == GA_ERR_OVERFLOW) first_err = rc;     // a failing GPU does not hold the others back
                break;
            }
        }
        for (uint32_t sh = 0; sh < n_sh; sh++) {
            Work &w = work[sh];
            if (!w.inflight) continue;
            w.inflight = false;
            const int rc = ga_wait_view(h->ctxs[sh], w.ticket, nullptr, nullptr, nullptr, nullptr);
            if (rc) { w.done = true; w.failed = true; if (!first_err || first_err == GA_ERR_OVERFLOW) first_err = rc; continue; }
            if (w.lo == w.batch.size()) w.done = true; else if (!w.done) busy = true;
        }
    }
    // chunks that were dropped never reach a ticket, so no wait releases their ring space: once a
    // GPU's whole prefix has been dealt with, everything up to its end is free
    for (uint32_t sh = 0; sh < n_sh; sh++)
        if (!work[sh].failed && work[sh].end_off) ga::ring_release_to(h->ctxs[sh], work[sh].end_off);
    if (n_msgs) *n_msgs = total - dropped - kept_back;
    if (first_err == GA_ERR_OVERFLOW && dropped)
        set_error("ga_hub_flush: dropped %u chunk(s) of relays whose sample arena is full; everything else was ingested", dropped);
    return first_err;
    GA_GUARD_END
}

// performArbitration :505-587 for a set of windows at once.
static int arbitrate_windows(ga_hub *h, const std::vector<uint64_t> &ids)
{
    if (ids.empty()) return GA_OK;
    if (h->ctxs.empty()) { set_error("arbitration needs a GPU context"); return GA_ERR_STATE; }
    std::vector<uint32_t> seg(1, 0), sid, shard, win_idx;
    std::vector<uint8_t> st_in, st_out;
    std::vector<double> scores;
    std::vector<Window *> ws;
    for (uint64_t id : ids) {
        auto it = h->windows.find(id);
        if (it == h->windows.end() || !it->second.is_active) continue;    // :508-511 already arbitrated
        Window &w = it->second;
        for (auto &rid : w.relays) {
            Relay *rp = find_relay(h, rid.c_str());
            if (!rp) {
                // the relay was cleaned up (cleanupRelay :916-923) while its window was open: its stream
                // slot is gone, so it takes part as a member that is not contending
                sid.push_back(0); shard.push_back(0); st_in.push_back(GA_RELAY_CANCELLED); scores.push_back(0.0);
                continue;
            }
            Relay &r = *rp;
            sid.push_back(r.slot);
            shard.push_back(r.shard);
            st_in.push_back(static_cast<uint8_t>(r.status.load()));
            scores.push_back(r.signal_strength);                          // untouched for skipped relays (:522-524)
        }
        seg.push_back(static_cast<uint32_t>(sid.size()));
        ws.push_back(&w);
    }
    if (ws.empty()) return GA_OK;
    win_idx.resize(ws.size());
    st_out.resize(sid.size());
    // one GPU: the context's own call; several: one fused launch per GPU, windows may span them
    int rc = h->cluster
        ? ga_cluster_arbitrate(h->cluster, seg.data(), static_cast<uint32_t>(ws.size()), shard.data(), sid.data(),
                               st_in.data(), win_idx.data(), scores.data(), st_out.data())
        : ga_arbitrate(h->ctxs[0], seg.data(), static_cast<uint32_t>(ws.size()), sid.data(), st_in.data(),
                       win_idx.data(), scores.data(), st_out.data());
    if (rc) return rc;
    for (size_t k = 0; k < ws.size(); k++) {
        Window &w = *ws[k];
        for (uint32_t m = seg[k]; m < seg[k + 1]; m++) {
            Relay *rp = find_relay(h, w.relays[m - seg[k]].c_str());
            if (!rp) continue;                                            // cleaned up meanwhile: nothing to update
            Relay &r = *rp;
            r.signal_strength = scores[m];                                // :528
            r.status = st_out[m];                                         // :553-571
            if (st_out[m] == GA_RELAY_CANCELLED) r.cancel_closed = true;  // :561 close(relay.CancelChannel)
        }
        w.winner = win_idx[k] == 0xffffffffu ? std::string() : w.relays[win_idx[k] - seg[k]];   // :550
        w.is_active = false;                                              // :551
        auto gw = h->group_window.find(w.group);
        if (gw != h->group_window.end() && gw->second == w.id) h->group_window.erase(gw);        // :584-586
    }
    return GA_OK;
}

extern "C" int ga_hub_perform_arbitration(ga_hub *h, uint64_t window_id)
{
    GA_GUARD_BEGIN
    if (!h) { set_error("ga_hub_perform_arbitration: NULL hub"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    if (h->windows.find(window_id) == h->windows.end()) { set_error("ga_hub_perform_arbitration: unknown window"); return GA_ERR_STATE; }
    return arbitrate_windows(h, {window_id});
    GA_GUARD_END
}

// runArbitrationTimer :494-502 for all groups: arbitrate every active window whose duration elapsed.
extern "C" int ga_hub_poll(ga_hub *h, uint64_t now_us, uint32_t *n_closed)
{
    GA_GUARD_BEGIN
    if (!h) { set_error("ga_hub_poll: NULL hub"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    std::vector<uint64_t> due;
    for (auto &kv : h->windows)
        if (kv.second.is_active && now_us >= kv.second.start_us && now_us - kv.second.start_us >= kv.second.duration_us) due.push_back(kv.first);
    if (n_closed) *n_closed = static_cast<uint32_t>(due.size());
    return arbitrate_windows(h, due);
    GA_GUARD_END
}

// The same, also telling WHICH windows were closed, so that the caller can fetch all their winners
// at once (ga_hub_fetch_winner_wavs).  At most cap ids are written; *n_closed is the full count.
extern "C" int ga_hub_poll_windows(ga_hub *h, uint64_t now_us, uint64_t *window_ids, uint32_t cap, uint32_t *n_closed)
{
    GA_GUARD_BEGIN
    if (!h || (cap && !window_ids)) { set_error("ga_hub_poll_windows: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    std::vector<uint64_t> due;
    for (auto &kv : h->windows)
        if (kv.second.is_active && now_us >= kv.second.start_us && now_us - kv.second.start_us >= kv.second.duration_us) due.push_back(kv.first);
    if (n_closed) *n_closed = static_cast<uint32_t>(due.size());
    for (size_t k = 0; k < due.size() && k < cap; k++) window_ids[k] = due[k];
    return arbitrate_windows(h, due);
    GA_GUARD_END
}

// processWinningRelayAudio (:590-609) for the winners of many windows in one go: per GPU one
// ga_fetch_wavs — every winner's WakeWordSignal ++ SpeechAudio as the WAV file the STT request
// carries (stt_client.go:365-398) plus, if audio_hashes is given, the voice-event fingerprint
// "sha256:<hex>" of those samples (events/voice_event.go:84-88; GA_AUDIO_HASH_LEN bytes each).
// wav_ptrs[k] / wav_bytes[k] describe window k's file inside C-owned pinned memory (valid until the
// next call); a window without a winner gets NULL / 0 / "".
extern "C" int ga_hub_fetch_winner_wavs(ga_hub *h, const uint64_t *window_ids, uint32_t n, uint32_t sample_rate,
                                        const uint8_t **wav_ptrs, uint64_t *wav_bytes, char *audio_hashes)
{
    GA_GUARD_BEGIN
    if (!h || (n && (!window_ids || !wav_ptrs || !wav_bytes))) { set_error("ga_hub_fetch_winner_wavs: bad argument"); return GA_ERR_ARG; }
    if (h->ctxs.empty()) { set_error("ga_hub_fetch_winner_wavs: this hub has no GPU context"); return GA_ERR_STATE; }
    std::vector<std::vector<uint32_t>> slots(h->n_shards), which(h->n_shards);
    {
        std::lock_guard<std::mutex> lk(h->mu);
        for (uint32_t k = 0; k < n; k++) {
            wav_ptrs[k] = nullptr; wav_bytes[k] = 0;
            if (audio_hashes) audio_hashes[static_cast<size_t>(k) * GA_AUDIO_HASH_LEN] = 0;
            auto it = h->windows.find(window_ids[k]);
            if (it == h->windows.end()) { set_error("ga_hub_fetch_winner_wavs: unknown window %llu", (unsigned long long)window_ids[k]); return GA_ERR_STATE; }
            if (it->second.winner.empty()) continue;
            Relay *r = find_relay(h, it->second.winner.c_str());
            if (!r) continue;                                     // the winner was cleaned up meanwhile
            slots[r->shard].push_back(r->slot);
            which[r->shard].push_back(k);
        }
    }
    for (uint32_t sh = 0; sh < h->n_shards; sh++) {
        const uint32_t m = static_cast<uint32_t>(slots[sh].size());
        if (!m) continue;
        std::vector<uint64_t> off(m + 1);
        std::vector<uint8_t> dg(audio_hashes ? 32ull * m : 0);
        const uint8_t *base = nullptr;
        int rc = ga_fetch_wavs(h->ctxs[sh], slots[sh].data(), m, sample_rate, &base, off.data(), audio_hashes ? dg.data() : nullptr);
        if (rc) return rc;
        for (uint32_t j = 0; j < m; j++) {
            const uint32_t k = which[sh][j];
            wav_ptrs[k] = base + off[j];
            const uint8_t *hdr = base + off[j];
            wav_bytes[k] = 8ull + (static_cast<uint64_t>(hdr[4]) | (static_cast<uint64_t>(hdr[5]) << 8) |
                                   (static_cast<uint64_t>(hdr[6]) << 16) | (static_cast<uint64_t>(hdr[7]) << 24));   // RIFF size + 8
            if (audio_hashes) ga_format_audio_hash(dg.data() + 32ull * j, audio_hashes + static_cast<size_t>(k) * GA_AUDIO_HASH_LEN);
        }
    }
    return GA_OK;
    GA_GUARD_END
}

extern "C" int ga_hub_window_info(ga_hub *h, uint64_t window_id, int *is_active, char *winner_id, size_t winner_cap,
                                  uint32_t *n_relays)
{
    GA_GUARD_BEGIN
    if (!h) { set_error("ga_hub_window_info: NULL hub"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    auto it = h->windows.find(window_id);
    if (it == h->windows.end()) { set_error("ga_hub_window_info: unknown window"); return GA_ERR_STATE; }
    if (is_active) *is_active = it->second.is_active ? 1 : 0;
    if (n_relays) *n_relays = static_cast<uint32_t>(it->second.relays.size());
    if (winner_id && winner_cap) {
        std::strncpy(winner_id, it->second.winner.c_str(), winner_cap - 1);
        winner_id[winner_cap - 1] = 0;
    }
    return GA_OK;
    GA_GUARD_END
}

// as.arbitrationWindow for a group: 0 = nil.
extern "C" int ga_hub_active_window(ga_hub *h, uint32_t group, uint64_t *window_id)
{
    GA_GUARD_BEGIN
    if (!h || !window_id) { set_error("ga_hub_active_window: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    auto gw = h->group_window.find(group);
    *window_id = gw == h->group_window.end() ? 0 : gw->second;
    return GA_OK;
    GA_GUARD_END
}

extern "C" int ga_hub_relay_info(ga_hub *h, const char *relay_id, int *status, double *signal_strength,
                                 int *cancel_closed, uint32_t *slot)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_relay_info: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_relay_info: unknown relay %s", relay_id); return GA_ERR_STATE; }
    if (status) *status = r->status.load();
    if (signal_strength) *signal_strength = r->signal_strength;
    if (cancel_closed) *cancel_closed = r->cancel_closed ? 1 : 0;
    if (slot) *slot = r->slot;
    return GA_OK;
    GA_GUARD_END
}

// isRelayActive :880-913: 1 / 0; a relay that is not in activeStreams is inactive (:884-888).
extern "C" int ga_hub_is_relay_active(ga_hub *h, const char *relay_id)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r || !r->in_active_streams.load()) return 0;
    const int st = r->status.load();
    return (st == GA_RELAY_WINNER || st == GA_RELAY_CONNECTED || st == GA_RELAY_CONTENDING) ? 1 : 0;
    GA_GUARD_END
}

// What the reference's tests do by writing the map directly (collision_detection_test.go:195-216,
// :361-369): put a relay into activeStreams with a given status.
extern "C" int ga_hub_set_relay_status(ga_hub *h, const char *relay_id, int status)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id || status < GA_RELAY_CONNECTED || status > GA_RELAY_CANCELLED) { set_error("ga_hub_set_relay_status: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_set_relay_status: unknown relay %s", relay_id); return GA_ERR_STATE; }
    r->status = status;
    r->in_active_streams = true;
    return GA_OK;
    GA_GUARD_END
}

// Request completion as the reference's tests spell it (collision_detection_test.go:361-369):
// drop the group's window if this relay won it, and return the relay to Connected.
extern "C" int ga_hub_complete_request(ga_hub *h, const char *relay_id)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_complete_request: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_complete_request: unknown relay %s", relay_id); return GA_ERR_STATE; }
    auto gw = h->group_window.find(r->group);
    if (gw != h->group_window.end() && h->windows[gw->second].winner == r->id) h->group_window.erase(gw);
    r->status = GA_RELAY_CONNECTED;
    return GA_OK;
    GA_GUARD_END
}

// cleanupRelay :916-923: remove from activeStreams; the stream slot and its GPU buffers are freed.
extern "C" int ga_hub_cleanup_relay(ga_hub *h, const char *relay_id)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_cleanup_relay: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    auto it = h->by_id.find(relay_id);
    if (it == h->by_id.end()) return GA_OK;                  // nonexistent relay: not an error (:916-923, test :316)
    Relay &r = h->table[it->second];
    const uint32_t slot = r.slot, shard = r.shard;
    // chunks of the relay that were not flushed yet go with it — the slot may get a new owner before
    // the next tick: the flush drops queued chunks whose generation is not the slot's current one
    // (their ring space is released with the rest of the flush's prefix)
    r.gen.store(r.gen.load() + 1, std::memory_order_release);           // odd -> even: free; stale handles are refused
    r.alive = false;
    r.in_active_streams.store(false);
    r.id.clear();
    h->by_id.erase(it);
    h->free_slots[shard].push_back(slot);
    return h->ctxs.empty() ? GA_OK : ga_stream_reset(h->ctxs[shard], slot);
    GA_GUARD_END
}

extern "C" int ga_hub_active_streams(ga_hub *h, uint32_t *n)
{
    GA_GUARD_BEGIN
    if (!h || !n) { set_error("ga_hub_active_streams: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    uint32_t c = 0;
    for (auto &kv : h->by_id) c += h->table[kv.second].in_active_streams.load() ? 1u : 0u;
    *n = c;
    return GA_OK;
    GA_GUARD_END
}

// ---------------------------------------------------------------- f32 -> WAV for the winner
// float32ToWAV, internal/llm/stt_client.go:365-398: 44-byte RIFF header (format 3 IEEE float,
// mono, 32 bits) followed by the little-endian f32 samples.  The samples already sit in HBM in
// exactly that byte order, so the payload is two device->host copies (wake-word region, speech
// region) straight into the caller's buffer behind the header — no per-sample allocation (:391-396).
extern "C" int ga_fetch_wav(ga_ctx *ctx, uint32_t stream_id, uint32_t sample_rate, uint8_t *out, size_t cap, size_t *n_bytes)
{
    GA_GUARD_BEGIN
    if (!ctx || !n_bytes) { set_error("ga_fetch_wav: bad argument"); return GA_ERR_ARG; }
    uint64_t n = 0, wake = 0;
    int rc = ga_stream_info(ctx, stream_id, &n, &wake);
    if (rc) return rc;
    const uint64_t data_size = n * 4, total = 44 + data_size;
    *n_bytes = static_cast<size_t>(total);
    if (!out) return GA_OK;                                       // size query
    if (cap < total || data_size > 0xffffffd0ull) { set_error("ga_fetch_wav: buffer of %zu bytes, need %llu", cap, (unsigned long long)total); return GA_ERR_ARG; }
    auto le32 = [](uint8_t *p, uint32_t v) { p[0] = v & 255; p[1] = (v >> 8) & 255; p[2] = (v >> 16) & 255; p[3] = v >> 24; };
    std::memcpy(out, "RIFF", 4);  le32(out + 4, static_cast<uint32_t>(36 + data_size));   // :375-376
    std::memcpy(out + 8, "WAVEfmt ", 8);  le32(out + 16, 16);                             // :377-379
    out[20] = 3; out[21] = 0; out[22] = 1; out[23] = 0;                                   // :380-381
    le32(out + 24, sample_rate);  le32(out + 28, sample_rate * 4);                        // :382-384
    out[32] = 4; out[33] = 0; out[34] = 32; out[35] = 0;                                  // :385-386
    std::memcpy(out + 36, "data", 4);  le32(out + 40, static_cast<uint32_t>(data_size));  // :387-388
    if (n == 0) return GA_OK;
    return ga_fetch_samples(ctx, stream_id, 0, n, reinterpret_cast<float *>(out + 44));
    GA_GUARD_END
}

// The winner's audio as the STT request body, from whichever GPU holds it.
extern "C" int ga_hub_fetch_wav(ga_hub *h, const char *relay_id, uint32_t sample_rate, uint8_t *out, size_t cap, size_t *n_bytes)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_fetch_wav: bad argument"); return GA_ERR_ARG; }
    ga_ctx *ctx = nullptr;
    uint32_t slot = 0;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        Relay *r = find_relay(h, relay_id);
        if (!r) { set_error("ga_hub_fetch_wav: unknown relay %s", relay_id); return GA_ERR_STATE; }
        if (h->ctxs.empty()) { set_error("ga_hub_fetch_wav: this hub has no GPU context"); return GA_ERR_STATE; }
        ctx = h->ctxs[r->shard]; slot = r->slot;
    }
    return ga_fetch_wav(ctx, slot, sample_rate, out, cap, n_bytes);
    GA_GUARD_END
}

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
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "common.cuh"

namespace {

struct Relay {                       // RelayStream :44-53
    std::string id;
    uint32_t group = 0;
    uint32_t shard = 0;              // which GPU's context holds the relay's audio
    uint32_t slot = 0;               // ga_ctx stream slot there: WakeWordSignal / SpeechAudio live on the GPU
    int status = GA_RELAY_CONNECTED; // :56-63
    double signal_strength = 0.0;    // :50
    bool cancel_closed = false;      // close(relay.CancelChannel) :561
    bool in_active_streams = false;  // as.activeStreams[relayID] exists (:91)
    bool first_chunk = true;
    uint64_t window = 0;             // window it last joined (0 = none)
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

}  // namespace

struct ga_hub {
    // no context: bookkeeping only; one: single GPU; several (with their cluster): one hub process
    // driving a multi-GPU box, relays placed by collision group, windows may span GPUs
    std::vector<ga_ctx *> ctxs;
    ga_cluster *cluster = nullptr;
    uint64_t window_duration_us = 300000;                 // arbitrationWindowDuration, 300 ms (:243)
    std::unordered_map<std::string, Relay> relays;
    std::map<uint64_t, Window> windows;                   // every window ever opened and not yet dropped
    std::unordered_map<uint32_t, uint64_t> group_window;  // as.arbitrationWindow per group (absent = nil)
    std::vector<std::vector<uint32_t>> free_slots;        // per shard
    std::vector<uint32_t> next_slot;
    uint32_t max_slots = 0;                               // stream slots per shard
    uint64_t next_window = 1;
    std::vector<std::vector<ga_msg>> pending;             // per shard: chunks pushed since the last flush
    std::mutex mu;
};

using namespace ga;

static Relay *find_relay(ga_hub *h, const char *id)
{
    auto it = h->relays.find(id);
    return it == h->relays.end() ? nullptr : &it->second;
}

extern "C" int ga_hub_create(ga_ctx *ctx, uint32_t max_streams, uint64_t window_duration_us, ga_hub **out)
{
    GA_GUARD_BEGIN
    // ctx may be NULL: the registry and the window bookkeeping then work on their own (that is how
    // the CPU-only tests run them); everything that needs samples reports GA_ERR_STATE
    if (!out || max_streams == 0) { set_error("ga_hub_create: bad argument"); return GA_ERR_ARG; }
    ga_hub *h = new (std::nothrow) ga_hub;
    if (!h) { set_error("ga_hub_create: out of memory"); return GA_ERR_NOMEM; }
    if (ctx) h->ctxs.push_back(ctx);
    h->max_slots = max_streams;
    h->free_slots.resize(1); h->next_slot.assign(1, 0); h->pending.resize(1);
    if (window_duration_us) h->window_duration_us = window_duration_us;
    *out = h;
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
    ga_hub *h = new (std::nothrow) ga_hub;
    if (!h) { set_error("ga_hub_create_cluster: out of memory"); return GA_ERR_NOMEM; }
    for (uint32_t i = 0; i < n; i++) h->ctxs.push_back(ga_cluster_context(cluster, i));
    h->cluster = cluster;
    h->max_slots = max_streams_per_shard;
    h->free_slots.resize(n); h->next_slot.assign(n, 0); h->pending.resize(n);
    if (window_duration_us) h->window_duration_us = window_duration_us;
    *out = h;
    return GA_OK;
    GA_GUARD_END
}

extern "C" void ga_hub_destroy(ga_hub *h) { delete h; }

// A relay device opened its gRPC stream (StreamAudio :926).  It is known to the hub but not yet
// in activeStreams — the reference only adds it there when it starts or joins a window.
static int connect_relay(ga_hub *h, const char *relay_id, uint32_t group, uint32_t shard, uint32_t *slot)
{
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) {
        uint32_t s;
        auto &fs = h->free_slots[shard];
        if (!fs.empty()) { s = fs.back(); fs.pop_back(); }
        else if (h->next_slot[shard] < h->max_slots) s = h->next_slot[shard]++;
        else { set_error("ga_hub_connect: all %u stream slots of shard %u are in use", h->max_slots, shard); return GA_ERR_OVERFLOW; }
        Relay nr;
        nr.id = relay_id; nr.slot = s; nr.shard = shard;
        r = &h->relays.emplace(relay_id, nr).first->second;
    }
    r->group = group;                 // a known relay keeps its shard and slot: its audio lives there
    r->first_chunk = true;
    if (slot) *slot = r->slot;
    return GA_OK;
}

extern "C" int ga_hub_connect(ga_hub *h, const char *relay_id, uint32_t group, uint32_t *slot)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_connect: bad argument"); return GA_ERR_ARG; }
    // a collision group lives on shard group % nGPU, so its windows normally stay on one GPU
    const uint32_t n = static_cast<uint32_t>(h->free_slots.size());
    return connect_relay(h, relay_id, group, group % n, slot);
    GA_GUARD_END
}

// A relay whose audio has to live on a given GPU (it connected through another front end, or the
// group's shard is full): its group's windows then span GPUs.
extern "C" int ga_hub_connect_on(ga_hub *h, const char *relay_id, uint32_t group, uint32_t shard, uint32_t *slot)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id) { set_error("ga_hub_connect_on: bad argument"); return GA_ERR_ARG; }
    if (shard >= h->free_slots.size()) { set_error("ga_hub_connect_on: shard %u of %zu", shard, h->free_slots.size()); return GA_ERR_ARG; }
    return connect_relay(h, relay_id, group, shard, slot);
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
    if (now_us - w.start_us > w.duration_us) return false;    // :467-470 elapsed > WindowDuration
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

// One received AudioChunk: the body of the StreamAudio loop, :939-1041.  keep_stream = 0 is the
// reference's `return nil` (the handler ends and the relay's stream is closed).
extern "C" int ga_hub_on_chunk(ga_hub *h, const char *relay_id, const uint8_t *audio, uint32_t nbytes,
                               int is_wake_word, int is_end_of_speech, uint64_t now_us, int *keep_stream)
{
    GA_GUARD_BEGIN
    if (!h || !relay_id || (nbytes && !audio)) { set_error("ga_hub_on_chunk: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    Relay *r = find_relay(h, relay_id);
    if (!r) { set_error("ga_hub_on_chunk: relay %s is not connected", relay_id); return GA_ERR_STATE; }
    if (h->ctxs.empty()) { set_error("ga_hub_on_chunk: this hub has no GPU context"); return GA_ERR_STATE; }
    if (keep_stream) *keep_stream = 1;
    bool ingest = false;
    if (is_wake_word) {                                           // :962
        ingest = true;                                            // :964-965 convert + append to wakeWordBuffer
        if (h->group_window.find(r->group) == h->group_window.end()) {
            start_window(h, r, now_us);                           // :972-980
        } else if (!join_window(h, r, now_us)) {                  // :983-995 too late
            if (keep_stream) *keep_stream = 0;
            return GA_OK;
        }
    } else {
        ingest = r->in_active_streams;                            // :1000-1004: only relays in activeStreams keep speech
    }
    if (ingest) {
        void *dst = nullptr;
        uint64_t off = 0;
        int rc = ga_ring_reserve(h->ctxs[r->shard], nbytes, &dst, &off);
        if (rc) return rc;
        if (nbytes) std::memcpy(dst, audio, nbytes);
        ga_msg m{};
        m.ring_offset = off; m.nbytes = nbytes; m.stream_id = r->slot;
        m.flags = (is_wake_word ? GA_MSG_WAKE : 0u) | (is_end_of_speech ? GA_MSG_EOS : 0u) |
                  (r->first_chunk ? GA_MSG_FIRST : 0u);
        r->first_chunk = false;
        h->pending[r->shard].push_back(m);
    }
    if (is_end_of_speech) {                                       // :1008-1041
        const bool active = r->in_active_streams && (r->status == GA_RELAY_WINNER || r->status == GA_RELAY_CONNECTED ||
                                                     r->status == GA_RELAY_CONTENDING);
        if (!active && keep_stream) *keep_stream = 0;             // :1018-1021
    }
    return GA_OK;
    GA_GUARD_END
}

// The batch tick: everything pushed since the last flush goes through ONE ga_submit per GPU, all
// GPUs working at the same time.
extern "C" int ga_hub_flush(ga_hub *h, uint32_t *n_msgs)
{
    GA_GUARD_BEGIN
    if (!h) { set_error("ga_hub_flush: NULL hub"); return GA_ERR_ARG; }
    if (h->ctxs.empty()) { set_error("ga_hub_flush: this hub has no GPU context"); return GA_ERR_STATE; }
    struct Work {
        std::vector<ga_msg> batch;
        std::vector<size_t> cuts;     // ends of the sub-batches
        size_t next = 0, lo = 0;      // next sub-batch to submit, first message not yet submitted
        uint64_t ticket = 0;
        bool inflight = false, done = false;
    };
    const uint32_t n_sh = static_cast<uint32_t>(h->ctxs.size());
    std::vector<Work> work(n_sh);
    {
        std::lock_guard<std::mutex> lk(h->mu);   // only to take the queues: chunks keep arriving while the GPUs work
        for (uint32_t sh = 0; sh < n_sh; sh++) work[sh].batch.swap(h->pending[sh]);
    }
    uint32_t total = 0;
    for (auto &w : work) {
        total += static_cast<uint32_t>(w.batch.size());
        w.done = w.batch.empty();
        // A relay that finished an utterance and started the next one within one tick has chunks
        // of both in the queue; ga_submit wants the restart in a later batch than the old chunks,
        // so the queue is cut there (rare: normally one submit per GPU).
        bool any_first = false;
        for (size_t k = 1; k < w.batch.size() && !any_first; k++) any_first = (w.batch[k].flags & GA_MSG_FIRST) != 0;
        if (any_first) {
            std::unordered_set<uint32_t> seen;
            for (size_t k = 0; k < w.batch.size(); k++) {
                if ((w.batch[k].flags & GA_MSG_FIRST) && seen.count(w.batch[k].stream_id)) { w.cuts.push_back(k); seen.clear(); }
                seen.insert(w.batch[k].stream_id);
            }
        }
        w.cuts.push_back(w.batch.size());
    }
    int first_err = GA_OK;
    uint32_t lost = 0;
    for (bool busy = true; busy;) {
        busy = false;
        for (uint32_t sh = 0; sh < n_sh; sh++) {            // round k: the k-th sub-batch of every GPU
            Work &w = work[sh];
            if (w.done) continue;
            const size_t hi = w.cuts[w.next];
            const int rc = ga_submit(h->ctxs[sh], w.batch.data() + w.lo, static_cast<uint32_t>(hi - w.lo), &w.ticket);
            if (rc) {                                       // nothing of [lo, end) was queued: it stays pending, in order
                std::lock_guard<std::mutex> lk(h->mu);
                h->pending[sh].insert(h->pending[sh].begin(), w.batch.begin() + w.lo, w.batch.end());
                lost += static_cast<uint32_t>(w.batch.size() - w.lo);
                w.done = true;
                if (!first_err) first_err = rc;             // a failing GPU does not hold the others back
                continue;
            }
            w.inflight = true;
            w.lo = hi;
            w.next++;
        }
        for (uint32_t sh = 0; sh < n_sh; sh++) {
            Work &w = work[sh];
            if (!w.inflight) continue;
            w.inflight = false;
            const int rc = ga_wait_view(h->ctxs[sh], w.ticket, nullptr, nullptr, nullptr, nullptr);
            if (rc) { w.done = true; if (!first_err) first_err = rc; continue; }
            if (w.next == w.cuts.size()) w.done = true; else busy = true;
        }
    }
    if (n_msgs) *n_msgs = total - lost;
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
            auto rit = h->relays.find(rid);
            if (rit == h->relays.end()) {
                // the relay was cleaned up (cleanupRelay :916-923) while its window was open: its stream
                // slot is gone, so it takes part as a member that is not contending
                sid.push_back(0); shard.push_back(0); st_in.push_back(GA_RELAY_CANCELLED); scores.push_back(0.0);
                continue;
            }
            Relay &r = rit->second;
            sid.push_back(r.slot);
            shard.push_back(r.shard);
            st_in.push_back(static_cast<uint8_t>(r.status));
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
            auto rit = h->relays.find(w.relays[m - seg[k]]);
            if (rit == h->relays.end()) continue;                         // cleaned up meanwhile: nothing to update
            Relay &r = rit->second;
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
        if (kv.second.is_active && now_us - kv.second.start_us >= kv.second.duration_us) due.push_back(kv.first);
    if (n_closed) *n_closed = static_cast<uint32_t>(due.size());
    return arbitrate_windows(h, due);
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
    if (status) *status = r->status;
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
    if (!r || !r->in_active_streams) return 0;
    return (r->status == GA_RELAY_WINNER || r->status == GA_RELAY_CONNECTED || r->status == GA_RELAY_CONTENDING) ? 1 : 0;
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
    auto it = h->relays.find(relay_id);
    if (it == h->relays.end()) return GA_OK;                 // nonexistent relay: not an error (:916-923, test :316)
    const uint32_t slot = it->second.slot, shard = it->second.shard;
    h->relays.erase(it);
    h->free_slots[shard].push_back(slot);
    // chunks of the relay that were not flushed yet go with it: the slot may get a new owner before
    // the next tick (their ring space is released with the next ticket that is waited)
    auto &q = h->pending[shard];
    q.erase(std::remove_if(q.begin(), q.end(), [slot](const ga_msg &m) { return m.stream_id == slot; }), q.end());
    return h->ctxs.empty() ? GA_OK : ga_stream_reset(h->ctxs[shard], slot);
    GA_GUARD_END
}

extern "C" int ga_hub_active_streams(ga_hub *h, uint32_t *n)
{
    GA_GUARD_BEGIN
    if (!h || !n) { set_error("ga_hub_active_streams: bad argument"); return GA_ERR_ARG; }
    std::lock_guard<std::mutex> lk(h->mu);
    uint32_t c = 0;
    for (auto &kv : h->relays) c += kv.second.in_active_streams ? 1u : 0u;
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

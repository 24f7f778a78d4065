// hub_realtime.cpp — the drop-in path under load, through the C ABI only.
//
// T "relay" threads stand for the StreamAudio goroutines of grpc-go (internal/grpc/audio_service.go:926):
// each owns streams/T relays and pushes one 20 ms chunk per relay per tick through
// ga_hub_on_chunk_h (wake-word chunks first, then speech, end of speech on the last chunk, then the
// relay reconnects and the next utterance starts).  ONE batching thread runs the tick:
// ga_hub_flush (ring -> GPU -> result records) and ga_hub_poll_windows + ga_hub_fetch_winner_wavs
// for the windows that closed (runArbitrationTimer :494-502, processWinningRelayAudio :590-609).
//
//   --paced 1   real time: tick k happens at k * 20 ms of wall time.  Reports per-tick latency (from the
//               tick's scheduled time to its records being on the host) and window-close latency.
//   --paced 0   as fast as it goes: reports sustained chunks per second.
// Prints one JSON object.  Build: see tools/build_tools.sh (g++, links libgpuaudio.so).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../include/gpuaudio.h"

using clk = std::chrono::steady_clock;
static double now_s() { return std::chrono::duration<double>(clk::now().time_since_epoch()).count(); }

#define CHECK(call)                                                                          \
    do {                                                                                     \
        int _rc = (call);                                                                    \
        if (_rc != GA_OK) {                                                                  \
            fprintf(stderr, "%s failed (%d): %s\n", #call, _rc, ga_last_error());            \
            exit(2);                                                                         \
        }                                                                                    \
    } while (0)

static double pct(std::vector<double> v, double p)
{
    if (v.empty()) return 0.0;
    std::sort(v.begin(), v.end());
    size_t i = static_cast<size_t>(p * (v.size() - 1) + 0.5);
    return v[std::min(i, v.size() - 1)];
}

int main(int argc, char **argv)
{
    uint32_t streams = 8192, threads = 8, group = 8, utt_chunks = 250, wake_chunks = 15, device = 0;
    double seconds = 5.0;
    int paced = 1, fetch = 1;
    for (int i = 1; i + 1 < argc; i += 2) {
        std::string k = argv[i];
        const char *v = argv[i + 1];
        if (k == "--streams") streams = atoi(v);
        else if (k == "--threads") threads = atoi(v);
        else if (k == "--seconds") seconds = atof(v);
        else if (k == "--paced") paced = atoi(v);
        else if (k == "--fetch") fetch = atoi(v);
        else if (k == "--device") device = atoi(v);
        else if (k == "--utt-chunks") utt_chunks = atoi(v);
        else { fprintf(stderr, "unknown option %s\n", k.c_str()); return 2; }
    }
    const uint32_t chunk_bytes = 640, ms = 320;
    const uint64_t tick_us = 20000;
    if (threads > streams) threads = streams;

    ga_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.device = device; cfg.max_streams = streams; cfg.max_batch_msgs = streams * 4; cfg.max_inflight = 2;
    cfg.ring_bytes = static_cast<uint64_t>(streams) * chunk_bytes * 16;
    cfg.arena_samples = static_cast<uint64_t>(utt_chunks) * ms; cfg.wake_samples_max = static_cast<uint64_t>(wake_chunks) * ms;
    cfg.mode = GA_DO_ALL; cfg.input_format = GA_FMT_MONO16K;
    ga_ctx *ctx = nullptr;
    CHECK(ga_init(&cfg, &ctx));
    ga_hub *hub = nullptr;
    CHECK(ga_hub_create(ctx, streams, 300000, &hub));

    // per relay one 20 ms chunk, louder for relay 0 of its group so that the winner is known
    std::vector<int16_t> pcm(static_cast<size_t>(streams) * ms);
    for (uint32_t s = 0; s < streams; s++)
        for (uint32_t k = 0; k < ms; k++) {
            const int amp = 12000 - 1200 * static_cast<int>(s % group);
            const int tri = static_cast<int>((k * 37u + s * 11u) % 200u) - 100;
            pcm[static_cast<size_t>(s) * ms + k] = static_cast<int16_t>(amp * tri / 100);
        }
    std::vector<std::string> names(streams);
    std::vector<uint64_t> handles(streams);
    for (uint32_t s = 0; s < streams; s++) {
        names[s] = "relay-" + std::to_string(s);
        CHECK(ga_hub_connect(hub, names[s].c_str(), s / group, nullptr));
        CHECK(ga_hub_relay_handle(hub, names[s].c_str(), &handles[s]));
    }

    std::atomic<bool> stop{false};
    std::vector<std::atomic<uint64_t>> done_tick(threads);      // ticks each relay thread has completed
    for (auto &d : done_tick) d.store(0);
    std::atomic<uint64_t> pushed{0}, backpressure{0};
    const double t0 = now_s();

    auto relay_thread = [&](uint32_t t) {
        // groups start their utterances spread over the first 50 ticks, so windows do not all close at once
        for (uint64_t k = 0; !stop.load(std::memory_order_relaxed); k++) {
            if (paced) {
                const double due = t0 + static_cast<double>(k * tick_us) * 1e-6;
                double w = due - now_s();
                if (w > 0) std::this_thread::sleep_for(std::chrono::duration<double>(w));
            }
            uint64_t n = 0;
            // a thread owns whole collision groups, so the relays of one window share one clock
            for (uint32_t s = 0; s < streams; s++) {
                if ((s / group) % threads != t) continue;
                const uint32_t phase = (s / group) % 50u;
                if (k < phase) continue;
                const uint64_t m = (k - phase) % utt_chunks;
                if (m == 0 && k != phase) {                        // next utterance: a new StreamAudio call
                    CHECK(ga_hub_connect(hub, names[s].c_str(), s / group, nullptr));       // same slot, same handle
                }
                int keep = 1;
                for (;;) {
                    int rc = ga_hub_on_chunk_h(hub, handles[s], reinterpret_cast<const uint8_t *>(&pcm[static_cast<size_t>(s) * ms]),
                                               chunk_bytes, m < wake_chunks, m + 1 == utt_chunks, k * tick_us, &keep);
                    if (rc == GA_OK) break;
                    if (rc != GA_ERR_RING_FULL) { fprintf(stderr, "on_chunk_h: %s\n", ga_last_error()); exit(2); }
                    backpressure.fetch_add(1, std::memory_order_relaxed);
                    if (stop.load(std::memory_order_relaxed)) return;
                    std::this_thread::yield();
                }
                n++;
            }
            pushed.fetch_add(n, std::memory_order_relaxed);
            done_tick[t].store(k + 1, std::memory_order_release);
        }
    };

    std::vector<double> tick_lat, flush_ms, close_ms, fetch_ms;
    double ramp_t = 0.0, last_flush_t = 0.0;
    uint64_t ramp_ingested = 0;
    uint64_t ingested = 0, windows_closed = 0, wav_bytes_total = 0, ticks_done = 0, winners_wrong = 0;
    std::vector<uint64_t> closed(streams / group + 1);
    std::vector<const uint8_t *> wav_ptrs(closed.size());
    std::vector<uint64_t> wav_sizes(closed.size());
    std::vector<char> hashes(closed.size() * GA_AUDIO_HASH_LEN);

    auto batcher = [&]() {
        uint64_t k = 0;                                            // next tick to flush
        while (!stop.load(std::memory_order_relaxed)) {
            uint64_t ready = ~0ull;
            for (auto &d : done_tick) ready = std::min(ready, d.load(std::memory_order_acquire));
            if (paced && ready <= k) { std::this_thread::sleep_for(std::chrono::microseconds(50)); continue; }
            const double a = now_s();
            uint32_t n = 0;
            int rc = ga_hub_flush(hub, &n);
            if (rc != GA_OK && rc != GA_ERR_OVERFLOW) { fprintf(stderr, "flush: %s\n", ga_last_error()); exit(2); }
            const double b = now_s();
            ingested += n;
            if (n) flush_ms.push_back((b - a) * 1e3);
            if (paced) {
                for (; k < ready; k++) tick_lat.push_back((b - (t0 + static_cast<double>(k * tick_us) * 1e-6)) * 1e3);
            } else {
                k = ready == ~0ull ? k : ready;
            }
            ticks_done = k;
            if (k >= 50 && ramp_t == 0.0) { ramp_t = b; ramp_ingested = ingested; }   // every group has started by tick 50
            last_flush_t = b;
            // the window timer: virtual time of the slowest relay thread
            uint32_t nc = 0;
            const double c = now_s();
            CHECK(ga_hub_poll_windows(hub, (k ? k - 1 : 0) * tick_us, closed.data(), static_cast<uint32_t>(closed.size()), &nc));
            if (nc) {
                const double d = now_s();
                close_ms.push_back((d - c) * 1e3);
                windows_closed += nc;
                if (fetch) {
                    CHECK(ga_hub_fetch_winner_wavs(hub, closed.data(), nc, 16000, wav_ptrs.data(), wav_sizes.data(), hashes.data()));
                    fetch_ms.push_back((now_s() - d) * 1e3);
                    for (uint32_t i = 0; i < nc; i++) wav_bytes_total += wav_sizes[i];
                }
                for (uint32_t i = 0; i < nc; i++) {                // relay 0 of the group must have won
                    char win[64]; int active = 0; uint32_t nr = 0;
                    CHECK(ga_hub_window_info(hub, closed[i], &active, win, sizeof win, &nr));
                    const uint32_t s = static_cast<uint32_t>(atoi(win + 6));
                    if (s % group != 0) winners_wrong++;
                }
            }
            if (!paced && n == 0) std::this_thread::yield();
        }
    };

    std::vector<std::thread> th;
    for (uint32_t t = 0; t < threads; t++) th.emplace_back(relay_thread, t);
    std::thread bt(batcher);
    std::this_thread::sleep_for(std::chrono::duration<double>(seconds));
    stop.store(true);
    for (auto &t : th) t.join();
    bt.join();
    const double wall = now_s() - t0;
    const double steady = (ramp_t > 0.0 && last_flush_t > ramp_t) ? (ingested - ramp_ingested) / (last_flush_t - ramp_t) : 0.0;
    uint32_t n = 0;
    ga_hub_flush(hub, &n);
    ingested += n;

    printf("{\"tool\": \"hub_realtime\", \"paced\": %d, \"streams\": %u, \"relay_threads\": %u, \"seconds\": %.3f, "
           "\"chunks_pushed\": %llu, \"chunks_ingested\": %llu, \"chunks_per_s\": %.0f, \"steady_chunks_per_s\": %.0f, \"audio_s_per_s\": %.0f, "
           "\"ticks\": %llu, \"backpressure_retries\": %llu, "
           "\"tick_latency_ms\": {\"p50\": %.3f, \"p99\": %.3f, \"max\": %.3f, \"n\": %zu}, "
           "\"flush_ms\": {\"p50\": %.3f, \"p99\": %.3f, \"n\": %zu}, "
           "\"window_close_ms\": {\"p50\": %.3f, \"p99\": %.3f, \"n\": %zu, \"windows\": %llu, \"wrong_winners\": %llu}, "
           "\"winner_fetch_ms\": {\"p50\": %.3f, \"p99\": %.3f, \"n\": %zu, \"wav_bytes\": %llu}}\n",
           paced, streams, threads, wall, (unsigned long long)pushed.load(), (unsigned long long)ingested,
           ingested / wall, steady, ingested * 0.02 / wall, (unsigned long long)ticks_done, (unsigned long long)backpressure.load(),
           pct(tick_lat, 0.5), pct(tick_lat, 0.99), pct(tick_lat, 1.0), tick_lat.size(),
           pct(flush_ms, 0.5), pct(flush_ms, 0.99), flush_ms.size(),
           pct(close_ms, 0.5), pct(close_ms, 0.99), close_ms.size(), (unsigned long long)windows_closed,
           (unsigned long long)winners_wrong,
           pct(fetch_ms, 0.5), pct(fetch_ms, 0.99), fetch_ms.size(), (unsigned long long)wav_bytes_total);
    ga_hub_destroy(hub);
    ga_shutdown(ctx);
    return winners_wrong ? 1 : 0;
}

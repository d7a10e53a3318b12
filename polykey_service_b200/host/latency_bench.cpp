// latency_bench.cpp — per-request routing latency through the batching front-end.
//
//   latency_bench <callers> <seconds> <max_batch> <max_wait_us> <payload_bytes> [n_backends [idle_flush_us]]
//   latency_bench open <requests_per_s> <seconds> <max_batch> <max_wait_us> <payload_bytes> [generators [idle_flush_us]]
//
// K caller threads play the reference's goroutine-per-RPC model (cmd/polykey/main.go:74-77):
// each calls BatchingService::ExecuteTool in a closed loop — enqueue, block until the batch it
// landed in has been routed on the GPU, receive the decision, build the response with the inner
// MockService — and records the wall time of every call.  Prints one JSON object: throughput,
// latency percentiles (enqueue -> decision + response), and how the batches formed.  The same
// loop against the bare MockService (no routing stage) is reported as the reference-faithful
// latency floor.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <thread>

#include "batching_service.hpp"

using namespace polykey;
using Clock = std::chrono::steady_clock;

static double pct(std::vector<float>& v, double p)
{
    if (v.empty()) return 0;
    const size_t k = std::min(v.size() - 1, (size_t)(p * v.size()));
    std::nth_element(v.begin(), v.begin() + k, v.end());
    return v[k];
}

struct RunResult { std::vector<float> lat_us; double seconds = 0; uint64_t errors = 0; };

static RunResult run(Service& svc, unsigned callers, double seconds, size_t payload_bytes)
{
    static const char* names[] = {"example_tool", "struct_tool", "file_tool", "other_tool"};
    std::vector<std::vector<float>> lat(callers);
    std::atomic<bool> stop{false};
    std::atomic<uint64_t> errors{0};
    std::vector<std::thread> th;
    const auto t0 = Clock::now();
    for (unsigned t = 0; t < callers; ++t) th.emplace_back([&, t] {
        std::string text;
        while (text.size() + 8 < payload_bytes) text += "word" + std::to_string((t * 31 + text.size()) % 997) + " ";
        Struct params{{"prompt", Value(text)}};
        Context ctx; ctx.now_unix = 1700000000;
        lat[t].reserve(1 << 16);
        for (unsigned i = 0; !stop.load(std::memory_order_relaxed); ++i) {
            const std::string name = names[(t + i) & 3];
            const auto a = Clock::now();
            Result r = svc.ExecuteTool(ctx, name, &params, nullptr, nullptr);
            const auto b = Clock::now();
            if (r.error || !r.resp) { ++errors; continue; }
            lat[t].push_back(std::chrono::duration<float, std::micro>(b - a).count());
        }
    });
    std::this_thread::sleep_for(std::chrono::duration<double>(seconds));
    stop.store(true);
    for (auto& x : th) x.join();
    RunResult out;
    out.seconds = std::chrono::duration<double>(Clock::now() - t0).count();
    out.errors = errors.load();
    for (auto& v : lat) out.lat_us.insert(out.lat_us.end(), v.begin(), v.end());
    return out;
}

// Open loop (SURVEY §8d iii): requests ARRIVE on a Poisson schedule of the given rate whether or
// not earlier ones have been answered — what a server sees from many independent clients.  Each
// generator thread owns a ring of AsyncRequests: it enqueues every arrival whose scheduled time has
// come (BatchingService::EnqueueAsync, non-blocking) and polls the outstanding ones in order;
// latency = completion observed - SCHEDULED arrival, so queueing behind a slow stage counts.
// A stand-in device for measuring the FRONT-END alone (PK_NULL_BACKEND_US=<µs>): every batch
// "completes" that many µs after its submit, records zeroed.  Not a routing implementation.
class NullBackend final : public RouteBackend {
public:
    explicit NullBackend(double us) : lat_(std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double, std::micro>(us))) {}
    uint32_t slots() const override { return 3; }
    void* alloc_host(size_t bytes) override { bufs_.emplace_back(new uint8_t[bytes]); return bufs_.back().get(); }
    int submit(uint32_t slot, const uint8_t*, const uint64_t*, const uint32_t*, const uint8_t*, const uint32_t*,
               uint32_t n, pk_decision* out) override
    {
        std::memset(out, 0, (size_t)n * sizeof(pk_decision));
        due_[slot] = Clock::now() + lat_;
        return 0;
    }
    int wait(uint32_t slot) override { while (Clock::now() < due_[slot]) {} return 0; }
    std::string error_string(int) override { return "null"; }
private:
    Clock::duration lat_;
    Clock::time_point due_[3];
    std::vector<std::unique_ptr<uint8_t[]>> bufs_;
};

struct OpenResult { std::vector<float> lat_us; double seconds = 0; uint64_t offered = 0, dropped = 0, errors = 0; };

static OpenResult run_open(BatchingService& svc, double rate, double seconds, size_t payload_bytes, unsigned gens)
{
    static const char* names[] = {"example_tool", "struct_tool", "file_tool", "other_tool"};
    std::vector<OpenResult> part(gens);
    std::vector<std::thread> th;
    const auto t0 = Clock::now();
    for (unsigned g = 0; g < gens; ++g) th.emplace_back([&, g] {
        constexpr size_t kRing = 1 << 15;
        std::vector<AsyncRequest> ring(kRing);
        std::vector<Clock::time_point> due(kRing);
        std::string text;
        while (text.size() + 8 < payload_bytes) text += "word" + std::to_string((g * 31 + text.size()) % 997) + " ";
        const std::string json = to_json(Struct{{"prompt", Value(text)}});
        uint64_t x = 0x9E3779B97F4A7C15ull * (g + 1);                 // xorshift: exponential gaps
        auto gap = [&] {
            x ^= x << 13; x ^= x >> 7; x ^= x << 17;
            const double u = ((x >> 11) + 1) * (1.0 / 9007199254740993.0);
            return std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double>(-std::log(u) * gens / rate));
        };
        OpenResult& out = part[g];
        out.lat_us.reserve(1 << 20);
        size_t head = 0, tail = 0;                                    // [tail, head) outstanding
        const auto end = Clock::now() + std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double>(seconds));
        auto next = Clock::now() + gap();
        for (;;) {
            const auto now = Clock::now();
            if (now >= end && tail == head) break;
            // every arrival that is due — in bursts of at most 32, so that a generator that has fallen
            // behind still picks records up between bursts (which is what frees arenas) instead of
            // enqueueing until the front-end pushes back
            for (int burst = 0; burst < 32 && now < end && next <= now; ++burst) {
                if (head - tail == kRing) { ++out.offered; ++out.dropped; next += gap(); continue; }   // generator out of tickets
                AsyncRequest& r = ring[head % kRing];
                due[head % kRing] = next;
                // the producer's copy of its request into the pinned arena happens in here
                const BatchingService::Enqueue e = svc.EnqueueAsync(r, names[head & 3], json);
                if (e == BatchingService::Enqueue::kBusy) break;      // back-pressure: the arrival stays due (its latency keeps
                                                                      // counting from its scheduled time); go pick records up
                ++out.offered;
                if (e == BatchingService::Enqueue::kAccepted) ++head; else ++out.errors;
                next += gap();
            }
            while (tail < head && ring[tail % kRing].done()) {        // completions, in arrival order
                if (ring[tail % kRing].rc() != 0) ++out.errors;
                else out.lat_us.push_back(std::chrono::duration<float, std::micro>(Clock::now() - due[tail % kRing]).count());
                ++tail;
            }
        }
    });
    for (auto& x : th) x.join();
    OpenResult out;
    out.seconds = std::chrono::duration<double>(Clock::now() - t0).count();
    for (auto& p : part) {
        out.lat_us.insert(out.lat_us.end(), p.lat_us.begin(), p.lat_us.end());
        out.offered += p.offered; out.dropped += p.dropped; out.errors += p.errors;
    }
    return out;
}

static int main_open(int argc, char** argv)
{
    if (argc < 7) { std::fprintf(stderr, "usage: %s open requests_per_s seconds max_batch max_wait_us payload_bytes [generators [idle_flush_us]]\n", argv[0]); return 2; }
    const double rate = std::atof(argv[2]), seconds = std::atof(argv[3]);
    BatchingConfig cfg;
    cfg.max_batch = (uint32_t)std::atoi(argv[4]);
    cfg.max_wait_us = (uint32_t)std::atoi(argv[5]);
    const size_t payload_bytes = (size_t)std::atol(argv[6]);
    const unsigned gens = argc > 7 ? (unsigned)std::atoi(argv[7]) : 1;
    if (argc > 8) cfg.idle_flush_us = (uint32_t)std::atoi(argv[8]);
    if (argc > 9) cfg.eager_flush = std::atoi(argv[9]) != 0;
    cfg.max_payload_bytes = (uint64_t)cfg.max_batch * ((payload_bytes + 64 + 15) & ~size_t(15));
    pk_table table;
    pk_default_table(&table);
    // PK_BACKENDS=N: one backend (flusher, table replica) per GPU 0..N-1; requests shard by id
    // PK_SHARDS_PER_GPU=K: K backends (handle, flusher thread, arenas) on every GPU — K open batches per GPU
    // take reservations in parallel, which is what the front-end scales with once one batch's
    // reservation word is the bottleneck
    const unsigned per_gpu = std::getenv("PK_SHARDS_PER_GPU") ? (unsigned)std::max(1, std::atoi(std::getenv("PK_SHARDS_PER_GPU"))) : 1u;
    const unsigned nb = per_gpu * (std::getenv("PK_BACKENDS") ? (unsigned)std::max(1, std::atoi(std::getenv("PK_BACKENDS"))) : 1u);
    std::vector<std::unique_ptr<RouteBackend>> devs;
    std::vector<RouteBackend*> bes;
    for (unsigned g = 0; g < nb; ++g) {
        if (const char* e = std::getenv("PK_NULL_BACKEND_US")) {
            devs.emplace_back(new NullBackend(std::atof(e)));
            if (g == 0) std::fprintf(stderr, "NULL backend: front-end measurement only, nothing is routed\n");
        } else {
            auto* cb = new CudaBackend((int)(g / per_gpu), table, cfg.max_batch, cfg.max_payload_bytes, 3);
            devs.emplace_back(cb);
            if (!cb->ok()) { std::fprintf(stderr, "no CUDA backend on device %u (status %d): there is no CPU fallback\n", g, cb->status()); return 1; }
        }
        bes.push_back(devs.back().get());
    }
    auto inner = NewMockService();
    BatchingService svc(inner.get(), bes, cfg);
    run_open(svc, std::min(rate, 2e5), 0.3, payload_bytes, gens);     // warm
    const BatchingStats before = svc.stats();
    OpenResult r = run_open(svc, rate, seconds, payload_bytes, gens);
    const BatchingStats st = svc.stats();
    svc.Shutdown();
    const double n = (double)r.lat_us.size();
    const uint64_t batches = st.batches - before.batches, reqs = st.requests - before.requests;
    std::printf("{\"mode\": \"open-loop\", \"shards_per_gpu\": %u, \"backends\": %u, \"offered_per_s\": %.0f, \"generators\": %u, \"max_batch\": %u, \"max_wait_us\": %u, \"idle_flush_us\": %u, \"eager\": %d, "
                "\"payload_bytes\": %zu, \"completed_per_s\": %.1f, \"dropped\": %llu, \"errors\": %llu, "
                "\"latency_us\": {\"p50\": %.1f, \"p90\": %.1f, \"p99\": %.1f, \"p999\": %.1f, \"max\": %.1f}, \"batches\": %llu, \"mean_batch\": %.1f}\n",
                per_gpu, nb, rate, gens, cfg.max_batch, cfg.max_wait_us, cfg.idle_flush_us, (int)cfg.eager_flush, payload_bytes, n / r.seconds,
                (unsigned long long)r.dropped, (unsigned long long)r.errors,
                pct(r.lat_us, 0.50), pct(r.lat_us, 0.90), pct(r.lat_us, 0.99), pct(r.lat_us, 0.999),
                r.lat_us.empty() ? 0.0 : *std::max_element(r.lat_us.begin(), r.lat_us.end()),
                (unsigned long long)batches, batches ? (double)reqs / batches : 0.0);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc > 1 && std::string(argv[1]) == "open") return main_open(argc, argv);
    if (argc < 6) { std::fprintf(stderr, "usage: %s callers seconds max_batch max_wait_us payload_bytes [n_backends]\n", argv[0]); return 2; }
    const unsigned callers = (unsigned)std::atoi(argv[1]);
    const double seconds = std::atof(argv[2]);
    BatchingConfig cfg;
    cfg.max_batch = (uint32_t)std::atoi(argv[3]);
    cfg.max_wait_us = (uint32_t)std::atoi(argv[4]);
    const size_t payload_bytes = (size_t)std::atol(argv[5]);
    const unsigned nb = argc > 6 ? (unsigned)std::atoi(argv[6]) : 1;
    if (argc > 7) cfg.idle_flush_us = (uint32_t)std::atoi(argv[7]);
    if (argc > 8) cfg.eager_flush = std::atoi(argv[8]) != 0;
    cfg.max_payload_bytes = (uint64_t)cfg.max_batch * ((payload_bytes + 64 + 15) & ~size_t(15));

    pk_table table;
    pk_default_table(&table);
    std::vector<std::unique_ptr<CudaBackend>> owned;
    std::vector<RouteBackend*> bes;
    for (unsigned g = 0; g < nb; ++g) {
        owned.emplace_back(new CudaBackend((int)g, table, cfg.max_batch, cfg.max_payload_bytes, 3));
        if (!owned.back()->ok()) { std::fprintf(stderr, "no CUDA backend on device %u (status %d): there is no CPU fallback\n", g, owned.back()->status()); return 1; }
        bes.push_back(owned.back().get());
    }
    auto inner = NewMockService();
    RunResult base = run(*inner, std::min(callers, 16u), std::min(seconds, 1.0), payload_bytes);
    BatchingService svc(inner.get(), bes, cfg);
    run(svc, callers, 0.3, payload_bytes);                      // warm
    RunResult r = run(svc, callers, seconds, payload_bytes);
    const BatchingStats st = svc.stats();
    svc.Shutdown();
    const double n = (double)r.lat_us.size();
    std::printf("{\"callers\": %u, \"backends\": %u, \"max_batch\": %u, \"max_wait_us\": %u, \"eager\": %d, \"payload_bytes\": %zu, "
                "\"requests\": %.0f, \"requests_per_s\": %.1f, \"latency_us\": {\"p50\": %.1f, \"p90\": %.1f, \"p99\": %.1f, \"max\": %.1f}, "
                "\"batches\": %llu, \"mean_batch\": %.1f, \"max_batch_seen\": %llu, \"flush_full\": %llu, \"flush_deadline\": %llu, \"flush_idle\": %llu, \"errors\": %llu, "
                "\"mock_only\": {\"requests_per_s\": %.1f, \"latency_us\": {\"p50\": %.2f, \"p99\": %.2f}}}\n",
                callers, nb, cfg.max_batch, cfg.max_wait_us, (int)cfg.eager_flush, payload_bytes, n, n / r.seconds,
                pct(r.lat_us, 0.50), pct(r.lat_us, 0.90), pct(r.lat_us, 0.99), r.lat_us.empty() ? 0.0 : *std::max_element(r.lat_us.begin(), r.lat_us.end()),
                (unsigned long long)st.batches, st.batches ? (double)st.requests / st.batches : 0.0, (unsigned long long)st.max_batch_seen,
                (unsigned long long)st.flush_full, (unsigned long long)st.flush_deadline, (unsigned long long)st.flush_idle, (unsigned long long)(st.errors + r.errors),
                base.lat_us.size() / base.seconds, pct(base.lat_us, 0.50), pct(base.lat_us, 0.99));
    return 0;
}

// latency_bench.cpp — per-request routing latency through the batching front-end.
//
//   latency_bench <callers> <seconds> <max_batch> <max_wait_us> <payload_bytes> [n_backends]
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
#include <cstdio>
#include <cstdlib>
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

int main(int argc, char** argv)
{
    if (argc < 6) { std::fprintf(stderr, "usage: %s callers seconds max_batch max_wait_us payload_bytes [n_backends]\n", argv[0]); return 2; }
    const unsigned callers = (unsigned)std::atoi(argv[1]);
    const double seconds = std::atof(argv[2]);
    BatchingConfig cfg;
    cfg.max_batch = (uint32_t)std::atoi(argv[3]);
    cfg.max_wait_us = (uint32_t)std::atoi(argv[4]);
    const size_t payload_bytes = (size_t)std::atol(argv[5]);
    const unsigned nb = argc > 6 ? (unsigned)std::atoi(argv[6]) : 1;
    if (argc > 7) cfg.idle_flush_us = (uint32_t)std::atoi(argv[7]);
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
    std::printf("{\"callers\": %u, \"backends\": %u, \"max_batch\": %u, \"max_wait_us\": %u, \"payload_bytes\": %zu, "
                "\"requests\": %.0f, \"requests_per_s\": %.1f, \"latency_us\": {\"p50\": %.1f, \"p90\": %.1f, \"p99\": %.1f, \"max\": %.1f}, "
                "\"batches\": %llu, \"mean_batch\": %.1f, \"max_batch_seen\": %llu, \"flush_full\": %llu, \"flush_deadline\": %llu, \"flush_idle\": %llu, \"errors\": %llu, "
                "\"mock_only\": {\"requests_per_s\": %.1f, \"latency_us\": {\"p50\": %.2f, \"p99\": %.2f}}}\n",
                callers, nb, cfg.max_batch, cfg.max_wait_us, payload_bytes, n, n / r.seconds,
                pct(r.lat_us, 0.50), pct(r.lat_us, 0.90), pct(r.lat_us, 0.99), r.lat_us.empty() ? 0.0 : *std::max_element(r.lat_us.begin(), r.lat_us.end()),
                (unsigned long long)st.batches, st.batches ? (double)st.requests / st.batches : 0.0, (unsigned long long)st.max_batch_seen,
                (unsigned long long)st.flush_full, (unsigned long long)st.flush_deadline, (unsigned long long)st.flush_idle, (unsigned long long)(st.errors + r.errors),
                base.lat_us.size() / base.seconds, pct(base.lat_us, 0.50), pct(base.lat_us, 0.99));
    return 0;
}

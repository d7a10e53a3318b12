// host_service_test.cpp — CPU tests of the C++ host mirror (polykey_service_b200/host/):
// MockService and Server against the literals of internal/service/mock.go:22-67 and
// internal/server/server.go:27-43, and BatchingService (the MPSC batcher at the
// service.Service seam) driven by many caller threads with a CPU stand-in backend that
// answers from the oracle.  Built and run by tests/test_host_service.py.  Test-only code.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <thread>

#include "../polykey_service_b200/host/batching_service.hpp"

extern "C" int pk_ref_route_batch(const pk_table*, const uint8_t*, const uint64_t*, const uint32_t*,
                                  const uint8_t*, const uint32_t*, uint32_t, pk_decision*);
extern "C" int pk_ref_default_table(pk_table*);

using namespace polykey;

static int g_fail = 0;
#define CHECK(c) do { if (!(c)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); ++g_fail; } } while (0)

// CPU stand-in for a routing device: same submit/wait contract, answers from the oracle.
class OracleBackend final : public RouteBackend {
public:
    explicit OracleBackend(uint32_t slots) : slots_(slots) { pk_ref_default_table(&table_); }
    uint32_t slots() const override { return slots_; }
    void* alloc_host(size_t bytes) override { bufs_.emplace_back(new uint8_t[bytes]); return bufs_.back().get(); }
    int submit(uint32_t, const uint8_t* payload, const uint64_t* pay_off, const uint32_t* pay_len,
               const uint8_t* names, const uint32_t* name_off, uint32_t n, pk_decision* out) override
    {
        ++batches;
        requests += n;
        if (fail_after >= 0 && batches > (uint64_t)fail_after) return PK_ERR_CUDA;
        return pk_ref_route_batch(&table_, payload, pay_off, pay_len, names, name_off, n, out);
    }
    int wait(uint32_t) override { return PK_OK; }
    std::string error_string(int s) override { return "oracle backend status " + std::to_string(s); }
    std::atomic<uint64_t> batches{0}, requests{0};
    int fail_after = -1;
    pk_table table_;
private:
    uint32_t slots_;
    std::vector<std::unique_ptr<uint8_t[]>> bufs_;
};

struct FailingService final : Service {
    Result ExecuteTool(Context&, const std::string&, const Struct*, const std::string*, const Metadata*) override
    { return Result{nullptr, std::string("boom")}; }
};

static pk_decision oracle_one(const pk_table& t, const std::string& name, const std::string& payload)
{
    std::vector<uint8_t> arena(payload.size() + 16, 0xEE);
    std::memcpy(arena.data(), payload.data(), payload.size());
    uint64_t off[2] = {0, (payload.size() + 15) & ~size_t(15)};
    uint32_t len[1] = {(uint32_t)payload.size()};
    uint32_t noff[2] = {0, (uint32_t)name.size()};
    pk_decision d{};
    int rc = pk_ref_route_batch(&t, arena.data(), off, len, (const uint8_t*)name.data(), noff, 1, &d);
    CHECK(rc == 0);
    return d;
}

static void test_mock_service()
{
    auto svc = NewMockService();
    Context ctx; ctx.now_unix = 1700000000;
    Result r = svc->ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr);      // mock.go:33-36
    CHECK(!r.error && r.resp && r.resp->status.code == 200 && r.resp->status.message == "Tool executed successfully");
    CHECK(std::get<std::string>(r.resp->output) == "Mock execution of example_tool at 2023-11-14T22:13:20Z");
    r = svc->ExecuteTool(ctx, "struct_tool", nullptr, nullptr, nullptr);               // mock.go:37-51
    const Struct& s = std::get<Struct>(r.resp->output);
    CHECK(to_json(s) == "{\"data\":{\"count\":42,\"processed\":true},\"result\":\"success\",\"timestamp\":1700000000}");
    r = svc->ExecuteTool(ctx, "file_tool", nullptr, nullptr, nullptr);                 // mock.go:52-59
    const File& f = std::get<File>(r.resp->output);
    CHECK(f.file_name == "example.txt" && f.mime_type == "text/plain" && f.content == "This is mock file content");
    r = svc->ExecuteTool(ctx, "Example_tool", nullptr, nullptr, nullptr);              // mock.go:60-64: not an error
    CHECK(!r.error && r.resp->status.code == 200 && std::get<std::string>(r.resp->output) == "Unknown tool: Example_tool");
    CHECK(!r.resp->decision.has_value());
}

static void test_server()
{
    auto svc = NewMockService();
    std::vector<std::string> log;
    Server srv(svc.get(), &log);
    Context ctx; ctx.now_unix = 1;
    // the dev client's request, cmd/dev_client/main.go:247-258
    ExecuteToolRequest req;
    req.tool_name = "example_tool";
    req.parameters = Struct{{"example_param", Value("value")}, {"timestamp", Value((int64_t)1700000000)}};
    req.secret_id = "secret-123";
    req.metadata = Metadata{{{"client_version", "1.0.0"}, {"request_source", "dev_client"}}};
    Result r = srv.ExecuteTool(ctx, req);
    CHECK(!r.error && r.resp && r.resp->status.code == 200);
    CHECK(log.size() == 1 && log[0] == "ExecuteTool called tool_name=example_tool has_parameters=true has_secret_id=true has_metadata=true");
    FailingService bad;
    Server srv2(&bad, &log);
    r = srv2.ExecuteTool(ctx, req);                                                    // server.go:37-40
    CHECK(r.error && *r.error == "boom" && !r.resp);
    CHECK(log.back() == "Service ExecuteTool failed error=boom");
}

static void test_logging_interceptor()
{
    // cmd/polykey/main.go:25-52
    std::vector<std::string> log;
    LoggingInterceptor li(&log);
    auto svc = NewMockService();
    Context ctx; ctx.now_unix = 1;
    Result ok = li.Intercept("/polykey.v2.PolykeyService/ExecuteTool",
                             [&] { return svc->ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr); });
    CHECK(!ok.error && ok.resp && log.size() == 2);
    CHECK(log[0] == "{\"level\":\"INFO\",\"msg\":\"gRPC call received\",\"method\":\"/polykey.v2.PolykeyService/ExecuteTool\"}");
    CHECK(log[1].rfind("{\"level\":\"INFO\",\"msg\":\"gRPC call finished\",\"method\":\"/polykey.v2.PolykeyService/ExecuteTool\",\"duration\":\"", 0) == 0);
    CHECK(log[1].find("\"code\":\"OK\"}") != std::string::npos);
    FailingService bad;
    Result err = li.Intercept("/polykey.v2.PolykeyService/ExecuteTool",
                              [&] { return bad.ExecuteTool(ctx, "x", nullptr, nullptr, nullptr); });
    CHECK(err.error && *err.error == "boom" && log.size() == 4);                       // error passes through (:50)
    CHECK(log[3].find("\"level\":\"ERROR\"") != std::string::npos && log[3].find("\"code\":\"Unknown\"") != std::string::npos);
    Result health = li.Intercept("/grpc.health.v1.Health/Check",
                                 [&] { return svc->ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr); });
    CHECK(!health.error && log.size() == 4);                                          // health checks are not logged (:29-31)
    CHECK(LoggingInterceptor::format_duration(17) == "17ns" && LoggingInterceptor::format_duration(340000) == "340µs");
    CHECK(LoggingInterceptor::format_duration(2250000) == "2.25ms" && LoggingInterceptor::format_duration(1500000000) == "1.5s");
}

static void test_batching(uint32_t n_gpus, uint32_t threads, uint32_t per_thread, uint32_t max_batch,
                          bool cuda = false, bool hold = false)
{
    auto inner = NewMockService();
    std::vector<std::unique_ptr<OracleBackend>> owned;
    std::vector<std::unique_ptr<CudaBackend>> cuda_owned;
    std::vector<RouteBackend*> bes;
    pk_table table; pk_ref_default_table(&table);
    for (uint32_t g = 0; g < n_gpus; ++g) {
        owned.emplace_back(new OracleBackend(2));                                    // also the table holder
        if (cuda) {
            cuda_owned.emplace_back(new CudaBackend(0, table, max_batch, 1 << 20, 2));   // handles share GPU 0
            CHECK(cuda_owned.back()->ok());
            bes.push_back(cuda_owned.back().get());
        } else {
            bes.push_back(owned.back().get());
        }
    }
    BatchingConfig cfg;
    cfg.max_batch = max_batch; cfg.max_wait_us = 300; cfg.max_payload_bytes = 1 << 20;
    std::vector<std::string> batch_log;
    cfg.batch_log = &batch_log;
    cfg.eager_flush = !hold;                     // hold: keep the batch open for max_wait / idle_flush
    BatchingService svc(inner.get(), bes, cfg);
    CHECK(svc.healthy());
    static const char* names[] = {"example_tool", "struct_tool", "file_tool", "nope", "", "example_tool "};
    std::atomic<uint32_t> bad{0};
    std::vector<std::thread> th;
    for (uint32_t t = 0; t < threads; ++t) th.emplace_back([&, t] {
        Context ctx; ctx.now_unix = 1700000000;
        for (uint32_t i = 0; i < per_thread; ++i) {
            const std::string name = names[(t + i) % 6];
            std::string text;
            for (uint32_t k = 0; k < (t * 7 + i * 13) % 200; ++k) text += "word" + std::to_string(k) + " ";
            Struct params{{"prompt", Value(text)}, {"n", Value((int)i)}};
            const bool with_params = (i % 5) != 4;
            Result r = svc.ExecuteTool(ctx, name, with_params ? &params : nullptr, nullptr, nullptr);
            if (r.error || !r.resp || !r.resp->decision) { ++bad; continue; }
            const pk_decision want = oracle_one(owned[0]->table_, name, with_params ? to_json(params) : std::string());
            if (std::memcmp(&want, &*r.resp->decision, sizeof want) != 0) ++bad;
            if (r.resp->status.code != 200) ++bad;                                   // inner service untouched
            if (name == "file_tool" && !std::holds_alternative<File>(r.resp->output)) ++bad;
        }
    });
    for (auto& x : th) x.join();
    CHECK(bad.load() == 0);
    const BatchingStats st = svc.stats();
    CHECK(st.requests == (uint64_t)threads * per_thread);
    CHECK(st.max_batch_seen <= max_batch && st.errors == 0);
    if (threads >= 8) CHECK(st.batches < st.requests);                               // coalescing happened
    if (!cuda) {
        uint64_t seen = 0;
        for (auto& b : owned) { CHECK(b->requests > 0); seen += b->requests; }        // every shard got work
        // (a payload larger than its slot of the arena also occupies the indices it runs over, as empty
        // placeholder requests: the device sees those too)
        CHECK(seen >= st.requests && (seen == st.requests || (1u << 20) / max_batch < 4096));
    }
    svc.Shutdown();
    // batch log: ONE record per flush (not three per request), and the records add up to the stream
    CHECK(batch_log.size() == st.batches);
    uint64_t logged = 0, cls_sum = 0;
    for (const std::string& line : batch_log) {
        CHECK(line.rfind("{\"level\":\"INFO\",\"msg\":\"routed batch\",\"gpu\":", 0) == 0 && line.back() == '}');
        const size_t r = line.find("\"requests\":"), c = line.find("\"classes\":[");
        CHECK(r != std::string::npos && c != std::string::npos && line.find("\"trigger\":\"") != std::string::npos &&
              line.find("\"duration\":\"") != std::string::npos && line.find("\"no_provider\":") != std::string::npos);
        if (r == std::string::npos || c == std::string::npos) continue;
        logged += std::strtoull(line.c_str() + r + 11, nullptr, 10);
        const char* q = line.c_str() + c + 11;
        for (int k = 0; k < 4; ++k) { char* end; cls_sum += std::strtoull(q, &end, 10); q = end + 1; }
    }
    CHECK(logged == st.requests && cls_sum == st.requests);
    Context ctx;
    Result r = svc.ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr);
    CHECK(r.error && !r.resp);                                                       // after Shutdown: error, not a hang
    std::printf("batching %s gpus=%u threads=%u: %llu requests in %llu batches (full %llu, deadline %llu, idle %llu, max %llu)\n",
                cuda ? "cuda" : "cpu-stand-in", n_gpus, threads, (unsigned long long)st.requests, (unsigned long long)st.batches,
                (unsigned long long)st.flush_full, (unsigned long long)st.flush_deadline,
                (unsigned long long)st.flush_idle, (unsigned long long)st.max_batch_seen);
}

// EnqueueAsync: the non-blocking half an event-loop server (or an open-loop load generator) uses.
static void test_async()
{
    auto inner = NewMockService();
    OracleBackend be0(2), be1(2);
    BatchingConfig cfg; cfg.max_batch = 64; cfg.max_wait_us = 200; cfg.max_payload_bytes = 1 << 20;
    BatchingService svc(inner.get(), {&be0, &be1}, cfg);
    static const char* names[] = {"example_tool", "struct_tool", "file_tool", "nope"};
    using Enq = BatchingService::Enqueue;
    std::vector<AsyncRequest> ring(512);
    std::vector<std::string> payloads(ring.size());
    uint32_t bad = 0, busy = 0;
    for (int round = 0; round < 3; ++round) {                                        // requests are reusable
        size_t polled = 0;                                                           // [0, polled) have been picked up
        auto poll_one = [&] {
            while (!ring[polled].done()) std::this_thread::yield();
            const pk_decision want = oracle_one(be0.table_, names[polled & 3], payloads[polled]);
            if (ring[polled].rc() != 0 || std::memcmp(&want, &ring[polled].decision(), sizeof want) != 0) ++bad;
            ++polled;
        };
        for (size_t i = 0; i < ring.size(); ++i) {
            payloads[i] = "{\"prompt\":\"";
            for (size_t k = 0; k < (i * 7 + round) % 90; ++k) payloads[i] += "tok" + std::to_string(k) + " ";
            payloads[i] += "\"}";
            CHECK(ring[i].done());
            for (;;) {
                const Enq e = svc.EnqueueAsync(ring[i], names[i & 3], payloads[i]);
                if (e == Enq::kAccepted) break;
                CHECK(e == Enq::kBusy);                                              // never refused
                if (e != Enq::kBusy) break;
                ++busy;
                if (polled < i) poll_one();                                          // picking records up is what frees arenas
                else std::this_thread::yield();                                      // (busy for a moment: the flusher is between two batches)
            }
        }
        while (polled < ring.size()) poll_one();
    }
    CHECK(bad == 0);
    CHECK(be0.requests > 0 && be1.requests > 0 && be0.requests + be1.requests == 3 * ring.size());
    std::printf("async: 3 x %zu requests, %u enqueue attempts met back-pressure\n", ring.size(), busy);
    AsyncRequest big;
    CHECK(svc.EnqueueAsync(big, "example_tool", std::string((2 << 20), 'x')) == Enq::kRefused && big.done());   // over the arena: refused
    AsyncRequest longname;
    CHECK(svc.EnqueueAsync(longname, std::string(300, 'n'), "{}") == Enq::kRefused && longname.done());
    svc.Shutdown();
    AsyncRequest late;
    CHECK(svc.EnqueueAsync(late, "example_tool", "{}") == Enq::kRefused && late.done());                       // after Shutdown: refused
}

// One client's oversized tool name must not take the service down (it did: the device rejected the
// whole batch and health went NOT_SERVING for good).  A name longer than the routing stage accepts
// equals no table entry, so it is the reference's default branch — 200, "Unknown tool: ..." —
// answered by the inner service without a decision; every other request is routed as usual.
static void test_oversized_name()
{
    auto inner = NewMockService();
    OracleBackend be(2);
    BatchingConfig cfg; cfg.max_batch = 16; cfg.max_wait_us = 100;
    BatchingService svc(inner.get(), {&be}, cfg);
    Context ctx; ctx.now_unix = 1;
    const std::string huge(100000, 'x');
    Result r = svc.ExecuteTool(ctx, huge, nullptr, nullptr, nullptr);
    CHECK(!r.error && r.resp && r.resp->status.code == 200 && !r.resp->decision.has_value());
    CHECK(std::get<std::string>(r.resp->output) == "Unknown tool: " + huge);            // mock.go:60-64
    const std::string edge(cfg.max_name_bytes_per_request, 'y');                        // the longest routed name
    r = svc.ExecuteTool(ctx, edge, nullptr, nullptr, nullptr);
    CHECK(!r.error && r.resp && r.resp->decision.has_value() && (r.resp->decision->flags & PK_FLAG_UNKNOWN_TOOL));
    // names of 57..256 bytes travel in the arena's side array, not in the request's record: same decision as the oracle's
    for (size_t len : {size_t(56), size_t(57), size_t(64), size_t(100), size_t(256)}) {
        std::string nm(len, 'a');
        for (size_t i = 0; i < len; ++i) nm[i] = (char)('a' + (i * 7 + len) % 26);
        Struct params{{"prompt", Value(std::string("route me ") + std::to_string(len))}};
        r = svc.ExecuteTool(ctx, nm, &params, nullptr, nullptr);
        const pk_decision want = oracle_one(be.table_, nm, to_json(params));
        CHECK(!r.error && r.resp && r.resp->decision.has_value() && std::memcmp(&want, &*r.resp->decision, sizeof want) == 0);
    }
    r = svc.ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr);
    CHECK(!r.error && r.resp && r.resp->decision.has_value() && r.resp->decision->tool_class == 0);
    CHECK(svc.healthy() && svc.stats().errors == 0);
}

// A batch the device REFUSES (argument / capacity) fails its own requests and nothing else:
// health stays SERVING and the next batch is routed.
static void test_rejected_batch_keeps_health()
{
    struct Picky final : RouteBackend {
        OracleBackend inner{2};
        int reject_next = 1;
        uint32_t slots() const override { return 2; }
        void* alloc_host(size_t b) override { return inner.alloc_host(b); }
        int submit(uint32_t s, const uint8_t* p, const uint64_t* po, const uint32_t* pl, const uint8_t* nm, const uint32_t* no,
                   uint32_t n, pk_decision* out) override
        { if (reject_next > 0) { --reject_next; return PK_ERR_CAPACITY; } return inner.submit(s, p, po, pl, nm, no, n, out); }
        int wait(uint32_t) override { return PK_OK; }
        std::string error_string(int s) override { return "picky " + std::to_string(s); }
    } be;
    auto inner = NewMockService();
    BatchingConfig cfg; cfg.max_batch = 8;
    BatchingService svc(inner.get(), {&be}, cfg);
    Context ctx;
    Result r = svc.ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr);
    CHECK(r.error && !r.resp && svc.healthy());                                         // this request failed, the service did not
    r = svc.ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr);
    CHECK(!r.error && r.resp && r.resp->decision.has_value() && svc.healthy() && svc.stats().errors == 1);
}

static void test_device_error_path()
{
    auto inner = NewMockService();
    OracleBackend be(2);
    be.fail_after = 0;                                                               // every submit fails
    BatchingConfig cfg; cfg.max_batch = 4; cfg.max_wait_us = 100;
    std::vector<std::string> batch_log;
    cfg.batch_log = &batch_log;
    BatchingService svc(inner.get(), {&be}, cfg);
    Context ctx;
    Result r = svc.ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr);
    CHECK(r.error && r.error->rfind("routing failed: ", 0) == 0 && !r.resp);          // plain error -> gRPC Unknown
    CHECK(batch_log.size() == 1 && batch_log[0].rfind("{\"level\":\"ERROR\",\"msg\":\"routed batch\"", 0) == 0 &&
          batch_log[0].find("\"error\":\"routing failed: ") != std::string::npos);
    CHECK(!svc.healthy());                                                           // -> health NOT_SERVING
    r = svc.ExecuteTool(ctx, "example_tool", nullptr, nullptr, nullptr);
    CHECK(r.error.has_value());
}

static void test_cuda_backend_fails_loudly()
{
    pk_table t; pk_ref_default_table(&t);
    CudaBackend be(0, t, 64, 1 << 16, 1);
    if (be.ok()) { std::printf("cuda backend present (GPU box)\n"); return; }
    CHECK(be.status() == PK_ERR_NO_DEVICE);                                          // no CPU fallback
    uint64_t off[2] = {0, 16}; uint32_t len[1] = {1}, noff[2] = {0, 0}; uint8_t p[16] = {0}; pk_decision d;
    CHECK(be.submit(0, p, off, len, p, noff, 1, &d) == PK_ERR_NO_DEVICE);
}

int main(int argc, char** argv)
{
    if (argc > 1 && std::strcmp(argv[1], "cuda") == 0) {
        // GPU box: the same batcher checks, answered by libpkroute.so through CudaBackend
        test_batching(1, 16, 120, 64, true);
        test_batching(2, 32, 60, 8, true);
        test_batching(1, 16, 60, 64, true, true);
        test_batching(1, 16, 60, 4096, true);        // 256-byte slots: payloads span several, placeholder requests reach the kernel
        std::printf(g_fail ? "HOST TESTS FAILED (%d)\n" : "HOST CUDA TESTS OK\n", g_fail);
        return g_fail ? 1 : 0;
    }
    test_mock_service();
    test_server();
    test_logging_interceptor();
    test_batching(1, 1, 40, 64);
    test_batching(1, 16, 60, 4);
    test_batching(1, 16, 120, 32);
    test_batching(2, 16, 120, 64);
    test_batching(4, 32, 60, 4096);
    test_batching(1, 16, 60, 32, false, true);
    test_batching(2, 32, 60, 8, false, true);
    test_async();
    test_oversized_name();
    test_rejected_batch_keeps_health();
    test_device_error_path();
    test_cuda_backend_fails_loudly();
    std::printf(g_fail ? "HOST TESTS FAILED (%d)\n" : "HOST TESTS OK\n", g_fail);
    return g_fail ? 1 : 0;
}

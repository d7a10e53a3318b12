// batching_service.hpp — the batching front-end that sits at the reference's plugin seam.
//
// `BatchingService` implements the same one-method interface as the reference's
// service.Service (internal/service/service.go:13-15) and is handed to the server exactly
// where NewMockService() is today (cmd/polykey/main.go:85-86):
//
//     inner   := service.NewMockService()
//     batched := NewBatchingService(inner, backends, cfg)      // <- new
//     srv     := server.NewServer(batched)
//
// grpc-go calls ExecuteTool from one goroutine (here: thread) per in-flight RPC, concurrently
// and without a limit (main.go:74-77), so this is a multi-producer / single-consumer design:
// callers enqueue and block on their own request; ONE flusher thread per GPU coalesces the
// queue into the flat arenas of include/pk_route.h (size- or deadline-triggered), crosses the
// C ABI once per batch (pk_submit / pk_wait, `pipeline_depth` batches in flight), hands each
// caller its 24-byte decision, and the caller then delegates to the inner service for the
// response body — so the gRPC surface and mock.go stay unchanged.
//
// Requests shard across GPUs by request id (id % n_gpus); there is no exchange between GPUs.
// A device error is returned to every request of the failed batch as a plain error (surfaced
// verbatim by server.go:37-40, i.e. gRPC Unknown) and flips healthy() to false so main can
// set health to NOT_SERVING (main.go:93-94).  Shutdown() drains the queues before
// GracefulStop (main.go:118-119).
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>

#include "service.hpp"

namespace polykey {

// One routing device.  CudaBackend binds libpkroute.so; tests inject a CPU stand-in.
class RouteBackend {
public:
    virtual ~RouteBackend() = default;
    virtual uint32_t slots() const = 0;
    // memory the flusher packs batches into (pinned for the CUDA backend)
    virtual void* alloc_host(size_t bytes) = 0;
    virtual int submit(uint32_t slot, const uint8_t* payload, const uint64_t* pay_off,
                       const uint32_t* pay_len, const uint8_t* names, const uint32_t* name_off,
                       uint32_t n, pk_decision* out) = 0;
    virtual int wait(uint32_t slot) = 0;
    virtual std::string error_string(int status) = 0;
};

// libpkroute.so through its C ABI (include/pk_route.h).  No CPU fallback: construction
// fails (ok() == false, status() != 0) when there is no CUDA device.
class CudaBackend final : public RouteBackend {
public:
    CudaBackend(int device, const pk_table& table, uint32_t max_requests, uint64_t max_payload_bytes,
                uint32_t n_slots);
    ~CudaBackend() override;
    bool ok() const { return h_ != nullptr; }
    int status() const { return status_; }
    uint32_t slots() const override { return n_slots_; }
    void* alloc_host(size_t bytes) override;
    int submit(uint32_t slot, const uint8_t* payload, const uint64_t* pay_off, const uint32_t* pay_len,
               const uint8_t* names, const uint32_t* name_off, uint32_t n, pk_decision* out) override;
    int wait(uint32_t slot) override;
    std::string error_string(int status) override;
private:
    pk_handle* h_ = nullptr;
    int status_ = 0;
    uint32_t n_slots_ = 0;
};

struct BatchingConfig {
    uint32_t max_batch = 4096;             // requests per flush
    uint32_t max_wait_us = 200;            // deadline from the first queued request
    uint32_t idle_flush_us = 20;           // flush earlier once no request has arrived for this long
                                           // (a closed population of callers is all queued by then)
    // Work-conserving flush: whenever a slot is free, route what is queued NOW instead of holding
    // the batch open for max_wait / idle_flush; batches then grow by themselves as load rises,
    // because requests pile up while all slots are busy.  Under open-loop arrivals this is what
    // keeps latency at the service time (at 1e5 requests/s the idle window almost never closes —
    // a gap of 20 µs between arrivals is rare — and p50 was 104 µs against 40).
    bool eager_flush = true;
    // Waiting without the kernel's help.  A futex sleep costs 10-70 µs to get out of (wake-up
    // latency, timer slack) — more than the GPU takes to route a small batch — so short waits poll:
    uint32_t flusher_spin_us = 100;        // an idle flusher polls its queue this long before it sleeps
    uint32_t caller_spin_us = 150;         // a caller polls for its decision this long before it sleeps ...
    uint32_t caller_spin_max_waiters = 8;  // ... unless more callers than this are already waiting (they would
                                           // only steal cores from each other and from the flusher)
    uint64_t max_payload_bytes = 32u << 20;
    uint32_t max_name_bytes_per_request = 64;
    // One summary record per routed batch: the reference writes three
    // synchronous log lines per REQUEST (main.go:33, 44-48; server.go:28-33), which dominate its
    // CPU path; a batching front-end can say the same once per flush.  JSON lines in the style of
    // the reference's slog handler minus the timestamp; nullptr = no batch log.
    std::vector<std::string>* batch_log = nullptr;
};

struct BatchingStats {
    uint64_t requests = 0, batches = 0, flush_full = 0, flush_deadline = 0, flush_idle = 0, flush_eager = 0,
             max_batch_seen = 0, errors = 0;
};

class BatchingService;

// One request handed to the routing stage without blocking (BatchingService::EnqueueAsync): an
// event-loop server — or the open-loop load generator of latency_bench — owns the storage, polls
// done(), then reads rc() / decision().  Reusable once done.  Must not move or die while in flight.
class AsyncRequest {
public:
    bool done() const { return done_.load(std::memory_order_acquire); }
    int rc() const { return rc_; }                          // valid once done(): 0 or a PK_ERR_* / backend status
    const pk_decision& decision() const { return decision_; }
private:
    friend class BatchingService;
    std::string name_, payload_;
    pk_decision decision_{};
    int rc_ = 0;
    std::atomic<bool> done_{true};
};

class BatchingService final : public Service {
public:
    // backends: one per GPU; not owned.
    BatchingService(Service* inner, std::vector<RouteBackend*> backends, BatchingConfig cfg);
    ~BatchingService() override;

    Result ExecuteTool(Context& ctx, const std::string& toolName, const Struct* parameters,
                       const std::string* secretId, const Metadata* metadata) override;

    // Non-blocking form of the routing half of ExecuteTool: queues (toolName, payload) on the shard
    // its request id selects and returns.  false = not accepted (shut down, unhealthy, too large).
    bool EnqueueAsync(AsyncRequest& r, std::string toolName, std::string payload);

    void Shutdown();                        // drain + join; further calls fail with an error
    bool healthy() const { return healthy_.load(); }
    BatchingStats stats() const;

private:
    struct Pending {
        const std::string* name;
        std::string payload;
        pk_decision decision{};
        int rc = 0;
        std::atomic<bool> done{false};      // set (release) under Shard::done_mu as the flusher's LAST access
        AsyncRequest* async = nullptr;      // set: the result goes there and this Pending is the service's to delete
    };
    struct Arena {                          // one per backend slot
        uint8_t* payload = nullptr;
        pk_decision* out = nullptr;
        std::vector<uint64_t> pay_off;
        std::vector<uint32_t> pay_len, name_off;
        std::vector<uint8_t> names;
        std::vector<Pending*> reqs;
        bool inflight = false;
        uint64_t bytes = 0;                 // payload bytes of the batch in flight
        const char* trigger = "";           // "full" | "idle" | "deadline" | "eager"
        int64_t t_submit_ns = 0;
    };
    // Producers never meet the flusher on a lock: a caller appends to ONE of kInboxes small
    // vectors (picked per thread), the flusher swaps each out in O(1) into its private backlog.
    // With one queue and one mutex, a few hundred thousand enqueues per second from several
    // threads starved the flusher (and each other) on the futex behind that mutex.
    struct Inbox {
        std::mutex mu;
        std::vector<Pending*> v;
        bool closed = false;                // set by the exiting flusher: pushes are refused from then on
    };
    static constexpr unsigned kInboxes = 8;
    struct Shard {
        RouteBackend* be = nullptr;
        Inbox inbox[kInboxes];
        std::deque<Pending*> backlog;       // flusher-private: drained, not yet in a batch
        std::atomic<uint32_t> queued{0};    // pushed and not yet in a batch (inboxes + backlog)
        std::atomic<uint32_t> waiting{0};   // callers blocked in ExecuteTool on this shard
        std::atomic<bool> sleeping{false};  // the flusher is (about to be) in its futex sleep
        std::mutex mu;                      // guards st, and the flusher's sleep
        std::condition_variable cv;
        std::vector<Arena> arenas;
        std::thread th;
        BatchingStats st;
        // Completion: a whole batch is marked done under done_mu and announced with ONE
        // notify_all (a futex broadcast) instead of one wake-up call per request.  The waiter,
        // whose stack owns the Pending, re-checks under done_mu, so it cannot return and destroy
        // the object while the flusher still touches it.
        std::mutex done_mu;
        std::condition_variable done_cv;
    };
    bool push(Shard* s, Pending* p);        // false: the shard is closed
    size_t drain(Shard* s);
    bool close_inboxes(Shard* s);
    void flusher(Shard* s);
    void complete(Shard* s, Arena& a, int rc);
    void log_batch(const Shard* s, const Arena& a, int rc);

    Service* inner_;
    BatchingConfig cfg_;
    std::vector<std::unique_ptr<Shard>> shards_;
    std::atomic<uint64_t> next_id_{0};
    std::atomic<bool> stopping_{false}, healthy_{true};
    std::mutex log_mu_;                     // flushers of different GPUs share the sink
};

}  // namespace polykey

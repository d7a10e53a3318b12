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
// and without a limit (main.go:74-77).  Every caller packs ITS OWN request: it takes the next
// record index of the open batch with one fetch_add, copies its payload into that index's slot of
// the page-locked arena and its tool name into the index's record, and commits by storing its
// length.  The ONE flusher thread per GPU never touches payload bytes: it seals the open batch
// (size-, deadline- or work-conserving trigger), opens the next arena so producers carry on, waits
// for the last commits while it spreads the records into the arrays of include/pk_route.h, crosses
// the C ABI once (pk_submit / pk_wait, `pipeline_depth` batches in flight) and publishes the whole
// batch with one store; each caller then reads its own 24-byte decision out of the record array
// and delegates to the inner service for the response body — so the gRPC surface and mock.go stay
// unchanged.  (Round 1 packed on the flusher thread: one memcpy per request on one core capped the
// seam at 3e6 requests/s.)
//
// Requests shard across GPUs by request id (id % n_gpus); there is no exchange between GPUs.
// A device error is returned to every request of the failed batch as a plain error (surfaced
// verbatim by server.go:37-40, i.e. gRPC Unknown); a CUDA / device failure also flips healthy()
// to false so main can set health to NOT_SERVING (main.go:93-94) — a rejected batch (bad
// argument, capacity) fails its own requests only.  Shutdown() drains before GracefulStop
// (main.go:118-119).
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
    // memory the producers pack batches into (pinned for the CUDA backend)
    virtual void* alloc_host(size_t bytes) = 0;
    virtual int submit(uint32_t slot, const uint8_t* payload, const uint64_t* pay_off,
                       const uint32_t* pay_len, const uint8_t* names, const uint32_t* name_off,
                       uint32_t n, pk_decision* out) = 0;
    virtual int wait(uint32_t slot) = 0;
    virtual std::string error_string(int status) = 0;
    // is this status a failure of the device (health goes NOT_SERVING), as opposed to a batch
    // the device refused (only its requests fail)?
    virtual bool is_device_failure(int status) const { return status == PK_ERR_CUDA || status == PK_ERR_NO_DEVICE; }
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
    uint32_t max_batch = 4096;             // requests per flush (<= 65536)
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
    uint32_t flusher_spin_us = 100;        // an idle flusher polls its open batch this long before it sleeps
    uint32_t caller_spin_us = 150;         // a caller polls for its decision this long before it sleeps ...
    uint32_t caller_spin_max_waiters = 8;  // ... unless more callers than this are already waiting (they would
                                           // only steal cores from each other and from the flusher)
    uint64_t max_payload_bytes = 32u << 20;   // payload arena per batch; a request's slot is max_payload_bytes / max_batch
                                              // (rounded down to 16) — a larger payload takes several slots
    // A request's tool name may be at most this long to be ROUTED.  A longer name cannot equal any
    // table entry (PK_MAX_TOOL_NAME is 64), so it is the switch's default branch whatever its
    // bytes (mock.go:60-64): ExecuteTool answers it through the inner service without a routing
    // decision instead of letting one client's oversized name fail a whole batch.
    uint32_t max_name_bytes_per_request = 256;
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

namespace detail {
// One batch under construction / in flight: the flat arrays of include/pk_route.h in memory the
// backend allocated, plus the words producers and the flusher meet on.
//
// Layout.  Request i of a batch owns a fixed SLOT of the payload arena, payload[i * slot_bytes ..)
// (the C ABI allows gaps between requests; the device reads only the bytes a request has), and ONE
// 64-byte record of its own — length, name length and the tool name itself when it has at most 56
// bytes (longer ones go to a side array).  So a reservation is nothing but taking the next index:
// ONE fetch_add on the batch's counter, which cannot fail and retry under contention the way a
// compare-and-swap on packed byte counts did (that word was the front-end's ceiling, and a
// metastable one: 4e6 or 9e6 requests/s from run to run), and everything a producer writes
// afterwards lies in cache lines no other producer of the batch touches.  A payload larger than
// one slot takes the k slots it covers; the k - 1 indices after it become empty placeholder requests.
struct Arena {
    struct alignas(64) Rec {
        uint32_t len;                      // payload bytes; kUncommitted until the producer has copied them, and again
                                           // once the caller has picked its record up (each caller marks its own)
        uint32_t name_len;                 // kPlaceholder: an index covered by the big payload before it
        uint8_t  name[56];                 // the tool name, if it fits
    };
    static constexpr uint32_t kInlineName = 56;
    // ---- written by producers ----
    uint8_t* payload = nullptr;            // page-locked
    uint8_t* long_names = nullptr;         // name i at long_names[i * name_stride ..) when longer than kInlineName
    pk_decision* out = nullptr;            // page-locked; the device's records
    std::vector<Rec> rec;
    // ---- written by the flusher (the arrays of include/pk_route.h) ----
    std::vector<uint64_t> pay_off;
    std::vector<uint32_t> pay_len;
    std::vector<uint32_t> name_off;
    std::vector<uint8_t>  names;           // packed
    // ---- reservation: [sealed:1][slots taken:63] ----
    alignas(64) std::atomic<uint64_t> state{kSealed};
    // How many slots were taken successfully, published by whoever closed the batch: the first to
    // set the sealed bit if the batch had not overflowed by then, else the one producer whose
    // request was the first not to fit (requests after it add to the counter but never fit).
    std::atomic<uint32_t> final_n{kUnset};
    // ---- completion ----
    alignas(64) std::atomic<uint32_t> gen_done{0};   // generation whose records are in `out`
    int rc = 0;                            // status of generation gen_done (written before gen_done)
    uint32_t gen = 0;                      // generation being filled / in flight (flusher writes, before opening)
    // ---- flusher-private ----
    uint32_t n = 0;
    uint32_t picked_up = 0;                // records [0, picked_up) of the completed batch have been picked up
    uint64_t bytes = 0;
    const char* trigger = "";              // "full" | "idle" | "deadline" | "eager"
    int64_t t_submit_ns = 0;

    static constexpr uint64_t kSealed = 1ull << 63;
    static constexpr uint32_t kUncommitted = 0xFFFFFFFFu, kPlaceholder = 0xFFFFFFFFu, kUnset = 0xFFFFFFFFu;
    static uint64_t count_of(uint64_t st) { return st & ~kSealed; }
};
}  // namespace detail

// One request handed to the routing stage without blocking (BatchingService::EnqueueAsync): an
// event-loop server — or the open-loop load generator of latency_bench — owns the storage, polls
// done() until it is true, then reads rc() / decision().  Reusable once done.  The owner MUST poll
// an accepted request to completion (the destructor does, if need be): the batch's arena is
// recycled only after every request of it has picked up its record.  Must not move while in flight,
// and must be done before the service is destroyed.
class AsyncRequest {
public:
    AsyncRequest() = default;
    AsyncRequest(const AsyncRequest&) = delete;
    AsyncRequest& operator=(const AsyncRequest&) = delete;
    ~AsyncRequest();
    bool done();                                            // polls; picks the record up the first time it is there
    int rc() const { return rc_; }                          // valid once done(): 0 or a PK_ERR_* / backend status
    const pk_decision& decision() const { return decision_; }
private:
    friend class BatchingService;
    detail::Arena* arena_ = nullptr;                        // set while in flight
    uint32_t idx_ = 0, gen_ = 0, slots_ = 1;
    pk_decision decision_{};
    int rc_ = 0;
};

class BatchingService final : public Service {
public:
    // backends: one per shard; not owned.  Usually one per GPU; two backends on the same GPU give it two
    // open batches, i.e. two reservation counters for its producers to meet on — one GPU then takes
    // 2.05e7 requests/s of 2 KB instead of 1.5e7 (the C ABI underneath: 2.4e7), at the price of a
    // second polling flusher thread.
    BatchingService(Service* inner, std::vector<RouteBackend*> backends, BatchingConfig cfg);
    ~BatchingService() override;

    Result ExecuteTool(Context& ctx, const std::string& toolName, const Struct* parameters,
                       const std::string* secretId, const Metadata* metadata) override;

    // Non-blocking form of the routing half of ExecuteTool: copies (toolName, payload) into the open
    // batch of the shard its request id selects and returns at once.
    //   kAccepted  poll r.done()
    //   kBusy      every arena of the shard is in flight or still holds records nobody has picked up:
    //              poll your outstanding requests (that is what frees arenas) and try again —
    //              back-pressure, never a wait inside this call
    //   kRefused   shut down, unhealthy, r still in flight, or payload / name too large
    enum class Enqueue { kAccepted, kBusy, kRefused };
    Enqueue EnqueueAsync(AsyncRequest& r, const std::string& toolName, const std::string& payload);

    void Shutdown();                        // drain + join; further calls fail with an error
    bool healthy() const { return healthy_.load(); }
    BatchingStats stats() const;

private:
    using Arena = detail::Arena;
    struct Ticket { Arena* a = nullptr; uint32_t idx = 0, gen = 0, slots = 1; };
    static constexpr uint32_t kNone = 0xFFFFFFFFu;
    struct Shard {
        RouteBackend* be = nullptr;
        std::vector<std::unique_ptr<Arena>> arenas;
        alignas(64) std::atomic<uint32_t> open{kNone};      // index of the arena accepting reservations
        std::atomic<uint32_t> open_seq{0};                  // bumped at every change of `open` (producers wait on it)
        std::atomic<bool> closed{false};                    // the flusher has exited: reservations are refused
        alignas(64) std::atomic<uint32_t> waiting{0};       // callers blocked in ExecuteTool on this shard
        std::atomic<bool> sleeping{false};                  // the flusher is (about to be) in its futex sleep
        std::mutex mu;                                      // guards st, and the flusher's sleep
        std::condition_variable cv;
        std::thread th;
        BatchingStats st;
    };
    enum class Reserve { kOk, kBusy, kClosed };
    Reserve reserve(Shard* s, const std::string& name, const void* payload, size_t len, Ticket& t, bool block);
    void wake_flusher(Shard* s);
    static uint64_t& thread_request_id();
    static uint64_t next_request_id();
    void flusher(Shard* s);
    void open_arena(Shard* s, uint32_t slot);
    void complete(Shard* s, Arena& a, int rc);
    void log_batch(const Shard* s, const Arena& a, int rc);

    Service* inner_;
    BatchingConfig cfg_;
    uint64_t slot_bytes_ = 0;               // payload slot of one request (a multiple of 16)
    uint32_t name_stride_ = 0;              // bytes per request in Arena::long_names
    std::vector<std::unique_ptr<Shard>> shards_;
    std::atomic<bool> stopping_{false}, healthy_{true};
    std::mutex log_mu_;                     // flushers of different GPUs share the sink
};

}  // namespace polykey

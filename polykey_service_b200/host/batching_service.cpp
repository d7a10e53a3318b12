// batching_service.cpp — see batching_service.hpp.
#include "batching_service.hpp"

#include <algorithm>
#include <chrono>
#include <cstring>

namespace polykey {

namespace {
inline uint64_t round16(uint64_t x) { return (x + 15u) & ~uint64_t(15); }
using Clock = std::chrono::steady_clock;
inline void cpu_relax()
{
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#else
    std::this_thread::yield();
#endif
}
}  // namespace

// ---------------------------------------------------------------- CudaBackend ------------
CudaBackend::CudaBackend(int device, const pk_table& table, uint32_t max_requests,
                         uint64_t max_payload_bytes, uint32_t n_slots)
    : n_slots_(n_slots)
{
    pk_config cfg{};
    cfg.max_requests = max_requests;
    cfg.n_slots = n_slots;
    cfg.max_payload_bytes = max_payload_bytes;
    cfg.max_name_bytes = (uint64_t)max_requests * 256u;
    status_ = pk_init(device, &table, &cfg, &h_);
    if (status_ != PK_OK) h_ = nullptr;
}

CudaBackend::~CudaBackend() { if (h_) pk_destroy(h_); }
void* CudaBackend::alloc_host(size_t bytes) { return h_ ? pk_host_alloc(h_, bytes) : nullptr; }

int CudaBackend::submit(uint32_t slot, const uint8_t* payload, const uint64_t* pay_off,
                        const uint32_t* pay_len, const uint8_t* names, const uint32_t* name_off,
                        uint32_t n, pk_decision* out)
{
    if (!h_) return PK_ERR_NO_DEVICE;
    return pk_submit(h_, slot, payload, pay_off, pay_len, names, name_off, n, out);
}

int CudaBackend::wait(uint32_t slot) { return h_ ? pk_wait(h_, slot) : PK_ERR_NO_DEVICE; }

std::string CudaBackend::error_string(int status)
{
    std::string s = pk_strerror(status);
    if (status == PK_ERR_CUDA && h_) s += " [cudaError " + std::to_string(pk_last_cuda_error(h_)) + "]";
    return s;
}

// ---------------------------------------------------------------- BatchingService --------
BatchingService::BatchingService(Service* inner, std::vector<RouteBackend*> backends, BatchingConfig cfg)
    : inner_(inner), cfg_(cfg)
{
    for (RouteBackend* be : backends) {
        auto s = std::make_unique<Shard>();
        s->be = be;
        s->arenas.resize(be->slots());
        for (Arena& a : s->arenas) {
            a.payload = static_cast<uint8_t*>(be->alloc_host(cfg_.max_payload_bytes + 64));
            a.out = static_cast<pk_decision*>(be->alloc_host((size_t)cfg_.max_batch * sizeof(pk_decision)));
            a.pay_off.resize(cfg_.max_batch + 1);
            a.pay_len.resize(cfg_.max_batch);
            a.name_off.resize(cfg_.max_batch + 1);
            a.reqs.reserve(cfg_.max_batch);
            if (!a.payload || !a.out) healthy_.store(false);
        }
        shards_.push_back(std::move(s));
    }
    if (shards_.empty()) healthy_.store(false);
    for (auto& s : shards_) s->th = std::thread(&BatchingService::flusher, this, s.get());
}

BatchingService::~BatchingService() { Shutdown(); }

void BatchingService::Shutdown()
{
    if (stopping_.exchange(true)) {
        for (auto& s : shards_) if (s->th.joinable()) s->th.join();
        return;
    }
    for (auto& s : shards_) {
        { std::lock_guard<std::mutex> g(s->mu); }
        s->cv.notify_all();
    }
    for (auto& s : shards_) if (s->th.joinable()) s->th.join();
}

BatchingStats BatchingService::stats() const
{
    BatchingStats t;
    for (const auto& s : shards_) {
        std::lock_guard<std::mutex> g(s->mu);
        t.requests += s->st.requests; t.batches += s->st.batches; t.flush_full += s->st.flush_full;
        t.flush_deadline += s->st.flush_deadline; t.flush_idle += s->st.flush_idle; t.flush_eager += s->st.flush_eager;
        t.errors += s->st.errors;
        if (s->st.max_batch_seen > t.max_batch_seen) t.max_batch_seen = s->st.max_batch_seen;
    }
    return t;
}

bool BatchingService::push(Shard* s, Pending* p)
{
    static std::atomic<unsigned> next_stripe{0};
    static thread_local const unsigned stripe = next_stripe.fetch_add(1) % kInboxes;
    Inbox& in = s->inbox[stripe];
    {
        std::lock_guard<std::mutex> g(in.mu);
        if (in.closed) return false;
        in.v.push_back(p);
    }
    s->queued.fetch_add(1);                                  // seq_cst: pairs with the flusher's sleeping / queued check
    if (s->sleeping.load()) {                                // a polling flusher needs no wake-up call
        { std::lock_guard<std::mutex> g(s->mu); }            // it is either before its predicate check or already waiting
        s->cv.notify_one();
    }
    return true;
}

// Flusher only: move everything the inboxes hold into the backlog.  O(1) per inbox under its lock.
size_t BatchingService::drain(Shard* s)
{
    static thread_local std::vector<Pending*> tmp;
    size_t got = 0;
    for (Inbox& in : s->inbox) {
        {
            std::lock_guard<std::mutex> g(in.mu);
            if (in.v.empty()) continue;
            tmp.swap(in.v);
        }
        s->backlog.insert(s->backlog.end(), tmp.begin(), tmp.end());
        got += tmp.size();
        tmp.clear();
    }
    return got;
}

// Flusher only, when stopping with nothing left: refuse further pushes.  false = something
// slipped in meanwhile; drain it and try again.
bool BatchingService::close_inboxes(Shard* s)
{
    for (Inbox& in : s->inbox) {
        std::lock_guard<std::mutex> g(in.mu);
        if (!in.v.empty()) return false;
        in.closed = true;
    }
    return true;
}

Result BatchingService::ExecuteTool(Context& ctx, const std::string& toolName, const Struct* parameters,
                                    const std::string* secretId, const Metadata* metadata)
{
    if (stopping_.load() || shards_.empty())
        return Result{nullptr, std::string("routing stage is shut down")};
    if (!healthy_.load())
        return Result{nullptr, std::string("routing stage is unhealthy")};

    Pending p;
    p.name = &toolName;
    if (parameters) p.payload = to_json(*parameters);       // the request's prompt payload bytes
    if (round16(p.payload.size()) > cfg_.max_payload_bytes)
        return Result{nullptr, std::string("request payload exceeds the routing stage's arena")};

    const uint64_t id = next_id_.fetch_add(1);
    Shard* s = shards_[id % shards_.size()].get();          // shard by request id, no exchange
    if (!push(s, &p)) return Result{nullptr, std::string("routing stage is shut down")};
    // Wait for the decision: poll first while few callers are waiting (the batch comes back in
    // tens of µs), then sleep.  `done` is the flusher's last access to p, so returning on it is safe.
    const uint32_t ahead = s->waiting.fetch_add(1, std::memory_order_relaxed);
    if (ahead < cfg_.caller_spin_max_waiters && cfg_.caller_spin_us) {
        const auto until = Clock::now() + std::chrono::microseconds(cfg_.caller_spin_us);
        while (!p.done.load(std::memory_order_acquire) && Clock::now() < until) cpu_relax();
    }
    if (!p.done.load(std::memory_order_acquire)) {
        std::unique_lock<std::mutex> w(s->done_mu);         // block this RPC's thread only
        s->done_cv.wait(w, [&] { return p.done.load(std::memory_order_acquire); });
    }
    s->waiting.fetch_sub(1, std::memory_order_relaxed);

    if (p.rc != 0) {
        // plain error: server.go:37-40 returns it verbatim (gRPC Unknown)
        return Result{nullptr, "routing failed: " + s->be->error_string(p.rc)};
    }
    Result r = inner_->ExecuteTool(ctx, toolName, parameters, secretId, metadata);
    if (r.resp) r.resp->decision = p.decision;
    return r;
}

// {"level":"INFO","msg":"routed batch","gpu":0,"requests":4096,"payload_bytes":8388608,
//  "trigger":"full","duration":"212.4µs","classes":[1024,1023,1025,1024],"no_provider":0}
// classes = decisions per tool class 0..3 (mock.go:32-64's four arms); on a failed batch the level
// is ERROR and "error" carries the text every request of the batch is answered with.
void BatchingService::log_batch(const Shard* s, const Arena& a, int rc)
{
    if (!cfg_.batch_log) return;
    const int64_t now = std::chrono::duration_cast<std::chrono::nanoseconds>(Clock::now().time_since_epoch()).count();
    size_t gpu = 0;
    while (gpu < shards_.size() && shards_[gpu].get() != s) ++gpu;
    std::string line = std::string("{\"level\":\"") + (rc ? "ERROR" : "INFO") + "\",\"msg\":\"routed batch\",\"gpu\":" +
                       std::to_string(gpu) + ",\"requests\":" + std::to_string(a.reqs.size()) + ",\"payload_bytes\":" +
                       std::to_string(a.bytes) + ",\"trigger\":\"" + a.trigger + "\",\"duration\":\"" +
                       LoggingInterceptor::format_duration(now - a.t_submit_ns) + "\"";
    if (rc) {
        line += ",\"error\":\"routing failed: " + s->be->error_string(rc) + "\"";
    } else {
        uint64_t cls[4] = {0, 0, 0, 0}, none = 0;
        for (size_t i = 0; i < a.reqs.size(); ++i) {
            ++cls[a.out[i].tool_class & 3u];
            if (a.out[i].flags & PK_FLAG_NO_PROVIDER) ++none;
        }
        line += ",\"classes\":[" + std::to_string(cls[0]) + "," + std::to_string(cls[1]) + "," + std::to_string(cls[2]) + "," +
                std::to_string(cls[3]) + "],\"no_provider\":" + std::to_string(none);
    }
    line += "}";
    std::lock_guard<std::mutex> g(log_mu_);
    cfg_.batch_log->push_back(std::move(line));
}

bool BatchingService::EnqueueAsync(AsyncRequest& r, std::string toolName, std::string payload)
{
    if (stopping_.load() || shards_.empty() || !healthy_.load()) return false;
    if (round16(payload.size()) > cfg_.max_payload_bytes) return false;
    r.name_ = std::move(toolName);
    r.payload_ = std::move(payload);
    r.rc_ = 0;
    r.done_.store(false, std::memory_order_relaxed);
    Pending* p = new Pending;
    p->name = &r.name_;
    p->payload.swap(r.payload_);
    p->async = &r;
    const uint64_t id = next_id_.fetch_add(1);
    Shard* s = shards_[id % shards_.size()].get();
    if (!push(s, p)) { r.payload_.swap(p->payload); delete p; r.done_.store(true); return false; }
    return true;
}

void BatchingService::complete(Shard* s, Arena& a, int rc)
{
    if (rc == 0) rc = s->be->wait((uint32_t)(&a - s->arenas.data()));
    log_batch(s, a, rc);
    if (rc != 0) {
        healthy_.store(false);
        std::lock_guard<std::mutex> g(s->mu);
        ++s->st.errors;
    }
    {
        std::lock_guard<std::mutex> w(s->done_mu);
        for (size_t i = 0; i < a.reqs.size(); ++i) {
            Pending* p = a.reqs[i];
            if (AsyncRequest* r = p->async) {
                r->rc_ = rc;
                if (rc == 0) r->decision_ = a.out[i];
                r->payload_.swap(p->payload);                // hand the buffer back for reuse
                delete p;
                r->done_.store(true, std::memory_order_release);   // r may be reused from here on
                continue;
            }
            p->rc = rc;
            if (rc == 0) p->decision = a.out[i];
            p->done.store(true, std::memory_order_release);  // p may be gone from here on
        }
    }
    s->done_cv.notify_all();
    a.reqs.clear();
    a.inflight = false;
}

void BatchingService::flusher(Shard* s)
{
    const size_t depth = s->arenas.size();
    std::deque<size_t> inflight;                            // submitted slots, oldest first
    std::vector<size_t> free_slots;
    for (size_t i = depth; i-- > 0;) free_slots.push_back(i);

    auto finish_oldest = [&] {
        const size_t slot = inflight.front();
        inflight.pop_front();
        complete(s, s->arenas[slot], 0);
        free_slots.push_back(slot);
    };

    for (;;) {
        drain(s);
        if (s->backlog.empty()) {
            if (!inflight.empty()) { finish_oldest(); continue; }   // nothing new: finish the oldest batch now
            // closed-loop callers come back within microseconds of their last answer: poll for
            // them before paying for a futex sleep and its wake-up
            const auto spin_until = Clock::now() + std::chrono::microseconds(cfg_.flusher_spin_us);
            while (s->queued.load(std::memory_order_acquire) == 0 && !stopping_.load() && Clock::now() < spin_until)
                cpu_relax();
            if (s->queued.load() == 0) {
                if (stopping_.load()) {
                    if (close_inboxes(s)) return;
                    continue;
                }
                std::unique_lock<std::mutex> lk(s->mu);
                s->sleeping.store(true);                     // seq_cst, then the predicate: see push()
                s->cv.wait(lk, [&] { return s->queued.load() != 0 || stopping_.load(); });
                s->sleeping.store(false);
            }
            continue;
        }
        // a batch is open: take it now (work-conserving), or hold it for more / the deadline of its first request
        const auto deadline = Clock::now() + std::chrono::microseconds(cfg_.max_wait_us);
        const auto idle = std::chrono::microseconds(cfg_.idle_flush_us ? cfg_.idle_flush_us : cfg_.max_wait_us);
        bool full = s->backlog.size() >= cfg_.max_batch, went_idle = false;
        const bool eager = cfg_.eager_flush && !full;       // a busy GPU makes the next batch bigger by itself
        while (!eager && !full && !stopping_.load()) {
            const size_t before = s->backlog.size();
            const uint32_t seen = s->queued.load(std::memory_order_acquire);
            const auto until = std::min(deadline, Clock::now() + idle);
            if (idle <= std::chrono::microseconds(100)) {
                // a short window is polled: a timed futex wait of 20 µs returns after 70 (timer slack)
                while (s->queued.load(std::memory_order_acquire) == seen && Clock::now() < until) cpu_relax();
            } else {
                std::unique_lock<std::mutex> lk(s->mu);
                s->sleeping.store(true);
                s->cv.wait_until(lk, until, [&] { return s->queued.load() != seen || stopping_.load(); });
                s->sleeping.store(false);
            }
            drain(s);
            full = s->backlog.size() >= cfg_.max_batch;
            if (full) break;
            const auto now = Clock::now();
            if (now >= deadline) break;
            if (s->backlog.size() == before && now >= until) { went_idle = true; break; }   // nothing new arrived
        }
        if (free_slots.empty()) {                           // free the slot we are about to reuse
            finish_oldest();
            drain(s);                                       // and take along what arrived meanwhile
        }
        const size_t slot = free_slots.back();
        free_slots.pop_back();
        Arena& a = s->arenas[slot];
        uint64_t at = 0;
        uint32_t n = 0;
        a.bytes = 0;
        while (!s->backlog.empty() && n < cfg_.max_batch) {
            Pending* p = s->backlog.front();
            const uint64_t sz = round16(p->payload.size());
            if (at + sz > cfg_.max_payload_bytes) break;    // arena full: the rest goes to the next batch
            s->backlog.pop_front();
            a.pay_off[n] = at;
            a.pay_len[n] = (uint32_t)p->payload.size();
            a.bytes += p->payload.size();
            a.reqs.push_back(p);
            at += sz;
            ++n;
        }
        s->queued.fetch_sub(n);
        const bool was_full = n >= cfg_.max_batch;          // (an eager batch may have filled while a slot was awaited)
        a.trigger = (full || was_full) ? "full" : eager ? "eager" : went_idle ? "idle" : "deadline";
        a.t_submit_ns = std::chrono::duration_cast<std::chrono::nanoseconds>(Clock::now().time_since_epoch()).count();
        {
            std::lock_guard<std::mutex> g(s->mu);
            ++s->st.batches;
            s->st.requests += n;
            if (full || was_full) ++s->st.flush_full; else if (eager) ++s->st.flush_eager;
            else if (went_idle) ++s->st.flush_idle; else ++s->st.flush_deadline;
            if (n > s->st.max_batch_seen) s->st.max_batch_seen = n;
        }

        // pack outside the lock: one memcpy per request into the (pinned) arena — gRPC buffers
        // live on the caller's heap and cannot be DMA sources
        a.names.clear();
        for (uint32_t i = 0; i < n; ++i) {
            const Pending* p = a.reqs[i];
            if (!p->payload.empty()) std::memcpy(a.payload + a.pay_off[i], p->payload.data(), p->payload.size());
            a.name_off[i] = (uint32_t)a.names.size();
            a.names.insert(a.names.end(), p->name->begin(), p->name->end());
        }
        a.pay_off[n] = at;
        a.name_off[n] = (uint32_t)a.names.size();
        const int rc = s->be->submit((uint32_t)slot, a.payload, a.pay_off.data(), a.pay_len.data(),
                                     a.names.data(), a.name_off.data(), n, a.out);
        if (rc != 0) {
            complete(s, a, rc);                             // fail exactly this batch's requests
            free_slots.push_back(slot);
        } else {
            a.inflight = true;
            inflight.push_back(slot);
        }
    }
}

}  // namespace polykey

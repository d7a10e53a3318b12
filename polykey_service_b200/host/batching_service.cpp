// batching_service.cpp — see batching_service.hpp.
#include "batching_service.hpp"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>

namespace polykey {

#if defined(__x86_64__)
#include <emmintrin.h>
#endif

namespace {
inline uint64_t round16(uint64_t x) { return (x + 15u) & ~uint64_t(15); }

// A request's payload into its slot of the page-locked arena.  Nobody on the CPU reads these bytes
// again (the GPU pulls them over PCIe), so whole 16-byte pieces go out with non-temporal stores:
// no read-for-ownership of the destination lines, no cache pollution — at 1e7 requests/s of 2 KB
// the ordinary copy's extra read of the arena is 20 GB/s of memory traffic.  dst is 16-byte aligned
// (slots are).  The caller's release store of the length is preceded by the sfence below.
inline void copy_to_arena(uint8_t* dst, const void* src, size_t len)
{
#if defined(__x86_64__)
    if (len >= 256) {
        const uint8_t* s = static_cast<const uint8_t*>(src);
        const size_t whole = len & ~size_t(15);
        for (size_t i = 0; i < whole; i += 16)
            _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i), _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + i)));
        if (len > whole) std::memcpy(dst + whole, s + whole, len - whole);
        _mm_sfence();                                        // the streamed bytes are visible before the commit
        return;
    }
#endif
    std::memcpy(dst, src, len);
}
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
    cfg.max_name_bytes = std::max<uint64_t>((uint64_t)max_requests * 256u, 1u << 20);   // >= the front-end's name arena
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
using detail::Arena;

BatchingService::BatchingService(Service* inner, std::vector<RouteBackend*> backends, BatchingConfig cfg)
    : inner_(inner), cfg_(cfg)
{
    cfg_.max_batch = std::max<uint32_t>(1u, std::min<uint32_t>(cfg_.max_batch, 65536u));
    cfg_.max_name_bytes_per_request = std::max<uint32_t>(64u, std::min<uint32_t>(cfg_.max_name_bytes_per_request, 4096u));
    slot_bytes_ = std::max<uint64_t>(16u, (cfg_.max_payload_bytes / cfg_.max_batch) & ~uint64_t(15));
    cfg_.max_payload_bytes = slot_bytes_ * cfg_.max_batch;   // what a single request can be, too (it then takes every slot)
    name_stride_ = (cfg_.max_name_bytes_per_request + 15u) & ~15u;
    for (RouteBackend* be : backends) {
        auto s = std::make_unique<Shard>();
        s->be = be;
        // more arenas than device slots: an arena stays taken after its batch has completed until
        // every caller has picked up its record, and that must not hold a device slot hostage
        const uint32_t n_arenas = 2u * be->slots() + 1u;
        for (uint32_t k = 0; k < n_arenas; ++k) {
            auto a = std::make_unique<Arena>();
            a->payload = static_cast<uint8_t*>(be->alloc_host(cfg_.max_payload_bytes + 64));
            a->out = static_cast<pk_decision*>(be->alloc_host((size_t)cfg_.max_batch * sizeof(pk_decision)));
            a->long_names = static_cast<uint8_t*>(std::malloc((size_t)cfg_.max_batch * name_stride_));
            a->pay_off.resize(cfg_.max_batch + 1);
            a->pay_len.resize(cfg_.max_batch);
            a->name_off.resize(cfg_.max_batch + 1);
            a->names.resize((size_t)cfg_.max_batch * cfg_.max_name_bytes_per_request);
            a->rec.resize(cfg_.max_batch);
            for (Arena::Rec& r : a->rec) { r.len = Arena::kUncommitted; r.name_len = 0; }
            if (!a->payload || !a->out || !a->long_names) healthy_.store(false);
            s->arenas.push_back(std::move(a));
        }
        shards_.push_back(std::move(s));
    }
    if (shards_.empty()) healthy_.store(false);
    if (!healthy_.load()) {                                  // nothing will be routed: refuse everything
        for (auto& s : shards_) s->closed.store(true);
        return;
    }
    for (auto& s : shards_) s->th = std::thread(&BatchingService::flusher, this, s.get());
}

BatchingService::~BatchingService()
{
    Shutdown();
    for (auto& s : shards_)
        for (auto& a : s->arenas) std::free(a->long_names);
}

void BatchingService::Shutdown()
{
    stopping_.store(true);
    for (auto& s : shards_) wake_flusher(s.get());
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

// Request ids are per calling thread — thread k's are k, k + 2^40, ... in the id space, i.e. its own
// round robin over the shards starting at shard k — so that taking one is not a read-modify-write on
// a line every caller of the service shares (with one batch counter per shard already being the
// front-end's ceiling, a second global counter per request was measurable).
uint64_t& BatchingService::thread_request_id()
{
    static std::atomic<uint64_t> next_thread{0};
    thread_local uint64_t id = next_thread.fetch_add(1, std::memory_order_relaxed);
    return id;
}
uint64_t BatchingService::next_request_id() { return thread_request_id()++; }

void BatchingService::wake_flusher(Shard* s)
{
    if (!s->sleeping.load()) return;                        // seq_cst: pairs with the flusher's sleeping / state check
    { std::lock_guard<std::mutex> g(s->mu); }               // it is either before its predicate check or already waiting
    s->cv.notify_one();
}

// Producer side: take the next record index of the open batch (k of them when the payload spans k
// slots) with one fetch_add, copy the request in, commit.  Callers have checked that the request
// fits an EMPTY batch.  When no batch is open (every arena busy) a blocking caller waits for one;
// the others get kBusy.
//
// A fetch_add cannot be taken back, so a request that does not fit has still moved the counter.
// That is harmless because nothing fits after it either (the counter only grows), and because the
// number of slots that WERE taken is published once, by exactly one party: the producer whose
// request was the first not to fit (it alone sees a counter still within the batch), or, if the
// batch never overflowed, whoever set the sealed bit first (it sees the counter at that instant).
BatchingService::Reserve BatchingService::reserve(Shard* s, const std::string& name, const void* payload, size_t len,
                                                  Ticket& t, bool block)
{
    const uint64_t k = len <= slot_bytes_ ? 1u : (len + slot_bytes_ - 1u) / slot_bytes_;
    const uint64_t cap = cfg_.max_batch;
    for (;;) {
        const uint32_t seq = s->open_seq.load(std::memory_order_acquire);
        const uint32_t slot = s->open.load(std::memory_order_acquire);
        if (slot != kNone) {
            Arena& a = *s->arenas[slot];
            const uint64_t seen = a.state.load(std::memory_order_relaxed);
            if (!(seen & Arena::kSealed) && seen + k <= cap) {           // do not touch a batch that cannot take us
                const uint64_t old = a.state.fetch_add(k);               // seq_cst: see wake_flusher()
                if (!(old & Arena::kSealed)) {
                    if (old + k <= cap) {
                        // slots [old, old + k) are ours; the arena cannot be recycled before we commit
                        const uint32_t i = (uint32_t)old;
                        t.a = &a; t.idx = i; t.gen = a.gen; t.slots = (uint32_t)k;
                        if (len) copy_to_arena(a.payload + (uint64_t)i * slot_bytes_, payload, len);
                        Arena::Rec& r = a.rec[i];
                        r.name_len = (uint32_t)name.size();
                        if (!name.empty())
                            std::memcpy(name.size() <= Arena::kInlineName ? r.name : a.long_names + (size_t)i * name_stride_,
                                        name.data(), name.size());
                        for (uint64_t j = 1; j < k; ++j) {                 // the indices our payload runs over: empty requests
                            a.rec[i + j].name_len = Arena::kPlaceholder;
                            std::atomic_ref<uint32_t>(a.rec[i + j].len).store(0u, std::memory_order_release);
                        }
                        std::atomic_ref<uint32_t>(r.len).store((uint32_t)len, std::memory_order_release);   // commit
                        if (i == 0) wake_flusher(s);                      // a batch has been started
                        return Reserve::kOk;
                    }
                    if (old <= cap) {
                        // the first request that did not fit: `old` slots were taken, close the batch
                        a.final_n.store((uint32_t)old, std::memory_order_release);
                        a.state.fetch_or(Arena::kSealed);
                        wake_flusher(s);
                    }
                }
            } else if (!(seen & Arena::kSealed) && seen <= cap) {
                // no room for this request in what is left of the batch: close it for the flusher
                const uint64_t old = a.state.fetch_or(Arena::kSealed);
                if (!(old & Arena::kSealed) && old <= cap) a.final_n.store((uint32_t)old, std::memory_order_release);
                wake_flusher(s);
            }
        }
        if (s->closed.load(std::memory_order_acquire)) return Reserve::kClosed;
        // every arena is busy, or the open one has just been sealed: wait for the next (back-pressure)
        for (int spin = 0; spin < 400 && s->open_seq.load(std::memory_order_acquire) == seq; ++spin) cpu_relax();
        if (s->open_seq.load(std::memory_order_acquire) != seq) continue;
        if (!block) return Reserve::kBusy;
        if (!s->closed.load(std::memory_order_acquire)) s->open_seq.wait(seq, std::memory_order_acquire);
    }
}

Result BatchingService::ExecuteTool(Context& ctx, const std::string& toolName, const Struct* parameters,
                                    const std::string* secretId, const Metadata* metadata)
{
    if (stopping_.load() || shards_.empty())
        return Result{nullptr, std::string("routing stage is shut down")};
    if (!healthy_.load())
        return Result{nullptr, std::string("routing stage is unhealthy")};
    // A name this long equals no table entry: it is the switch's default branch (mock.go:60-64)
    // whatever the routing stage would say, so answer it without one.
    if (toolName.size() > cfg_.max_name_bytes_per_request)
        return inner_->ExecuteTool(ctx, toolName, parameters, secretId, metadata);

    std::string payload;
    if (parameters) payload = to_json(*parameters);         // the request's prompt payload bytes
    if (round16(payload.size()) > cfg_.max_payload_bytes || payload.size() >= Arena::kUncommitted)
        return Result{nullptr, std::string("request payload exceeds the routing stage's arena")};

    const uint64_t id = next_request_id();
    Shard* s = shards_[id % shards_.size()].get();          // shard by request id, no exchange
    Ticket t;
    if (reserve(s, toolName, payload.data(), payload.size(), t, true) != Reserve::kOk)
        return Result{nullptr, std::string("routing stage is shut down")};
    // Wait for the batch: poll first while few callers are waiting (it comes back in tens of µs),
    // then sleep on the arena's generation word.
    Arena& a = *t.a;
    const uint32_t ahead = s->waiting.fetch_add(1, std::memory_order_relaxed);
    if (ahead < cfg_.caller_spin_max_waiters && cfg_.caller_spin_us) {
        const auto until = Clock::now() + std::chrono::microseconds(cfg_.caller_spin_us);
        while (a.gen_done.load(std::memory_order_acquire) != t.gen && Clock::now() < until) cpu_relax();
    }
    for (;;) {
        const uint32_t g = a.gen_done.load(std::memory_order_acquire);
        if (g == t.gen) break;
        a.gen_done.wait(g, std::memory_order_acquire);       // block this RPC's thread only
    }
    s->waiting.fetch_sub(1, std::memory_order_relaxed);
    const int rc = a.rc;
    pk_decision decision{};
    if (rc == 0) decision = a.out[t.idx];
    for (uint32_t j = t.slots; j-- > 0;)                     // picked up (placeholders too): our last access
        std::atomic_ref<uint32_t>(a.rec[t.idx + j].len).store(Arena::kUncommitted, std::memory_order_release);

    if (rc != 0) {
        // plain error: server.go:37-40 returns it verbatim (gRPC Unknown)
        return Result{nullptr, "routing failed: " + s->be->error_string(rc)};
    }
    Result r = inner_->ExecuteTool(ctx, toolName, parameters, secretId, metadata);
    if (r.resp) r.resp->decision = decision;
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
    uint64_t requests = 0;
    for (size_t i = 0; i < a.n; ++i) requests += a.rec[i].name_len != Arena::kPlaceholder;
    std::string line = std::string("{\"level\":\"") + (rc ? "ERROR" : "INFO") + "\",\"msg\":\"routed batch\",\"gpu\":" +
                       std::to_string(gpu) + ",\"requests\":" + std::to_string(requests) + ",\"payload_bytes\":" +
                       std::to_string(a.bytes) + ",\"trigger\":\"" + a.trigger + "\",\"duration\":\"" +
                       LoggingInterceptor::format_duration(now - a.t_submit_ns) + "\"";
    if (rc) {
        line += ",\"error\":\"routing failed: " + s->be->error_string(rc) + "\"";
    } else {
        uint64_t cls[4] = {0, 0, 0, 0}, none = 0;
        for (size_t i = 0; i < a.n; ++i) {
            if (a.rec[i].name_len == Arena::kPlaceholder) continue;      // an index a bigger payload runs over
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

BatchingService::Enqueue BatchingService::EnqueueAsync(AsyncRequest& r, const std::string& toolName, const std::string& payload)
{
    if (r.arena_) return Enqueue::kRefused;                  // still in flight
    if (stopping_.load() || shards_.empty() || !healthy_.load()) return Enqueue::kRefused;
    if (round16(payload.size()) > cfg_.max_payload_bytes || payload.size() >= Arena::kUncommitted) return Enqueue::kRefused;
    if (toolName.size() > cfg_.max_name_bytes_per_request) return Enqueue::kRefused;
    // the id (hence the shard) is consumed only once the request is in: a busy shard must not skew the others
    uint64_t& local = thread_request_id();
    Shard* s = shards_[local % shards_.size()].get();
    Ticket t;
    switch (reserve(s, toolName, payload.data(), payload.size(), t, false)) {
    case Reserve::kBusy:   return Enqueue::kBusy;
    case Reserve::kClosed: return Enqueue::kRefused;
    case Reserve::kOk:     break;
    }
    ++local;
    r.rc_ = 0;
    r.idx_ = t.idx;
    r.slots_ = t.slots;
    r.gen_ = t.gen;
    r.arena_ = t.a;
    return Enqueue::kAccepted;
}

bool AsyncRequest::done()
{
    Arena* a = arena_;
    if (!a) return true;
    if (a->gen_done.load(std::memory_order_acquire) != gen_) return false;
    rc_ = a->rc;
    if (rc_ == 0) decision_ = a->out[idx_];
    arena_ = nullptr;
    for (uint32_t j = slots_; j-- > 0;)                      // picked up (placeholders too): our last access
        std::atomic_ref<uint32_t>(a->rec[idx_ + j].len).store(Arena::kUncommitted, std::memory_order_release);
    return true;
}

AsyncRequest::~AsyncRequest()
{
    while (!done()) std::this_thread::yield();
}

// Flusher only.  The batch's records are in a.out (or rc says why not): publish the whole batch
// with one store; its callers pick their records up themselves.
void BatchingService::complete(Shard* s, Arena& a, int rc)
{
    log_batch(s, a, rc);
    if (rc != 0) {
        if (s->be->is_device_failure(rc)) healthy_.store(false);
        std::lock_guard<std::mutex> g(s->mu);
        ++s->st.errors;
    }
    a.rc = rc;
    a.picked_up = 0;
    a.gen_done.store(a.gen, std::memory_order_release);
    a.gen_done.notify_all();
}

// Flusher only: make arena `slot` the one producers reserve in.
void BatchingService::open_arena(Shard* s, uint32_t slot)
{
    Arena& a = *s->arenas[slot];                             // every rec[i].len reads kUncommitted again: its callers saw to that
    a.n = 0;
    ++a.gen;
    a.final_n.store(Arena::kUnset, std::memory_order_relaxed);
    a.state.store(0, std::memory_order_seq_cst);             // publishes gen and final_n
    s->open.store(slot, std::memory_order_release);
    s->open_seq.fetch_add(1, std::memory_order_release);
    s->open_seq.notify_all();
}

void BatchingService::flusher(Shard* s)
{
    struct Flight { uint32_t arena, be_slot; };
    std::deque<Flight> inflight;                            // submitted, oldest first
    std::deque<uint32_t> draining;                          // completed; callers are still picking records up
    std::vector<uint32_t> free_arenas, free_be;
    for (uint32_t i = (uint32_t)s->arenas.size(); i-- > 0;) free_arenas.push_back(i);
    for (uint32_t i = s->be->slots(); i-- > 0;) free_be.push_back(i);

    auto reap = [&] {
        for (size_t k = 0; k < draining.size();) {
            Arena& d = *s->arenas[draining[k]];
            while (d.picked_up < d.n &&
                   std::atomic_ref<uint32_t>(d.rec[d.picked_up].len).load(std::memory_order_acquire) == Arena::kUncommitted)
                ++d.picked_up;
            if (d.picked_up == d.n) {
                free_arenas.push_back(draining[k]);
                draining.erase(draining.begin() + (ptrdiff_t)k);
            } else {
                ++k;
            }
        }
    };
    auto finish_oldest = [&] {
        const Flight f = inflight.front();
        inflight.pop_front();
        Arena& a = *s->arenas[f.arena];
        complete(s, a, s->be->wait(f.be_slot));
        free_be.push_back(f.be_slot);
        draining.push_back(f.arena);
    };
    // An arena nobody is reading any more.  Finishes batches (blocks on the device) if that is what
    // it takes; producers keep filling the open batch meanwhile.
    auto take_free_arena = [&]() -> uint32_t {
        for (;;) {
            reap();
            if (!free_arenas.empty()) { const uint32_t k = free_arenas.back(); free_arenas.pop_back(); return k; }
            if (!inflight.empty()) { finish_oldest(); continue; }
            cpu_relax();                                    // all completed, some records not yet picked up
        }
    };

    // Close batch `a` to further reservations; returns how many slots were taken (see reserve()).
    auto seal = [&](Arena& a) -> uint32_t {
        const uint64_t old = a.state.fetch_or(Arena::kSealed);
        if (!(old & Arena::kSealed) && Arena::count_of(old) <= cfg_.max_batch) {
            a.final_n.store((uint32_t)old, std::memory_order_relaxed);          // we closed it, and it had not overflowed
            return (uint32_t)old;
        }
        uint32_t n;                                         // closed (or overflowed) by a producer, who says how many
        while ((n = a.final_n.load(std::memory_order_acquire)) == Arena::kUnset) cpu_relax();
        return n;
    };

    uint32_t open = take_free_arena();
    open_arena(s, open);
    auto first_seen = Clock::time_point{};                  // when the flusher first saw the open batch non-empty
    bool have_first = false;
    for (;;) {
        Arena& a = *s->arenas[open];
        uint64_t st = a.state.load(std::memory_order_seq_cst);
        if (Arena::count_of(st) == 0 && !(st & Arena::kSealed)) {
            have_first = false;
            if (!inflight.empty()) { finish_oldest(); continue; }   // nothing new: finish the oldest batch now
            reap();
            if (stopping_.load()) {
                // shut the door; a producer that slipped in first still gets its batch
                if (seal(a) == 0) break;
                st = a.state.load();
            } else {
                // closed-loop callers come back within microseconds of their last answer: poll for
                // them before paying for a futex sleep and its wake-up
                const auto spin_until = Clock::now() + std::chrono::microseconds(cfg_.flusher_spin_us);
                while (a.state.load(std::memory_order_acquire) == 0 && !stopping_.load() && Clock::now() < spin_until)
                    cpu_relax();
                if (a.state.load() == 0 && !stopping_.load()) {
                    std::unique_lock<std::mutex> lk(s->mu);
                    s->sleeping.store(true);                 // seq_cst, then the predicate: see wake_flusher()
                    s->cv.wait(lk, [&] { return a.state.load() != 0 || stopping_.load(); });
                    s->sleeping.store(false);
                }
                continue;
            }
        }
        if (!have_first) { first_seen = Clock::now(); have_first = true; }
        // a batch is open: take it now (work-conserving), or hold it for more / the deadline of its first request
        bool full = (st & Arena::kSealed) != 0 || Arena::count_of(st) >= cfg_.max_batch, went_idle = false;
        const bool eager = cfg_.eager_flush && !full;       // a busy GPU makes the next batch bigger by itself
        if (!eager && !full && !stopping_.load()) {
            const auto deadline = first_seen + std::chrono::microseconds(cfg_.max_wait_us);
            const auto idle = std::chrono::microseconds(cfg_.idle_flush_us ? cfg_.idle_flush_us : cfg_.max_wait_us);
            for (;;) {
                const uint64_t seen = a.state.load(std::memory_order_acquire);
                const auto until = std::min(deadline, Clock::now() + idle);
                if (idle <= std::chrono::microseconds(100)) {
                    // a short window is polled: a timed futex wait of 20 µs returns after 70 (timer slack)
                    while (a.state.load(std::memory_order_acquire) == seen && Clock::now() < until) cpu_relax();
                } else {
                    std::unique_lock<std::mutex> lk(s->mu);
                    s->sleeping.store(true);
                    s->cv.wait_until(lk, until, [&] {
                        const uint64_t v = a.state.load();
                        return (v & Arena::kSealed) != 0 || Arena::count_of(v) >= cfg_.max_batch || stopping_.load();
                    });
                    s->sleeping.store(false);
                }
                st = a.state.load(std::memory_order_acquire);
                full = (st & Arena::kSealed) != 0 || Arena::count_of(st) >= cfg_.max_batch;
                if (full || stopping_.load()) break;
                const auto now = Clock::now();
                if (now >= deadline) break;
                if (st == seen && now >= until) { went_idle = true; break; }   // nothing new arrived
            }
        }
        // the arena the NEXT batch opens in, before this one is sealed: producers never find the door shut
        // for longer than it takes to swap two words
        const uint32_t next = take_free_arena();
        const uint32_t n = seal(a);
        full = full || n >= cfg_.max_batch;
        open_arena(s, next);                                // producers carry on in the next arena
        have_first = false;
        if (n == 0) {                                       // sealed empty (shutdown race): nothing to route
            free_arenas.push_back(open);
            open = next;
            continue;
        }
        // The last producers may still be copying: wait for their commits, spreading the records into
        // the arrays the C ABI takes as they come: request i sits in slot i of the payload arena (an
        // index a bigger payload runs over starts where that payload ends, with length 0), names packed.
        a.bytes = 0;
        uint32_t real = 0, names_at = 0;
        uint64_t prev_end = 0;
        for (uint32_t i = 0; i < n; ++i) {
            Arena::Rec& rec = a.rec[i];
            uint32_t len;
            while ((len = std::atomic_ref<uint32_t>(rec.len).load(std::memory_order_acquire)) == Arena::kUncommitted) cpu_relax();
            const uint64_t off = std::max<uint64_t>((uint64_t)i * slot_bytes_, prev_end);
            a.pay_off[i] = off;
            a.pay_len[i] = len;
            a.name_off[i] = names_at;
            prev_end = off + round16(len);
            a.bytes += len;
            if (rec.name_len == Arena::kPlaceholder) continue;
            ++real;
            if (rec.name_len) {
                std::memcpy(a.names.data() + names_at,
                            rec.name_len <= Arena::kInlineName ? rec.name : a.long_names + (size_t)i * name_stride_, rec.name_len);
                names_at += rec.name_len;
            }
        }
        a.pay_off[n] = std::max<uint64_t>((uint64_t)n * slot_bytes_, prev_end);
        a.name_off[n] = names_at;
        a.n = n;
        a.trigger = full ? "full" : eager ? "eager" : went_idle ? "idle" : "deadline";
        a.t_submit_ns = std::chrono::duration_cast<std::chrono::nanoseconds>(Clock::now().time_since_epoch()).count();
        {
            std::lock_guard<std::mutex> g(s->mu);
            ++s->st.batches;
            s->st.requests += real;
            if (full) ++s->st.flush_full; else if (eager) ++s->st.flush_eager;
            else if (went_idle) ++s->st.flush_idle; else ++s->st.flush_deadline;
            if (real > s->st.max_batch_seen) s->st.max_batch_seen = real;
        }
        if (free_be.empty()) finish_oldest();               // a device slot for this batch
        const uint32_t be_slot = free_be.back();
        free_be.pop_back();
        const int rc = s->be->submit(be_slot, a.payload, a.pay_off.data(), a.pay_len.data(), a.names.data(),
                                     a.name_off.data(), n, a.out);
        if (rc != 0) {
            complete(s, a, rc);                             // fail exactly this batch's requests
            free_be.push_back(be_slot);
            draining.push_back(open);
        } else {
            inflight.push_back(Flight{open, be_slot});
        }
        open = next;
    }
    // stopping, the open batch sealed empty: refuse further reservations, finish what is in flight
    s->closed.store(true, std::memory_order_release);
    s->open.store(kNone, std::memory_order_release);
    s->open_seq.fetch_add(1, std::memory_order_release);
    s->open_seq.notify_all();
    while (!inflight.empty()) finish_oldest();
    // give callers a moment to pick their last records up (an abandoned AsyncRequest must not hang shutdown)
    const auto give_up = Clock::now() + std::chrono::seconds(5);
    for (reap(); !draining.empty() && Clock::now() < give_up; reap()) std::this_thread::yield();
}

}  // namespace polykey

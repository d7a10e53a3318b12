// doorbell_experiment.cu — ONE measured experiment for the launch floor (VERDICT r1 "next" #7,
// SURVEY §7 step 4): what does a request cost when no kernel is launched for it at all?
//
// A RESIDENT kernel (one CTA) polls a doorbell word in mapped page-locked host memory.  The host
// writes a request into a mapped arena, rings the doorbell, and spins on a completion word the
// kernel writes back; the kernel routes the request with the SAME device code as the product
// (this file includes pk_kernels.cu: route_body, latency mode).  Compared with the product's
// path for the same request: pk_submit + pk_wait, in place (one launch + one event wait).
//
// This is an experiment, not a product path: nothing in libpkroute.so uses it.  The kernel has a
// watchdog (it exits after `idle_ms` without a doorbell and after `life_ms` in any case), the host
// checks every record against pk_route_batch, and the run is wrapped in `timeout` by its caller.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/doorbell_experiment \
//        tools/doorbell_experiment.cu -Lpolykey_service_b200 -lpkroute -Xlinker -rpath=$PWD/polykey_service_b200
#include <chrono>
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>
#include <atomic>

#include "../polykey_service_b200/csrc/pk_kernels.cu"

namespace {

struct Doorbell {
    volatile uint32_t seq;        // host -> device: request number, 1, 2, ...
    volatile uint32_t n;          // requests of this ring (<= 8)
    volatile uint32_t stop;       // host -> device: leave
    uint32_t pad0[13];
    volatile uint32_t done;       // device -> host: last request number finished
    volatile uint32_t exited;     // device -> host: 1 = stop seen, 2 = watchdog
    uint32_t pad1[14];
};

__device__ __forceinline__ uint64_t gtime_ns()
{
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// One CTA, resident.  Thread 0 polls; the CTA routes up to 8 requests per ring (a warp each).
__global__ void __launch_bounds__(pk::kThreads) resident_router(Doorbell* db, pk::RouteParams P, const __grid_constant__ pk::DevTable T,
                                                                uint64_t idle_ns, uint64_t life_ns)
{
    __shared__ uint32_t s_seq, s_n, s_exit;
    uint32_t seen = 0;
    const uint64_t born = gtime_ns();
    for (;;) {
        if (threadIdx.x == 0) {
            const uint64_t t0 = gtime_ns();
            uint32_t s, why = 0;
            while ((s = db->seq) == seen) {
                if (db->stop) { why = 1; break; }
                const uint64_t now = gtime_ns();
                if (now - t0 > idle_ns || now - born > life_ns) { why = 2; break; }     // watchdog
            }
            __threadfence_system();                       // the request's bytes were written before the doorbell
            s_seq = s; s_n = db->n; s_exit = why;
        }
        __syncthreads();
        if (s_exit) {
            if (threadIdx.x == 0) { db->exited = s_exit; __threadfence_system(); }
            return;
        }
        pk::RouteParams Q = P;
        Q.n = s_n;
        pk::route_body<32, true, 1>(Q, T);                // latency mode: a warp scans and finalises a request
        __threadfence_system();                           // the records are in host memory ...
        __syncthreads();
        seen = s_seq;
        if (threadIdx.x == 0) { db->done = seen; __threadfence_system(); }              // ... before the completion word
        __syncthreads();
    }
}

double pct(std::vector<double>& v, double p)
{
    std::sort(v.begin(), v.end());
    return v[std::min(v.size() - 1, (size_t)(p * v.size()))];
}

#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e__)); return 1; } } while (0)

}  // namespace

int main(int argc, char** argv)
{
    const int iters = argc > 1 ? std::atoi(argv[1]) : 20000;
    const uint32_t plen = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 2048;
    using Clock = std::chrono::steady_clock;
    CK(cudaSetDevice(0));
    pk_table table;
    pk_default_table(&table);
    pk::DevTable dt;
    pk::make_dev_table(table, dt);
    pk::DevTable* d_table;
    CK(cudaMalloc(&d_table, sizeof dt));
    CK(cudaMemcpy(d_table, &dt, sizeof dt, cudaMemcpyHostToDevice));
    pk::SegScratch* d_scratch;
    CK(cudaMalloc(&d_scratch, 8 * sizeof(pk::SegScratch)));
    CK(cudaMemset(d_scratch, 0, 8 * sizeof(pk::SegScratch)));

    // mapped page-locked memory: a ring of request slots (so the device never re-reads an address it
    // has seen recently), the request's metadata, the record, the doorbell
    const uint32_t kSlots = 4096, slot_bytes = (plen + 15u) & ~15u;
    uint8_t* arena; uint64_t* pay_off; uint32_t *pay_len, *name_off; uint8_t* names; pk_decision* out; Doorbell* db;
    CK(cudaHostAlloc((void**)&arena, (size_t)kSlots * slot_bytes, cudaHostAllocMapped));
    CK(cudaHostAlloc((void**)&pay_off, 64, cudaHostAllocMapped));
    CK(cudaHostAlloc((void**)&pay_len, 64, cudaHostAllocMapped));
    CK(cudaHostAlloc((void**)&name_off, 64, cudaHostAllocMapped));
    CK(cudaHostAlloc((void**)&names, 64, cudaHostAllocMapped));
    CK(cudaHostAlloc((void**)&out, 8 * sizeof(pk_decision), cudaHostAllocMapped));
    CK(cudaHostAlloc((void**)&db, sizeof(Doorbell), cudaHostAllocMapped));
    std::memset((void*)db, 0, sizeof(Doorbell));
    db->n = 1;                                             // one request per ring
    static const char* kName = "example_tool";
    std::memcpy(names, kName, 12);
    name_off[0] = 0; name_off[1] = 12;
    pay_len[0] = plen;

    // the product path, for the expected records and for the comparison: pk_submit + pk_wait, in place
    pk_config cfg{8, 1, (uint64_t)kSlots * slot_bytes, 1024};
    pk_handle* h = nullptr;
    if (pk_init(0, &table, &cfg, &h) != PK_OK) { std::fprintf(stderr, "pk_init failed\n"); return 1; }
    std::vector<pk_decision> want(kSlots);
    std::vector<uint8_t> text((size_t)kSlots * slot_bytes);
    for (size_t i = 0; i < text.size(); ++i) text[i] = (uint8_t)("route the request to a provider, 42 times. "[(i * 7 + i / 2048) % 43]);
    std::memcpy(arena, text.data(), text.size());
    uint64_t off2[2]; uint32_t noff2[2] = {0, 12};
    pk_decision* pinned_out = (pk_decision*)pk_host_alloc(h, sizeof(pk_decision));
    std::vector<double> lat_call;
    for (uint32_t k = 0; k < kSlots; ++k) {
        off2[0] = (uint64_t)k * slot_bytes; off2[1] = off2[0] + slot_bytes;
        const auto a = Clock::now();
        if (pk_submit(h, 0, arena, off2, pay_len, names, noff2, 1, pinned_out) != PK_OK || pk_wait(h, 0) != PK_OK) { std::fprintf(stderr, "pk_submit failed\n"); return 1; }
        lat_call.push_back(std::chrono::duration<double, std::micro>(Clock::now() - a).count());
        want[k] = *pinned_out;
    }
    const uint64_t direct = pk_direct_count(h);

    pk::RouteParams P{};
    P.payload = arena; P.pay_off = pay_off; P.pay_len = pay_len; P.names = names; P.name_off = name_off;
    P.extra = nullptr; P.table = d_table; P.scratch = d_scratch; P.out = out; P.n = 1; P.n_extra = 0;
    P.reqs_per_cta = 8; P.n_req_ctas = 1; P.one = 1u; P.pdl = 0;
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    resident_router<<<1, pk::kThreads, 0, st>>>(db, P, dt, 2000000000ull /* idle: 2 s */, 30000000000ull /* life: 30 s */);
    CK(cudaGetLastError());

    std::vector<double> lat;
    lat.reserve(iters);
    int bad = 0;
    for (int it = 0; it < iters; ++it) {
        const uint32_t k = (uint32_t)it % kSlots;
        const auto a = Clock::now();
        pay_off[0] = (uint64_t)k * slot_bytes;             // (the payload bytes are in the ring already; a server would
        pay_off[1] = pay_off[0] + slot_bytes;              //  memcpy them here — the same 2 KB copy on either path)
        std::atomic_thread_fence(std::memory_order_release);
        db->seq = (uint32_t)it + 1u;                       // ring
        while (db->done != (uint32_t)it + 1u) {            // wait for the record
            if (db->exited) { std::fprintf(stderr, "resident kernel left early (%u) at %d\n", db->exited, it); return 1; }
        }
        std::atomic_thread_fence(std::memory_order_acquire);
        const pk_decision got = out[0];
        lat.push_back(std::chrono::duration<double, std::micro>(Clock::now() - a).count());
        if (std::memcmp(&got, &want[k], sizeof got) != 0) ++bad;
    }
    db->stop = 1;
    CK(cudaStreamSynchronize(st));
    std::vector<double> warm(lat.begin() + std::min<size_t>(lat.size() / 10, 1000), lat.end());
    std::vector<double> warm_call(lat_call.begin() + 200, lat_call.end());
    std::printf("{\"payload_bytes\": %u, \"iterations\": %d, \"records_differing_from_pk_submit\": %d, "
                "\"doorbell_us\": {\"p50\": %.2f, \"p99\": %.2f}, \"pk_submit_wait_us\": {\"p50\": %.2f, \"p99\": %.2f, \"in_place\": %s}, "
                "\"kernel_exit\": %u}\n",
                plen, iters, bad, pct(warm, 0.5), pct(warm, 0.99), pct(warm_call, 0.5), pct(warm_call, 0.99),
                direct >= kSlots ? "true" : "false", db->exited);
    pk_destroy(h);
    return bad ? 2 : 0;
}

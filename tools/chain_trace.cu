// chain_trace.cu — when does each launch of a short PDL chain start and end?  (diagnostic)
// Same device code as the product (includes pk_kernels.cu); every launch records the earliest CTA
// start and the latest CTA end (%globaltimer) into a trace, so the fill / drain of a K-launch chain
// can be read off:   build/chain_trace [K=20] [R=64] [chains=1]
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>

#include "../polykey_service_b200/csrc/pk_kernels.cu"

namespace {
struct Span { unsigned long long start, end; };

__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(pk::kThreads, PK_MIN_CTAS)
traced(const pk::RouteParams P, const __grid_constant__ pk::DevTable T, Span* span)
{
    if (threadIdx.x == 0) atomicMin(&span->start, gtime());
    if (P.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    pk::route_body<32, false, 1>(P, T);
    if (P.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0) atomicMax(&span->end, gtime());
}
#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e__)); return 1; } } while (0)
}  // namespace

int main(int argc, char** argv)
{
    const int K = argc > 1 ? std::atoi(argv[1]) : 20;
    const uint32_t R = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 64u;
    const int chains = argc > 3 ? std::atoi(argv[3]) : 1;
    const uint32_t n = 4096, plen = 2048;
    const int ring = 33;
    CK(cudaSetDevice(0));
    pk_table table; pk_default_table(&table);
    pk::DevTable dt; pk::make_dev_table(table, dt);
    pk::DevTable* d_table; CK(cudaMalloc(&d_table, sizeof dt)); CK(cudaMemcpy(d_table, &dt, sizeof dt, cudaMemcpyHostToDevice));
    // one set of metadata (identical for every batch), `ring` payload arenas
    std::vector<uint64_t> off(n + 1); std::vector<uint32_t> len(n, plen), noff(n + 1);
    std::vector<uint8_t> names((size_t)n * 12);
    for (uint32_t i = 0; i <= n; ++i) { off[i] = (uint64_t)i * plen; noff[i] = i * 12; }
    for (uint32_t i = 0; i < n; ++i) std::memcpy(&names[(size_t)i * 12], "example_tool", 12);
    uint64_t* d_off; uint32_t *d_len, *d_noff; uint8_t* d_names; pk_decision* d_out; pk::SegScratch* d_scr; Span* d_span;
    CK(cudaMalloc(&d_off, (n + 1) * 8)); CK(cudaMalloc(&d_len, n * 4)); CK(cudaMalloc(&d_noff, (n + 1) * 4));
    CK(cudaMalloc(&d_names, names.size())); CK(cudaMalloc(&d_out, (size_t)ring * n * sizeof(pk_decision)));
    CK(cudaMalloc(&d_scr, n * sizeof(pk::SegScratch))); CK(cudaMemset(d_scr, 0, n * sizeof(pk::SegScratch)));
    CK(cudaMalloc(&d_span, K * sizeof(Span)));
    CK(cudaMemcpy(d_off, off.data(), (n + 1) * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_len, len.data(), n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_noff, noff.data(), (n + 1) * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_names, names.data(), names.size(), cudaMemcpyHostToDevice));
    std::vector<uint8_t*> arena(ring);
    std::vector<uint8_t> text((size_t)n * plen);
    for (size_t i = 0; i < text.size(); ++i) text[i] = (uint8_t)("route the request to a provider, 42 times. "[(i * 7 + i / 2048) % 43]);
    for (int j = 0; j < ring; ++j) { CK(cudaMalloc(&arena[j], text.size())); CK(cudaMemcpy(arena[j], text.data(), text.size(), cudaMemcpyHostToDevice)); }

    std::vector<cudaStream_t> st(chains);
    for (auto& s : st) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaEvent_t fork, join[16];
    CK(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
    for (auto& e : join) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    auto launch = [&](int k, int pos, cudaStream_t s) {
        pk::RouteParams P{};
        P.payload = arena[pos % ring]; P.pay_off = d_off; P.pay_len = d_len; P.names = d_names; P.name_off = d_noff;
        P.extra = nullptr; P.table = d_table; P.scratch = d_scr; P.out = d_out + (size_t)(pos % ring) * n; P.n = n; P.n_extra = 0;
        P.reqs_per_cta = R; P.n_req_ctas = (P.n + R - 1u) / R; P.one = 1u; P.pdl = 1u;
        cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3((n + R - 1) / R); cfg.blockDim = dim3(pk::kThreads); cfg.stream = s;
        cudaLaunchAttribute attr{}; attr.id = cudaLaunchAttributeProgrammaticStreamSerialization; attr.val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = &attr; cfg.numAttrs = 1;
        return cudaLaunchKernelEx(&cfg, traced, P, dt, d_span + k);
    };
    int pos = 0;
    for (int rep = 0; rep < 4; ++rep) {
        std::vector<Span> init(K, Span{~0ull, 0ull});
        CK(cudaMemcpy(d_span, init.data(), K * sizeof(Span), cudaMemcpyHostToDevice));
        cudaGraph_t graph; cudaGraphExec_t exec;
        CK(cudaStreamBeginCapture(st[0], cudaStreamCaptureModeThreadLocal));
        CK(cudaEventRecord(fork, st[0]));
        for (int b = 1; b < chains; ++b) CK(cudaStreamWaitEvent(st[b], fork, 0));
        for (int k = 0; k < K; ++k) CK(launch(k, pos + k, st[((pos + k) % ring) % chains]));
        for (int b = 1; b < chains; ++b) { CK(cudaEventRecord(join[b], st[b])); CK(cudaStreamWaitEvent(st[0], join[b], 0)); }
        CK(cudaStreamEndCapture(st[0], &graph));
        CK(cudaGraphInstantiate(&exec, graph, 0));
        CK(cudaGraphUpload(exec, st[0]));
        CK(cudaStreamSynchronize(st[0]));
        CK(cudaGraphLaunch(exec, st[0]));
        CK(cudaStreamSynchronize(st[0]));
        pos += K;
        std::vector<Span> sp(K);
        CK(cudaMemcpy(sp.data(), d_span, K * sizeof(Span), cudaMemcpyDeviceToHost));
        if (rep < 3) continue;                       // the last repetition is the one printed
        unsigned long long t0 = ~0ull, t1 = 0;
        for (auto& x : sp) { t0 = std::min(t0, x.start); t1 = std::max(t1, x.end); }
        std::printf("K=%d R=%u chains=%d: first start -> last end %.2f us = %.3f us per launch\n", K, R, chains, (t1 - t0) / 1e3, (t1 - t0) / 1e3 / K);
        for (int k = 0; k < K; ++k)
            std::printf("  launch %2d  start %7.2f  end %7.2f  (%.2f us)\n", k, (sp[k].start - t0) / 1e3, (sp[k].end - t0) / 1e3, (sp[k].end - sp[k].start) / 1e3);
        cudaGraphExecDestroy(exec); cudaGraphDestroy(graph);
    }
    return 0;
}

// call_latency.cpp — what one pk_route_batch / pk_submit + pk_wait costs from C++ (no Python, no
// front-end): median µs for a batch of n requests x B bytes in page-locked buffers, in place and
// with copies, and the split between pk_submit (CPU: validate, pack metadata, launch) and pk_wait.
//   g++ -O2 -std=c++17 tools/call_latency.cpp -Iinclude -Lpolykey_service_b200 -lpkroute -o build/call_latency
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pk_route.h"

using Clock = std::chrono::steady_clock;
static double us(Clock::time_point a, Clock::time_point b) { return std::chrono::duration<double, std::micro>(b - a).count(); }
static double med(std::vector<double>& v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; }

int main(int argc, char** argv)
{
    const uint32_t n = argc > 1 ? (uint32_t)std::atoi(argv[1]) : 1;
    const uint32_t B = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 2048;
    pk_table t; pk_default_table(&t);
    pk_config cfg{}; cfg.max_requests = n; cfg.n_slots = 1; cfg.max_payload_bytes = (uint64_t)n * ((B + 15) & ~15u) + 64; cfg.max_name_bytes = 64ull * n;
    pk_handle* h = nullptr;
    if (pk_init(0, &t, &cfg, &h) != PK_OK) { std::fprintf(stderr, "pk_init failed (no GPU?)\n"); return 1; }
    uint8_t* pay = (uint8_t*)pk_host_alloc(h, cfg.max_payload_bytes);
    pk_decision* out = (pk_decision*)pk_host_alloc(h, sizeof(pk_decision) * n);
    std::vector<uint64_t> off(n + 1); std::vector<uint32_t> len(n), noff(n + 1); std::vector<uint8_t> names;
    for (uint32_t i = 0; i < n; ++i) {
        off[i] = (uint64_t)i * ((B + 15) & ~15u); len[i] = B; noff[i] = (uint32_t)names.size();
        const char* nm = i % 3 ? "example_tool" : "file_tool"; names.insert(names.end(), nm, nm + std::strlen(nm));
        for (uint32_t k = 0; k < B; ++k) pay[off[i] + k] = (uint8_t)("the quick brown fox "[(k + i) % 20]);
    }
    off[n] = (uint64_t)n * ((B + 15) & ~15u); noff[n] = (uint32_t)names.size();
    for (int mode = 0; mode < 2; ++mode) {
        pk_set_direct_bytes(h, mode == 0 ? (16u << 20) : 0);
        std::vector<double> tot, sub, wai;
        for (int r = 0; r < 3000; ++r) {
            const auto a = Clock::now();
            int rc = pk_submit(h, 0, pay, off.data(), len.data(), names.data(), noff.data(), n, out);
            const auto b = Clock::now();
            rc |= pk_wait(h, 0);
            const auto c = Clock::now();
            if (rc) { std::fprintf(stderr, "rc %d\n", rc); return 1; }
            if (r >= 500) { tot.push_back(us(a, c)); sub.push_back(us(a, b)); wai.push_back(us(b, c)); }
        }
        std::printf("{\"n\": %u, \"payload_bytes\": %u, \"path\": \"%s\", \"call_us\": %.2f, \"submit_us\": %.2f, \"wait_us\": %.2f, \"direct_batches\": %llu}\n",
                    n, B, mode == 0 ? "in place" : "copies", med(tot), med(sub), med(wai), (unsigned long long)pk_direct_count(h));
    }
    pk_destroy(h);
    return 0;
}

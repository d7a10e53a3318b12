// pk_kernels.cu — the fused scan + hash + select kernel for sm_100a.
//
// One pass over the payload bytes does all three steps of the routing decision:
//   (1) SIMD-in-word byte classification / token estimate         (SPEC §B)
//   (2) lane-striped route-key hash                                (SPEC §C)
//   (3) tool-class match against the device-resident table — the batched form of the
//       `switch toolName` at internal/service/mock.go:32-64 — and provider selection
//       (SPEC §A, §D).
//
// Shape of a CTA (256 threads = 8 warps), which owns R consecutive requests:
//   phase 1  warp-cooperative scan.  A group of G lanes (G = 32, or 16 / 8 when every
//            payload of the batch fits one step of G 16-byte blocks) owns one request; lane l
//            of the group loads block (G*i + l) with one 128-bit streaming load, so a warp
//            reads 512 contiguous bytes per step, four steps in flight.  Each lane keeps its
//            own hash stripe and token counter and parks them in shared memory — no
//            cross-lane traffic at all in the scan (a letter run is cut at every 8th offset,
//            and every block starts at one).
//   phase 2  one THREAD per request: sums the G partials, adds the tool-name terms, runs the
//            tool switch and the provider choice and stores the 24-byte record.  32 requests
//            per warp instruction instead of one.
// Requests longer than one 16 KiB segment are split: the extra segments are scanned by
// separate warps (listed by the host), partials are folded into a 32-byte scratch entry with
// integer atomics (a commutative sum: any order gives the same bits) and the warp that
// arrives last finalises the request.
//
// HBM-bound integer/byte work: no tensor cores — nothing here is a contraction.
#include <algorithm>
#include <cstdlib>

#include "pk_kernels.cuh"
#include "pk_spec.cuh"

// Build-time tuning knobs (python -m polykey_service_b200.build passes the chosen values).
#ifndef PK_TMA
#define PK_TMA 0         // 1: segments >= 1 of long requests are staged through shared memory with
#endif                   //    cp.async.bulk + mbarrier (a 4-stage, 2 KiB ring per warp) instead of LDG.128
#ifndef PK_MIN_CTAS
#define PK_MIN_CTAS 5    // __launch_bounds__ minimum resident CTAs per SM, full-warp kernels (G == 32)
#endif
#ifndef PK_MIN_CTAS_TINY
#define PK_MIN_CTAS_TINY 8   // ... lane-group kernels (G < 32): latency-bound, more resident warps win
#endif
#ifndef PK_MIN_CTAS_TINY4
#define PK_MIN_CTAS_TINY4 6  // ... the 4-lane x 4-block variant: four 128-bit loads in flight per lane need the registers
#endif
#ifndef PK_UNROLL
#define PK_UNROLL 4      // 128-bit loads in flight per lane
#endif

namespace pk {
namespace {

constexpr int      kWarpsPerCta = 8;
constexpr int      kThreads     = kWarpsPerCta * 32;
constexpr int      kUnroll      = PK_UNROLL;  // 128-bit loads in flight per lane
constexpr unsigned kFull        = 0xFFFFFFFFu;
constexpr int      kPartialSlots = 32 * 65;   // >= G * (R + 1) for every (G, R) launch_route picks
constexpr uint32_t kTinyWarpMax = 8192;       // launch_route: batches of tiny payloads up to this size take a warp per request
                                              // (serial µs, packed vs warp each: 1 request 4.3 / 2.2, 1,024: 6.1 / 2.9,
                                              // 4,096: 6.3 / 3.8, 8,192: 6.5 / 5.6, 16,384: 7.7 / 13.2)
constexpr uint32_t kNameStage   = 3072;       // bytes of tool names a CTA stages in shared memory

__device__ __forceinline__ uint4 ldg_stream(const uint4* p)
{
    uint4 r;
    asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
        : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ uint64_t warp_sum64(uint64_t v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(kFull, v, d);
    return v;
}

// Scan one 16 KiB segment of one payload with a full warp.  pay points at the payload's first
// byte (16-byte aligned).  Returns this lane's stripe digest (0 if it saw no block) and its
// token count.
// A warp's private staging ring for the cp.async.bulk variant: kTmaStages chunks of 2 KiB (one
// chunk = four warp-steps) and one mbarrier per chunk.
#ifndef PK_TMA_STAGES
#define PK_TMA_STAGES 4
#endif
#ifndef PK_TMA_MIN_CTAS
#define PK_TMA_MIN_CTAS 3
#endif
constexpr int      kTmaStages = PK_TMA_STAGES;
constexpr uint32_t kTmaChunk  = 32u * kUnroll * kBlockBytes;       // 2 KiB
struct TmaRing { uint8_t* stage; uint64_t* bar; };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tma_issue(const TmaRing& r, uint32_t s, const uint8_t* src)
{
    const uint32_t bar = smem_u32(r.bar + s), dst = smem_u32(r.stage + s * kTmaChunk);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(kTmaChunk) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(kTmaChunk), "r"(bar) : "memory");
}

__device__ __forceinline__ void tma_wait(const TmaRing& r, uint32_t s, uint32_t parity)
{
    const uint32_t bar = smem_u32(r.bar + s);
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}

template <bool kTma = false>
__device__ __forceinline__ void scan_segment(const uint8_t* __restrict__ pay, uint32_t len,
                                             uint32_t seg, uint32_t lane, uint32_t one,
                                             uint64_t& digest, uint32_t& tok, TmaRing ring = TmaRing{nullptr, nullptr})
{
    const uint32_t nblk   = (len >> 4) + ((len & 15u) ? 1u : 0u);
    const bool     tiny   = nblk <= kTinyBlocks;                    // SPEC v3: 8 stripes instead of 32 (see the remainder loop)
    const uint32_t first  = seg * kSegBlocks;
    const uint32_t nb     = min(kSegBlocks, nblk - first);          // blocks of this segment
    const uint32_t nwhole = min(kSegBlocks, (len >> 4) - first);    // ... lying wholly inside the payload
    const uint4*   base   = reinterpret_cast<const uint4*>(pay) + first + lane;

    Stripe st;
    stripe_init(st, seg * kStripes + lane);
    uint32_t tok128 = 0;
    uint32_t b0 = 0;

    // main loop: groups of four whole warp-steps (2 KiB), no predicates
    const uint32_t ngroups = nwhole / (32u * kUnroll);
    auto load_group = [&](uint4 (&x)[kUnroll], uint32_t g) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) x[u] = ldg_stream(base + g * (32u * kUnroll) + u * 32);
    };
    auto consume_group = [&](const uint4 (&x)[kUnroll]) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const Words4 w = {{x[u].x, x[u].y, x[u].z, x[u].w}};
            stripe_round(st, w);
            scan_block(w, one, tok128);
        }
    };
    if (kTma) {
        // chunks of 2 KiB land in the warp's ring by bulk async copy; lane 0 keeps kTmaStages of
        // them in flight, every lane then reads its 16-byte block of each step with LDS.128
        const uint8_t* src = pay + (size_t)first * kBlockBytes;
        if (lane == 0)
            for (uint32_t c = 0; c < ngroups && c < (uint32_t)kTmaStages; ++c) tma_issue(ring, c, src + (size_t)c * kTmaChunk);
        for (uint32_t c = 0; c < ngroups; ++c) {
            const uint32_t s = c % kTmaStages;
            tma_wait(ring, s, (c / kTmaStages) & 1u);
            uint4 x[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u)
                x[u] = *reinterpret_cast<const uint4*>(ring.stage + s * kTmaChunk + u * 512 + lane * 16);
            consume_group(x);
            __syncwarp();                                   // every lane has read stage s
            if (lane == 0 && c + kTmaStages < ngroups) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic reads before async write
                tma_issue(ring, s, src + (size_t)(c + kTmaStages) * kTmaChunk);
            }
        }
    } else {
        for (uint32_t g = 0; g < ngroups; ++g) {
            uint4 x[kUnroll];
            load_group(x, g);
            consume_group(x);
        }
    }
    b0 = ngroups * (32u * kUnroll);
    // remainder: ragged steps, the payload's last (partial) block
    for (; b0 < nb; b0 += 32u * kUnroll) {
        uint4 x[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const uint32_t k = b0 + u * 32 + lane;
            x[u] = (k < nb) ? ldg_stream(base + b0 + u * 32) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            if (b0 + u * 32 >= nb) break;                                 // warp-uniform
            const uint32_t k = b0 + u * 32 + lane;
            Words4 w = {{x[u].x, x[u].y, x[u].z, x[u].w}};
            if (k < nb) {
                const uint32_t valid = len - (first + k) * kBlockBytes;   // >= 1
                if (valid < kBlockBytes) mask_tail(w, valid);
                if (!tiny) stripe_round(st, w);
            }
            if (tiny) {                          // warp-uniform; only u == 0 gets here (nb <= 16)
                // SPEC v3: a payload of at most 16 blocks has 8 stripes — lane l < 8 folds block l and
                // then block l + 8, which lane l + 8 loaded and masked (the words move in place, after
                // the lane has counted its own block's tokens: no extra registers)
                if (lane < kTinyStripes && lane < nb) stripe_round(st, w);
                scan_block(w, one, tok128);
#pragma unroll
                for (int j = 0; j < 4; ++j) w.w[j] = __shfl_down_sync(0xFFFFFFFFu, w.w[j], kTinyStripes);
                if (lane + kTinyStripes < nb && lane < kTinyStripes) stripe_round(st, w);
            } else {
                scan_block(w, one, tok128);      // an all-zero block counts nothing
            }
        }
    }
    digest = (lane < (tiny ? min(nb, kTinyStripes) : nb)) ? stripe_digest(st) : 0ull;
    tok    = tok128 >> 7;
}

// A payload of at most G * B blocks scanned by a group of G lanes: lane `lig` owns blocks
// lig, lig + G, ... (B of them at most), each block its own stripe (block k of a payload this
// short is stripe k).  x[] are the lane's (already loaded) blocks; blocks past the payload's last
// hold zeros and are skipped.
template <int G, int B>
__device__ __forceinline__ void scan_tiny(const uint4 (&x)[B], uint32_t len, uint32_t lig, uint32_t one,
                                          uint64_t& digest, uint32_t& tok)
{
    // SPEC v3: a payload of at most 16 blocks has kTinyStripes = 8 stripes, block k in stripe k mod 8.
    // This lane's blocks are lig, lig + G, ...: they fall into kPer = 8 / G stripes (two for four
    // lanes, one for eight), which the lane keeps and finalises — half as many digests as blocks
    // once a payload is longer than 128 bytes.
    static_assert(G == 4 || G == 8, "lane groups of 4 or 8: a stripe never spans two lanes");
    constexpr int kPer = (int)kTinyStripes / G;
    const uint32_t nblk = (len + 15u) >> 4;                            // <= G * B
    uint32_t tok128 = 0;
    Stripe st[kPer];
#pragma unroll
    for (int b = 0; b < B; ++b) {
        const uint32_t k = lig + (uint32_t)b * G;
        if (b < kPer) stripe_init(st[b], k);                           // (k < 8 here: the stripe's own number)
        if (k < nblk) {
            Words4 w = {{x[b].x, x[b].y, x[b].z, x[b].w}};
            const uint32_t valid = len - k * kBlockBytes;
            if (valid < kBlockBytes) mask_tail(w, valid);
            stripe_round(st[b % kPer], w);
            scan_block(w, one, tok128);
        }
    }
    digest = 0ull;
#pragma unroll
    for (int q = 0; q < kPer && q < B; ++q)
        if (lig + (uint32_t)q * G < nblk) digest += stripe_digest(st[q]);
    tok = tok128 >> 7;
}

// One copy of the sequential decision code per kernel (it is long; inlining it at every call
// site only bloats the instruction cache footprint).
__device__ __noinline__ void decide_call(const DevTable* tbl, const uint8_t* name, uint32_t name_len,
                                         uint32_t pay_len, uint64_t payload_sum, uint32_t tok,
                                         pk_decision* out)
{
    decide(tbl, name, name_len, pay_len, payload_sum, tok, out);
}

// Latency mode (small batches): the warp that scanned a request finalises it on the spot,
// lane-parallel — name bytes one per lane, the tool switch as a warp vote, the provider rule as
// a warp prefix sum — with no shared memory and no CTA barrier on the path.  Same results as
// decide(): the hash sum is commutative, so name terms can join the lanes' digests before the
// reduction, and the first lane whose running weight exceeds x is the first such provider.
//
// Everything that does not depend on the payload is done BEFORE the scan (warp_prologue), while
// the payload's first 2 KiB are on their way to L2: the name bytes are two dependent loads away,
// the table rows one each, and after the scan nothing but arithmetic and one store is left.
struct WarpPrologue {
    uint64_t name_digest;     // this lane's share of the name terms of the hash
    uint32_t nlen;
    uint32_t cls;             // tool class; bit 31 set = unknown tool (default branch)
    uint32_t max_tokens, weight, tool_mask;   // provider row `lane` (weight 0 past n_providers)
};

__device__ __forceinline__ WarpPrologue warp_prologue(const RouteParams& P, uint32_t noff, uint32_t nlen,
                                                      uint32_t lane)
{
    const DevTable* __restrict__ tbl = P.table;
    const uint8_t* __restrict__ nm = P.names + noff;
    WarpPrologue w;
    w.nlen = nlen;
    // independent table loads first, so they are in flight together with the name bytes
    const uint32_t nt = __ldg(&tbl->t.n_tools);
    const uint32_t np = __ldg(&tbl->t.n_providers);
    uint32_t cls = __ldg(&tbl->t.default_class);
    uint4 row = make_uint4(0u, 0u, 0u, 0u);
    if (lane < np) row = __ldg(reinterpret_cast<const uint4*>(&tbl->t.providers[lane]));
    const uint32_t my_tool_len = lane < nt ? __ldg(&tbl->t.tools[lane].name_len) : 0xFFFFFFFFu;
    // lane j holds word j of the name (bytes 4j .. 4j+3, zero padded): the same register feeds the
    // hash term and the comparison with the table's (word-aligned, zero-padded) names
    const uint32_t nwords = (nlen + 3u) >> 2;
    uint32_t wj = 0u;
    if (lane < nwords) {
#pragma unroll
        for (uint32_t k = 0; k < 4; ++k)
            if (4u * lane + k < nlen) wj |= (uint32_t)__ldg(nm + 4u * lane + k) << (8u * k);
    }
    w.max_tokens = row.x; w.weight = row.y; w.tool_mask = row.z;

    uint64_t d = 0;
    if (lane < nwords) d = name_term(lane, wj);
    for (uint32_t j = lane + 32u; j < nwords; j += 32) d += name_term(j, name_word(nm, nlen, j));   // names over 128 bytes
    w.name_digest = d;

    // tool switch (mock.go:32-64): whole-name byte equality, first match wins.  Lane t holds
    // entry t's length; candidates are the entries of the request's length, in table order
    // (a table name has at most 64 bytes = 16 words, so lanes 0..15 decide).
    uint32_t unknown = 0x80000000u;
    uint32_t cand = __ballot_sync(kFull, my_tool_len == nlen);
    while (cand != 0u) {
        const uint32_t t = (uint32_t)__ffs((int)cand) - 1u;
        cand &= cand - 1u;
        bool ok = true;
        if (lane < nwords && lane < PK_MAX_TOOL_NAME / 4u)
            ok = __ldg(reinterpret_cast<const uint32_t*>(tbl->t.tools[t].name) + lane) == wj;
        if (__all_sync(kFull, ok)) { cls = __ldg(&tbl->t.tools[t].class_id); unknown = 0u; break; }
    }
    w.cls = cls | unknown;
    return w;
}

__device__ __forceinline__ void finalize_warp(const RouteParams& P, uint32_t req, uint32_t len,
                                              uint64_t digest, uint32_t tok_lane, uint32_t lane,
                                              const WarpPrologue& w)
{
    const uint64_t total = warp_sum64(digest + w.name_digest);
    const uint32_t tok = __reduce_add_sync(kFull, tok_lane);
    const uint64_t h = route_hash_final(total, len, w.nlen);
    const uint32_t cls = w.cls & 0xFFu;
    const uint32_t unknown = w.cls >> 31;

    // provider: lane p holds provider p's eligible weight; inclusive prefix sum over the warp
    const bool elig = w.weight != 0u && ((w.tool_mask >> cls) & 1u) != 0u && tok <= w.max_tokens;
    uint32_t cum = elig ? w.weight : 0u;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(kFull, cum, d);
        if (lane >= (uint32_t)d) cum += up;
    }
    const uint32_t wtotal = __shfl_sync(kFull, cum, 31);
    uint32_t prov = PK_PROVIDER_NONE;
    if (wtotal != 0u) {
        const uint32_t x = __umulhi((uint32_t)(h >> 32), wtotal);
        prov = (uint32_t)__ffs((int)__ballot_sync(kFull, cum > x)) - 1u;
    }
    const uint32_t flags = (prov == PK_PROVIDER_NONE ? PK_FLAG_NO_PROVIDER : 0u) |
                           (len == 0u ? PK_FLAG_EMPTY_PAYLOAD : 0u) | (unknown ? PK_FLAG_UNKNOWN_TOOL : 0u);
    if (lane < 3) {                                // the 24-byte record: one coalesced store
        uint64_t word;
        if (lane == 0)      word = (uint64_t)len | ((uint64_t)tok << 32);
        else if (lane == 1) word = h;
        else                word = (uint64_t)(cls & 0xFFu) | ((uint64_t)(prov & 0xFFu) << 8) |
                                   ((uint64_t)flags << 16) | (200ull << 32);   // Status.Code 200, mock.go:26
        reinterpret_cast<uint64_t*>(P.out)[(size_t)req * 3 + lane] = word;
    }
}

// Fold one segment of a long request into its scratch entry; the warp that arrives last reads
// the totals (resetting the entry to zero for the next launch) and lane 0 finalises.
__device__ __forceinline__ void fold_long_segment(const RouteParams& P, uint32_t req, uint32_t len,
                                                  uint64_t digest, uint32_t tok, uint32_t lane)
{
    const uint32_t nseg = (len + kSegBytes - 1u) / kSegBytes;
    const uint64_t sum  = warp_sum64(digest);
    const uint32_t tsum = __reduce_add_sync(kFull, tok);
    if (lane != 0) return;
    SegScratch* s = P.scratch + req;
    atomicAdd(&s->hash_sum, (unsigned long long)sum);
    atomicAdd(&s->tok, tsum);
    __threadfence();
    if (atomicAdd(&s->arrived, 1u) != nseg - 1u) return;
    __threadfence();
    const uint64_t total = atomicExch(&s->hash_sum, 0ull);
    const uint32_t ttok  = atomicExch(&s->tok, 0u);
    atomicExch(&s->arrived, 0u);
    const uint32_t noff = P.name_off[req];
    decide_call(P.table, P.names + noff, P.name_off[req + 1] - noff, len, total, ttok, P.out + req);
}

// T is the routing table as a kernel parameter (constant bank: every thread reads the same rows, so
// its loads are uniform LDCs that cost no shared memory, no staging and no LSU slot); P.table is
// the same table in global memory for the few places that index it per lane.
template <int G, bool kWarpFinal, int B>
__device__ __forceinline__ void route_body(const RouteParams& P, const DevTable& T)
{
    static_assert(!kWarpFinal || G == 32, "latency mode finalises with the full warp");
    static_assert(G < 32 || B == 1, "B: blocks per lane of the lane-group kernels");
    constexpr uint32_t kRmax = G == 32 ? 64u : G == 16 ? 128u : 256u;   // launch_route's r_max
    static_assert(kRmax <= (uint32_t)kThreads, "one thread per request stages its offsets and finalises it");
    __shared__ uint64_t psum[kPartialSlots];     // [k][R+1]: lane-in-group k's digest for request i
    __shared__ uint16_t ptok[kPartialSlots];      // a lane sees at most 512 bytes of a segment: its count fits 16 bits
    __shared__ __align__(16) uint8_t s_names[kNameStage + 8];   // this CTA's tool names, when they fit (+8: decide reads whole words)
    __shared__ uint32_t s_noff[kRmax + 1];       // this CTA's name offsets
    __shared__ uint64_t s_meta[G < 32 ? kRmax : 1u];   // lane-group kernels: (payload offset / 16) << 12 | length, per request

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t wid  = threadIdx.x >> 5;
    const uint32_t R    = P.reqs_per_cta;
    const uint32_t n_req_ctas = P.n_req_ctas;                   // ceil(n / R), computed on the host

    if (G == 32 && blockIdx.x >= n_req_ctas) {       // (a lane-group batch has no request longer than 256 bytes)
        // ---- extra segments (>= 1) of long requests: one warp per (request, segment) ----
        const uint32_t task = (blockIdx.x - n_req_ctas) * kWarpsPerCta + wid;
        if (task >= P.n_extra) return;
        const uint2 e = __ldg(P.extra + task);
        const uint32_t len = __ldg(P.pay_len + e.x);
        uint64_t digest;
        uint32_t tok;
        scan_segment(P.payload + __ldg(P.pay_off + e.x), len, e.y, lane, P.one, digest, tok);
        fold_long_segment(P, e.x, len, digest, tok, lane);
        return;
    }

    const uint32_t base_req = blockIdx.x * R;
    const uint32_t r_here   = min(R, P.n - base_req);

    if (kWarpFinal) {
        // ---- latency mode: scan and finalise per warp; nothing shared, no barrier ----
        for (uint32_t i = wid; i < r_here; i += kWarpsPerCta) {
            const uint32_t req = base_req + i;
            const uint32_t len = __ldg(P.pay_len + req);
            const uint64_t off = __ldg(P.pay_off + req);
            const uint32_t noff = __ldg(P.name_off + req);
            const uint32_t nlen = __ldg(P.name_off + req + 1) - noff;
            // pull the head of the payload towards L2, then do all payload-independent work
            // (name bytes, tool switch, provider row) while that HBM fetch is under way
            if (lane < 16u && lane * 128u < len)
                asm volatile("prefetch.global.L2 [%0];" :: "l"(P.payload + off + lane * 128u));
            const WarpPrologue w = warp_prologue(P, noff, nlen, lane);
            uint64_t digest;
            uint32_t tok;
            scan_segment(P.payload + off, len, 0u, lane, P.one, digest, tok);
            if (len > kSegBytes) fold_long_segment(P, req, len, digest, tok, lane);
            else                 finalize_warp(P, req, len, digest, tok, lane, w);
        }
        return;
    }

    const uint32_t ld       = R + 1u;            // odd leading dimension: conflict-free columns
    constexpr uint32_t kGroups = 32u / G;
    const uint32_t grp = lane / G, lig = lane % G;

    // ---- prologue.  Phase 2 reads the name offsets and the tool names out of shared memory; the
    //      offsets are requested here (consumed after the barrier that follows the scan), the name
    //      bytes after the scan, so no warp stalls on a staging round trip before its first payload
    //      load is out. ----
    if (threadIdx.x < r_here) s_noff[threadIdx.x + 1] = __ldg(P.name_off + base_req + threadIdx.x + 1);
    if (threadIdx.x == 0) s_noff[0] = __ldg(P.name_off + base_req);
    if (G < 32) {
        // lane-group kernels: every request's length and offset in one coalesced sweep, so the scan
        // loop below reads them with one LDS instead of two dependent global loads and 64-bit
        // address arithmetic per request (every payload here is at most 256 bytes: 12 bits)
        if (threadIdx.x < r_here)
            s_meta[threadIdx.x] = ((__ldg(P.pay_off + base_req + threadIdx.x) >> 4) << 12) |
                                  (uint64_t)__ldg(P.pay_len + base_req + threadIdx.x);
        __syncthreads();
    }

    // ---- phase 1: scan (first segment of) each request ----
    if constexpr (G == 32) {
        // the next request's length and offset are loaded while this one is scanned
        uint32_t len_next = 0u;
        uint64_t off_next = 0ull;
        if (wid < r_here) {
            len_next = __ldg(P.pay_len + base_req + wid);
            off_next = __ldg(P.pay_off + base_req + wid);
        }
        for (uint32_t i = wid; i < r_here; i += kWarpsPerCta) {
            const uint32_t req = base_req + i;
            const uint32_t len = len_next;
            const uint8_t* pay = P.payload + off_next;
            if (i + kWarpsPerCta < r_here) {
                len_next = __ldg(P.pay_len + req + kWarpsPerCta);
                off_next = __ldg(P.pay_off + req + kWarpsPerCta);
            }
            uint64_t digest;
            uint32_t tok;
            scan_segment(pay, len, 0u, lane, P.one, digest, tok);
            if (len > kSegBytes) {               // warp-uniform: the whole warp owns this request
                fold_long_segment(P, req, len, digest, tok, lane);
                continue;
            }
            psum[lane * ld + i] = digest;
            ptok[lane * ld + i] = (uint16_t)tok;
        }
    } else {
        // tiny payloads: B blocks per lane (all loaded before any is used), one request per lane
        // group per pass.  Occupancy (8 CTAs per SM) hides the rest of the payload loads' latency.
        constexpr uint32_t kStride = kWarpsPerCta * kGroups;
        const uint4* const arena = reinterpret_cast<const uint4*>(P.payload);
#pragma unroll 1
        for (uint32_t i = wid * kGroups + grp; i < r_here; i += kStride) {
            const uint64_t m = s_meta[i];
            const uint32_t len = (uint32_t)m & 0xFFFu;
            const uint32_t nblk = (len + 15u) >> 4;
            const uint4* const src = arena + (m >> 12) + lig;
            uint4 x[B];
#pragma unroll
            for (int b = 0; b < B; ++b) {
                x[b] = make_uint4(0u, 0u, 0u, 0u);
                if (lig + (uint32_t)b * G < nblk) x[b] = ldg_stream(src + b * G);
            }
            uint64_t digest;
            uint32_t tok;
            scan_tiny<G, B>(x, len, lig, P.one, digest, tok);
            psum[lig * ld + i] = digest;
            ptok[lig * ld + i] = (uint16_t)tok;
        }
    }
    // ---- staging for phase 2 (its load latency hides behind the warps still scanning): this CTA's
    //      tool names as aligned words ----
    const uint32_t names_lo = __ldg(P.name_off + base_req), names_hi = __ldg(P.name_off + base_req + r_here);
    const uint32_t names_sz = names_hi - names_lo;
    const uint8_t* names_src = P.names + names_lo;
    const uint32_t names_head = (uint32_t)(reinterpret_cast<uintptr_t>(names_src) & 3u);   // s_names[names_head] = first name byte
    const bool names_staged = names_sz + 3u <= kNameStage;
    if (names_staged) {
        // whole aligned words, including the up to three bytes before the first and after the last
        // name that share a word with them (same word, same page: always readable, never used)
        const uint32_t* al = reinterpret_cast<const uint32_t*>(names_src - names_head);
        const uint32_t nw = (names_head + names_sz + 3u) >> 2;
        for (uint32_t w = threadIdx.x; w < nw; w += kThreads) reinterpret_cast<uint32_t*>(s_names)[w] = __ldg(al + w);
    }
    __syncthreads();

    // ---- phase 2: one thread per request ----
    if (const uint32_t t = threadIdx.x; t < r_here) {
        const uint32_t req = base_req + t;
        const uint32_t len = G < 32 ? ((uint32_t)s_meta[t] & 0xFFFu) : __ldg(P.pay_len + req);
        if (G == 32 && len > kSegBytes) return;          // finalised by its last segment
        uint64_t sum = 0;
        uint32_t tok = 0;
#pragma unroll
        for (int k = 0; k < G; ++k) {
            sum += psum[k * ld + t];
            tok += ptok[k * ld + t];
        }
        const uint32_t noff = s_noff[t];
        const uint32_t nlen = s_noff[t + 1] - noff;
        // names in shared memory (the usual case): decide() inlined, the name read as whole words,
        // the table out of the constant bank
        if (names_staged) decide<true>(&T, s_names + names_head + (noff - names_lo), nlen, len, sum, tok, P.out + req);
        else              decide_call(P.table, P.names + noff, nlen, len, sum, tok, P.out + req);
    }
}

// Programmatic dependent launch (P.pdl).  Batches are independent — a launch reads its own staged
// arenas and the table and writes its own records — so the next kernel of the stream may start
// while this one is still running: every CTA releases its dependents at once, and the launch
// latency, the two dependent HBM round trips (offsets, then bytes) and the tail of one batch
// overlap the body of its neighbours.  The wait for the predecessor is the LAST instruction,
// which keeps stream order for whatever follows (a copy out after kernel k sees kernels <= k
// complete) without holding anything back.  What two launches of the SAME slot share is that
// slot's long-request scratch, so launch_route never sets pdl for a batch with long requests —
// those are bandwidth-bound launches of hundreds of µs that have nothing to gain from overlap.
template <int G, bool kWarpFinal, int B>
__global__ void __launch_bounds__(kThreads, G == 32 ? PK_MIN_CTAS : B >= 4 ? PK_MIN_CTAS_TINY4 : PK_MIN_CTAS_TINY)
route_kernel(const RouteParams P, const __grid_constant__ DevTable T)
{
    if (P.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    route_body<G, kWarpFinal, B>(P, T);
    if (P.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
}

// Several batches per launch: CTA (x, y) is CTA x of batch y (FusedParams; launch_route_fused).
template <int G, bool kWarpFinal, int B>
__global__ void __launch_bounds__(kThreads, G == 32 ? PK_MIN_CTAS : B >= 4 ? PK_MIN_CTAS_TINY4 : PK_MIN_CTAS_TINY)
route_fused_kernel(const __grid_constant__ FusedParams M, const __grid_constant__ DevTable T)
{
    if (M.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const RouteParams& P = M.b[blockIdx.y];
    // the grid is as wide as the widest batch needs
    const uint32_t ctas = P.n_req_ctas + (P.n_extra + kWarpsPerCta - 1u) / kWarpsPerCta;
    if (blockIdx.x < ctas) route_body<G, kWarpFinal, B>(P, T);
    if (M.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
}

#if PK_TMA
// Segments >= 1 of long requests, one warp per (request, segment), the bytes staged through a
// per-warp shared-memory ring by cp.async.bulk (SASS: UBLKCP) and consumed with LDS.128.
constexpr size_t kTmaSmemBytes = (size_t)kWarpsPerCta * kTmaStages * kTmaChunk + (size_t)kWarpsPerCta * kTmaStages * 8;

__global__ void __launch_bounds__(kThreads, PK_TMA_MIN_CTAS) segment_kernel_tma(const RouteParams P)
{
    extern __shared__ __align__(128) uint8_t tma_smem[];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t wid  = threadIdx.x >> 5;
    TmaRing ring;
    ring.stage = tma_smem + (size_t)wid * kTmaStages * kTmaChunk;
    ring.bar   = reinterpret_cast<uint64_t*>(tma_smem + (size_t)kWarpsPerCta * kTmaStages * kTmaChunk) + wid * kTmaStages;
    if (lane == 0) {
        for (int s = 0; s < kTmaStages; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(ring.bar + s)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    const uint32_t task = blockIdx.x * kWarpsPerCta + wid;
    if (task >= P.n_extra) return;
    const uint2 e = __ldg(P.extra + task);
    const uint32_t len = __ldg(P.pay_len + e.x);
    uint64_t digest;
    uint32_t tok;
    scan_segment<true>(P.payload + __ldg(P.pay_off + e.x), len, e.y, lane, P.one, digest, tok, ring);
    fold_long_segment(P, e.x, len, digest, tok, lane);
}
#endif

uint32_t pow2_floor(uint32_t x)
{
    uint32_t p = 1;
    while (p * 2u <= x) p *= 2u;
    return p;
}

}  // namespace

namespace {
template <int G, bool kWarpFinal, int B = 1>
cudaError_t launch_one(const RouteParams& p, const DevTable& table, uint32_t ctas, cudaStream_t stream)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(ctas);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = p.pdl ? 1u : 0u;
    return cudaLaunchKernelEx(&cfg, route_kernel<G, kWarpFinal, B>, p, table);
}
}  // namespace

namespace {
struct Shape {
    int G, B;
    bool warp_final;
    uint32_t ctas;
};

// Kernel shape and requests per CTA for one batch (sets p.reqs_per_cta, p.one, p.pdl).  `n_total`:
// the requests of the whole launch (the batch itself, or every batch of a fused launch) — what the
// "enough CTAs to fill the machine" rule counts.
Shape choose_shape(RouteParams& p, uint32_t max_len, int num_sms, bool& pdl, uint64_t n_total)
{
    pdl = pdl && p.n_extra == 0u;       // long requests share their slot's scratch between launches: no overlap
    // lanes per request: a full warp, or a half / quarter warp when every payload fits one step
    // Payloads of at most 256 bytes (16 blocks): FOUR lanes per request, each taking 1, 2 or 4 blocks
    // (all loaded before any is used), so a warp pass covers eight requests.  Against 16 lanes x 1
    // block (round 1) and 8 x 2 the loop overhead per request and the partial sums phase 2 adds up
    // halve each time; the four loads in flight per lane cost registers (40: six CTAs per SM instead
    // of eight), which the measurement pays for: 65,536 x 256 B under the bench's issue 5.01 µs per
    // batch at 4 x 4 against 5.15 at 8 x 2 (PK_TINY_SHAPE=8 selects the older shape).
    static const char* const t_env = std::getenv("PK_TINY_SHAPE");
    const int t_shape = t_env ? std::atoi(t_env) : 4;
    int G = 32, B = 1;
    if (max_len <= 16u * kBlockBytes) {
        const uint32_t nblk_max = (max_len + kBlockBytes - 1u) / kBlockBytes;
        G = 4; B = nblk_max <= 4u ? 1 : nblk_max <= 8u ? 2 : 4;
        if (t_shape == 8) { G = 8; B = nblk_max <= 8u ? 1 : 2; }   // (16 lanes x 1 block went with SPEC v3: a stripe would span two lanes)
    }
    // Few tiny requests: a warp each (latency mode below) beats packing them 2 or 4 to a warp —
    // the packed path pays a CTA barrier and a second phase that only pay off when warps are scarce.
    static const char* const g_env = std::getenv("PK_TINY_WARP_MAX");      // tuning aid: largest such batch
    const uint32_t tiny_warp_max = g_env ? (uint32_t)std::atoi(g_env) : kTinyWarpMax;
    if (G < 32 && !pdl && p.n <= tiny_warp_max) G = 32;
    const uint32_t r_min = kWarpsPerCta * (32u / G);          // one request per lane group per pass
    const uint32_t r_max = G == 32 ? 64u : G == 16 ? 128u : 256u;   // G * (R + 1) <= kPartialSlots
    // enough CTAs to fill the machine (4 per SM resident), then as many requests per CTA as
    // possible so that phase 2 runs with full warps.  (Sizing R so that the grid is a whole number
    // of waves of resident CTAs — 20 fused batches of 4,096 are 1.73 waves at R = 64, two full ones
    // at R = 56 — was measured and changes nothing at K = 20, and costs 3 % in the steady state and
    // on 256-byte payloads: the kernel is issue-bound, a thinner last wave simply runs faster.)
    uint32_t R = pow2_floor((uint32_t)std::min<uint64_t>(n_total / (uint64_t)(num_sms * 4) + 1u, 1u << 20));
    // A stream of batches issued with programmatic dependent launch wants the opposite shape: few,
    // fat CTAs per batch, so that many batches are resident at once and one's latency chain hides
    // behind the others' streaming (measured, C1: 1.94 µs per batch at R = 64 against 2.86 at
    // R = 8; alone, the R = 64 batch takes 15 µs).
    // No more than 256 KiB of payload per CTA, so long payloads keep the thin shape.
    // (A fused launch already holds many batches: it keeps the fill-the-machine shape, the same R
    // for every batch in it.)
    if (pdl && n_total == p.n) R = std::min(pow2_floor(p.n / 16u + 1u), pow2_floor((256u << 10) / (max_len ? max_len : 1u) + 1u));
    static const char* const r_env = std::getenv("PK_REQS_PER_CTA");     // tuning aid
    if (r_env && std::atoi(r_env) > 0) R = ((uint32_t)std::atoi(r_env) + 3u) & ~3u;
    R = R < r_min ? r_min : R > r_max ? r_max : R;
    p.reqs_per_cta = R;
    p.n_req_ctas = (p.n + R - 1u) / R;
    p.one = 1u;
    p.pdl = pdl ? 1u : 0u;
    const uint32_t ctas = p.n_req_ctas + (p.n_extra + kWarpsPerCta - 1u) / kWarpsPerCta;
    // latency mode while one request per warp already fills the machine's resident CTA slots
    static const char* const m_env = std::getenv("PK_LATENCY_MODE");     // tuning aid: 0 / 1 forces
    bool warp_final = G == 32 && R == r_min;
    if (m_env && G == 32) warp_final = std::atoi(m_env) != 0;
    return Shape{G, B, warp_final, ctas};
}
}  // namespace

cudaError_t launch_route(RouteParams p, const DevTable& table, uint32_t max_len, int num_sms, cudaStream_t stream, bool pdl)
{
    if (p.n == 0) return cudaSuccess;
#if PK_TMA
    const uint32_t n_extra_tma = p.n_extra;
    if (n_extra_tma) pdl = false;                   // (as choose_shape would have decided)
    p.n_extra = 0;                                  // the main kernel scans first segments only
#endif
    const Shape sh = choose_shape(p, max_len, num_sms, pdl, p.n);
    const int G = sh.G, B = sh.B;
    const bool warp_final = sh.warp_final;
    const uint32_t ctas = sh.ctas;
    cudaError_t le;
    switch (G) {
    case 8:  le = B == 2 ? launch_one<8, false, 2>(p, table, ctas, stream) : launch_one<8, false, 1>(p, table, ctas, stream); break;
    case 4:  le = B == 4 ? launch_one<4, false, 4>(p, table, ctas, stream)
                : B == 2 ? launch_one<4, false, 2>(p, table, ctas, stream) : launch_one<4, false, 1>(p, table, ctas, stream); break;
    default: le = warp_final ? launch_one<32, true>(p, table, ctas, stream) : launch_one<32, false>(p, table, ctas, stream); break;
    }
    if (le != cudaSuccess) return le;
#if PK_TMA
    if (n_extra_tma != 0) {
        static const cudaError_t attr = cudaFuncSetAttribute(segment_kernel_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTmaSmemBytes);
        if (attr != cudaSuccess) return attr;
        p.n_extra = n_extra_tma;
        segment_kernel_tma<<<(n_extra_tma + kWarpsPerCta - 1u) / kWarpsPerCta, kThreads, kTmaSmemBytes, stream>>>(p);
    }
#endif
    return cudaGetLastError();
}

namespace {
template <int G, bool kWarpFinal, int B = 1>
cudaError_t launch_fused(const FusedParams& f, const DevTable& table, uint32_t ctas, cudaStream_t stream)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(ctas, f.m);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = f.pdl ? 1u : 0u;
    return cudaLaunchKernelEx(&cfg, route_fused_kernel<G, kWarpFinal, B>, f, table);
}
}  // namespace

cudaError_t launch_route_fused(const RouteParams* batches, const uint32_t* max_len, uint32_t m, const DevTable& table,
                               int num_sms, cudaStream_t stream, bool pdl, bool* launched)
{
    *launched = false;
#if PK_TMA
    return cudaSuccess;                             // (the TMA variant keeps its two-kernel form)
#endif
    if (m == 0 || m > kMaxFused) return cudaSuccess;
    FusedParams f{};
    uint64_t n_total = 0;
    for (uint32_t k = 0; k < m; ++k) {
        if (batches[k].n == 0) return cudaSuccess;
        n_total += batches[k].n;
        if (batches[k].n_extra) pdl = false;        // one batch with long requests: the launch waits for its predecessor
    }
    Shape sh{};
    uint32_t ctas = 0;
    for (uint32_t k = 0; k < m; ++k) {
        f.b[k] = batches[k];
        bool pdl_k = pdl;
        const Shape s = choose_shape(f.b[k], max_len[k], num_sms, pdl_k, n_total);
        if (k == 0) sh = s;
        else if (s.G != sh.G || s.B != sh.B || s.warp_final != sh.warp_final) return cudaSuccess;
        ctas = std::max(ctas, s.ctas);
    }
    f.m = m;
    f.pdl = pdl ? 1u : 0u;
    cudaError_t le;
    switch (sh.G) {
    case 8:  le = sh.B == 2 ? launch_fused<8, false, 2>(f, table, ctas, stream) : launch_fused<8, false, 1>(f, table, ctas, stream); break;
    case 4:  le = sh.B == 4 ? launch_fused<4, false, 4>(f, table, ctas, stream)
                : sh.B == 2 ? launch_fused<4, false, 2>(f, table, ctas, stream) : launch_fused<4, false, 1>(f, table, ctas, stream); break;
    default: le = sh.warp_final ? launch_fused<32, true>(f, table, ctas, stream) : launch_fused<32, false>(f, table, ctas, stream); break;
    }
    if (le != cudaSuccess) return le;
    *launched = true;
    return cudaGetLastError();
}

}  // namespace pk

// pk_gcm.cu — batched AES-256-GCM on the GPU and its C ABI (declared in include/pk_gcm.h).
//
// The batched form of BatchEncrypt / BatchDecrypt, internal/adapters/security/cipher.go:110-141:
// the reference walks the batch sequentially and rebuilds the AES key schedule and the GCM state
// for every message (cipher.go:35-43 inside the loops at :116-122 and :133-139).  Here the key
// schedule, the T-table and the GHASH table are built once per key on the host; one GPU thread
// then seals or opens one message (messages here are secrets: tens of bytes to a few KB; the
// longer ones get a warp each, gcm_long_kernel), with
// the key's tables staged in shared memory (Te0 and Te2 replicated per bank).  Integer/LDS-bound work —
// no tensor cores, and not HBM-bound.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <new>

#include "../../include/pk_gcm.h"
#include "pk_gcm.cuh"

namespace {

using pkgcm::KeyContext;
constexpr int kThreads = 256;          // largest block; small batches use narrower blocks (see run())

// Shared memory.  gcm_kernel (one thread per message; latency-bound, wants occupancy): static,
// 41 KB -> 5 CTAs of 256 threads per SM
//   the key's round keys and GHASH tables (pkgcm::GcmTables, 8.8 KB) and Te0 replicated once per
//   bank (32 KB: entry x of lane l at word 32·x + l), so the AES lookups of a warp never collide.
// gcm_long_kernel (one warp per message): dynamic, 104.5 KB -> 2 CTAs per SM (of up to 16 warps)
//   [0, 64 KB)   Te0 | Te2 replicated per bank, pkgcm::TeDual's layout
//   then         the round keys (GcmTables::rk), the round keys rotated left by 8 (TeDual::rkr),
//                the 4-bit tables of H^1..H^32 (KeyContext::pw4, 8 KB: a lane closes with its own
//                power, once per message — bit-serial, that product was a quarter of the kernel's
//                instructions on 4 KiB messages, r02 ncu)
//   then         the table of H^32 — the lanes' Horner step — replicated once per 16-byte bank
//                group, as in the lane-group kernel below (32 KB; unreplicated, 39M of the
//                kernel's 133M shared wavefronts were bank conflicts on it, r02 ncu)
constexpr uint32_t kTabBytes = 256u * 256u;
constexpr uint32_t kLongRkAt = kTabBytes, kLongRkrAt = kLongRkAt + 256u, kLongPwAt = kLongRkrAt + 256u;
constexpr uint32_t kLongHrepAt = kLongPwAt + 32u * 16u * (uint32_t)sizeof(pkgcm::HEntry);
constexpr uint32_t kSmemBytes = kLongHrepAt + 8u * 256u * (uint32_t)sizeof(pkgcm::HEntry);
constexpr uint32_t kLongCtasPerSm = 2;
constexpr int kLongThreads = 512;      // largest block of gcm_long_kernel
static_assert(kLongHrepAt % 128 == 0, "long-kernel shared layout");

__device__ __forceinline__ void stage_key(pkgcm::GcmTables* kc, const KeyContext* __restrict__ kc_g)
{
    for (uint32_t i = threadIdx.x; i < sizeof(pkgcm::GcmTables) / 16; i += blockDim.x)
        reinterpret_cast<uint4*>(kc)[i] = __ldg(reinterpret_cast<const uint4*>(&kc_g->g) + i);
}

__device__ __forceinline__ void stage_single(pkgcm::GcmTables& kc, uint32_t* te_rep, const KeyContext* __restrict__ kc_g)
{
    stage_key(&kc, kc_g);
    for (uint32_t x = threadIdx.x; x < 256; x += blockDim.x) {
        const uint32_t t = __ldg(&kc_g->te[0][x]);
        const uint4 t4 = make_uint4(t, t, t, t);
#pragma unroll
        for (int q = 0; q < 8; ++q) reinterpret_cast<uint4*>(te_rep + x * 32)[q] = t4;
    }
    __syncthreads();
}

// mode 0: seal   in = plaintext arena,  out = sealed arena (nonce || ct || tag per message)
// mode 1: open   in = sealed arena,     out = plaintext arena; status[i] = 0 ok / 1 auth / 2 short
__global__ void __launch_bounds__(kThreads) gcm_kernel(const KeyContext* __restrict__ kc_g,
                                                       const uint8_t* __restrict__ in, const uint64_t* __restrict__ in_off,
                                                       const uint32_t* __restrict__ in_len, const uint8_t* __restrict__ nonces,
                                                       uint8_t* __restrict__ out, const uint64_t* __restrict__ out_off,
                                                       uint32_t* __restrict__ status, uint32_t first, uint32_t n, int mode,
                                                       uint32_t long_bytes)
{
    __shared__ pkgcm::GcmTables kc_s;
    __shared__ uint32_t te_rep[256 * 32];
    stage_single(kc_s, te_rep, kc_g);
    const pkgcm::GcmTables* kc = &kc_s;
    const pkgcm::TeReplicated te{te_rep, threadIdx.x & 31u};

    const uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;      // messages [first, n)
    if (i >= n) return;
    const uint32_t len = in_len[i];
    if (len >= long_bytes) return;                        // gcm_long_kernel's message
    const uint8_t* src = in + in_off[i];
    uint8_t* dst = out + out_off[i];
    uint8_t tag[16];
    if (mode == 0) {
        const uint8_t* nonce = nonces + (size_t)i * pkgcm::kNonceBytes;
        for (uint32_t b = 0; b < pkgcm::kNonceBytes; ++b) dst[b] = nonce[b];           // cipher.go:51-52
        pkgcm::gcm_crypt_message(kc, te, nonce, src, dst + pkgcm::kNonceBytes, len, false, tag);
        for (uint32_t b = 0; b < pkgcm::kTagBytes; ++b) dst[pkgcm::kNonceBytes + len + b] = tag[b];
    } else {
        if (len < pkgcm::kOverhead) {                     // shorter than nonce + tag: cipher.go:69-71 / Open fails
            status[i] = len < pkgcm::kNonceBytes ? 2u : 1u;
            return;
        }
        const uint32_t clen = len - pkgcm::kOverhead;
        pkgcm::gcm_crypt_message(kc, te, src, src + pkgcm::kNonceBytes, dst, clen, true, tag);
        uint32_t diff = 0;
        for (uint32_t b = 0; b < pkgcm::kTagBytes; ++b) diff |= (uint32_t)(tag[b] ^ src[pkgcm::kNonceBytes + clen + b]);
        if (diff != 0) {                                  // authentication failed: release no plaintext
            for (uint32_t b = 0; b < clen; ++b) dst[b] = 0;
            status[i] = 1u;
        } else {
            status[i] = 0u;
        }
    }
}

// kGroupLanes lanes per SHORT message (the messages gcm_long_kernel does not take) — the default
// for them since round 2; gcm_kernel above (one thread per message) stays for comparison
// (PK_GCM_GROUP=0).  65,536 secrets are 65,536 threads there: 14 warps per SM, each on one serial
// chain of 17 dependent AES blocks and table walks; here they are four times as many threads on
// chains a quarter as long.  Its time goes to shared-memory wavefronts first (the AES and GHASH
// table lookups) and instruction issue second, so the layout is chosen for the lookups.  Dynamic
// shared memory, 108.5 KB -> 2 CTAs of 512 threads per SM (which is also what its registers allow):
//   [0, 64 KB)   Te0 | Te2 replicated per bank, pkgcm::TeDual's layout (conflict-free, and half the
//                ALU instructions per AES round of the single replicated table)
//   then         the head of pkgcm::GroupTables: round keys and the tables of H^1..H^(G-1), the
//                lanes' closing powers (12.2 KB; one multiplication in five for a 256-byte secret)
//   then         the round keys rotated left by 8 (TeDual::rkr)
//   then         the table of H^G — the Horner step, four multiplications in five — replicated once
//                per 16-byte bank group (32 KB: entry e of copy r at (8·e + r)·16, lane l reads copy
//                l mod 8): the eight lanes a 128-bit shared load serves per wavefront never collide.
//                Unreplicated, a lookup costs 10 wavefronts instead of 4 (r02 ncu: 7.0M of 19.8M
//                load wavefronts were bank conflicts, all of them here).
constexpr int kGroupThreads = 512;
constexpr uint32_t kGroupCtasPerSm = 2;
constexpr uint32_t kGroupHeadBytes = (uint32_t)(sizeof(pkgcm::GroupTables) - sizeof(pkgcm::HEntry) * 256u);
constexpr uint32_t kGroupRkrAt = kTabBytes + kGroupHeadBytes;
constexpr uint32_t kGroupHrepAt = kGroupRkrAt + 256u;
constexpr uint32_t kGroupSmemBytes = kGroupHrepAt + 8u * 256u * (uint32_t)sizeof(pkgcm::HEntry);
static_assert(kGroupHeadBytes % 16 == 0 && kGroupHrepAt % 128 == 0, "group-kernel shared layout");

// The GHASH tables of gcm_group_pass in the layout above.
struct GroupTablesSmem {
    const uint8_t* hp;          // tables of H^1..H^(G-1), 4 KB each, plain
    const uint8_t* hrep_lane;   // replicated table of H^G + this lane's copy offset
    __device__ __forceinline__ void horner(uint64_t& yh, uint64_t& yl) const { pkgcm::ghash_mul_smem(hrep_lane, 7u, yh, yl); }
    __device__ __forceinline__ void close(uint32_t e, uint64_t& yh, uint64_t& yl) const
    {
        const bool top = e == (uint32_t)pkgcm::kGroupLanes;      // (one lane of every group: no divergence)
        pkgcm::ghash_mul_smem(top ? hrep_lane : hp + (e - 1u) * 4096u, top ? 7u : 4u, yh, yl);
    }
};

__global__ void __launch_bounds__(kGroupThreads) gcm_group_kernel(const KeyContext* __restrict__ kc_g,
                                                                  const uint8_t* __restrict__ in, const uint64_t* __restrict__ in_off,
                                                                  const uint32_t* __restrict__ in_len, const uint8_t* __restrict__ nonces,
                                                                  uint8_t* __restrict__ out, const uint64_t* __restrict__ out_off,
                                                                  uint32_t* __restrict__ status, uint32_t first, uint32_t n, int mode,
                                                                  uint32_t long_bytes)
{
    constexpr uint32_t G = pkgcm::kGroupLanes;
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t* rk = reinterpret_cast<const uint32_t*>(smem + kTabBytes);       // GroupTables::rk
    uint32_t* rkr = reinterpret_cast<uint32_t*>(smem + kGroupRkrAt);
    for (uint32_t i = threadIdx.x; i < kGroupHeadBytes / 16; i += blockDim.x)
        reinterpret_cast<uint4*>(smem + kTabBytes)[i] = __ldg(reinterpret_cast<const uint4*>(&kc_g->gg) + i);
    for (uint32_t i = threadIdx.x; i < 60; i += blockDim.x) {
        const uint32_t k = __ldg(&kc_g->gg.rk[i]);
        rkr[i] = (k << 8) | (k >> 24);
    }
    // a thread fills a 128-byte row with eight copies of one value — a T-table word, or an entry of
    // the H^G table (row 8·e .. 8·e+7 of 16 bytes).  It starts at chunk (row mod 8), so the eight lanes
    // a 128-bit store serves per wavefront are on eight different bank groups (all starting at chunk 0
    // put every lane on the same banks: 1.8M of the kernel's 2.2M store wavefronts).
    // (One loop of this shape on purpose: with a second staging loop, or one store per thread and
    // iteration, ptxas keeps the shared base in a vector register instead of a uniform one and every
    // AES lookup below pays an extra add — 66 instead of 50 instructions per round.)
    for (uint32_t x = threadIdx.x; x < 768; x += blockDim.x) {       // x = 2·entry + table; from 512: entry of H^G
        uint4 t4;
        uint32_t at = x * 128u;
        if (x < 512) {
            const uint32_t t = __ldg(&kc_g->te[(x & 1u) * 2u][x >> 1]);
            t4 = make_uint4(t, t, t, t);
        } else {
            t4 = __ldg(reinterpret_cast<const uint4*>(&kc_g->gg.hp[G - 1][x - 512u]));
            at += kGroupHrepAt - 512u * 128u;
        }
#pragma unroll
        for (uint32_t q = 0; q < 8; ++q) reinterpret_cast<uint4*>(smem + at)[(q + x) & 7u] = t4;
    }
    __syncthreads();
    const GroupTablesSmem gt{smem + kTabBytes + 256u, smem + kGroupHrepAt + (threadIdx.x & 7u) * 16u};
    const pkgcm::TeDual te{smem, rkr, (threadIdx.x & 31u) * 4u};

    // The grid is resident (the host launches at most what fits on the SMs at once) and a warp walks
    // UNITS of 32 / G consecutive messages, unit u = k · (warps in the grid) + warp · gridDim + CTA:
    // warp-major, so when the units do not divide evenly every CTA — every SM — gets the same number
    // of second helpings.  (With one CTA per 128 messages, 65,536 secrets were 1.73 waves of CTAs:
    // the SMs sat idle for 18% of the kernel waiting for the last wave, r02 ncu sm__cycles_active.)
    // Every lane of a warp stays to the end of a unit: the group's lanes meet in shuffles.
    const uint32_t lane = threadIdx.x & 31u, lig = lane % G, lead = lane & ~(G - 1u);
    const uint32_t n_units = (n - first + 32u / G - 1u) / (32u / G);
    const uint32_t warp = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);     // (uniform by construction: see gcm_long_kernel)
    for (uint32_t u = warp * gridDim.x + blockIdx.x; u < n_units; u += (blockDim.x >> 5) * gridDim.x) {
        const uint32_t i = first + u * (32u / G) + lane / G;
        uint32_t len = i < n ? in_len[i] : 0u;
        bool active = i < n && len < long_bytes;                 // (the long ones are gcm_long_kernel's)
        const uint8_t* src = active ? in + in_off[i] : nullptr;
        uint8_t* dst = active ? out + out_off[i] : nullptr;
        if (active && mode != 0 && len < pkgcm::kOverhead) {     // shorter than nonce + tag: cipher.go:69-71 / Open fails
            if (lig == 0) status[i] = len < pkgcm::kNonceBytes ? 2u : 1u;
            active = false;
        }
        const uint32_t clen = mode == 0 ? len : len - pkgcm::kOverhead;   // (meaningful when active)
        const uint8_t* nonce = mode == 0 ? nonces + (size_t)i * pkgcm::kNonceBytes : src;
        uint64_t yh = 0, yl = 0;
        uint32_t j0[4] = {0u, 0u, 0u, 0u};
        if (active) {
            if (mode == 0 && lig == 0) {                                                    // cipher.go:51-52
                if (((uintptr_t)nonce & 3u) == 0)      // (dst is 4-byte aligned by contract)
                    for (uint32_t b = 0; b < pkgcm::kNonceBytes; b += 4) pkgcm::store32(dst + b, pkgcm::load32(nonce + b));
                else
                    for (uint32_t b = 0; b < pkgcm::kNonceBytes; ++b) dst[b] = nonce[b];
            }
            pkgcm::gcm_group_pass<(int)G>(rk, gt, te, nonce,
                                          mode == 0 ? src : src + pkgcm::kNonceBytes,
                                          mode == 0 ? dst + pkgcm::kNonceBytes : dst, clen, mode != 0, lig, yh, yl, j0);
        }
#pragma unroll
        for (uint32_t d = G / 2u; d > 0; d >>= 1) {
            yh ^= __shfl_xor_sync(0xFFFFFFFFu, yh, d);
            yl ^= __shfl_xor_sync(0xFFFFFFFFu, yl, d);
        }
        // the lane that owned the lengths block holds E_K(J0): it makes the tag
        const uint32_t owner = active ? ((clen + 15u) >> 4) % G : 0u;
        uint32_t ok = 1u;
        if (active && lig == owner) {
            const uint32_t tw[4] = {(uint32_t)(yh >> 32) ^ j0[0], (uint32_t)yh ^ j0[1], (uint32_t)(yl >> 32) ^ j0[2], (uint32_t)yl ^ j0[3]};
            // (the tag follows the ciphertext: word accesses when the message length is a multiple of 4)
            if (mode == 0) {
                uint8_t* tp = dst + pkgcm::kNonceBytes + len;
                if ((len & 3u) == 0)
                    for (uint32_t j = 0; j < 4; ++j) pkgcm::store32(tp + 4 * j, pkgcm::bswap32(tw[j]));
                else
                    for (uint32_t b = 0; b < pkgcm::kTagBytes; ++b) tp[b] = (uint8_t)(tw[b >> 2] >> (24u - 8u * (b & 3u)));
            } else {
                const uint8_t* tp = src + pkgcm::kNonceBytes + clen;
                uint32_t diff = 0;
                if ((clen & 3u) == 0)
                    for (uint32_t j = 0; j < 4; ++j) diff |= pkgcm::load32(tp + 4 * j) ^ pkgcm::bswap32(tw[j]);
                else
                    for (uint32_t b = 0; b < pkgcm::kTagBytes; ++b) diff |= (uint32_t)tp[b] ^ ((tw[b >> 2] >> (24u - 8u * (b & 3u))) & 0xFFu);
                ok = diff == 0u;
                status[i] = ok ? 0u : 1u;
            }
        }
        if (mode != 0) {
            ok = __shfl_sync(0xFFFFFFFFu, ok, lead + owner);
            if (active && !ok)                                   // authentication failed: release no plaintext
                for (uint32_t b = lig * 16u; b < clen; b += G * 16u)
                    for (uint32_t k = b; k < b + 16u && k < clen; ++k) dst[k] = 0;
        }
    }
}

// The GHASH tables of gcm_lane_pass in gcm_long_kernel's shared layout.
struct LaneTablesSmem {
    const uint8_t* hrep_lane;        // replicated table of H^32 + this lane's copy offset
    const pkgcm::HEntry* pw4;        // 4-bit tables of H^1..H^32, 16 entries each
    __device__ __forceinline__ void horner(uint64_t& yh, uint64_t& yl) const { pkgcm::ghash_mul_smem(hrep_lane, 7u, yh, yl); }
    __device__ __forceinline__ void close(uint32_t e, uint64_t& yh, uint64_t& yl) const { pkgcm::ghash_mul_table4(pw4 + (e - 1u) * 16u, yh, yl); }
};

// One WARP per long message (those the short-message kernels skip; long_idx lists them).  Lane l takes the
// 16-byte blocks l, l+32, ... — a warp reads and writes 512 consecutive bytes per step — and the
// lanes' GHASH terms (pkgcm::gcm_lane_pass) XOR together through shuffles.  The grid is resident and
// a warp walks the list warp-major (entry k · (warps in the grid) + warp · gridDim + CTA), like
// gcm_group_kernel: the tables are staged once per CTA, not once per eight messages.
__global__ void __launch_bounds__(kLongThreads) gcm_long_kernel(const KeyContext* __restrict__ kc_g,
                                                                const uint8_t* __restrict__ in, const uint64_t* __restrict__ in_off,
                                                                const uint32_t* __restrict__ in_len, const uint8_t* __restrict__ nonces,
                                                                uint8_t* __restrict__ out, const uint64_t* __restrict__ out_off,
                                                                uint32_t* __restrict__ status, const uint32_t* __restrict__ long_idx,
                                                                uint32_t first, uint32_t n_long, int mode)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t* rk_s = reinterpret_cast<uint32_t*>(smem + kLongRkAt);
    uint32_t* rkr = reinterpret_cast<uint32_t*>(smem + kLongRkrAt);
    for (uint32_t i = threadIdx.x; i < 60; i += blockDim.x) {
        const uint32_t k = __ldg(&kc_g->g.rk[i]);
        rk_s[i] = k;
        rkr[i] = (k << 8) | (k >> 24);
    }
    for (uint32_t i = threadIdx.x; i < 32 * 16; i += blockDim.x)
        reinterpret_cast<uint4*>(smem + kLongPwAt)[i] = __ldg(reinterpret_cast<const uint4*>(&kc_g->pw4[0][0]) + i);
    // (the staging loop of gcm_group_kernel, with its reasons: rotated chunks, one loop)
    for (uint32_t x = threadIdx.x; x < 768; x += blockDim.x) {       // x = 2·entry + table; from 512: entry of H^32
        uint4 t4;
        uint32_t at = x * 128u;
        if (x < 512) {
            const uint32_t t = __ldg(&kc_g->te[(x & 1u) * 2u][x >> 1]);
            t4 = make_uint4(t, t, t, t);
        } else {
            t4 = __ldg(reinterpret_cast<const uint4*>(&kc_g->g.ht32[x - 512u]));
            at += kLongHrepAt - 512u * 128u;
        }
#pragma unroll
        for (uint32_t q = 0; q < 8; ++q) reinterpret_cast<uint4*>(smem + at)[(q + x) & 7u] = t4;
    }
    __syncthreads();
    const uint32_t* rk = rk_s;
    const LaneTablesSmem ht{smem + kLongHrepAt + (threadIdx.x & 7u) * 16u, reinterpret_cast<const pkgcm::HEntry*>(smem + kLongPwAt)};
    const pkgcm::TeDual te{smem, rkr, (threadIdx.x & 31u) * 4u};
    const uint32_t lane = threadIdx.x & 31u;

    // (the warp's rank comes out of a shuffle so that the compiler KNOWS it is the same in every lane:
    // the loop, the message's pointers and the shared-memory base of every table lookup then live in
    // uniform registers — 40 registers instead of 63, and no extra add per AES lookup)
    const uint32_t warp = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);
    for (uint32_t w = first + warp * gridDim.x + blockIdx.x; w < n_long; w += (blockDim.x >> 5) * gridDim.x) {   // list entries [first, n_long)
        const uint32_t i = long_idx[w];
        const uint8_t* src = in + in_off[i];
        uint8_t* dst = out + out_off[i];
        const uint32_t len = in_len[i];                       // >= long_bytes > kOverhead
        const uint8_t* nonce = mode == 0 ? nonces + (size_t)i * pkgcm::kNonceBytes : src;
        const uint32_t clen = mode == 0 ? len : len - pkgcm::kOverhead;
        uint64_t yh, yl;
        uint32_t j0[4] = {0u, 0u, 0u, 0u};
        pkgcm::gcm_lane_pass(rk, ht, te, nonce, mode == 0 ? src : src + pkgcm::kNonceBytes,
                             mode == 0 ? dst + pkgcm::kNonceBytes : dst, clen, mode != 0, lane, yh, yl, j0);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            yh ^= __shfl_xor_sync(0xFFFFFFFFu, yh, d);
            yl ^= __shfl_xor_sync(0xFFFFFFFFu, yl, d);
        }
        // the lane that owned the lengths block holds E_K(J0): it makes the tag
        const uint32_t owner = ((clen + 15u) >> 4) & 31u;
        uint32_t ok = 1;
        if (lane == owner) {
            const uint32_t tw[4] = {(uint32_t)(yh >> 32) ^ j0[0], (uint32_t)yh ^ j0[1], (uint32_t)(yl >> 32) ^ j0[2], (uint32_t)yl ^ j0[3]};
            if (mode == 0) {
                for (uint32_t b = 0; b < pkgcm::kNonceBytes; ++b) dst[b] = nonce[b];
                for (uint32_t b = 0; b < pkgcm::kTagBytes; ++b) dst[pkgcm::kNonceBytes + len + b] = (uint8_t)(tw[b >> 2] >> (24u - 8u * (b & 3u)));
            } else {
                uint32_t diff = 0;
                for (uint32_t b = 0; b < pkgcm::kTagBytes; ++b)
                    diff |= (uint32_t)src[pkgcm::kNonceBytes + clen + b] ^ ((tw[b >> 2] >> (24u - 8u * (b & 3u))) & 0xFFu);
                ok = diff == 0;
                status[i] = ok ? 0u : 1u;
            }
        }
        if (mode != 0) {
            ok = __shfl_sync(0xFFFFFFFFu, ok, owner);
            if (!ok) {                                        // authentication failed: release no plaintext
                __syncwarp();
                for (uint32_t b = lane; b < clen; b += 32u) dst[b] = 0;
            }
        }
    }
}

}  // namespace

struct pk_gcm {
    int device = 0;
    int num_sms = 148;
    uint32_t max_messages = 0;
    uint64_t max_bytes = 0;
    KeyContext* d_kc = nullptr;
    uint8_t *d_in = nullptr, *d_out = nullptr, *d_nonce = nullptr;
    uint64_t *d_in_off = nullptr, *d_out_off = nullptr;
    uint32_t *d_len = nullptr, *d_status = nullptr, *d_long = nullptr;
    uint32_t* h_long = nullptr;                 // pinned: indices of the messages the warp kernel takes
    uint32_t* h_long_dev = nullptr;             // the same block as the GPU addresses it
    uint64_t direct_bytes = 0;                  // largest batch (arena bytes in + out) sealed / opened in place
    uint64_t direct_batches = 0;
    cudaStream_t stream = nullptr;              // copies in, and everything when a batch is not pipelined
    cudaStream_t s_run = nullptr, s_out = nullptr;   // kernels / copies out of a pipelined batch
    cudaEvent_t ev_in[8] = {}, ev_run[8] = {};  // chunk c: its input has landed / its kernels are done
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // bracket the kernels of the last call
    uint32_t chunks = 0;                        // 0 = choose per batch, 1 = never pipeline, 2..8 = this many
    int last_cuda = 0;
    uint64_t launches = 0;
};

namespace {

#define PKG_CUDA(h, call)                                   \
    do {                                                    \
        cudaError_t e__ = (call);                           \
        if (e__ != cudaSuccess) {                           \
            if (h) (h)->last_cuda = (int)e__;               \
            (void)cudaGetLastError();                       \
            return PK_GCM_ERR_CUDA;                         \
        }                                                   \
    } while (0)

bool is_pinned(const void* p)
{
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// The device-side address of page-locked host memory the GPU can read and write, else nullptr.
template <typename T>
T* device_view(T* p)
{
    if (!p) return nullptr;
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
    return static_cast<T*>(a.devicePointer);
}

// Smallest message (bytes, as passed in) the warp-per-message kernel takes for a batch of n.
// PK_GCM_LONG_BYTES overrides it (tools/bench_gcm.py sweeps it).
uint32_t long_threshold(uint32_t n)
{
    const char* e = std::getenv("PK_GCM_LONG_BYTES");
    if (e && *e) { const long forced = std::atol(e); return forced < 64 ? 64u : (uint32_t)forced; }
    // crossovers measured on B200 with both kernels as they are since round 2 (tools/bench_gcm.py "sweep",
    // profiles/README.md): the more messages there are, the longer the lane-group kernel keeps winning
    return n < 512u ? 256u : n < 2048u ? 512u : n < 8192u ? 1024u : n < 32768u ? 2048u : 8192u;
}

// Warps per CTA of gcm_long_kernel for a batch of n_long messages (narrow blocks spread a small
// batch over more SMs; PK_GCM_LONG_WARPS overrides), and the resident grid for it.
uint32_t long_warps(uint32_t n_long)
{
    static const char* const env = std::getenv("PK_GCM_LONG_WARPS");
    if (env && *env) { const long w = std::atol(env); return w < 1 ? 1u : w > 16 ? 16u : (uint32_t)w; }
    return n_long >= 148u * 2u * 16u ? 16u : n_long >= 148u * 2u * 8u ? 8u : n_long >= 148u * 2u ? 2u : 1u;
}
uint32_t long_grid(const pk_gcm* h, uint32_t n_long, uint32_t warps)
{
    const uint32_t ctas = (n_long + warps - 1) / warps, resident = (uint32_t)h->num_sms * kLongCtasPerSm;
    return ctas < resident ? ctas : resident;
}

int run(pk_gcm* h, int mode, const uint8_t* in, const uint64_t* in_off, const uint32_t* in_len,
        const uint8_t* nonces, uint32_t n, uint8_t* out, const uint64_t* out_off, uint32_t* status)
{
    if (!h) return PK_GCM_ERR_ARG;
    if (n == 0) return PK_GCM_OK;
    if (!in_off || !in_len || !out_off || !out || (mode == 0 && !nonces) || (mode == 1 && !status)) return PK_GCM_ERR_ARG;
    if (n > h->max_messages) return PK_GCM_ERR_CAPACITY;
    // validate layout: 16-byte aligned, non-overlapping slots in order; compute extents
    // Messages of long_bytes or more go to the warp-per-message kernel.  It does one extra field
    // multiplication per lane, so with enough messages to fill the machine one thread per message
    // is the cheaper form until the message is a few KB; with few messages the warp form is what
    // keeps the SMs busy.
    const uint32_t long_bytes = long_threshold(n);
    uint32_t n_long = 0;
    uint64_t in_end = 0, out_end = 0;
    for (uint32_t i = 0; i < n; ++i) {
        if (in_len[i] >= long_bytes) h->h_long[n_long++] = i;
        const uint64_t need_out = mode == 0 ? (uint64_t)in_len[i] + pkgcm::kOverhead
                                            : (in_len[i] >= pkgcm::kOverhead ? in_len[i] - pkgcm::kOverhead : 0);
        if (in_off[i] % 16 || out_off[i] % 16) return PK_GCM_ERR_ALIGN;
        if (in_off[i] < in_end || out_off[i] < out_end) return PK_GCM_ERR_OFFSETS;
        // checked before the lengths are added, so a huge offset cannot wrap past the capacity test
        if (in_off[i] > h->max_bytes || out_off[i] > h->max_bytes) return PK_GCM_ERR_CAPACITY;
        in_end = in_off[i] + in_len[i];
        out_end = out_off[i] + need_out;
    }
    if (in_end > h->max_bytes || out_end > h->max_bytes) return PK_GCM_ERR_CAPACITY;
    if (in_end && !in) return PK_GCM_ERR_ARG;

    PKG_CUDA(h, cudaSetDevice(h->device));
    // A small batch of long messages whose every array is page-locked is sealed / opened IN PLACE:
    // the warp-per-message kernel reads the caller's arena over PCIe, 512 contiguous bytes per
    // warp-step, and writes the result straight back — one launch and a synchronisation instead of
    // up to six stream operations around it (256 x 4 KiB: 0.071 ms against 0.135).  Not for the
    // thread-per-message kernel: its 4-byte accesses, a message apart, are hopeless over PCIe
    // (65,536 x 64 B: 3.2 ms against 0.33).
    if (n_long == n && in_end + out_end <= h->direct_bytes) {
        const uint8_t* v_in = in_end ? device_view(in) : h->d_in;
        uint8_t* v_out = out_end ? device_view(out) : h->d_out;
        const uint64_t* v_in_off = device_view(in_off);
        const uint64_t* v_out_off = device_view(out_off);
        const uint32_t* v_len = device_view(in_len);
        const uint8_t* v_nonce = mode == 0 ? device_view(nonces) : h->d_nonce;
        uint32_t* v_status = mode == 1 ? device_view(status) : h->d_status;
        if (v_in && v_out && v_in_off && v_out_off && v_len && v_nonce && v_status) {
            cudaStream_t st = h->stream;
            PKG_CUDA(h, cudaEventRecord(h->ev0, st));
            if (n_long < n) {
                const uint32_t threads = n >= 148u * 64u ? 64u : 32u;
                gcm_kernel<<<(n + threads - 1) / threads, threads, 0, st>>>(h->d_kc, v_in, v_in_off, v_len, v_nonce, v_out, v_out_off,
                                                                               v_status, 0u, n, mode, long_bytes);
                PKG_CUDA(h, cudaGetLastError());
                ++h->launches;
            }
            if (n_long) {
                const uint32_t warps = n_long >= 148u * 2u ? 2u : 1u;
                gcm_long_kernel<<<long_grid(h, n_long, warps), warps * 32u, kSmemBytes, st>>>(
                    h->d_kc, v_in, v_in_off, v_len, v_nonce, v_out, v_out_off, v_status, h->h_long_dev, 0u, n_long, mode);
                PKG_CUDA(h, cudaGetLastError());
                ++h->launches;
            }
            PKG_CUDA(h, cudaEventRecord(h->ev1, st));
            PKG_CUDA(h, cudaStreamSynchronize(st));
            ++h->direct_batches;
            if (mode == 1)
                for (uint32_t i = 0; i < n; ++i)
                    if (status[i] != 0) return PK_GCM_ERR_AUTH;
            return PK_GCM_OK;
        }
    }
    // A large batch is cut into chunks of whole messages: the copy in of chunk c+1, the kernels of
    // chunk c and the copy out of chunk c-1 run at the same time (PCIe is full duplex), on three
    // streams chained by events.  Slots are in increasing order, so a chunk is one byte range of
    // each arena and one range of the long-message list.
    // Measured (profiles/README.md §6): pays from about 4 MiB of traffic, and only with page-locked
    // arenas — a pageable copy is staged synchronously by the driver and chunking only adds calls.
    uint32_t chunks = h->chunks;
    if (chunks == 0) {
        const uint64_t traffic = in_end + out_end;
        chunks = traffic >= (64u << 20) ? 8u : traffic >= (8u << 20) ? 4u : traffic >= (4u << 20) ? 2u : 1u;
        if (chunks > 1 && !(is_pinned(in) && is_pinned(out))) chunks = 1;
    }
    if (chunks > 8u) chunks = 8u;
    if (chunks > n) chunks = n;
    const bool piped = chunks > 1;
    cudaStream_t s_in = h->stream, s_run = piped ? h->s_run : h->stream, s_out = piped ? h->s_out : h->stream;
    PKG_CUDA(h, cudaMemcpyAsync(h->d_in_off, in_off, (size_t)n * 8, cudaMemcpyHostToDevice, s_in));
    PKG_CUDA(h, cudaMemcpyAsync(h->d_out_off, out_off, (size_t)n * 8, cudaMemcpyHostToDevice, s_in));
    PKG_CUDA(h, cudaMemcpyAsync(h->d_len, in_len, (size_t)n * 4, cudaMemcpyHostToDevice, s_in));
    if (mode == 0) PKG_CUDA(h, cudaMemcpyAsync(h->d_nonce, nonces, (size_t)n * pkgcm::kNonceBytes, cudaMemcpyHostToDevice, s_in));
    if (n_long) PKG_CUDA(h, cudaMemcpyAsync(h->d_long, h->h_long, (size_t)n_long * 4, cudaMemcpyHostToDevice, s_in));
    uint32_t lo = 0, l_lo = 0;
    for (uint32_t c = 0; c < chunks; ++c) {
        // chunk c = messages [lo, hi): cut by bytes, so chunks carry equal copy time
        uint32_t hi = n;
        if (c + 1 < chunks) {
            const uint64_t cut = in_end / chunks * (c + 1);
            hi = lo;
            while (hi < n && in_off[hi] < cut) ++hi;
        }
        if (hi == lo) continue;
        const uint64_t in_lo = in_off[lo], in_hi = hi < n ? in_off[hi] : in_end;
        const uint64_t out_lo = out_off[lo], out_hi = hi < n ? out_off[hi] : out_end;
        uint32_t l_hi = l_lo;
        while (l_hi < n_long && h->h_long[l_hi] < hi) ++l_hi;
        if (in_hi > in_lo) PKG_CUDA(h, cudaMemcpyAsync(h->d_in + in_lo, in + in_lo, in_hi - in_lo, cudaMemcpyHostToDevice, s_in));
        if (piped) {
            PKG_CUDA(h, cudaEventRecord(h->ev_in[c], s_in));
            PKG_CUDA(h, cudaStreamWaitEvent(s_run, h->ev_in[c], 0));
        }
        if (lo == 0) PKG_CUDA(h, cudaEventRecord(h->ev0, s_run));
        const uint32_t cnt = hi - lo, cnt_long = l_hi - l_lo;
        if (cnt_long < cnt) {
            static const char* const g_env = std::getenv("PK_GCM_GROUP");     // "0": one thread per message (comparison)
            if (!(g_env && g_env[0] == '0')) {
                // kGroupLanes lanes per message; narrower blocks spread a small batch over more SMs
                const uint64_t lanes = (uint64_t)cnt * pkgcm::kGroupLanes;
                const uint32_t threads = lanes >= 148u * (uint64_t)kGroupThreads ? (uint32_t)kGroupThreads : lanes >= 148u * 128u ? 128u : 64u;
                const uint64_t resident = (uint64_t)h->num_sms * kGroupCtasPerSm;      // (shared memory allows two per SM)
                const uint64_t ctas = (lanes + threads - 1) / threads;
                gcm_group_kernel<<<(uint32_t)(ctas < resident ? ctas : resident), threads, kGroupSmemBytes, s_run>>>(
                    h->d_kc, h->d_in, h->d_in_off, h->d_len, h->d_nonce, h->d_out, h->d_out_off, h->d_status, lo, hi, mode, long_bytes);
            } else {
                // one thread per message: spread a small batch over more SMs with narrower blocks
                const uint32_t threads = cnt >= 148u * 2u * kThreads ? kThreads : cnt >= 148u * 64u ? 64u : 32u;
                gcm_kernel<<<(cnt + threads - 1) / threads, threads, 0, s_run>>>(h->d_kc, h->d_in, h->d_in_off, h->d_len, h->d_nonce,
                                                                                 h->d_out, h->d_out_off, h->d_status, lo, hi, mode, long_bytes);
            }
            PKG_CUDA(h, cudaGetLastError());
            ++h->launches;
        }
        if (cnt_long) {
            const uint32_t warps = long_warps(cnt_long);
            gcm_long_kernel<<<long_grid(h, cnt_long, warps), warps * 32u, kSmemBytes, s_run>>>(
                h->d_kc, h->d_in, h->d_in_off, h->d_len, h->d_nonce, h->d_out, h->d_out_off, h->d_status, h->d_long, l_lo, l_hi, mode);
            PKG_CUDA(h, cudaGetLastError());
            ++h->launches;
        }
        if (hi == n) PKG_CUDA(h, cudaEventRecord(h->ev1, s_run));
        if (piped) {
            PKG_CUDA(h, cudaEventRecord(h->ev_run[c], s_run));
            PKG_CUDA(h, cudaStreamWaitEvent(s_out, h->ev_run[c], 0));
        }
        if (out_hi > out_lo) PKG_CUDA(h, cudaMemcpyAsync(out + out_lo, h->d_out + out_lo, out_hi - out_lo, cudaMemcpyDeviceToHost, s_out));
        if (mode == 1) PKG_CUDA(h, cudaMemcpyAsync(status + lo, h->d_status + lo, (size_t)cnt * 4, cudaMemcpyDeviceToHost, s_out));
        lo = hi;
        l_lo = l_hi;
    }
    PKG_CUDA(h, cudaStreamSynchronize(s_out));
    if (piped) {                                           // all three idle before the handle is reused
        PKG_CUDA(h, cudaStreamSynchronize(s_run));
        PKG_CUDA(h, cudaStreamSynchronize(s_in));
    }
    if (mode == 1)
        for (uint32_t i = 0; i < n; ++i)
            if (status[i] != 0) return PK_GCM_ERR_AUTH;       // BatchDecrypt fails as a whole, cipher.go:135-137
    return PK_GCM_OK;
}

}  // namespace

extern "C" {

const char* pk_gcm_strerror(int status)
{
    switch (status) {
    case PK_GCM_OK:           return "ok";
    case PK_GCM_ERR_ARG:      return "invalid argument";
    case PK_GCM_ERR_ALIGN:    return "message offset is not a multiple of 16";
    case PK_GCM_ERR_OFFSETS:  return "message slots overlap or are out of order";
    case PK_GCM_ERR_CAPACITY: return "batch exceeds the capacity given to pk_gcm_init";
    case PK_GCM_ERR_CUDA:     return "CUDA runtime error";
    case PK_GCM_ERR_NO_DEVICE:return "no usable CUDA device";
    case PK_GCM_ERR_KEYLEN:   return "key length must be 32 bytes";
    case PK_GCM_ERR_AUTH:     return "failed to decrypt: message authentication failed";
    default:                  return "unknown pk_gcm status";
    }
}

int pk_gcm_init(int device, const uint8_t* key, uint32_t key_len, uint32_t max_messages,
                uint64_t max_bytes, pk_gcm** out)
{
    if (!out) return PK_GCM_ERR_ARG;
    *out = nullptr;
    if (!key || max_messages == 0) return PK_GCM_ERR_ARG;
    if (key_len != 32) return PK_GCM_ERR_KEYLEN;          // validateKey, cipher.go:15-19
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { (void)cudaGetLastError(); return PK_GCM_ERR_NO_DEVICE; }
    if (device < 0 || device >= count) return PK_GCM_ERR_NO_DEVICE;
    pk_gcm* h = new (std::nothrow) pk_gcm();
    if (!h) return PK_GCM_ERR_ARG;
    h->device = device;
    h->max_messages = max_messages;
    h->max_bytes = max_bytes;
    KeyContext kc;
    pkgcm::make_key_context(key, kc);
    const size_t n = max_messages;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gcm_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gcm_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGroupSmemBytes);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->s_run, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking);
    for (int c = 0; c < 8 && e == cudaSuccess; ++c) {
        e = cudaEventCreateWithFlags(&h->ev_in[c], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_run[c], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev1);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_kc, sizeof kc);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_kc, &kc, sizeof kc, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_in, max_bytes + 64);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_out, max_bytes + 64);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_nonce, n * pkgcm::kNonceBytes);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_in_off, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_out_off, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_len, n * 4);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_status, n * 4);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_long, n * 4);
    if (e == cudaSuccess) e = cudaMallocHost(&h->h_long, n * 4);
    if (e == cudaSuccess) e = cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->h_long_dev), h->h_long, 0);
    {
        const char* env = std::getenv("PK_GCM_DIRECT_BYTES");     // tuning aid; pk_gcm_set_direct_bytes at run time
        h->direct_bytes = env && *env ? (uint64_t)std::atoll(env) : (4u << 20);
    }
    std::memset(&kc, 0, sizeof kc);                       // do not leave key material on the stack
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        pk_gcm_destroy(h);
        return PK_GCM_ERR_CUDA;
    }
    *out = h;
    return PK_GCM_OK;
}

void pk_gcm_destroy(pk_gcm* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->s_run) cudaStreamSynchronize(h->s_run);
    if (h->s_out) cudaStreamSynchronize(h->s_out);
    if (h->d_kc) { cudaMemset(h->d_kc, 0, sizeof(KeyContext)); cudaFree(h->d_kc); }
    void* bufs[] = {h->d_in, h->d_out, h->d_nonce, h->d_in_off, h->d_out_off, h->d_len, h->d_status};
    for (void* b : bufs) if (b) cudaFree(b);
    if (h->d_long) cudaFree(h->d_long);
    if (h->h_long) cudaFreeHost(h->h_long);
    for (int c = 0; c < 8; ++c) {
        if (h->ev_in[c]) cudaEventDestroy(h->ev_in[c]);
        if (h->ev_run[c]) cudaEventDestroy(h->ev_run[c]);
    }
    if (h->s_run) cudaStreamDestroy(h->s_run);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    (void)cudaGetLastError();
    delete h;
}

int pk_gcm_seal_batch(pk_gcm* h, const uint8_t* plain, const uint64_t* plain_off, const uint32_t* plain_len,
                      const uint8_t* nonces, uint32_t n, uint8_t* sealed, const uint64_t* sealed_off)
{
    return run(h, 0, plain, plain_off, plain_len, nonces, n, sealed, sealed_off, nullptr);
}

int pk_gcm_open_batch(pk_gcm* h, const uint8_t* sealed, const uint64_t* sealed_off, const uint32_t* sealed_len,
                      uint32_t n, uint8_t* plain, const uint64_t* plain_off, uint32_t* status)
{
    return run(h, 1, sealed, sealed_off, sealed_len, nullptr, n, plain, plain_off, status);
}

void* pk_gcm_host_alloc(pk_gcm* h, size_t bytes)
{
    if (!h || bytes == 0) return nullptr;
    void* p = nullptr;
    if (cudaSetDevice(h->device) != cudaSuccess || cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        h->last_cuda = (int)cudaGetLastError();
        return nullptr;
    }
    return p;
}

void pk_gcm_host_free(pk_gcm* h, void* p)
{
    if (!h || !p) return;
    if (cudaFreeHost(p) != cudaSuccess) h->last_cuda = (int)cudaGetLastError();
}

int pk_gcm_set_direct_bytes(pk_gcm* h, uint64_t bytes)
{
    if (!h) return PK_GCM_ERR_ARG;
    h->direct_bytes = bytes;
    return PK_GCM_OK;
}

uint64_t pk_gcm_direct_count(const pk_gcm* h) { return h ? h->direct_batches : 0; }

int pk_gcm_set_chunks(pk_gcm* h, uint32_t chunks)
{
    if (!h || chunks > 8u) return PK_GCM_ERR_ARG;
    h->chunks = chunks;
    return PK_GCM_OK;
}

uint64_t pk_gcm_launch_count(const pk_gcm* h) { return h ? h->launches : 0; }

float pk_gcm_last_kernel_ms(pk_gcm* h)
{
    float ms = -1.f;
    if (!h || h->launches == 0) return ms;
    cudaSetDevice(h->device);
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) { (void)cudaGetLastError(); return -1.f; }
    return ms;
}
int pk_gcm_last_cuda_error(const pk_gcm* h) { return h ? h->last_cuda : 0; }

}  // extern "C"

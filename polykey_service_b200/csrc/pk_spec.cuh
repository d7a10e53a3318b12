// pk_spec.cuh — constants and per-block arithmetic of the routing SPEC (DESIGN.md "SPEC"),
// shared by every kernel.  The functions here are __host__ __device__ so that
// tests/test_swar_host.py can compile this header with g++ and check the SIMD-in-word byte
// classifier, the stripe round and the selection rule against the byte-at-a-time oracle
// without a GPU.
//
// Reference-pinned: only the tool-name switch (internal/service/mock.go:32-64).
// Everything else in this header is build-defined.
#pragma once
#include <stdint.h>

#include "../../include/pk_route.h"

#if defined(__CUDACC__)
#define PK_HD __host__ __device__ __forceinline__
#else
#define PK_HD static inline
#endif

namespace pk {

constexpr uint32_t kBlockBytes   = 16;      // one 16-byte block = one lane's vector load
constexpr uint32_t kStripes      = 32;      // stripes per segment = warp width
constexpr uint32_t kSegBlocks    = 1024;    // blocks per segment (16 KiB)
constexpr uint32_t kSegBytes     = kSegBlocks * kBlockBytes;
constexpr uint32_t kTinyBlocks   = 16;      // SPEC v3: a payload of at most 16 blocks (256 bytes) ...
constexpr uint32_t kTinyStripes  = 8;       // ... has 8 stripes (block k in stripe k mod 8), not 32
constexpr uint32_t kP1 = 0x9E3779B1u, kP2 = 0x85EBCA77u, kP5 = 0x165667B1u;
constexpr uint64_t kNameMul = 0xD6E8FEB86659FD93ull;
constexpr uint64_t kFmixC1 = 0xFF51AFD7ED558CCDull, kFmixC2 = 0xC4CEB9FE1A85EC53ull;

struct Words4 { uint32_t w[4]; };

PK_HD uint32_t rotl32(uint32_t x, uint32_t r)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(x, x, r);
#else
    return (x << r) | (x >> (32u - r));
#endif
}

// (cur << 8) | (prev >> 24): byte i of the result is byte i-1 of the 8-byte string prev:cur
PK_HD uint32_t shift_in_prev_byte(uint32_t prev, uint32_t cur)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(prev, cur, 8);
#else
    return (cur << 8) | (prev >> 24);
#endif
}

// adds 128 * (number of bytes of m whose bit 7 is set) to acc; m must be pre-masked to
// bit 7 of each byte.
PK_HD uint32_t add_flag_bytes(uint32_t m, uint32_t acc)
{
#if defined(__CUDA_ARCH__)
    return __dp4a(m, 0x01010101u, acc);
#else
    return acc + (m & 0xFFu) + ((m >> 8) & 0xFFu) + ((m >> 16) & 0xFFu) + (m >> 24);
#endif
}

PK_HD uint64_t fmix64(uint64_t x)
{
    x ^= x >> 33; x *= kFmixC1;
    x ^= x >> 33; x *= kFmixC2;
    x ^= x >> 33;
    return x;
}

// ---- SPEC §B, four bytes at a time ------------------------------------------------
// Byte classes: SEP = {0x00, 0x20}; LETTER = A-Z a-z; LEAD = 0xC0..0xFF; CONT = 0x80..0xBF;
// OTHER = every other ASCII byte (digits, punctuation, control characters, DEL).
// A token starts at every OTHER byte, every LEAD byte, and every LETTER that does not continue
// a letter run; runs are cut at every payload offset that is a multiple of 8.
//
// Every quantity is a "flag word": bit 7 of byte i carries the predicate for byte i of the
// word, the other bits are don't-care until the final mask.  On the device the five logic ops
// are explicit lop3 (the compiler otherwise re-associates them into seven) and the three range
// tests are mad.lo with a run-time multiplier `one` == 1: a*one + c is an IMAD on the FMA pipe
// where a + c would be an IADD on the ALU pipe, which is what balances the SM's two integer
// pipes (DESIGN.md "Kernels").
#if defined(__CUDA_ARCH__)
template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(d) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return d;
}
__device__ __forceinline__ uint32_t mad_lo(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
#else
template <int LUT>
static inline uint32_t lop3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d = 0;                                  // truth table: bit (a<<2 | b<<1 | c) of LUT
    for (int bit = 0; bit < 32; ++bit) {
        const uint32_t idx = (((a >> bit) & 1u) << 2) | (((b >> bit) & 1u) << 1) | ((c >> bit) & 1u);
        d |= ((uint32_t)(LUT >> idx) & 1u) << bit;
    }
    return d;
}
static inline uint32_t mad_lo(uint32_t a, uint32_t b, uint32_t c) { return a * b + c; }
#endif

struct WordScan { uint32_t t, lm; };

// t: the word with bits 5 and 7 of every byte cleared ("folded": A-Z and a-z -> 0x41..0x5A,
// NUL and space -> 0).  lm: ASCII-letter flags.
PK_HD WordScan fold_word(uint32_t x, uint32_t one)
{
    WordScan r;
    r.t = x & 0x5F5F5F5Fu;
    const uint32_t a1 = mad_lo(r.t, one, 0x3F3F3F3Fu);    // bit7: folded >= 0x41
    const uint32_t a2 = mad_lo(r.t, one, 0x25252525u);    // bit7: folded >= 0x5B
    r.lm = lop3<0x10>(a1, a2, x);                         // a1 & ~a2 & ~x : letter, and ASCII
    return r;
}

// Token-start flags of one word, pre-masked to bit 7.  pl: letter flags of the PREVIOUS byte of
// each byte, already cleared where a run is cut.
PK_HD uint32_t token_flags(uint32_t x, const WordScan& ws, uint32_t pl, uint32_t one)
{
    const uint32_t g = mad_lo(ws.t, one, 0x7F7F7F7Fu);    // bit7: folded != 0, i.e. not NUL / space
    const uint32_t a = lop3<0x70>(g, ws.lm, pl);          // g & ~(lm & pl): not a separator, not a continuing letter
    const uint32_t s = x + x;                             // bit7: bit 6 of the byte
    const uint32_t m = lop3<0xCA>(x, s, a);               // x ? s : a  -> byte >= 0xC0, or ASCII and a
    return m & 0x80808080u;
}

// One 16-byte block at payload offset 16k.  A letter run is cut at every offset that is a
// multiple of 8, i.e. at byte 0 of words 0 and 2 — so a block never needs its neighbour.
// Bytes past the payload end must already be zero (NUL is a separator: counts nothing).
PK_HD void scan_block(const Words4& b, uint32_t one, uint32_t& tok128)
{
    const WordScan s0 = fold_word(b.w[0], one), s1 = fold_word(b.w[1], one);
    const WordScan s2 = fold_word(b.w[2], one), s3 = fold_word(b.w[3], one);
    tok128 = add_flag_bytes(token_flags(b.w[0], s0, s0.lm << 8, one), tok128);
    tok128 = add_flag_bytes(token_flags(b.w[1], s1, shift_in_prev_byte(s0.lm, s1.lm), one), tok128);
    tok128 = add_flag_bytes(token_flags(b.w[2], s2, s2.lm << 8, one), tok128);
    tok128 = add_flag_bytes(token_flags(b.w[3], s3, shift_in_prev_byte(s2.lm, s3.lm), one), tok128);
}

// Zero the bytes of a block that lie at or past `valid` (0..16) bytes.
PK_HD void mask_tail(Words4& b, uint32_t valid)
{
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int nv = (int)valid - 4 * j;                 // valid bytes in word j
        const uint32_t m = nv >= 4 ? 0xFFFFFFFFu : nv <= 0 ? 0u : ((1u << (8 * nv)) - 1u);
        b.w[j] &= m;
    }
}

// ---- SPEC §C ----------------------------------------------------------------------
struct Stripe { uint32_t a[4]; };

PK_HD void stripe_init(Stripe& s, uint32_t q)
{
    const uint32_t base = q * kP5;
    s.a[0] = kP1 + kP2 + base;
    s.a[1] = kP2 + base;
    s.a[2] = base;
    s.a[3] = base - kP1;
}

PK_HD void stripe_round(Stripe& s, const Words4& b)
{
#pragma unroll
    for (int j = 0; j < 4; ++j) s.a[j] = rotl32(s.a[j] + b.w[j] * kP2, 13) * kP1;
}

PK_HD uint64_t stripe_digest(const Stripe& s)
{
    const uint32_t lo = rotl32(s.a[0], 1) + rotl32(s.a[1], 7) + rotl32(s.a[2], 12) + rotl32(s.a[3], 18);
    const uint32_t hi = rotl32(s.a[0], 24) + rotl32(s.a[1], 3) + rotl32(s.a[2], 19) + rotl32(s.a[3], 10);
    return fmix64(((uint64_t)hi << 32) | lo);
}

// SPEC §C, the tool name's share of the hash.  The name is read as little-endian 32-bit words
// (the last one zero padded; the true length enters route_hash_final); word j contributes
// mixn((j + 1) << 32 | word) — an injective 64-bit input, two multiplications — and the terms
// add up mod 2^64 like the payload's stripe digests, in any order.
PK_HD uint64_t name_term(uint32_t j, uint32_t word)
{
    uint64_t x = ((uint64_t)(j + 1u) << 32) | word;
    x *= kNameMul;
    x ^= x >> 32;
    x *= kFmixC1;
    return x;
}

// word j of a name: bytes 4j .. 4j+3, little endian, zero past name_len.  Byte loads: any
// pointer, any alignment.
PK_HD uint32_t name_word(const uint8_t* name, uint32_t name_len, uint32_t j)
{
    uint32_t w = 0;
#pragma unroll
    for (uint32_t k = 0; k < 4; ++k)
        if (4u * j + k < name_len) w |= (uint32_t)name[4u * j + k] << (8u * k);
    return w;
}

PK_HD uint64_t route_hash_final(uint64_t sum, uint32_t pay_len, uint32_t name_len)
{
    return fmix64(sum + ((uint64_t)name_len << 32) + (uint64_t)pay_len);
}

// ---- SPEC §A + §D: one request's decision from its scan results ---------------------
// The routing table as the kernels see it: the caller's pk_table plus, per tool, the sum of
// its name's hash terms, so that the switch compares 12 bytes (length + sum) per entry and
// touches name bytes only to confirm a hit.
struct alignas(16) DevTable {
    pk_table t;
    uint64_t tool_name_sum[PK_MAX_TOOLS];
};

PK_HD uint64_t name_sum(const uint8_t* name, uint32_t len)
{
    uint64_t s = 0;
    for (uint32_t j = 0; 4u * j < len; ++j) s += name_term(j, name_word(name, len, j));
    return s;
}

// Host side: the device copy is normalised — name bytes past name_len, tool rows past n_tools
// and provider rows past n_providers are zero — so that decide() may compare names as words
// and read whole groups of provider rows.
static inline void make_dev_table(const pk_table& t, DevTable& d)
{
    d = DevTable{};
    d.t.n_tools = t.n_tools;
    d.t.default_class = t.default_class;
    d.t.n_providers = t.n_providers;
    for (uint32_t k = 0; k < t.n_tools && k < PK_MAX_TOOLS; ++k) {
        const uint32_t len = t.tools[k].name_len <= PK_MAX_TOOL_NAME ? t.tools[k].name_len : PK_MAX_TOOL_NAME;
        for (uint32_t j = 0; j < len; ++j) d.t.tools[k].name[j] = t.tools[k].name[j];
        d.t.tools[k].name_len = t.tools[k].name_len;
        d.t.tools[k].class_id = t.tools[k].class_id;
        d.tool_name_sum[k] = name_sum(d.t.tools[k].name, len);
    }
    for (uint32_t p = 0; p < t.n_providers && p < PK_MAX_PROVIDERS; ++p) d.t.providers[p] = t.providers[p];
}

PK_HD bool provider_eligible(const pk_provider& e, uint32_t cls, uint32_t tok)
{
    return e.weight != 0u && ((e.tool_mask >> cls) & 1u) != 0u && tok <= e.max_tokens;
}

// Sequential (one thread per request): name terms of the hash, the tool switch
// (internal/service/mock.go:32-64: whole-name, case-sensitive byte equality, first match
// wins, no match = the default branch — a class, not an error), final hash, provider.
//   payload_sum   sum of the payload's stripe digests (SPEC §C)
//   tok           est_tokens (SPEC §B)
// Provider rule: among eligible providers (weight != 0, serves the tool class,
// tok <= max_tokens) pick with probability weight / total weight, deterministically from
// the top 32 bits of the route hash: x = (hi32(hash) * W) >> 32, first provider whose
// cumulative weight exceeds x.
//
// Written for instruction count, since 32 of these run per warp-instruction and the whole batch
// pays for them: the name is handled as 32-bit words (the hash terms and the comparison with the
// table, whose names are word-aligned and zero padded, use the same four registers); provider
// eligibility is evaluated ONCE per provider into a 32-bit mask that the pick then replays
// (the "first p whose running weight exceeds x" is the number of running weights <= x, since
// they only grow); rows past n_providers are zero in the device table and need no test.
//   kWordLoads   the name may be read as aligned 32-bit words, up to 7 bytes past its end (true
//                for names staged in shared memory, which is padded); otherwise byte loads
template <bool kWordLoads = false>
PK_HD void decide(const DevTable* __restrict__ tbl, const uint8_t* __restrict__ name,
                  uint32_t name_len, uint32_t pay_len, uint64_t payload_sum, uint32_t tok,
                  pk_decision* __restrict__ out)
{
    const uint32_t nwords = (name_len + 3u) >> 2;
    auto word = [&](uint32_t j) -> uint32_t {
#if defined(__CUDA_ARCH__)
        if (kWordLoads) {
            const uintptr_t a = reinterpret_cast<uintptr_t>(name);
            const uint32_t* al = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3)) + j;
            const uint32_t w = __funnelshift_r(al[0], al[1], (uint32_t)(a & 3u) * 8u);
            const uint32_t valid = name_len - 4u * j;                 // > 0: callers ask for j < nwords only
            return valid >= 4u ? w : (w & ((1u << (8u * valid)) - 1u));
        }
#endif
        return name_word(name, name_len, j);
    };
    uint32_t nw[4] = {0u, 0u, 0u, 0u};               // name bytes 0..15 as words, zero padded
    uint64_t nsum = 0;
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) {
        if (j < nwords) {
            nw[j] = word(j);
            nsum += name_term(j, nw[j]);
        }
    }
    for (uint32_t j = 4; j < nwords; ++j) nsum += name_term(j, word(j));

    uint32_t cls = tbl->t.default_class;
    uint32_t unknown = 1;
    const uint32_t nt = tbl->t.n_tools;
    for (uint32_t t = 0; t < nt; ++t) {
        const pk_tool& e = tbl->t.tools[t];
        if (e.name_len != name_len) continue;
        // the switch is exact byte equality.  pk_tool::name is 4-byte aligned and zero padded past
        // name_len (pk_init / pk_update_table see to that), so equal lengths and equal words are equal bytes.
        const uint32_t* tw = reinterpret_cast<const uint32_t*>(e.name);
        bool same = (tw[0] == nw[0]) & (tw[1] == nw[1]) & (tw[2] == nw[2]) & (tw[3] == nw[3]);
        for (uint32_t j = 4; same && j < nwords; ++j) same = tw[j] == word(j);      // names of 17..64 bytes
        if (same) { cls = e.class_id; unknown = 0; break; }
    }

    const uint64_t h = route_hash_final(payload_sum + nsum, pay_len, name_len);

    // pass 1: which providers are eligible (one bit each — there are at most 32) and their total weight
    const uint32_t np = tbl->t.n_providers;
    uint32_t elig = 0, total = 0;
#pragma unroll
    for (uint32_t u = 0; u < 8; ++u) {
        const pk_provider& e = tbl->t.providers[u];           // rows past n_providers: weight 0, not eligible
        const bool ok = provider_eligible(e, cls, tok);
        elig |= ok ? (1u << u) : 0u;
        total += ok ? e.weight : 0u;
    }
    for (uint32_t p = 8; p < np; ++p) {
        const pk_provider& e = tbl->t.providers[p];
        const bool ok = provider_eligible(e, cls, tok);
        elig |= ok ? (1u << p) : 0u;
        total += ok ? e.weight : 0u;
    }
    // pass 2: the first provider whose running eligible weight exceeds x = the number of running
    // weights <= x (they only grow)
    uint32_t prov = PK_PROVIDER_NONE;
    if (total != 0u) {
        const uint32_t x = (uint32_t)(((uint64_t)(uint32_t)(h >> 32) * (uint64_t)total) >> 32);   // < total
        uint32_t cum = 0, before = 0;
#pragma unroll
        for (uint32_t u = 0; u < 8; ++u) {
            cum += ((elig >> u) & 1u) ? tbl->t.providers[u].weight : 0u;
            before += cum <= x ? 1u : 0u;
        }
        for (uint32_t p = 8; before == p && p < np; ++p) {    // not among the first eight
            cum += ((elig >> p) & 1u) ? tbl->t.providers[p].weight : 0u;
            before += cum <= x ? 1u : 0u;
        }
        prov = before;                                        // < np: the running weight reaches total > x
    }

    const uint32_t flags = (prov == PK_PROVIDER_NONE ? PK_FLAG_NO_PROVIDER : 0u) |
                           (pay_len == 0u ? PK_FLAG_EMPTY_PAYLOAD : 0u) |
                           (unknown ? PK_FLAG_UNKNOWN_TOOL : 0u);
#if defined(__CUDA_ARCH__)
    // the 24-byte record as three 8-byte stores
    uint64_t* o = reinterpret_cast<uint64_t*>(out);
    o[0] = (uint64_t)pay_len | ((uint64_t)tok << 32);
    o[1] = h;
    o[2] = (uint64_t)(cls & 0xFFu) | ((uint64_t)(prov & 0xFFu) << 8) | ((uint64_t)flags << 16) |
           (200ull << 32);                          // Status.Code 200, mock.go:26
#else
    out->byte_len = pay_len;
    out->est_tokens = tok;
    out->route_hash = h;
    out->tool_class = (uint8_t)cls;
    out->provider = (uint8_t)prov;
    out->flags = (uint16_t)flags;
    out->status = 200;                              // Status.Code 200, mock.go:26
#endif
}

}  // namespace pk

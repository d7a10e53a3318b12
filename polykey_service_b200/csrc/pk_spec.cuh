// pk_spec.cuh — constants and per-block arithmetic of the routing SPEC (DESIGN.md "SPEC"),
// shared by every kernel.  The functions here are __host__ __device__ so that
// tests/test_swar_host.py can compile this header with g++ and check the SWAR byte
// classifier and the stripe round against the byte-at-a-time oracle without a GPU.
//
// Reference-pinned: only the tool-name switch (internal/service/mock.go:32-64), which
// lives in pk_kernels.cu.  Everything in this header is build-defined.
#pragma once
#include <stdint.h>

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
constexpr uint32_t kP1 = 0x9E3779B1u, kP2 = 0x85EBCA77u, kP3 = 0xC2B2AE3Du, kP5 = 0x165667B1u;
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
// Returns the letter mask of this word (bit 7 of byte i set iff byte i is an ASCII
// letter; other bits are don't-care) and adds 128 per token-starting byte to tok128.
//   x          4 payload bytes, little endian (byte i = payload[o+i]); bytes past the
//              payload end must already be zero (zero is a separator: counts nothing)
//   prev_lm    letter mask of the previous word (only its bit 31 is used)
//   brk        0x00000080 if o % 8 == 0 (a letter run is cut at every 8th offset), else 0
// Letter mask of a word: bit 7 of byte i set iff byte i is an ASCII letter (other bits
// are don't-care).  Depends on the word alone, so a lane can hand the mask of its last
// word to its neighbour before scanning.
PK_HD uint32_t letter_mask(uint32_t x)
{
    const uint32_t y  = (x & 0x7F7F7F7Fu) | 0x20202020u;
    const uint32_t a1 = y + 0x1F1F1F1Fu;          // bit7: folded byte >= 'a'
    const uint32_t a2 = y + 0x05050505u;          // bit7: folded byte >  'z'
    return a1 & ~a2 & ~x;
}

PK_HD uint32_t scan_word(uint32_t x, uint32_t prev_lm, uint32_t brk, uint32_t& tok128)
{
    const uint32_t g  = (x & 0x7F7F7F7Fu) + 0x5F5F5F5Fu;   // bit7: low 7 bits >= 0x21
    const uint32_t lm = letter_mask(x);
    const uint32_t pl = shift_in_prev_byte(prev_lm, lm);
    const uint32_t c  = lm & pl & ~brk;           // letter continuing a run
    const uint32_t s  = x << 1;                   // bit7: bit 6 of the byte
    const uint32_t m  = ((g & ~x & ~c) | (x & s)) & 0x80808080u;   // graphic ASCII not continuing, or UTF-8 lead
    tok128 = add_flag_bytes(m, tok128);
    return lm;
}

// One 16-byte block.  prev_lm: letter mask of the word before this block (bit 31).
// Returns the letter mask of the block's last word.
PK_HD uint32_t scan_block(const Words4& b, uint32_t prev_lm, uint32_t& tok128, uint32_t& orbits)
{
    uint32_t l0 = scan_word(b.w[0], prev_lm, 0x80u, tok128);
    uint32_t l1 = scan_word(b.w[1], l0, 0u, tok128);
    uint32_t l2 = scan_word(b.w[2], l1, 0x80u, tok128);
    uint32_t l3 = scan_word(b.w[3], l2, 0u, tok128);
    orbits |= (b.w[0] | b.w[1]) | (b.w[2] | b.w[3]);
    return l3;
}

// Is byte b an ASCII letter?  Returned in "letter mask" form (bit 31), for the byte that
// precedes a task's first block.
PK_HD uint32_t letter_mask_of_byte(uint32_t b)
{
    const uint32_t y = (b & 0x7Fu) | 0x20u;
    return (b < 0x80u && y >= 0x61u && y <= 0x7Au) ? 0x80000000u : 0u;
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

PK_HD uint64_t name_term(uint32_t j, uint32_t byte)
{
    uint32_t t = ((j << 8) | byte) + 1u;
    t *= kP1;
    t ^= t >> 15;
    t *= kP2;
    t ^= t >> 13;
    return (uint64_t)t * kNameMul;
}

PK_HD uint64_t route_hash_final(uint64_t sum, uint32_t pay_len, uint32_t name_len)
{
    return fmix64(sum + ((uint64_t)name_len << 32) + (uint64_t)pay_len);
}

// ---- SPEC §D ----------------------------------------------------------------------
PK_HD uint64_t provider_score(uint64_t route_hash, uint32_t salt, uint32_t weight)
{
    uint32_t r = ((uint32_t)route_hash ^ salt) * kP2 + (uint32_t)(route_hash >> 32);
    r ^= r >> 15;
    r *= kP3;
    r ^= r >> 13;
    return (uint64_t)r * (uint64_t)weight;
}

}  // namespace pk

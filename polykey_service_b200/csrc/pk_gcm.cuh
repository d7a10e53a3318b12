// pk_gcm.cuh — AES-256-GCM building blocks shared by the GPU kernels and the CPU unit test.
//
// Widening step: the batched form of
// internal/adapters/security/cipher.go:92-141 (Encrypt / Decrypt / BatchEncrypt / BatchDecrypt).
// The reference computes AES-256-GCM with Go's crypto/aes + crypto/cipher.NewGCM: a 12-byte nonce,
// no additional data, a 16-byte tag, output nonce || ciphertext || tag (cipher.go:51-52, 73-77).
// The algorithm is the published one (NIST SP 800-38D); parity is pinned by its known-answer tests
// and by an independent implementation (OpenSSL via Python `cryptography`), not by reference tests
// — the reference has none.
//
// What is here
//   * AES-256 encryption of one block with the four classic 1 KB T-tables (the last round's S-box
//     is a byte of Te0).  The tables and the key schedule are built on the HOST, once per key
//     (the reference rebuilds both per message, cipher.go:35-43 inside the loops at :116-122 /
//     :133-139).
//   * GHASH by Shoup's 8-bit table method: a 256-entry table of H·b per key (4 KB) plus the
//     256-entry reduction table, built on the host; 16 table steps per block.
//   * gcm_crypt_message(): CTR + GHASH over one message, sequentially — one GPU thread per message.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define PKG_HD __host__ __device__ __forceinline__
#define PKG_HDM __host__ __device__ __forceinline__      // member functions
#else
#define PKG_HD static inline
#define PKG_HDM inline
#endif

namespace pkgcm {

constexpr uint32_t kNonceBytes = 12;      // aesgcm.NonceSize(), cipher.go:45
constexpr uint32_t kTagBytes   = 16;
constexpr uint32_t kOverhead   = kNonceBytes + kTagBytes;

// One field element, big-endian halves; 16 bytes so a table lookup is one 128-bit load.
struct alignas(16) HEntry { uint64_t h, l; };

// Per-key tables, built by make_key_context() on the host.  GcmTables (8.8 KB) is what a kernel
// keeps in shared memory next to its bank-replicated copy of te[0].
struct GcmTables {
    uint32_t rk[60];            // AES-256 round keys, big-endian words, 15 × 4
    uint32_t pad[4];
    HEntry ht[256];             // Shoup table: ht[b] = H · b(x), H = E_K(0^128); entry 0x80 is H
    HEntry ht32[256];           // the same table for H^32 (the warp-per-message Horner step)
    HEntry pw[32];              // pw[e-1] = H^e, e = 1..32
};
// What the lane-group kernel (kGroupLanes lanes per short message) keeps in shared memory: the round
// keys and the 8-bit Shoup tables of H^1 .. H^G — H^G is the lanes' Horner step, H^e (e = 1..G) a
// lane's closing power, so every multiplication is a 16-step table walk (no bit-serial product).
constexpr int kGroupLanes = 4;
struct GroupTables {
    uint32_t rk[60];
    uint32_t pad[4];
    HEntry hp[kGroupLanes][256];    // hp[e-1] = table of H^e
};
struct KeyContext {
    uint32_t te[4][256];        // T-tables: te[0][x] = (2·S[x], S[x], S[x], 3·S[x]) big-endian; te[k] = te[0] rotated right by 8k
    GcmTables g;
    GroupTables gg;
    HEntry pw4[32][16];         // 4-bit Shoup tables of H^1..H^32: pw4[e-1][v] = H^e · (v << 4)(x) — the warp-per-message
                                // kernel's closing powers, a different one per lane (8 KB for all 32; ghash_mul_table4)
};
static_assert(sizeof(GroupTables) % 16 == 0, "GroupTables is staged as uint4");
static_assert(sizeof(GcmTables) % 16 == 0, "GcmTables is staged as uint4");
static_assert(sizeof(KeyContext) % 16 == 0, "KeyContext is staged as uint4");

PKG_HD uint32_t rotr32(uint32_t x, uint32_t r) { return (x >> r) | (x << (32u - r)); }
PKG_HD uint32_t load_be32(const uint8_t* p)
{
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
}
// 32-bit access to message bytes that are 4-byte aligned by contract
PKG_HD uint32_t load32(const uint8_t* p)
{
#if defined(__CUDA_ARCH__)
    return *reinterpret_cast<const uint32_t*>(p);
#else
    uint32_t w; memcpy(&w, p, 4); return w;
#endif
}
PKG_HD void store32(uint8_t* p, uint32_t w)
{
#if defined(__CUDA_ARCH__)
    *reinterpret_cast<uint32_t*>(p) = w;
#else
    memcpy(p, &w, 4);
#endif
}
PKG_HD uint32_t bswap32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(x, 0, 0x0123);
#else
    return (x >> 24) | ((x >> 8) & 0xFF00u) | ((x << 8) & 0xFF0000u) | (x << 24);
#endif
}

// T-table access.  Host / plain form: the four tables of the key context.
struct TeTables {
    const uint32_t (*te)[256];
    PKG_HDM uint32_t operator()(int k, uint32_t x) const { return te[k][x]; }
};
// Device form: ONE table, replicated 32 times in shared memory so that lane l always reads bank l
// (entry x of lane l at word 32·x + l).  A warp's 32 data-dependent lookups then never collide —
// with a single copy they serialise 3-4 deep, and the kernel is bound by exactly that.  The other
// three classic tables are byte rotations of the first.
struct TeReplicated {
    const uint32_t* rep;      // [256][32]
    uint32_t lane;
    PKG_HDM uint32_t operator()(int k, uint32_t x) const
    {
        const uint32_t t = rep[(x << 5) + lane];
        return k ? rotr32(t, 8u * (uint32_t)k) : t;
    }
};

// One AES-256 block encryption.  s[4]: state as big-endian column words, in place.
template <typename Te>
PKG_HD void aes256_encrypt_block(const Te& te, const uint32_t* __restrict__ rk, uint32_t s[4])
{
    uint32_t s0 = s[0] ^ rk[0], s1 = s[1] ^ rk[1], s2 = s[2] ^ rk[2], s3 = s[3] ^ rk[3];
#pragma unroll 1
    for (int r = 1; r < 14; ++r) {
        const uint32_t* k = rk + 4 * r;
        const uint32_t t0 = te(0, s0 >> 24) ^ te(1, (s1 >> 16) & 0xFF) ^ te(2, (s2 >> 8) & 0xFF) ^ te(3, s3 & 0xFF) ^ k[0];
        const uint32_t t1 = te(0, s1 >> 24) ^ te(1, (s2 >> 16) & 0xFF) ^ te(2, (s3 >> 8) & 0xFF) ^ te(3, s0 & 0xFF) ^ k[1];
        const uint32_t t2 = te(0, s2 >> 24) ^ te(1, (s3 >> 16) & 0xFF) ^ te(2, (s0 >> 8) & 0xFF) ^ te(3, s1 & 0xFF) ^ k[2];
        const uint32_t t3 = te(0, s3 >> 24) ^ te(1, (s0 >> 16) & 0xFF) ^ te(2, (s1 >> 8) & 0xFF) ^ te(3, s2 & 0xFF) ^ k[3];
        s0 = t0; s1 = t1; s2 = t2; s3 = t3;
    }
    // last round: SubBytes + ShiftRows + AddRoundKey; S[x] is byte 2 (bits 8..15) of Te0[x]
    const uint32_t* k = rk + 56;
#define PKG_SB(x) ((te(0, (x)) >> 8) & 0xFFu)
    s[0] = (PKG_SB(s0 >> 24) << 24 | PKG_SB((s1 >> 16) & 0xFF) << 16 | PKG_SB((s2 >> 8) & 0xFF) << 8 | PKG_SB(s3 & 0xFF)) ^ k[0];
    s[1] = (PKG_SB(s1 >> 24) << 24 | PKG_SB((s2 >> 16) & 0xFF) << 16 | PKG_SB((s3 >> 8) & 0xFF) << 8 | PKG_SB(s0 & 0xFF)) ^ k[1];
    s[2] = (PKG_SB(s2 >> 24) << 24 | PKG_SB((s3 >> 16) & 0xFF) << 16 | PKG_SB((s0 >> 8) & 0xFF) << 8 | PKG_SB(s1 & 0xFF)) ^ k[2];
    s[3] = (PKG_SB(s3 >> 24) << 24 | PKG_SB((s0 >> 16) & 0xFF) << 16 | PKG_SB((s1 >> 8) & 0xFF) << 8 | PKG_SB(s2 & 0xFF)) ^ k[3];
#undef PKG_SB
}

#if defined(__CUDACC__)
#ifndef PK_GCM_AES_UNROLL
#define PK_GCM_AES_UNROLL 1      // rounds per loop iteration of the device AES (2 measured: 0.089-0.093 against 0.085-0.087 ms)
#endif
constexpr int kAesUnroll = PK_GCM_AES_UNROLL;
// Device form used by the kernels: Te0 and Te2 (= Te0 rotated by 16) side by side, each replicated
// per bank — entry x is 256 bytes: 32 lanes × Te0[x], then 32 lanes × Te2[x] (64 KB in all).
//   * the address of a lookup is ONE byte-permute: (byte k of the state word) << 8 | lane·4,
//     and Te2 is the same address + 128 (an immediate);
//   * the two remaining rotations of a column are both by 8 and XOR is linear over rotations, so
//     column = Te0[a] ^ Te2[c] ^ rotr8(Te0[b] ^ Te2[d] ^ rotl8(rk)) — one rotation, three 3-input
//     logic ops.  `rkr` holds the round keys rotated left by 8.
// That is 28 ALU-pipe instructions per round against 56 for the single replicated table
// (shift + mask + lane-or per index, three rotations per column): the kernels are bound by that pipe.
struct TeDual {
    const uint8_t* tab;       // [256][2][32] words
    const uint32_t* rkr;      // rotl8 of the 60 round-key words
    uint32_t lane4;           // lane · 4
    __device__ __forceinline__ uint32_t at(uint32_t word, uint32_t k, uint32_t second) const
    {
        const uint32_t off = __byte_perm(word, lane4, 0x5504u | (k << 4));
        return *reinterpret_cast<const uint32_t*>(tab + off + second * 128u);
    }
    // TeTables-style access (last round helper, tag path): byte value -> rotated table word
    __device__ __forceinline__ uint32_t operator()(int k, uint32_t x) const
    {
        const uint32_t t = *reinterpret_cast<const uint32_t*>(tab + (x << 8) + lane4 + ((k & 2) ? 128u : 0u));
        return (k & 1) ? rotr32(t, 8u) : t;
    }
};

__device__ __forceinline__ void aes256_encrypt_block(const TeDual& te, const uint32_t* __restrict__ rk, uint32_t s[4])
{
    uint32_t s0 = s[0] ^ rk[0], s1 = s[1] ^ rk[1], s2 = s[2] ^ rk[2], s3 = s[3] ^ rk[3];
#pragma unroll kAesUnroll
    for (int r = 1; r < 14; ++r) {
        const uint4 kr = *reinterpret_cast<const uint4*>(te.rkr + 4 * r);
        const uint32_t u0 = te.at(s1, 2, 0) ^ te.at(s3, 0, 1) ^ kr.x;
        const uint32_t u1 = te.at(s2, 2, 0) ^ te.at(s0, 0, 1) ^ kr.y;
        const uint32_t u2 = te.at(s3, 2, 0) ^ te.at(s1, 0, 1) ^ kr.z;
        const uint32_t u3 = te.at(s0, 2, 0) ^ te.at(s2, 0, 1) ^ kr.w;
        const uint32_t t0 = te.at(s0, 3, 0) ^ te.at(s2, 1, 1) ^ __byte_perm(u0, 0, 0x0321);
        const uint32_t t1 = te.at(s1, 3, 0) ^ te.at(s3, 1, 1) ^ __byte_perm(u1, 0, 0x0321);
        const uint32_t t2 = te.at(s2, 3, 0) ^ te.at(s0, 1, 1) ^ __byte_perm(u2, 0, 0x0321);
        const uint32_t t3 = te.at(s3, 3, 0) ^ te.at(s1, 1, 1) ^ __byte_perm(u3, 0, 0x0321);
        s0 = t0; s1 = t1; s2 = t2; s3 = t3;
    }
    // last round: S[x] sits in bits 16..23 and 8..15 of Te0[x] and in bits 24..31 and 0..7 of Te2[x]
    const uint4 k = *reinterpret_cast<const uint4*>(rk + 56);
#define PKG_LAST(a, b, c, d)                                                                       \
    ((te.at(a, 3, 1) & 0xFF000000u) | (te.at(b, 2, 0) & 0x00FF0000u) | (te.at(c, 1, 0) & 0x0000FF00u) | \
     (te.at(d, 0, 1) & 0x000000FFu))
    s[0] = PKG_LAST(s0, s1, s2, s3) ^ k.x;
    s[1] = PKG_LAST(s1, s2, s3, s0) ^ k.y;
    s[2] = PKG_LAST(s2, s3, s0, s1) ^ k.z;
    s[3] = PKG_LAST(s3, s0, s1, s2) ^ k.w;
#undef PKG_LAST
}
#endif  // __CUDACC__

// (xh, xl) <- (xh, xl) · H in GF(2^128) (GCM bit order) with the key's 8-bit tables: walk the 16
// bytes of X from the last to the first; each step shifts the accumulator right by one byte,
// folds the byte that fell off back in at the top (r8), and adds H · (next byte of X).
PKG_HD HEntry load_entry(const HEntry* __restrict__ t, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    const uint4 v = *reinterpret_cast<const uint4*>(t + b);
    HEntry e;
    e.h = ((uint64_t)v.y << 32) | v.x;
    e.l = ((uint64_t)v.w << 32) | v.z;
    return e;
#else
    return t[b];
#endif
}

// X <- X · V for the element V whose 8-bit Shoup table is `t`: Horner over the 16 bytes of X from
// the last (highest powers of x) to the first.  Moving the accumulator right by 8 bits drops a
// byte r off the low end: its bit j (0 = least significant) is the coefficient of x^(135-j) =
// x^(7-j) · x^128, and x^128 = 1 + x + x^2 + x^7 — 0xE1 at the top — so the byte folds back as the
// carry-less product r ⊗ 0x1C2 placed at bits 48..63: four shifts, no table.
#if defined(__CUDACC__)
// Device form of the walk below for a table in SHARED memory, on 32-bit words and fully unrolled:
// one 128-bit shared load per step (on 64-bit halves the compiler splits it in two, twice the
// shared-memory wavefronts), constant byte positions, the fold as four shifts of one word.
//   tbl     entry 0, in shared memory (plus the caller's lane offset in a replicated table)
//   stride  log2 of the distance between entries in bytes: 4 plain, 7 replicated per bank group
__device__ __forceinline__ void ghash_mul_smem(const uint8_t* __restrict__ tbl, uint32_t stride, uint64_t& xh, uint64_t& xl)
{
    const uint32_t x[4] = {(uint32_t)xl, (uint32_t)(xl >> 32), (uint32_t)xh, (uint32_t)(xh >> 32)};
    uint4 e = *reinterpret_cast<const uint4*>(tbl + ((x[0] & 0xFFu) << stride));
    uint32_t z0 = e.z, z1 = e.w, z2 = e.x, z3 = e.y;           // HEntry {h, l}: x, y = h low, high; z, w = l low, high
#pragma unroll
    for (int i = 1; i < 16; ++i) {
        e = *reinterpret_cast<const uint4*>(tbl + (((x[i >> 2] >> (8 * (i & 3))) & 0xFFu) << stride));
        const uint32_t r = z0 << 24;                           // the byte that falls off, at the top of a word
        const uint32_t f = r ^ (r >> 1) ^ (r >> 2) ^ (r >> 7); // (rem<<8 ^ rem<<7 ^ rem<<6 ^ rem<<1) << 16
        z0 = __funnelshift_r(z0, z1, 8) ^ e.z;
        z1 = __funnelshift_r(z1, z2, 8) ^ e.w;
        z2 = __funnelshift_r(z2, z3, 8) ^ e.x;
        z3 = (z3 >> 8) ^ f ^ e.y;
    }
    xh = ((uint64_t)z3 << 32) | z2;
    xl = ((uint64_t)z1 << 32) | z0;
}
#endif

// On the device `t` must point into shared memory (every kernel stages its tables there).
PKG_HD void ghash_mul_table(const HEntry* __restrict__ t, uint64_t& xh, uint64_t& xl)
{
#if defined(__CUDA_ARCH__)
    ghash_mul_smem(reinterpret_cast<const uint8_t*>(t), 4u, xh, xl);
#else
    HEntry e = load_entry(t, (uint32_t)xl & 0xFFu);
    uint64_t zh = e.h, zl = e.l;
#pragma unroll 1
    for (int i = 1; i < 16; ++i) {
        const uint64_t word = i < 8 ? xl : xh;
        e = load_entry(t, (uint32_t)(word >> (8 * (i & 7))) & 0xFFu);
        const uint32_t rem = (uint32_t)zl & 0xFFu;
        const uint32_t fold = (rem << 1) ^ (rem << 6) ^ (rem << 7) ^ (rem << 8);
        zl = (zh << 56) | (zl >> 8);
        zh = (zh >> 8) ^ ((uint64_t)fold << 48) ^ e.h;
        zl ^= e.l;
    }
    xh = zh;
    xl = zl;
#endif
}
// The same product with a 4-BIT table t16[v] = V · (v << 4)(x) — sixteen entries of V's 8-bit table,
// 256 bytes per element instead of 4 KB — walking the 32 nibbles of X from the last to the first:
// a byte step of the walk above is two of these (low nibble, then high), because shifting right by
// 8 with its fold is shifting right by 4 with its fold, twice, and the table is linear in the byte.
// Twice the steps, so only where 4 KB per multiplier is too much: gcm_long_kernel's 32 closing powers.
PKG_HD void ghash_mul_table4(const HEntry* __restrict__ t16, uint64_t& xh, uint64_t& xl)
{
#if defined(__CUDA_ARCH__)
    const uint8_t* tbl = reinterpret_cast<const uint8_t*>(t16);             // shared memory, as for ghash_mul_table
    const uint32_t x[4] = {(uint32_t)xl, (uint32_t)(xl >> 32), (uint32_t)xh, (uint32_t)(xh >> 32)};
    uint4 e = *reinterpret_cast<const uint4*>(tbl + ((x[0] & 0xFu) << 4));
    uint32_t z0 = e.z, z1 = e.w, z2 = e.x, z3 = e.y;
#pragma unroll
    for (int i = 1; i < 32; ++i) {
        e = *reinterpret_cast<const uint4*>(tbl + (((x[i >> 3] >> (4 * (i & 7))) & 0xFu) << 4));
        const uint32_t r = z0 << 28;                           // the four bits that fall off, at the top of a word
        const uint32_t f = r ^ (r >> 1) ^ (r >> 2) ^ (r >> 7); // (rem<<8 ^ rem<<7 ^ rem<<6 ^ rem<<1) << 20
        z0 = __funnelshift_r(z0, z1, 4) ^ e.z;
        z1 = __funnelshift_r(z1, z2, 4) ^ e.w;
        z2 = __funnelshift_r(z2, z3, 4) ^ e.x;
        z3 = (z3 >> 4) ^ f ^ e.y;
    }
    xh = ((uint64_t)z3 << 32) | z2;
    xl = ((uint64_t)z1 << 32) | z0;
#else
    uint64_t zh = t16[xl & 0xFu].h, zl = t16[xl & 0xFu].l;
    for (int i = 1; i < 32; ++i) {
        const uint64_t word = i < 16 ? xl : xh;
        const HEntry e = t16[(word >> (4 * (i & 15))) & 0xFu];
        const uint64_t rem = zl & 0xFu;
        const uint64_t fold = (rem << 1) ^ (rem << 6) ^ (rem << 7) ^ (rem << 8);
        zl = (zh << 60) | (zl >> 4);
        zh = (zh >> 4) ^ (fold << 52) ^ e.h;
        zl ^= e.l;
    }
    xh = zh;
    xl = zl;
#endif
}
PKG_HD void ghash_mul(const GcmTables* __restrict__ kc, uint64_t& xh, uint64_t& xl)
{
    ghash_mul_table(kc->ht, xh, xl);
}

// Textbook GF(2^128) multiplication (SP 800-38D Algorithm 1), bit by bit: X <- X · P for an
// arbitrary P.  Slow (128 steps); used once per lane per long message, and by the unit test.
PKG_HD void ghash_mul_bitwise(uint64_t ph, uint64_t pl, uint64_t& xh, uint64_t& xl)
{
    uint64_t zh = 0, zl = 0, vh = ph, vl = pl;
#pragma unroll 1
    for (int i = 0; i < 128; ++i) {
        const uint64_t bit = i < 64 ? (xh >> (63 - i)) & 1u : (xl >> (127 - i)) & 1u;
        if (bit) { zh ^= vh; zl ^= vl; }
        const uint64_t carry = (vl & 1u) ? 0xE100000000000000ull : 0ull;
        vl = (vh << 63) | (vl >> 1);
        vh = (vh >> 1) ^ carry;
    }
    xh = zh; xl = zl;
}

// CTR-crypt one 16-byte (or final partial) block at offset `at` and return the CIPHERTEXT block
// as big-endian words, zero padded.
template <typename Te>
PKG_HD void ctr_block(const uint32_t* __restrict__ rk, const Te& te, uint32_t n0, uint32_t n1, uint32_t n2,
                      const uint8_t* __restrict__ in, uint8_t* __restrict__ out, uint32_t len, uint32_t at,
                      bool decrypt, uint32_t c[4])
{
    uint32_t ks[4] = {n0, n1, n2, 2u + (at >> 4)};
    aes256_encrypt_block(te, rk, ks);
    const uint32_t nb = len - at < 16 ? len - at : 16;
    c[0] = c[1] = c[2] = c[3] = 0;
    if (nb == 16) {
        uint32_t w[4], o[4];
#if defined(__CUDA_ARCH__)
        // one 128-bit access where the side is 16-byte aligned (the contract only promises 4; in a
        // sealed slot the ciphertext sits 12 bytes in, so one side of every call is the narrow one)
        if (((uintptr_t)(in + at) & 15u) == 0) {
            const uint4 v = *reinterpret_cast<const uint4*>(in + at);
            w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
        } else
#endif
        {
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = load32(in + at + 4 * j);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            o[j] = w[j] ^ bswap32(ks[j]);
            c[j] = bswap32(decrypt ? w[j] : o[j]);
        }
#if defined(__CUDA_ARCH__)
        if (((uintptr_t)(out + at) & 15u) == 0) {
            *reinterpret_cast<uint4*>(out + at) = make_uint4(o[0], o[1], o[2], o[3]);
        } else
#endif
        {
#pragma unroll
            for (int j = 0; j < 4; ++j) store32(out + at + 4 * j, o[j]);
        }
    } else {
        for (uint32_t b = 0; b < nb; ++b) {
            const uint8_t ib = in[at + b];
            const uint8_t ob = (uint8_t)(ib ^ (uint8_t)(ks[b >> 2] >> (24 - 8 * (b & 3))));
            out[at + b] = ob;
            c[b >> 2] |= (uint32_t)(decrypt ? ib : ob) << (24 - 8 * (b & 3));
        }
    }
}

// The 12 nonce bytes as three big-endian words: word loads when the nonce is 4-byte aligned (a
// sealed slot's always is; a caller's nonce array need not be), byte loads otherwise.
PKG_HD void load_nonce(const uint8_t* __restrict__ nonce, uint32_t& n0, uint32_t& n1, uint32_t& n2)
{
#if defined(__CUDA_ARCH__)
    if (((uintptr_t)nonce & 3u) == 0) {
        n0 = bswap32(load32(nonce)); n1 = bswap32(load32(nonce + 4)); n2 = bswap32(load32(nonce + 8));
        return;
    }
#endif
    n0 = load_be32(nonce); n1 = load_be32(nonce + 4); n2 = load_be32(nonce + 8);
}

// Warp-per-message form, one lane's share.  GHASH is X = sum over blocks B_b · H^(m-b), b = 0..m-1,
// where m = (ciphertext blocks) + 1 and the last block is the lengths block.  Lane l takes blocks
// l, l+32, ...: Horner with H^32 over its own blocks, then one multiplication by H^(m - last block)
// — the lanes' results XOR to X.  Returns this lane's term in (yh, yl); (0, 0) if it had no block.
// The GHASH tables arrive as a policy H (as for gcm_group_pass below): H.horner multiplies by H^32,
// H.close(e, ..) by H^e, e = 1..32.  The lane that owns the lengths block has no AES block to
// encrypt for it: it encrypts J0 = nonce || 1 instead and returns E_K(J0) in j0[], so the tag
// needs no serial AES block after the lanes have met.
struct LaneTablesPlain {
    const KeyContext* kc;
    PKG_HDM void horner(uint64_t& yh, uint64_t& yl) const { ghash_mul_table(kc->g.ht32, yh, yl); }
    PKG_HDM void close(uint32_t e, uint64_t& yh, uint64_t& yl) const { ghash_mul_table4(kc->pw4[e - 1u], yh, yl); }
};
template <typename Te, typename H>
PKG_HD void gcm_lane_pass(const uint32_t* __restrict__ rk, const H& ht, const Te& te, const uint8_t* __restrict__ nonce,
                          const uint8_t* __restrict__ in, uint8_t* __restrict__ out, uint32_t len,
                          bool decrypt, uint32_t lane, uint64_t& yh, uint64_t& yl, uint32_t j0[4])
{
    uint32_t n0, n1, n2;
    load_nonce(nonce, n0, n1, n2);
    const uint32_t nblk = (len + 15u) >> 4;
    const uint32_t m = nblk + 1u;
    yh = 0; yl = 0;
    uint32_t last = 0xFFFFFFFFu;
    for (uint32_t b = lane; b < m; b += 32u) {
        if (last != 0xFFFFFFFFu) ht.horner(yh, yl);
        if (b < nblk) {
            uint32_t c[4];
            ctr_block(rk, te, n0, n1, n2, in, out, len, b << 4, decrypt, c);
            yh ^= ((uint64_t)c[0] << 32) | c[1];
            yl ^= ((uint64_t)c[2] << 32) | c[3];
        } else {
            yl ^= (uint64_t)len * 8u;              // the lengths block: len(A) = 0 || len(C) in bits
            j0[0] = n0; j0[1] = n1; j0[2] = n2; j0[3] = 1u;
            aes256_encrypt_block(te, rk, j0);
        }
        last = b;
    }
    if (last != 0xFFFFFFFFu) ht.close(m - last, yh, yl);       // H^(m - last), 1..32
}

// Lane-group form for SHORT messages (a secret of 32-256 bytes is 2-16 blocks: one thread walking
// them serially leaves the SM's warp slots a fifth full and every thread on one dependent chain).
// G lanes share a message: lane l takes blocks l, l+G, ... of the m = (ciphertext blocks) + 1 GHASH
// blocks — Horner with H^G over its own blocks, closed with H^(m - its last block), every step a
// table walk — and the lanes' terms XOR to X.  The lane that owns the lengths block (it has no AES
// block to encrypt for it) encrypts J0 = nonce || 1 instead and returns E_K(J0) in j0[]: the work
// is even and the tag needs no serial tail.  (yh, yl) = this lane's term.
// The GHASH tables arrive as a policy H: H.horner(yh, yl) multiplies by H^G, H.close(e, yh, yl) by H^e
// (e = 1..G) — GroupTablesPlain below (host, unit test) or the kernel's shared-memory form.
struct GroupTablesPlain {
    const GroupTables* gt;
    PKG_HDM void horner(uint64_t& yh, uint64_t& yl) const { ghash_mul_table(gt->hp[kGroupLanes - 1], yh, yl); }
    PKG_HDM void close(uint32_t e, uint64_t& yh, uint64_t& yl) const { ghash_mul_table(gt->hp[e - 1u], yh, yl); }
};
template <int G, typename Te, typename H>
PKG_HD void gcm_group_pass(const uint32_t* __restrict__ rk, const H& ht, const Te& te, const uint8_t* __restrict__ nonce,
                           const uint8_t* __restrict__ in, uint8_t* __restrict__ out, uint32_t len,
                           bool decrypt, uint32_t lig, uint64_t& yh, uint64_t& yl, uint32_t j0[4])
{
    uint32_t n0, n1, n2;
    load_nonce(nonce, n0, n1, n2);
    const uint32_t nblk = (len + 15u) >> 4;
    const uint32_t m = nblk + 1u;
    yh = 0; yl = 0;
    uint32_t last = 0xFFFFFFFFu;
    for (uint32_t b = lig; b < m; b += (uint32_t)G) {
        if (last != 0xFFFFFFFFu) ht.horner(yh, yl);
        if (b < nblk) {
            uint32_t c[4];
            ctr_block(rk, te, n0, n1, n2, in, out, len, b << 4, decrypt, c);
            yh ^= ((uint64_t)c[0] << 32) | c[1];
            yl ^= ((uint64_t)c[2] << 32) | c[3];
        } else {
            yl ^= (uint64_t)len * 8u;              // the lengths block: len(A) = 0 || len(C) in bits
            j0[0] = n0; j0[1] = n1; j0[2] = n2; j0[3] = 1u;
            aes256_encrypt_block(te, rk, j0);
        }
        last = b;
    }
    if (last != 0xFFFFFFFFu) ht.close(m - last, yh, yl);       // H^(m - last), 1..G
}

// tag = X ^ E_K(nonce || 1), as 16 bytes
template <typename Te>
PKG_HD void gcm_tag(const GcmTables* __restrict__ kc, const Te& te, const uint8_t* __restrict__ nonce,
                    uint64_t xh, uint64_t xl, uint8_t tag[16])
{
    uint32_t j0[4] = {load_be32(nonce), load_be32(nonce + 4), load_be32(nonce + 8), 1u};
    aes256_encrypt_block(te, kc->rk, j0);
    const uint32_t t[4] = {(uint32_t)(xh >> 32) ^ j0[0], (uint32_t)xh ^ j0[1], (uint32_t)(xl >> 32) ^ j0[2], (uint32_t)xl ^ j0[3]};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        tag[4 * j] = (uint8_t)(t[j] >> 24); tag[4 * j + 1] = (uint8_t)(t[j] >> 16);
        tag[4 * j + 2] = (uint8_t)(t[j] >> 8); tag[4 * j + 3] = (uint8_t)t[j];
    }
}

// One message: CTR-mode encryption (or decryption) with counter blocks nonce || be32(2, 3, ...) and
// GHASH over the CIPHERTEXT, then the tag = GHASH ^ E_K(nonce || be32(1)).
//   in / out   message bytes (len), 4-byte aligned; may not alias unless identical
//   decrypt    false: in = plaintext, out = ciphertext;  true: in = ciphertext, out = plaintext
//   tag        receives the 16 tag bytes (computed over the ciphertext in both directions)
template <typename Te>
PKG_HD void gcm_crypt_message(const GcmTables* __restrict__ kc, const Te& te,
                              const uint8_t* __restrict__ nonce, const uint8_t* __restrict__ in,
                              uint8_t* __restrict__ out, uint32_t len, bool decrypt, uint8_t tag[16])
{
    const uint32_t n0 = load_be32(nonce), n1 = load_be32(nonce + 4), n2 = load_be32(nonce + 8);
    uint64_t xh = 0, xl = 0;                         // GHASH state, big-endian halves
    uint32_t ctr = 2;
    for (uint32_t at = 0; at < len; at += 16, ++ctr) {
        uint32_t ks[4] = {n0, n1, n2, ctr};
        aes256_encrypt_block(te, kc->rk, ks);
        const uint32_t nb = len - at < 16 ? len - at : 16;
        uint32_t c[4] = {0, 0, 0, 0};                // ciphertext block, big-endian words, zero padded
        if (nb == 16) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t w = load32(in + at + 4 * j);
                const uint32_t o = w ^ bswap32(ks[j]);
                store32(out + at + 4 * j, o);
                c[j] = bswap32(decrypt ? w : o);
            }
        } else {
            for (uint32_t b = 0; b < nb; ++b) {
                const uint8_t ib = in[at + b];
                const uint8_t ob = (uint8_t)(ib ^ (uint8_t)(ks[b >> 2] >> (24 - 8 * (b & 3))));
                out[at + b] = ob;
                c[b >> 2] |= (uint32_t)(decrypt ? ib : ob) << (24 - 8 * (b & 3));
            }
        }
        xh ^= ((uint64_t)c[0] << 32) | c[1];
        xl ^= ((uint64_t)c[2] << 32) | c[3];
        ghash_mul(kc, xh, xl);
    }
    xl ^= (uint64_t)len * 8u;                        // lengths block: len(A) = 0 || len(C) in bits
    ghash_mul(kc, xh, xl);
    uint32_t j0[4] = {n0, n1, n2, 1u};
    aes256_encrypt_block(te, kc->rk, j0);
    const uint32_t t[4] = {(uint32_t)(xh >> 32) ^ j0[0], (uint32_t)xh ^ j0[1], (uint32_t)(xl >> 32) ^ j0[2], (uint32_t)xl ^ j0[3]};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        tag[4 * j] = (uint8_t)(t[j] >> 24); tag[4 * j + 1] = (uint8_t)(t[j] >> 16);
        tag[4 * j + 2] = (uint8_t)(t[j] >> 8); tag[4 * j + 3] = (uint8_t)t[j];
    }
}

// ---------------------------------------------------------------- host-side set-up -----------
static inline uint8_t gf256_mul(uint8_t a, uint8_t b)
{
    uint8_t p = 0;
    for (int i = 0; i < 8; ++i) {
        if (b & 1) p ^= a;
        const uint8_t hi = a & 0x80;
        a = (uint8_t)(a << 1);
        if (hi) a ^= 0x1B;
        b >>= 1;
    }
    return p;
}

// S-box from its definition (multiplicative inverse in GF(2^8), then the affine map) — no table
// of 256 literals to mistype.
static inline void make_sbox(uint8_t sbox[256])
{
    for (int x = 0; x < 256; ++x) {
        uint8_t inv = 0;
        if (x) for (int y = 1; y < 256; ++y) if (gf256_mul((uint8_t)x, (uint8_t)y) == 1) { inv = (uint8_t)y; break; }
        uint8_t s = inv;
        for (int i = 1; i <= 4; ++i) s ^= (uint8_t)((inv << i) | (inv >> (8 - i)));
        sbox[x] = (uint8_t)(s ^ 0x63);
    }
}

// 8-bit Shoup table of a field element V: entry b is V · b(x) where the byte's most significant
// bit is x^0 (GCM's bit order): 0x80 -> V, 0x40 -> V·x, ... 0x01 -> V·x^7; the rest by linearity.
static inline void make_shoup_table(uint64_t vh, uint64_t vl, HEntry t[256])
{
    for (int i = 0; i < 256; ++i) { t[i].h = 0; t[i].l = 0; }
    t[0x80].h = vh; t[0x80].l = vl;
    for (int i = 0x40; i > 0; i >>= 1) {
        const uint64_t carry = (vl & 1u) ? 0xE100000000000000ull : 0ull;     // multiply by x
        vl = (vh << 63) | (vl >> 1);
        vh = (vh >> 1) ^ carry;
        t[i].h = vh; t[i].l = vl;
    }
    for (int i = 2; i <= 0x80; i *= 2)
        for (int j = 1; j < i; ++j) { t[i + j].h = t[i].h ^ t[j].h; t[i + j].l = t[i].l ^ t[j].l; }
}

// key: exactly 32 bytes (cipher.go:15-19 rejects anything else).
static inline void make_key_context(const uint8_t key[32], KeyContext& kc)
{
    uint8_t sbox[256];
    make_sbox(sbox);
    for (int x = 0; x < 256; ++x) {
        const uint8_t s = sbox[x];
        const uint32_t t = ((uint32_t)gf256_mul(s, 2) << 24) | ((uint32_t)s << 16) | ((uint32_t)s << 8) | gf256_mul(s, 3);
        for (int k = 0; k < 4; ++k) kc.te[k][x] = k ? rotr32(t, 8 * k) : t;
    }
    for (int k = 0; k < 4; ++k) kc.g.pad[k] = 0;
    // AES-256 key expansion (FIPS-197 §5.2, Nk = 8, Nr = 14)
    for (int i = 0; i < 8; ++i) kc.g.rk[i] = load_be32(key + 4 * i);
    uint8_t rcon = 1;
    for (int i = 8; i < 60; ++i) {
        uint32_t t = kc.g.rk[i - 1];
        if (i % 8 == 0) {
            t = (t << 8) | (t >> 24);
            t = ((uint32_t)sbox[t >> 24] << 24) | ((uint32_t)sbox[(t >> 16) & 0xFF] << 16) | ((uint32_t)sbox[(t >> 8) & 0xFF] << 8) | sbox[t & 0xFF];
            t ^= (uint32_t)rcon << 24;
            rcon = gf256_mul(rcon, 2);
        } else if (i % 8 == 4) {
            t = ((uint32_t)sbox[t >> 24] << 24) | ((uint32_t)sbox[(t >> 16) & 0xFF] << 16) | ((uint32_t)sbox[(t >> 8) & 0xFF] << 8) | sbox[t & 0xFF];
        }
        kc.g.rk[i] = kc.g.rk[i - 8] ^ t;
    }
    // H = E_K(0^128)
    uint32_t h[4] = {0, 0, 0, 0};
    aes256_encrypt_block(TeTables{kc.te}, kc.g.rk, h);
    const uint64_t Hh = ((uint64_t)h[0] << 32) | h[1], Hl = ((uint64_t)h[2] << 32) | h[3];
    make_shoup_table(Hh, Hl, kc.g.ht);
    // H^1 .. H^32 (warp-per-message path: Horner step H^32, per-lane closing power H^e)
    uint64_t ph = Hh, pl = Hl;
    for (int e = 1; e <= 32; ++e) {
        kc.g.pw[e - 1].h = ph; kc.g.pw[e - 1].l = pl;
        ghash_mul_bitwise(Hh, Hl, ph, pl);
    }
    make_shoup_table(kc.g.pw[31].h, kc.g.pw[31].l, kc.g.ht32);
    for (int e = 1; e <= 32; ++e) {                                  // 4-bit tables of the closing powers
        HEntry t[256];
        make_shoup_table(kc.g.pw[e - 1].h, kc.g.pw[e - 1].l, t);
        for (int v = 0; v < 16; ++v) kc.pw4[e - 1][v] = t[v << 4];
    }
    // the lane-group kernel's copy: round keys, tables of H^1 .. H^G
    for (int i = 0; i < 60; ++i) kc.gg.rk[i] = kc.g.rk[i];
    for (int k = 0; k < 4; ++k) kc.gg.pad[k] = 0;
    for (int e = 1; e <= kGroupLanes; ++e) make_shoup_table(kc.g.pw[e - 1].h, kc.g.pw[e - 1].l, kc.gg.hp[e - 1]);
}

}  // namespace pkgcm

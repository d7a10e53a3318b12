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

// Per-key tables, built by make_key_context() on the host.  GcmTables (4.9 KB) is what a kernel
// keeps in shared memory next to its bank-replicated copy of te[0].
struct GcmTables {
    uint32_t rk[60];            // AES-256 round keys, big-endian words, 15 × 4
    uint32_t pad[4];
    uint64_t hh[256], hl[256];  // Shoup table: (hh[b], hl[b]) = H · b(x), H = E_K(0^128); entry 0x80 is H
    uint16_t r8[256];           // reduction of the byte shifted out at the low end (goes to bits 48..63)
};
struct KeyContext {
    uint32_t te[4][256];        // T-tables: te[0][x] = (2·S[x], S[x], S[x], 3·S[x]) big-endian; te[k] = te[0] rotated right by 8k
    GcmTables g;
};
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

// (xh, xl) <- (xh, xl) · H in GF(2^128) (GCM bit order) with the key's 8-bit tables: walk the 16
// bytes of X from the last to the first; each step shifts the accumulator right by one byte,
// folds the byte that fell off back in at the top (r8), and adds H · (next byte of X).
PKG_HD void ghash_mul(const GcmTables* __restrict__ kc, uint64_t& xh, uint64_t& xl)
{
    uint32_t b = (uint32_t)xl & 0xFFu;
    uint64_t zh = kc->hh[b], zl = kc->hl[b];
#pragma unroll 1
    for (int i = 1; i < 16; ++i) {
        const uint64_t word = i < 8 ? xl : xh;
        b = (uint32_t)(word >> (8 * (i & 7))) & 0xFFu;
        const uint32_t rem = (uint32_t)zl & 0xFFu;
        zl = (zh << 56) | (zl >> 8);
        zh = (zh >> 8) ^ ((uint64_t)kc->r8[rem] << 48);
        zh ^= kc->hh[b];
        zl ^= kc->hl[b];
    }
    xh = zh;
    xl = zl;
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
    // H = E_K(0^128).  Table entry b is H · b(x) where the byte's most significant bit is x^0
    // (GCM's bit order): 0x80 -> H, 0x40 -> H·x, ... 0x01 -> H·x^7; the rest by linearity.
    uint32_t h[4] = {0, 0, 0, 0};
    aes256_encrypt_block(TeTables{kc.te}, kc.g.rk, h);
    uint64_t vh = ((uint64_t)h[0] << 32) | h[1], vl = ((uint64_t)h[2] << 32) | h[3];
    for (int i = 0; i < 256; ++i) { kc.g.hh[i] = 0; kc.g.hl[i] = 0; }
    kc.g.hh[0x80] = vh; kc.g.hl[0x80] = vl;
    for (int i = 0x40; i > 0; i >>= 1) {
        const uint64_t carry = (vl & 1u) ? 0xE100000000000000ull : 0ull;     // multiply by x
        vl = (vh << 63) | (vl >> 1);
        vh = (vh >> 1) ^ carry;
        kc.g.hh[i] = vh; kc.g.hl[i] = vl;
    }
    for (int i = 2; i <= 0x80; i *= 2)
        for (int j = 1; j < i; ++j) { kc.g.hh[i + j] = kc.g.hh[i] ^ kc.g.hh[j]; kc.g.hl[i + j] = kc.g.hl[i] ^ kc.g.hl[j]; }
    // byte r falls off the low end when the accumulator moves right by 8 bits: its bit j (0 = least
    // significant) is the coefficient of x^(135-j) = x^(7-j) · x^128, and x^128 = 1 + x + x^2 + x^7,
    // i.e. 0xE1 at the top — so bit j contributes 0xE100 >> (7 - j) to the top 16 bits.
    for (int r = 0; r < 256; ++r) {
        uint16_t v = 0;
        for (int j = 0; j < 8; ++j) if (r & (1 << j)) v ^= (uint16_t)(0xE100u >> (7 - j));
        kc.g.r8[r] = v;
    }
}

// Textbook GF(2^128) multiplication (SP 800-38D Algorithm 1), bit by bit — used only to check
// the table-driven ghash_mul in the unit test.
static inline void ghash_mul_bitwise(uint64_t hh, uint64_t hl, uint64_t& xh, uint64_t& xl)
{
    uint64_t zh = 0, zl = 0, vh = hh, vl = hl;
    for (int i = 0; i < 128; ++i) {
        const uint64_t bit = i < 64 ? (xh >> (63 - i)) & 1u : (xl >> (127 - i)) & 1u;
        if (bit) { zh ^= vh; zl ^= vl; }
        const uint64_t carry = (vl & 1u) ? 0xE100000000000000ull : 0ull;
        vl = (vh << 63) | (vl >> 1);
        vh = (vh >> 1) ^ carry;
    }
    xh = zh; xl = zl;
}

}  // namespace pkgcm

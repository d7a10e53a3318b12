// gcm_host_check.cpp — the AES-256-GCM building blocks of csrc/pk_gcm.cuh compiled for the host,
// so tests/test_gcm_cpu.py can check them against OpenSSL (Python `cryptography`) and the NIST
// known-answer tests without a GPU.  Test-only code.
#include <cstdint>
#include <cstring>

#include "../polykey_service_b200/csrc/pk_gcm.cuh"

using namespace pkgcm;

extern "C" void gcmh_seal(const uint8_t* key, const uint8_t* nonce, const uint8_t* pt, uint32_t len, uint8_t* out)
{
    KeyContext kc;
    make_key_context(key, kc);
    std::memcpy(out, nonce, 12);
    uint8_t tag[16];
    gcm_crypt_message(&kc.g, TeTables{kc.te}, nonce, pt, out + 12, len, false, tag);
    std::memcpy(out + 12 + len, tag, 16);
}

// returns 0 ok, 1 auth failure, 2 too short
extern "C" int gcmh_open(const uint8_t* key, const uint8_t* sealed, uint32_t len, uint8_t* pt)
{
    if (len < 12) return 2;
    if (len < 28) return 1;
    KeyContext kc;
    make_key_context(key, kc);
    uint8_t tag[16];
    gcm_crypt_message(&kc.g, TeTables{kc.te}, sealed, sealed + 12, pt, len - 28, true, tag);
    return std::memcmp(tag, sealed + len - 16, 16) == 0 ? 0 : 1;
}

extern "C" void gcmh_encrypt_block(const uint8_t* key, const uint8_t in[16], uint8_t out[16])
{
    KeyContext kc;
    make_key_context(key, kc);
    uint32_t s[4] = {load_be32(in), load_be32(in + 4), load_be32(in + 8), load_be32(in + 12)};
    aes256_encrypt_block(TeTables{kc.te}, kc.g.rk, s);
    for (int j = 0; j < 4; ++j) { out[4*j] = s[j] >> 24; out[4*j+1] = s[j] >> 16; out[4*j+2] = s[j] >> 8; out[4*j+3] = s[j]; }
}

// table-driven multiply vs the bit-by-bit textbook one; returns the number of mismatches
extern "C" int gcmh_check_ghash(const uint8_t* key, uint64_t seed, int trials)
{
    KeyContext kc;
    make_key_context(key, kc);
    int bad = 0;
    for (int t = 0; t < trials; ++t) {
        seed = seed * 6364136223846793005ull + 1442695040888963407ull; uint64_t a = seed;
        seed = seed * 6364136223846793005ull + 1442695040888963407ull; uint64_t b = seed;
        uint64_t xh = a, xl = b, yh = a, yl = b;
        ghash_mul(&kc.g, xh, xl);
        ghash_mul_bitwise(kc.g.ht[0x80].h, kc.g.ht[0x80].l, yh, yl);
        if (xh != yh || xl != yl) ++bad;
    }
    return bad;
}

// ghash_mul_table4 with the 4-bit table of H^e against the bit-serial product by H^e, e = 1..32
extern "C" int gcmh_check_table4(const uint8_t* key, uint64_t seed, int rounds)
{
    KeyContext kc;
    make_key_context(key, kc);
    int bad = 0;
    for (int r = 0; r < rounds; ++r)
        for (int e = 1; e <= 32; ++e) {
            seed = seed * 6364136223846793005ull + 1442695040888963407ull; const uint64_t a = seed;
            seed = seed * 6364136223846793005ull + 1442695040888963407ull; const uint64_t b = seed;
            uint64_t xh = a, xl = b, yh = a, yl = b;
            ghash_mul_table4(kc.pw4[e - 1], xh, xl);
            ghash_mul_bitwise(kc.g.pw[e - 1].h, kc.g.pw[e - 1].l, yh, yl);
            if (xh != yh || xl != yl) ++bad;
        }
    return bad;
}

// tag = X ^ E_K(J0), big-endian words to bytes
static void lanes_tag(uint64_t xh, uint64_t xl, const uint32_t j0[4], uint8_t tag[16])
{
    const uint32_t t[4] = {(uint32_t)(xh >> 32) ^ j0[0], (uint32_t)xh ^ j0[1], (uint32_t)(xl >> 32) ^ j0[2], (uint32_t)xl ^ j0[3]};
    for (int b = 0; b < 16; ++b) tag[b] = (uint8_t)(t[b >> 2] >> (24 - 8 * (b & 3)));
}

// The warp-per-message form run as a loop over 32 lanes on the host: must equal gcm_crypt_message.
extern "C" void gcmh_seal_lanes(const uint8_t* key, const uint8_t* nonce, const uint8_t* pt, uint32_t len, uint8_t* out)
{
    KeyContext kc;
    make_key_context(key, kc);
    std::memcpy(out, nonce, 12);
    uint64_t xh = 0, xl = 0;
    uint32_t j0[4] = {0, 0, 0, 0};
    for (uint32_t lane = 0; lane < 32; ++lane) {
        uint64_t yh, yl;
        uint32_t mine[4] = {0, 0, 0, 0};
        gcm_lane_pass(kc.g.rk, LaneTablesPlain{&kc}, TeTables{kc.te}, nonce, pt, out + 12, len, false, lane, yh, yl, mine);
        xh ^= yh; xl ^= yl;
        if (lane == ((len + 15u) >> 4) % 32u) std::memcpy(j0, mine, 16);      // the owner of the lengths block
    }
    uint8_t tag[16];
    lanes_tag(xh, xl, j0, tag);
    std::memcpy(out + 12 + len, tag, 16);
}

extern "C" int gcmh_open_lanes(const uint8_t* key, const uint8_t* sealed, uint32_t len, uint8_t* pt)
{
    if (len < 12) return 2;
    if (len < 28) return 1;
    KeyContext kc;
    make_key_context(key, kc);
    uint64_t xh = 0, xl = 0;
    uint32_t j0[4] = {0, 0, 0, 0};
    for (uint32_t lane = 0; lane < 32; ++lane) {
        uint64_t yh, yl;
        uint32_t mine[4] = {0, 0, 0, 0};
        gcm_lane_pass(kc.g.rk, LaneTablesPlain{&kc}, TeTables{kc.te}, sealed, sealed + 12, pt, len - 28, true, lane, yh, yl, mine);
        xh ^= yh; xl ^= yl;
        if (lane == ((len - 28u + 15u) >> 4) % 32u) std::memcpy(j0, mine, 16);
    }
    uint8_t tag[16];
    lanes_tag(xh, xl, j0, tag);
    return std::memcmp(tag, sealed + len - 16, 16) == 0 ? 0 : 1;
}

// The lane-group form for short messages (kGroupLanes lanes, Horner with H^G, table multiplications,
// E_K(J0) computed by the lane that owns the lengths block) run as a loop over the lanes on the host.
extern "C" int gcmh_group(const uint8_t* key, const uint8_t* nonce, const uint8_t* in, uint32_t len, int decrypt,
                          uint8_t* out, uint8_t* tag)
{
    KeyContext kc;
    make_key_context(key, kc);
    uint64_t xh = 0, xl = 0;
    uint32_t j0[4] = {0, 0, 0, 0};
    int owners = 0;
    for (uint32_t lig = 0; lig < (uint32_t)kGroupLanes; ++lig) {
        uint64_t yh, yl;
        uint32_t mine[4] = {0, 0, 0, 0};
        gcm_group_pass<kGroupLanes>(kc.gg.rk, GroupTablesPlain{&kc.gg}, TeTables{kc.te}, nonce, in, out, len, decrypt != 0, lig, yh, yl, mine);
        xh ^= yh; xl ^= yl;
        if (lig == ((len + 15u) >> 4) % (uint32_t)kGroupLanes) { std::memcpy(j0, mine, 16); ++owners; }
    }
    const uint32_t t[4] = {(uint32_t)(xh >> 32) ^ j0[0], (uint32_t)xh ^ j0[1], (uint32_t)(xl >> 32) ^ j0[2], (uint32_t)xl ^ j0[3]};
    for (int b = 0; b < 16; ++b) tag[b] = (uint8_t)(t[b >> 2] >> (24 - 8 * (b & 3)));
    return owners;
}

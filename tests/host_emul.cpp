// host_emul.cpp — CPU emulation of the warp algorithm in pk_kernels.cu, built by
// tests/test_swar_host.py with g++ from the SAME pk_spec.cuh the kernels include.
// It walks a (request, segment) task exactly as a 32-lane warp does — 16-byte block per
// lane, four steps per iteration, the rotating "last byte was a letter" hand-off, the
// tail mask, per-lane stripes, the commutative merge — so the SWAR classifier, the stripe
// geometry and the carry logic are checked against the byte-at-a-time oracle without a GPU.
// Test-only code: it is not part of libpkroute.so.
#include <cstdint>
#include <cstring>

#include "../include/pk_route.h"
#include "../polykey_service_b200/csrc/pk_spec.cuh"

using namespace pk;

namespace {
struct Partial { uint64_t sum; uint32_t tok; uint32_t orb; };

Partial scan_segment_emul(const uint8_t* pay, uint32_t len, uint32_t seg)
{
    const uint32_t nblk = (len >> 4) + ((len & 15u) ? 1u : 0u);
    const uint32_t first = seg * kSegBlocks;
    const uint32_t nb = (nblk - first) < kSegBlocks ? (nblk - first) : kSegBlocks;
    Stripe st[32];
    uint32_t tok128[32] = {0}, orb[32] = {0}, carry[32] = {0};
    for (uint32_t lane = 0; lane < 32; ++lane) stripe_init(st[lane], seg * kStripes + lane);
    if (first != 0) carry[0] = letter_mask_of_byte(pay[(size_t)first * 16 - 1]);
    for (uint32_t b0 = 0; b0 < nb; b0 += 32 * 4) {
        for (int u = 0; u < 4; ++u) {
            if (b0 + u * 32 >= nb) break;
            Words4 w[32];
            uint32_t l3[32];
            for (uint32_t lane = 0; lane < 32; ++lane) {
                const uint32_t k = b0 + u * 32 + lane;
                std::memset(&w[lane], 0, sizeof(Words4));
                if (k < nb) {
                    // the device reads a full 16-byte block even past the payload end
                    // (arena padding holds garbage there): emulate with 0xEE filler
                    uint8_t raw[16];
                    std::memset(raw, 0xEE, 16);
                    const size_t at = (size_t)(first + k) * 16;
                    const uint32_t valid = len - (first + k) * 16;
                    std::memcpy(raw, pay + at, valid < 16 ? valid : 16);
                    std::memcpy(&w[lane], raw, 16);
                    if (valid < 16) mask_tail(w[lane], valid);
                    stripe_round(st[lane], w[lane]);
                }
                l3[lane] = letter_mask(w[lane].w[3]);
            }
            for (uint32_t lane = 0; lane < 32; ++lane) {
                const uint32_t r = l3[(lane + 31u) & 31u];
                const uint32_t prev = lane == 0 ? carry[lane] : r;
                carry[lane] = r;
                scan_block(w[lane], prev, tok128[lane], orb[lane]);
            }
        }
    }
    Partial p{0, 0, 0};
    for (uint32_t lane = 0; lane < 32; ++lane) {
        if (lane < nb) p.sum += stripe_digest(st[lane]);
        p.tok += tok128[lane] >> 7;
        p.orb |= orb[lane];
    }
    return p;
}
}  // namespace

extern "C" void emul_route_one(const pk_table* t, const uint8_t* pay, uint32_t len,
                               const uint8_t* name, uint32_t nlen, pk_decision* d)
{
    const uint32_t nseg = len <= kSegBytes ? 1u : (len + kSegBytes - 1u) / kSegBytes;
    uint64_t sum = 0;
    uint32_t tok = 0, orb = 0;
    for (uint32_t g = nseg; g-- > 0;) {          // reverse order: the merge is commutative
        Partial p = scan_segment_emul(pay, len, g);
        sum += p.sum; tok += p.tok; orb |= p.orb;
    }
    for (uint32_t j = 0; j < nlen; ++j) sum += name_term(j, name[j]);
    uint32_t cls = t->default_class;
    bool unknown = true;
    for (uint32_t k = 0; k < t->n_tools; ++k) {
        if (t->tools[k].name_len != nlen) continue;
        if (std::memcmp(t->tools[k].name, name, nlen) == 0) { cls = t->tools[k].class_id; unknown = false; break; }
    }
    const uint64_t h = route_hash_final(sum, len, nlen);
    uint32_t prov = PK_PROVIDER_NONE;
    uint64_t best = 0;
    for (uint32_t p = 0; p < t->n_providers; ++p) {
        const pk_provider& e = t->providers[p];
        if (!(e.weight != 0 && ((e.tool_mask >> cls) & 1u) && tok <= e.max_tokens)) continue;
        const uint64_t sc = provider_score(h, e.salt, e.weight);
        if (prov == PK_PROVIDER_NONE || sc > best) { prov = p; best = sc; }
    }
    std::memset(d, 0, sizeof *d);
    d->byte_len = len; d->est_tokens = tok; d->route_hash = h;
    d->tool_class = (uint8_t)cls; d->provider = (uint8_t)prov;
    d->flags = (uint16_t)(((orb & 0x80808080u) ? PK_FLAG_NONASCII : 0) | (prov == PK_PROVIDER_NONE ? PK_FLAG_NO_PROVIDER : 0) |
                          (len == 0 ? PK_FLAG_EMPTY_PAYLOAD : 0) | (unknown ? PK_FLAG_UNKNOWN_TOOL : 0));
    d->status = 200;
}

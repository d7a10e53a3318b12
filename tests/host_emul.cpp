// host_emul.cpp — CPU emulation of the warp algorithm in pk_kernels.cu, built by
// tests/test_swar_host.py with g++ from the SAME pk_spec.cuh the kernels include.
// It walks a (request, segment) task as the warp does — one 16-byte block per lane per step,
// the tail mask, per-lane stripes and token counters, the commutative merge, then the same
// decide() the device runs — so the SIMD-in-word classifier, the stripe geometry and the
// selection rule are checked against the byte-at-a-time oracle without a GPU.
// Test-only code: it is not part of libpkroute.so.
#include <cstdint>
#include <cstring>

#include "../include/pk_route.h"
#include "../polykey_service_b200/csrc/pk_spec.cuh"

using namespace pk;

namespace {
struct Partial { uint64_t sum; uint32_t tok; };

// one (request, segment) task walked the way scan_segment() in pk_kernels.cu walks it:
// lane l owns blocks 32*i + l of the segment, with its own stripe and token counter
Partial scan_segment_emul(const uint8_t* pay, uint32_t len, uint32_t seg, uint32_t lanes)
{
    const uint32_t nblk = (len >> 4) + ((len & 15u) ? 1u : 0u);
    const uint32_t first = seg * kSegBlocks;
    const uint32_t nb = (nblk - first) < kSegBlocks ? (nblk - first) : kSegBlocks;
    Partial p{0, 0};
    for (uint32_t lane = 0; lane < lanes; ++lane) {
        Stripe st;
        stripe_init(st, seg * kStripes + lane);
        uint32_t tok128 = 0;
        for (uint32_t k = lane; k < nb; k += lanes) {
            // the device reads a full 16-byte block even past the payload end (arena padding
            // holds garbage there): emulate with 0xEE filler, then mask as the kernel does
            uint8_t raw[16];
            std::memset(raw, 0xEE, 16);
            const uint32_t valid = len - (first + k) * 16;
            std::memcpy(raw, pay + (size_t)(first + k) * 16, valid < 16 ? valid : 16);
            Words4 w;
            std::memcpy(&w, raw, 16);
            if (valid < 16) mask_tail(w, valid);
            stripe_round(st, w);
            scan_block(w, 1u, tok128);
        }
        if (lane < nb) p.sum += stripe_digest(st);
        p.tok += tok128 >> 7;
    }
    return p;
}
}  // namespace

// lanes = 32 (full warp) or 16 / 8 (the tiny-payload lane groups; valid when the payload has
// at most `lanes` blocks, where both walks visit the same (stripe, block) pairs)
extern "C" void emul_route_one(const pk_table* t, const uint8_t* pay, uint32_t len,
                               const uint8_t* name, uint32_t nlen, uint32_t lanes, pk_decision* d)
{
    const uint32_t nseg = len <= kSegBytes ? 1u : (len + kSegBytes - 1u) / kSegBytes;
    uint64_t sum = 0;
    uint32_t tok = 0;
    for (uint32_t g = nseg; g-- > 0;) {          // reverse order: the merge is commutative
        Partial p = scan_segment_emul(pay, len, g, lanes);
        sum += p.sum; tok += p.tok;
    }
    std::memset(d, 0, sizeof *d);
    DevTable dt;
    make_dev_table(*t, dt);
    decide(&dt, name, nlen, len, sum, tok, d);
}

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

// one (request, segment) task walked the way the kernels walk it.  A block k of the segment belongs
// to stripe k % width (width = 8 for payloads of at most 16 blocks, else 32) and is folded by lane
// k % min(lanes, width): the full warp's lane l owns blocks l, l + 32, ... (scan_segment), and for a
// short payload lanes 0..7 fold blocks l and l + 8; a lane group of 4 or 8 lanes (scan_tiny) gives
// lane l the blocks l, l + G, ... and with them the one or two stripes those fall into.  Each lane
// keeps its own stripes and token counter; the merge is a commutative sum.
Partial scan_segment_emul(const uint8_t* pay, uint32_t len, uint32_t seg, uint32_t lanes)
{
    const uint32_t nblk = (len >> 4) + ((len & 15u) ? 1u : 0u);
    const uint32_t first = seg * kSegBlocks;
    const uint32_t nb = (nblk - first) < kSegBlocks ? (nblk - first) : kSegBlocks;
    const uint32_t width = nblk <= kTinyBlocks ? kTinyStripes : kStripes;
    const uint32_t owners = lanes < width ? lanes : width;
    Partial p{0, 0};
    for (uint32_t lane = 0; lane < owners; ++lane) {
        Stripe st[2];
        bool used[2] = {false, false};
        uint32_t tok128 = 0;
        for (uint32_t k = lane; k < nb; k += owners) {
            // the device reads a full 16-byte block even past the payload end (arena padding
            // holds garbage there): emulate with 0xEE filler, then mask as the kernel does
            uint8_t raw[16];
            std::memset(raw, 0xEE, 16);
            const uint32_t valid = len - (first + k) * 16;
            std::memcpy(raw, pay + (size_t)(first + k) * 16, valid < 16 ? valid : 16);
            Words4 w;
            std::memcpy(&w, raw, 16);
            if (valid < 16) mask_tail(w, valid);
            const uint32_t s = k % width;                      // the block's stripe
            const uint32_t slot = (s / owners) & 1u;           // a lane owns width / owners <= 2 stripes
            if (!used[slot]) { stripe_init(st[slot], seg * kStripes + s); used[slot] = true; }
            stripe_round(st[slot], w);
            scan_block(w, 1u, tok128);
        }
        for (int slot = 0; slot < 2; ++slot)
            if (used[slot]) p.sum += stripe_digest(st[slot]);
        p.tok += tok128 >> 7;
    }
    return p;
}
}  // namespace

// lanes = 32 (full warp) or 4 / 8 (the tiny-payload lane groups: payloads of at most 16 blocks)
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

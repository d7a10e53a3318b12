// pk_kernels.cuh — launch interface between the C-ABI shim (pk_abi.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pk_route.h"
#include "pk_spec.cuh"

namespace pk {

// Per-request accumulator used only by requests longer than one 16 KiB segment, whose
// segments are scanned by different warps.  All-zero between launches: the warp that
// finalises a request reads its entry with atomicExch(…, 0).
struct __align__(32) SegScratch {
    unsigned long long hash_sum;   // sum of stripe digests so far
    uint32_t tok;                  // est_tokens partial sum
    uint32_t arrived;              // segments done
    uint32_t pad[4];
};
static_assert(sizeof(SegScratch) == 32, "SegScratch is 32 bytes");

struct RouteParams {
    const uint8_t*  payload;     // arena
    const uint64_t* pay_off;     // n+1, each a multiple of 16
    const uint32_t* pay_len;     // n
    const uint8_t*  names;
    const uint32_t* name_off;    // n+1
    const uint2*    extra;       // (request, segment) for segments >= 1 of long requests
    const DevTable* table;       // the routing table in global memory (launch_route also passes it by value)
    SegScratch*     scratch;     // n entries, zero
    pk_decision*    out;         // n
    uint32_t        n;
    uint32_t        n_extra;
    uint32_t        reqs_per_cta;   // R, a multiple of (warps per CTA) x (requests per warp pass): set by launch_route
    uint32_t        n_req_ctas;     // ceil(n / R): set by launch_route (CTAs past it take the long requests' extra segments)
    uint32_t        one;            // the integer 1 (see pk_spec.cuh: keeps adds on the FMA pipe)
    uint32_t        pdl;            // launched with programmatic stream serialization: set by launch_route
};

// Chooses the lane-group width, requests per CTA and grid, and launches the fused
// scan+hash+select kernel on `stream`.  max_len: the longest payload of the batch.
// pdl: launch with programmatic stream serialization (programmatic dependent launch) — the grid
// may start while the previous kernel of the stream is still running; see route_kernel.
// table: the host copy of *p.table — it also travels as a kernel parameter (constant bank).
cudaError_t launch_route(RouteParams p, const DevTable& table, uint32_t max_len, int num_sms, cudaStream_t stream, bool pdl = false);

// Several staged batches in ONE launch (grid.y = batch): what a router with a queue of batches does
// in throughput mode — one launch latency, one ramp and one tail for all of them.  The batches'
// parameter blocks travel by value (constant bank), like the table.  All batches of a launch must
// resolve to the same kernel shape (lane-group width, blocks per lane, latency mode), which they do
// when their longest payloads fall in the same class; otherwise *launched = false, nothing was
// issued and the caller launches them one by one.  No batch may appear twice.
constexpr uint32_t kMaxFused = 32;
struct FusedParams {
    RouteParams b[kMaxFused];
    uint32_t    m;              // batches in this launch
    uint32_t    pdl;
};
cudaError_t launch_route_fused(const RouteParams* batches, const uint32_t* max_len, uint32_t m, const DevTable& table,
                               int num_sms, cudaStream_t stream, bool pdl, bool* launched);

}  // namespace pk

// pk_kernels.cuh — launch interface between the C-ABI shim (pk_abi.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pk_route.h"

namespace pk {

// Per-request accumulator used only by requests longer than one 16 KiB segment, whose
// segments are scanned by different warps.  All-zero between launches: the warp that
// finalises a request reads its entry with atomicExch(…, 0).
struct __align__(32) SegScratch {
    unsigned long long hash_sum;
    uint32_t tok;        // est_tokens partial sum
    uint32_t orbits;     // OR of all payload words
    uint32_t arrived;    // segments done
    uint32_t pad[3];
};
static_assert(sizeof(SegScratch) == 32, "SegScratch is 32 bytes");

struct RouteParams {
    const uint8_t*  payload;     // arena
    const uint64_t* pay_off;     // n+1, each a multiple of 16
    const uint32_t* pay_len;     // n
    const uint8_t*  names;
    const uint32_t* name_off;    // n+1
    const uint2*    extra;       // (request, segment) for segments >= 1 of long requests
    const pk_table* table;       // device-resident routing table
    SegScratch*     scratch;     // n entries, zero
    pk_decision*    out;         // n
    uint32_t        n;
    uint32_t        n_extra;
};

// Chooses grid/block and launches the fused scan+hash+select kernel on `stream`.
cudaError_t launch_route(const RouteParams& p, int num_sms, cudaStream_t stream);

}  // namespace pk

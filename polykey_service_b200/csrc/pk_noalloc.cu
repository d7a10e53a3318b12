// pk_noalloc.cu — pk_route_device_noalloc(): the routing kernel for callers that must not
// allocate.
//
// Who needs this.  Every other entry point works on buffers that pk_init() obtained with
// cudaMalloc / cudaHostAlloc.  Some hosts of this library cannot do that:
//   * a process whose GPU memory is owned outright by something else — an inference engine
//     that reserved all of HBM for weights and KV cache at start-up, or a pool allocator
//     with a fixed budget.  A cudaMalloc from a side library either fails there or eats
//     memory the owner has already promised away; such a caller carves its request arenas
//     out of its own reservation and can hand us pointers, but nothing more;
//   * a latency-critical flusher thread: cudaMalloc and cudaFree synchronise the whole
//     device, so one allocation on the routing path stalls every stream of the process.
// For them this file holds the only device memory the kernel needs beyond its inputs and
// outputs — the per-request scratch entries that long (> 16 KiB) requests are merged in,
// and the 1.8 KB routing table — in file-scope __device__ storage that exists from the
// moment the library is loaded.  The caller passes device pointers for everything else.
//
// Cost of that storage: PK_NOALLOC_MAX_REQUESTS * 32 bytes = 2 MiB (plus the table) on
// EVERY GPU the process touches with this library loaded, held until the process exits,
// whether or not this entry point is ever called.
#include <cuda_runtime.h>

#include <mutex>

#include "../../include/pk_route.h"
#include "pk_kernels.cuh"
#include "pk_spec.cuh"

namespace {

// The scratch the long-request path of route_kernel accumulates in: one 32-byte SegScratch
// per request.  In host code these names are only handles for cudaGetSymbolAddress — they
// are NOT device addresses and must never be passed to a kernel directly.
static __device__ __align__(256) unsigned char g_scratch[(size_t)PK_NOALLOC_MAX_REQUESTS * sizeof(pk::SegScratch)];
static __device__ __align__(256) pk::DevTable g_table;

constexpr int kMaxDevices = 64;

struct DeviceState {
    bool ready = false;
    pk::SegScratch* scratch = nullptr;    // device address of g_scratch on this device
    pk::DevTable*   table = nullptr;      // device address of g_table on this device
    int             num_sms = 0;
};

std::mutex  g_mu;
DeviceState g_dev[kMaxDevices];

// Once per device, with that device current: resolve the device addresses of the two
// file-scope arrays.  (A __device__ variable has one instance per device; the address
// differs from device to device, hence the per-device table.)
int setup_current_device(int dev, DeviceState** out)
{
    std::lock_guard<std::mutex> g(g_mu);
    DeviceState& s = g_dev[dev];
    if (!s.ready) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return PK_ERR_CUDA; }
        s.scratch = static_cast<pk::SegScratch*>(p);
        e = cudaGetSymbolAddress(&p, g_table);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return PK_ERR_CUDA; }
        s.table = static_cast<pk::DevTable*>(p);
        e = cudaDeviceGetAttribute(&s.num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return PK_ERR_CUDA; }
        s.ready = true;
    }
    *out = &s;
    return PK_OK;
}

}  // namespace

extern "C" int pk_route_device_noalloc(const pk_table* table,
                                              const uint8_t* d_payload, const uint64_t* d_pay_off,
                                              const uint32_t* d_pay_len, const uint8_t* d_names,
                                              const uint32_t* d_name_off, const uint32_t* d_extra,
                                              uint32_t n_extra, uint32_t n, uint32_t max_payload_len,
                                              pk_decision* d_out, void* stream)
{
    if (n == 0) return PK_OK;
    if (!table || !d_payload || !d_pay_off || !d_pay_len || !d_name_off || !d_out) return PK_ERR_ARG;
    if (n_extra != 0 && !d_extra) return PK_ERR_ARG;
    if (table->n_tools > PK_MAX_TOOLS || table->n_providers > PK_MAX_PROVIDERS || table->default_class >= 32)
        return PK_ERR_ARG;
    for (uint32_t k = 0; k < table->n_tools; ++k)
        if (table->tools[k].name_len > PK_MAX_TOOL_NAME || table->tools[k].class_id >= 32) return PK_ERR_ARG;
    for (uint32_t p = 0; p < table->n_providers; ++p)
        if (table->providers[p].weight > PK_MAX_WEIGHT) return PK_ERR_ARG;
    // the static scratch has room for PK_NOALLOC_MAX_REQUESTS requests and no more
    if (n > PK_NOALLOC_MAX_REQUESTS) return PK_ERR_CAPACITY;

    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return PK_ERR_NO_DEVICE; }
    if (dev < 0 || dev >= kMaxDevices) return PK_ERR_NO_DEVICE;
    DeviceState* ds = nullptr;
    int rc = setup_current_device(dev, &ds);
    if (rc) return rc;

    cudaStream_t st = static_cast<cudaStream_t>(stream);

    // Table: host -> the static device copy.  The source is pageable stack memory, which
    // cudaMemcpyAsync stages before it returns, so `dt` may go out of scope afterwards.  (This is
    // also why the header forbids calling under stream capture: a recorded copy would read `dt`
    // long after this frame is gone.)
    pk::DevTable dt;
    pk::make_dev_table(*table, dt);
    if (cudaMemcpyAsync(ds->table, &dt, sizeof dt, cudaMemcpyHostToDevice, st) != cudaSuccess) {
        (void)cudaGetLastError();
        return PK_ERR_CUDA;
    }
    // Scratch entries must read zero when the kernel starts.  A completed launch leaves them
    // zero, but the array is shared by every call and keeps whatever an earlier, failed or
    // overlapping, call left behind — so clear the part this call can touch.  Only requests
    // with extra segments ever use their entry.
    if (n_extra != 0 &&
        cudaMemsetAsync(ds->scratch, 0, (size_t)n * sizeof(pk::SegScratch), st) != cudaSuccess) {
        (void)cudaGetLastError();
        return PK_ERR_CUDA;
    }

    pk::RouteParams P{};
    P.payload  = d_payload;
    P.pay_off  = d_pay_off;
    P.pay_len  = d_pay_len;
    P.names    = d_names;
    P.name_off = d_name_off;
    P.extra    = reinterpret_cast<const uint2*>(d_extra);
    P.table    = ds->table;
    P.scratch  = ds->scratch;          // the resolved device address, never the symbol itself
    P.out      = d_out;
    P.n        = n;
    P.n_extra  = n_extra;
    if (pk::launch_route(P, dt, max_payload_len, ds->num_sms, st) != cudaSuccess) {
        (void)cudaGetLastError();
        return PK_ERR_CUDA;
    }
    return PK_OK;
}

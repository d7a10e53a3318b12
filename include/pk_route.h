/*
 * pk_route.h — C ABI of the B200 batched request-routing stage (libpkroute.so).
 *
 * This is the drop-in boundary for the ExecuteTool hot path of
 * spounge-ai/polykey-service.  The reference has no FFI of its own (it is pure Go,
 * CGO_ENABLED=0, Makefile:69,71); the seam this library plugs into is the one-method
 * plugin interface
 *
 *     service.Service.ExecuteTool(ctx, toolName, parameters, secretId, metadata)
 *                                              (internal/service/service.go:13-15)
 *
 * injected at server.NewServer(polykeySvc) (cmd/polykey/main.go:85-86) and called
 * at internal/server/server.go:36.  A Go `BatchingService` decorator coalesces the
 * concurrent per-RPC calls into flat byte arenas and crosses cgo ONCE per batch
 * through the entry points below (binding shown in INTEGRATION.md).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross the boundary.
 *   - every function returns 0 (PK_OK) or a negative pk_status; nothing aborts or
 *     throws.  CUDA runtime failures are folded into PK_ERR_CUDA and the numeric
 *     cudaError_t is kept per handle (pk_last_cuda_error) — the Go side maps any
 *     non-zero code to a gRPC status exactly as server.go:37-40 passes an error
 *     through, and flips health to NOT_SERVING (cmd/polykey/main.go:93-94).
 *   - the caller owns every buffer it passes; the library owns device memory,
 *     streams and events.  Pointers must stay valid until the call (or, for
 *     pk_submit, the matching pk_wait) returns.
 *   - a handle may be used by ONE thread at a time (the batcher's flusher);
 *     different handles (one per GPU) may be used concurrently.
 *   - there is NO CPU fallback inside this library.
 *
 * Batch layout (flat SoA; see DESIGN.md "Data layout")
 *   payload   : one contiguous byte arena; request i owns
 *               payload[pay_off[i] .. pay_off[i] + pay_len[i]) where
 *               pay_len[i] = pay_end(i) - pay_off[i].  pay_off has n+1 entries;
 *               every pay_off[i] MUST be a multiple of 16 (PK_ALIGN); the bytes of
 *               request i end at pay_off[i] + pay_len[i] <= pay_off[i+1] and the
 *               gap up to pay_off[i+1] is padding whose content is ignored.
 *               Because of the padding the true length travels separately in
 *               pay_len[] (u32, one per request).
 *   names     : packed tool-name bytes, request i owns
 *               names[name_off[i] .. name_off[i+1]); no alignment requirement.
 *   out       : n fixed-size decision records (24 bytes each).
 */
#ifndef PK_ROUTE_H
#define PK_ROUTE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PK_API __attribute__((visibility("default")))
#else
#define PK_API
#endif

#define PK_ABI_VERSION        1u
#define PK_ALIGN              16u      /* payload start alignment in the arena      */
#define PK_MAX_TOOLS          16u
#define PK_MAX_TOOL_NAME      64u
#define PK_MAX_PROVIDERS      32u
#define PK_MAX_NAME_LEN       0xFFFFFFu /* a request's tool name, bytes              */
#define PK_PROVIDER_NONE      0xFFu
#define PK_MAX_WEIGHT         (1u << 20) /* provider weight upper bound              */
#define PK_NOALLOC_MAX_REQUESTS 65536u  /* batch limit of pk_route_device_noalloc    */

/* decision flags */
                                       /* 0x0001u is reserved (always 0)            */
#define PK_FLAG_NO_PROVIDER   0x0002u  /* no provider eligible; provider = 0xFF     */
#define PK_FLAG_EMPTY_PAYLOAD 0x0004u  /* pay_len == 0                              */
#define PK_FLAG_UNKNOWN_TOOL  0x0008u  /* name matched no table entry: the default
                                          branch of the switch, mock.go:60-64 —
                                          still status 200, never an error          */

typedef enum pk_status {
    PK_OK              =  0,
    PK_ERR_ARG         = -1,   /* null pointer / bad count / bad table              */
    PK_ERR_ALIGN       = -2,   /* a pay_off[i] is not a multiple of PK_ALIGN        */
    PK_ERR_OFFSETS     = -3,   /* offsets not monotone, or pay_len overruns a slot  */
    PK_ERR_CAPACITY    = -4,   /* batch exceeds the limits given to pk_init         */
    PK_ERR_CUDA        = -5,   /* CUDA runtime error; see pk_last_cuda_error        */
    PK_ERR_NO_DEVICE   = -6,   /* no usable CUDA device / device ordinal invalid    */
    PK_ERR_BUSY        = -7,   /* slot has a submit in flight (missing pk_wait)     */
    PK_ERR_STATE       = -8    /* call not valid in this state (e.g. wait w/o submit)*/
} pk_status;

/* One entry of the tool table: the `case "...":` labels of the switch at
 * internal/service/mock.go:32-59.  Matching is whole-name, case-sensitive byte
 * equality (Go string switch semantics); the first matching entry wins. */
typedef struct pk_tool {
    uint8_t  name[PK_MAX_TOOL_NAME];
    uint32_t name_len;
    uint32_t class_id;            /* < 32 */
} pk_tool;

/* One provider row of the device-resident routing table (build-defined: the
 * reference has no provider list — SURVEY.md §0.3, §8 a6). */
typedef struct pk_provider {
    uint32_t max_tokens;          /* eligible iff est_tokens <= max_tokens          */
    uint32_t weight;              /* share of traffic among the eligible providers;
                                     0 disables the provider; <= PK_MAX_WEIGHT      */
    uint32_t tool_mask;           /* bit c set: serves tool class c                 */
    uint32_t reserved;            /* ignored                                        */
} pk_provider;

typedef struct pk_table {
    uint32_t    n_tools;          /* <= PK_MAX_TOOLS                                */
    uint32_t    default_class;    /* class of the `default:` branch, mock.go:60     */
    uint32_t    n_providers;      /* <= PK_MAX_PROVIDERS                            */
    uint32_t    reserved;
    pk_tool     tools[PK_MAX_TOOLS];
    pk_provider providers[PK_MAX_PROVIDERS];
} pk_table;

/* 24-byte decision record, one per request. */
typedef struct pk_decision {
    uint32_t byte_len;            /* payload length                                 */
    uint32_t est_tokens;          /* token-count estimate (SPEC §B)                 */
    uint64_t route_hash;          /* 64-bit route key hash (SPEC §C)                */
    uint8_t  tool_class;          /* mock.go:32-64                                  */
    uint8_t  provider;            /* index into table.providers, or 0xFF            */
    uint16_t flags;               /* PK_FLAG_*                                      */
    uint32_t status;              /* always 200: Status.Code, mock.go:25-28         */
} pk_decision;

typedef struct pk_config {
    uint32_t max_requests;        /* per batch                                      */
    uint32_t n_slots;             /* independent in-flight batches (>=1)            */
    uint64_t max_payload_bytes;   /* arena bytes per batch (incl. padding)          */
    uint64_t max_name_bytes;      /* name arena bytes per batch                     */
} pk_config;

typedef struct pk_handle pk_handle;

/* Fills *t with the reference table: example_tool→0, struct_tool→1, file_tool→2,
 * default→3 (mock.go:33,37,52,60) and 8 default providers.               */
PK_API int  pk_default_table(pk_table* t);

/* Replaces server start-up wiring at cmd/polykey/main.go:85-86: creates the CUDA
 * context on `device`, uploads the routing table, allocates every device buffer,
 * stream and event for cfg->n_slots batches.  Nothing is allocated afterwards.   */
PK_API int  pk_init(int device, const pk_table* tbl, const pk_config* cfg, pk_handle** out);

/* Teardown hook for cmd/polykey/main.go:118-119 (before GracefulStop returns).    */
PK_API void pk_destroy(pk_handle* h);

/* Re-uploads the routing table (stream-ordered after all submitted work).         */
PK_API int  pk_update_table(pk_handle* h, const pk_table* tbl);

/* THE hot call — the batched replacement for N calls of
 * MockService.ExecuteTool's routing decision (internal/service/mock.go:22-67):
 * H2D copies, the fused scan+hash+select kernel, D2H copy of the records, all on
 * slot 0's stream, synchronous.  Host pointers.                                  */
PK_API int  pk_route_batch(pk_handle* h,
                    const uint8_t* payload, const uint64_t* pay_off /* n+1 */,
                    const uint32_t* pay_len /* n */,
                    const uint8_t* names, const uint32_t* name_off /* n+1 */,
                    uint32_t n, pk_decision* out /* n */);

/* Batches up to a size skip the copies.  When the bytes of a batch's payloads (gaps between
 * requests not counted) + its metadata come to at most `bytes` (default 16 MiB; 0 turns the path
 * off), and `payload` and `out` are page-locked host memory the GPU can address (pk_host_alloc, or any
 * cudaHostAlloc / cudaHostRegister'ed block), pk_submit / pk_route_batch launch the kernel
 * directly on the caller's buffers: it reads the bytes over PCIe and writes the records into
 * `out` — one launch instead of copy, copy, launch, copy.  Same records; the buffers must stay
 * untouched until pk_wait returns, as in the copying path.  The path also needs the first byte
 * the batch uses (payload + pay_off[0]) on a 16-byte boundary and `out` on an 8-byte one; a
 * batch that is not so aligned is simply copied.  pk_direct_count: batches so routed. */
PK_API int      pk_set_direct_bytes(pk_handle* h, uint64_t bytes);
PK_API uint64_t pk_direct_count(const pk_handle* h);

/* Asynchronous form of the same call for a pipelined flusher: pk_submit enqueues
 * H2D → kernel → D2H on slot `slot`'s stream and returns; pk_wait blocks until
 * that slot's records are in `out`.  Use pinned buffers (pk_host_alloc) or the
 * copies degrade to synchronous staging inside the CUDA driver.                 */
PK_API int  pk_submit(pk_handle* h, uint32_t slot,
               const uint8_t* payload, const uint64_t* pay_off, const uint32_t* pay_len,
               const uint8_t* names, const uint32_t* name_off,
               uint32_t n, pk_decision* out);
PK_API int  pk_wait(pk_handle* h, uint32_t slot);

/* Device-resident form (used to time the kernels alone, inputs already in HBM):
 * pk_stage copies a batch into slot `slot` and returns when it is resident;
 * pk_launch runs the kernels over the staged batch (no copies, asynchronous on the
 * slot's stream); pk_launch_many issues `count` launches back to back, walking the
 * slot list round robin — launch k uses slots[(p + k) % n_slots_in_list], where p is where
 * the handle's previous pk_launch_many call stopped, so that a short call does not start on
 * the batches the call before it left in L2 — all on ONE stream (slot slots[0]'s), either
 * as `count` plain launches or (use_graph == 1) as one CUDA graph of `count` kernel
 * nodes in a chain, or (use_graph = k >= 2) as a graph of k parallel chains — a launch runs in
 * chain (its position in the slot list) % k, so a slot is always in the same chain — so that
 * independent batches may overlap on the GPU; it brackets them with
 * CUDA events on that stream, waits, and reports the device time in *elapsed_ms;
 * use_graph | PK_LAUNCH_PDL issues the same launches with programmatic stream serialization
 * (programmatic dependent launch): still one stream / one chain, but a kernel may start while
 * its predecessor drains — batches are independent; a batch with long requests, which owns
 * per-request scratch in its slot, is launched without the attribute (it starts after its
 * predecessor completed).  Stream order for what follows is kept.
 * use_graph | PK_LAUNCH_FUSE(m), 2 <= m <= 32 and m <= n_slots_in_list: every m consecutive
 * launches of the walk become ONE kernel launch over m staged batches (grid.y = batch) — what a
 * router with a queue of ready batches does in throughput mode: one launch latency, one ramp and
 * one tail for all of them; the records are the same as from m launches.  Batches whose longest
 * payloads fall in different kernel classes (<= 256 bytes, or longer) are launched one by one
 * instead.  Fused launches run in one chain (a graph width is ignored); PK_LAUNCH_PDL applies
 * between them.  pk_launch_count counts kernel launches, so it shows which happened.
 * pk_fetch copies the slot's records out (synchronous).                               */
#define PK_LAUNCH_PDL 0x100
#define PK_LAUNCH_FUSE(m) (((m) & 0xFF) << 12)
PK_API int  pk_stage(pk_handle* h, uint32_t slot,
              const uint8_t* payload, const uint64_t* pay_off, const uint32_t* pay_len,
              const uint8_t* names, const uint32_t* name_off, uint32_t n);
PK_API int  pk_launch(pk_handle* h, uint32_t slot);
PK_API int  pk_launch_many(pk_handle* h, const uint32_t* slots, uint32_t n_slots_in_list,
                    uint32_t count, int use_graph, float* elapsed_ms /* may be NULL */);
PK_API int  pk_fetch(pk_handle* h, uint32_t slot, pk_decision* out /* staged n */);
PK_API int  pk_sync(pk_handle* h);                 /* waits for every slot's stream        */

/* The cudaStream_t of a slot, as an opaque pointer, so a caller can record its own
 * CUDA events on the launching stream (bench.py does).                            */
PK_API void* pk_slot_stream(pk_handle* h, uint32_t slot);

/* Number of kernel launches this handle has issued (bench.py's gpu_launches).     */
PK_API uint64_t pk_launch_count(const pk_handle* h);

/* Pinned host memory owned by the C side: gRPC buffers live on the Go heap and
 * cannot be DMA sources, so the batcher copies each request into an arena from
 * here (SURVEY.md §7 "Pinned host memory from Go").                              */
PK_API void* pk_host_alloc(pk_handle* h, size_t bytes);
PK_API void  pk_host_free(pk_handle* h, void* p);

/* Measurement aid (no counterpart in the reference): how fast can THIS GPU pull page-locked host
 * memory with plain 128-bit loads from a kernel — the ceiling the in-place path of pk_submit
 * runs under.  `host` must be page-locked, GPU-addressable and 16-byte aligned (pk_host_alloc);
 * `reps` back-to-back read passes over [host, host + bytes) on slot 0's stream are timed with
 * CUDA events; *elapsed_ms = their device time.  Slot 0 must be idle.                       */
PK_API int pk_probe_host_read(pk_handle* h, const void* host, size_t bytes, uint32_t reps, float* elapsed_ms);

/* pk_route_device_noalloc — routing for callers that cannot allocate GPU memory.
 * (No counterpart in the reference: it has no device code at all.)
 *
 * Everything above runs on buffers pk_init() allocates.  That does not work for a host
 * process whose GPU memory already belongs to something else (an inference engine that
 * reserved all of HBM at start-up, a fixed-budget pool), nor on a thread that must never
 * hit the device-wide synchronisation cudaMalloc / cudaFree imply.  This entry point makes
 * no allocation of any kind, on the device or on the host, and needs no handle: the caller
 * passes DEVICE pointers to a batch it has already placed in its own memory and to the
 * record array to fill; the only other device memory the kernel needs — one 32-byte scratch
 * entry per request, used by requests longer than 16 KiB, and the routing table — lives in
 * static storage inside the library: PK_NOALLOC_MAX_REQUESTS * 32 bytes = 2 MiB per GPU
 * that loads the library, held until the process exits.
 *
 *   table            HOST pointer; copied into the library's static device copy each call
 *   d_payload .. d_name_off   DEVICE pointers, same layout and alignment rules as
 *                    pk_route_batch (pay_off[i] % 16 == 0, offsets relative to d_payload /
 *                    d_names); d_payload must be 16-byte aligned and readable up to the
 *                    end of the last request's final 16-byte block
 *   d_extra, n_extra DEVICE array of n_extra (request, segment) uint32 pairs: one pair for
 *                    every segment s >= 1 of every request longer than 16384 bytes
 *                    (segment s covers bytes [16384*s, 16384*(s+1)) ); NULL / 0 if no request
 *                    is that long.  The list must match d_pay_len exactly — it is not
 *                    checked, the lengths being on the device.
 *   n                requests; more than PK_NOALLOC_MAX_REQUESTS is refused with
 *                    PK_ERR_CAPACITY (the static scratch has no room for them)
 *   max_payload_len  an upper bound of the payload lengths (0xFFFFFFFF if unknown)
 *   d_out            DEVICE pointer to n records
 *   stream           cudaStream_t (as void*) of the CURRENT device; the work is enqueued
 *                    there and the call returns without waiting — synchronise the stream
 *                    before reading d_out
 *
 * The first call on each device resolves the static storage's device address
 * (cudaGetSymbolAddress).  Never call this entry point on a stream that is being captured —
 * not the first call, not a later one: every call copies the table from a variable on its own
 * stack, which is gone by the time anything recorded would run.
 * All calls share the one static scratch array and table copy: two calls whose GPU work
 * can overlap (different streams, different threads) race on them — serialise them on one
 * stream.  The part of the scratch a call can touch is cleared by that call before its
 * kernel runs, so a call is not affected by what earlier calls left behind.             */
PK_API int pk_route_device_noalloc(const pk_table* table,
                                   const uint8_t* d_payload, const uint64_t* d_pay_off,
                                   const uint32_t* d_pay_len, const uint8_t* d_names,
                                   const uint32_t* d_name_off, const uint32_t* d_extra,
                                   uint32_t n_extra, uint32_t n, uint32_t max_payload_len,
                                   pk_decision* d_out, void* stream);

PK_API const char* pk_strerror(int status);
PK_API int         pk_last_cuda_error(const pk_handle* h);   /* cudaError_t of last failure */
PK_API uint32_t    pk_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PK_ROUTE_H */

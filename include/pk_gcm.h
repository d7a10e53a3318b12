/*
 * pk_gcm.h — C ABI of the batched AES-256-GCM stage (part of libpkroute.so).
 *
 * Widening step after the routing path.  Replaces the loops of
 *   security.BatchEncrypt / security.BatchDecrypt   internal/adapters/security/cipher.go:110-141
 * (and, for a batch of one, Encrypt / Decrypt, cipher.go:92-104) with one GPU launch per batch.
 * Same cryptography — AES-256-GCM, 12-byte nonce, no additional data, 16-byte tag — and the same
 * wire format: sealed message = nonce(12) || ciphertext || tag(16)   (cipher.go:51-52, 73-77).
 *
 * Differences a caller must know:
 *   - nonces are an INPUT of pk_gcm_seal_batch (the reference draws them from crypto/rand,
 *     cipher.go:25-31, which makes its output unrepeatable); the host mirror draws them from the OS;
 *   - the key schedule and GCM tables are built once per handle, not once per message;
 *   - messages travel in flat arenas with 16-byte aligned slots, like the routing batches;
 *   - BatchDecrypt's fail-fast error becomes: PK_GCM_ERR_AUTH for the batch, plus a per-message
 *     status so the caller can tell which one failed; no plaintext is released for a failed one.
 * There is no CPU fallback.  A handle is used by one thread at a time.
 */
#ifndef PK_GCM_H
#define PK_GCM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PK_GCM_API __attribute__((visibility("default")))
#else
#define PK_GCM_API
#endif

#define PK_GCM_NONCE_BYTES 12u
#define PK_GCM_TAG_BYTES   16u
#define PK_GCM_OVERHEAD    28u

typedef enum pk_gcm_status {
    PK_GCM_OK            =  0,
    PK_GCM_ERR_ARG       = -1,
    PK_GCM_ERR_ALIGN     = -2,   /* a slot offset is not a multiple of 16                     */
    PK_GCM_ERR_OFFSETS   = -3,   /* slots overlap or are not in increasing order              */
    PK_GCM_ERR_CAPACITY  = -4,
    PK_GCM_ERR_CUDA      = -5,
    PK_GCM_ERR_NO_DEVICE = -6,
    PK_GCM_ERR_KEYLEN    = -9,   /* validateKey, cipher.go:15-19: key must be 32 bytes        */
    PK_GCM_ERR_AUTH      = -10   /* at least one message failed to open (cipher.go:78-81)     */
} pk_gcm_status;

typedef struct pk_gcm pk_gcm;

/* Builds the per-key context (cipher.go:35-43, done once) and allocates device buffers for
 * batches of up to max_messages messages and max_bytes arena bytes (input and output each). */
PK_GCM_API int  pk_gcm_init(int device, const uint8_t* key, uint32_t key_len, uint32_t max_messages,
                            uint64_t max_bytes, pk_gcm** out);
PK_GCM_API void pk_gcm_destroy(pk_gcm* h);

/* BatchEncrypt (cipher.go:110-124).  Message i: plaintext plain[plain_off[i] .. +plain_len[i]),
 * nonce nonces[12*i .. 12*i+12); its sealed form, plain_len[i] + 28 bytes, is written at
 * sealed[sealed_off[i]].  Offsets are multiples of 16 and slots must not overlap. Synchronous. */
PK_GCM_API int  pk_gcm_seal_batch(pk_gcm* h, const uint8_t* plain, const uint64_t* plain_off,
                                  const uint32_t* plain_len, const uint8_t* nonces, uint32_t n,
                                  uint8_t* sealed, const uint64_t* sealed_off);

/* BatchDecrypt (cipher.go:126-141).  status[i]: 0 opened, 1 authentication failed ("failed to
 * decrypt"), 2 shorter than a nonce ("ciphertext too short", cipher.go:69-71).  Returns
 * PK_GCM_ERR_AUTH if any status is non-zero; plaintext of a failed message is zeroed.
 * The kernels decrypt while they authenticate (one pass over the ciphertext) and wipe the output
 * of a message whose tag does not match before the call returns — so `plain` must not be read by
 * anyone until the call has returned.  That matters for the in-place path (every array page-locked,
 * small batch: the kernel writes straight into the caller's `plain`), where for the duration of the
 * call unauthenticated plaintext sits in host memory another thread could see; with the copying
 * path it only ever exists in the library's device buffer.  A caller that cannot guarantee that
 * turns the in-place path off (pk_gcm_set_direct_bytes(h, 0)).                                  */
PK_GCM_API int  pk_gcm_open_batch(pk_gcm* h, const uint8_t* sealed, const uint64_t* sealed_off,
                                  const uint32_t* sealed_len, uint32_t n, uint8_t* plain,
                                  const uint64_t* plain_off, uint32_t* status);

/* How a batch is pipelined.  0 (default): batches of 4 MiB and more in page-locked arenas are cut
 * into 2, 4 or 8 chunks of whole messages so that copy in, kernels and copy out overlap; 1: never
 * (one copy in, the kernels, one copy out — what pk_gcm_last_kernel_ms needs to time the kernels
 * alone); 2..8: that many.                                                                        */
PK_GCM_API int   pk_gcm_set_chunks(pk_gcm* h, uint32_t chunks);

/* Small batches of long messages are sealed / opened in place.  When the arena bytes in + out come
 * to at most `bytes` (default 4 MiB; 0 turns it off), every message is one the warp-per-message
 * kernel takes (256 bytes or more for batches under 8,192 messages), and EVERY array of the call —
 * arenas, offsets, lengths, nonces, status — is page-locked host memory the GPU can address
 * (pk_gcm_host_alloc), no copy is issued: the kernel works on the caller's buffers over PCIe.
 * pk_gcm_direct_count: batches so done.                                                          */
PK_GCM_API int      pk_gcm_set_direct_bytes(pk_gcm* h, uint64_t bytes);
PK_GCM_API uint64_t pk_gcm_direct_count(const pk_gcm* h);

/* Page-locked host memory for the arenas (plain, sealed, offsets, nonces): copies to and from
 * pinned memory run at PCIe rate; pageable memory is staged by the driver at about a third of it.
 * Any host pointer is accepted by the calls above; these are the fast ones.  NULL on failure.     */
PK_GCM_API void* pk_gcm_host_alloc(pk_gcm* h, size_t bytes);
PK_GCM_API void  pk_gcm_host_free(pk_gcm* h, void* p);

PK_GCM_API const char* pk_gcm_strerror(int status);
PK_GCM_API uint64_t    pk_gcm_launch_count(const pk_gcm* h);
/* Device time from the first to the last kernel of the last seal/open call (CUDA events on their
 * stream), ms; < 0 if none.  With a pipelined batch it includes waiting for the copies in.        */
PK_GCM_API float       pk_gcm_last_kernel_ms(pk_gcm* h);
PK_GCM_API int         pk_gcm_last_cuda_error(const pk_gcm* h);

#ifdef __cplusplus
}
#endif
#endif /* PK_GCM_H */

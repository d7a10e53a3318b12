/*
 * pk_aes_ref.c — CPU baseline for the batched AES-256-GCM stage: OpenSSL (libcrypto, EVP).
 * TEST INFRASTRUCTURE ONLY, like the rest of oracle/ (separate library: libpkaesref.so, so that
 * the routing oracle does not depend on libcrypto).
 *
 * Two shapes of the same work, both multi-threaded over the batch:
 *   per_item_setup = 1   what the reference does: a fresh cipher (key schedule + GCM state) for
 *                        every message — aes.NewCipher + cipher.NewGCM inside the loops of
 *                        BatchEncrypt / BatchDecrypt, internal/adapters/security/cipher.go:35-43,
 *                        116-122, 133-139
 *   per_item_setup = 0   the key set once per thread, only the nonce per message
 * Output format nonce(12) || ciphertext || tag(16), cipher.go:51-52.
 */
#include <openssl/evp.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    const uint8_t* key; const uint8_t* plain; const uint64_t* plain_off; const uint32_t* plain_len;
    const uint8_t* nonces; uint8_t* sealed; const uint64_t* sealed_off; uint32_t lo, hi; int per_item_setup; int rc;
} job_t;

static void* seal_worker(void* arg)
{
    job_t* j = (job_t*)arg;
    EVP_CIPHER_CTX* ctx = EVP_CIPHER_CTX_new();
    if (!j->per_item_setup) EVP_EncryptInit_ex(ctx, EVP_aes_256_gcm(), NULL, j->key, NULL);
    for (uint32_t i = j->lo; i < j->hi; ++i) {
        const uint8_t* nonce = j->nonces + (size_t)i * 12;
        uint8_t* out = j->sealed + j->sealed_off[i];
        int n = 0, m = 0;
        if (j->per_item_setup) {
            EVP_CIPHER_CTX_reset(ctx);
            EVP_EncryptInit_ex(ctx, EVP_aes_256_gcm(), NULL, j->key, nonce);
        } else {
            EVP_EncryptInit_ex(ctx, NULL, NULL, NULL, nonce);
        }
        memcpy(out, nonce, 12);
        if (j->plain_len[i]) EVP_EncryptUpdate(ctx, out + 12, &n, j->plain + j->plain_off[i], (int)j->plain_len[i]);
        EVP_EncryptFinal_ex(ctx, out + 12 + n, &m);
        EVP_CIPHER_CTX_ctrl(ctx, EVP_CTRL_GCM_GET_TAG, 16, out + 12 + j->plain_len[i]);
    }
    EVP_CIPHER_CTX_free(ctx);
    return NULL;
}

int pk_aes_ref_seal_batch(const uint8_t* key, const uint8_t* plain, const uint64_t* plain_off,
                          const uint32_t* plain_len, const uint8_t* nonces, uint32_t n, uint8_t* sealed,
                          const uint64_t* sealed_off, uint32_t nthreads, int per_item_setup)
{
    if (n == 0) return 0;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    if (nthreads > n) nthreads = n;
    pthread_t th[256];
    job_t jobs[256];
    for (uint32_t k = 0; k < nthreads; ++k) {
        jobs[k] = (job_t){key, plain, plain_off, plain_len, nonces, sealed, sealed_off,
                          (uint32_t)(((uint64_t)n * k) / nthreads), (uint32_t)(((uint64_t)n * (k + 1)) / nthreads), per_item_setup, 0};
        if (nthreads == 1) seal_worker(&jobs[0]);
        else pthread_create(&th[k], NULL, seal_worker, &jobs[k]);
    }
    if (nthreads > 1) for (uint32_t k = 0; k < nthreads; ++k) pthread_join(th[k], NULL);
    return 0;
}

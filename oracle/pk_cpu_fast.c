/*
 * pk_cpu_fast.c — an efficient CPU port of the same routing SPEC, used ONLY as the timed CPU
 * baseline (bench.py cpu_baseline / --impl reference).  TEST INFRASTRUCTURE, like the rest of
 * oracle/.  The checker stays pk_oracle.c (byte at a time, obviously correct); this file is
 * checked against it in tests/test_oracle_golden.py.  It is what a careful C programmer would
 * write for one core: a 64-bit SIMD-in-word scan (SPEC §B), the 32 stripe states of a
 * segment kept hot and walked block-major (SPEC §C), 4-byte loads.  No SIMD intrinsics; gcc -O3
 * may vectorise what it can.  The tool switch is internal/service/mock.go:32-64.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/pk_route.h"

#define P1 0x9E3779B1u
#define P2 0x85EBCA77u
#define P5 0x165667B1u
#define NAME_MUL 0xD6E8FEB86659FD93ull

static inline uint32_t rotl32(uint32_t x, unsigned r) { return (x << r) | (x >> (32u - r)); }
static inline uint64_t fmix64(uint64_t x)
{
    x ^= x >> 33; x *= 0xFF51AFD7ED558CCDull;
    x ^= x >> 33; x *= 0xC4CEB9FE1A85EC53ull;
    x ^= x >> 33;
    return x;
}

/* SPEC §B eight bytes at a time (64-bit SIMD-in-word).  A letter run is cut at every offset
 * j % 8 == 0, i.e. at byte 0 of every 8-byte word, so words are independent of each other. */
#define K64(b) (0x0101010101010101ull * (uint64_t)(b))
static inline uint32_t scan_word64(uint64_t x)
{
    const uint64_t t  = x & K64(0x5F);            /* folded: bits 5 and 7 cleared */
    const uint64_t a1 = t + K64(0x3F);            /* bit7: folded >= 0x41 */
    const uint64_t a2 = t + K64(0x25);            /* bit7: folded >= 0x5B */
    const uint64_t g  = t + K64(0x7F);            /* bit7: folded != 0 (not NUL / space) */
    const uint64_t lm = a1 & ~a2 & ~x;            /* ASCII letter */
    const uint64_t c  = lm & (lm << 8);           /* letter preceded by a letter in this word */
    const uint64_t m  = ((g & ~x & ~c) | (x & (x << 1))) & K64(0x80);
    return (uint32_t)__builtin_popcountll(m);
}

static uint32_t scan_tokens(const uint8_t* p, uint32_t len, int* nonascii)
{
    uint32_t tok = 0, j = 0;
    for (; j + 8 <= len; j += 8) {
        uint64_t x;
        memcpy(&x, p + j, 8);
        tok += scan_word64(x);
    }
    if (j < len) {
        uint64_t x = 0;                            /* zero bytes are separators: count nothing */
        memcpy(&x, p + j, len - j);
        tok += scan_word64(x);
    }
    *nonascii = 0;
    return tok;
}

static uint64_t payload_sum(const uint8_t* p, uint32_t len)
{
    uint32_t nblk = (len >> 4) + ((len & 15u) ? 1u : 0u);
    uint64_t sum = 0;
    for (uint32_t first = 0, seg = 0; first < nblk; first += 1024u, ++seg) {
        uint32_t nb = nblk - first < 1024u ? nblk - first : 1024u;
        uint32_t a[32][4];
        uint32_t ns = nb < 32u ? nb : 32u;
        for (uint32_t s = 0; s < ns; ++s) {
            uint32_t base = (seg * 32u + s) * P5;
            a[s][0] = P1 + P2 + base; a[s][1] = P2 + base; a[s][2] = base; a[s][3] = base - P1;
        }
        const uint32_t full = (len >> 4) - first < nb ? (len >> 4) - first : nb;   /* whole blocks */
        for (uint32_t r = 0; r < full; ++r) {
            uint32_t w[4];
            memcpy(w, p + (uint64_t)(first + r) * 16u, 16);
            uint32_t* st = a[r & 31u];
            for (int j = 0; j < 4; ++j) st[j] = rotl32(st[j] + w[j] * P2, 13) * P1;
        }
        if (full < nb) {                                   /* the payload's last, partial block */
            uint32_t w[4] = {0, 0, 0, 0};
            memcpy(w, p + (uint64_t)(first + full) * 16u, len & 15u);
            uint32_t* st = a[full & 31u];
            for (int j = 0; j < 4; ++j) st[j] = rotl32(st[j] + w[j] * P2, 13) * P1;
        }
        for (uint32_t s = 0; s < ns; ++s) {
            uint32_t lo = rotl32(a[s][0], 1) + rotl32(a[s][1], 7) + rotl32(a[s][2], 12) + rotl32(a[s][3], 18);
            uint32_t hi = rotl32(a[s][0], 24) + rotl32(a[s][1], 3) + rotl32(a[s][2], 19) + rotl32(a[s][3], 10);
            sum += fmix64(((uint64_t)hi << 32) | lo);
        }
    }
    return sum;
}

static void route_one(const pk_table* t, const uint8_t* p, uint32_t len, const uint8_t* name,
                      uint32_t nlen, pk_decision* d)
{
    int nonascii = 0, unknown = 1;
    uint32_t cls = t->default_class;
    for (uint32_t k = 0; k < t->n_tools; ++k)                         /* mock.go:32-64 */
        if (t->tools[k].name_len == nlen && memcmp(t->tools[k].name, name, nlen) == 0) {
            cls = t->tools[k].class_id; unknown = 0; break;
        }
    uint32_t tok = scan_tokens(p, len, &nonascii);
    uint64_t s = payload_sum(p, len);
    for (uint32_t j = 0; j < nlen; ++j) {
        uint32_t x = ((j << 8) | name[j]) + 1u;
        x *= P1; x ^= x >> 15; x *= P2; x ^= x >> 13;
        s += (uint64_t)x * NAME_MUL;
    }
    uint64_t h = fmix64(s + ((uint64_t)nlen << 32) + len);
    uint64_t total = 0;
    for (uint32_t q = 0; q < t->n_providers; ++q) {
        const pk_provider* e = &t->providers[q];
        if (e->weight != 0 && ((e->tool_mask >> cls) & 1u) && tok <= e->max_tokens) total += e->weight;
    }
    uint32_t best = PK_PROVIDER_NONE;
    if (total) {
        uint64_t x = ((h >> 32) * total) >> 32, running = 0;
        for (uint32_t q = 0; q < t->n_providers; ++q) {
            const pk_provider* e = &t->providers[q];
            if (!(e->weight != 0 && ((e->tool_mask >> cls) & 1u) && tok <= e->max_tokens)) continue;
            running += e->weight;
            if (running > x) { best = q; break; }
        }
    }
    d->byte_len = len; d->est_tokens = tok; d->route_hash = h;
    d->tool_class = (uint8_t)cls; d->provider = (uint8_t)best;
    (void)nonascii;
    d->flags = (uint16_t)((best == PK_PROVIDER_NONE ? PK_FLAG_NO_PROVIDER : 0) |
                          (len == 0 ? PK_FLAG_EMPTY_PAYLOAD : 0) | (unknown ? PK_FLAG_UNKNOWN_TOOL : 0));
    d->status = 200;                                                  /* mock.go:25-28 */
}

typedef struct {
    const pk_table* t; const uint8_t* payload; const uint64_t* pay_off; const uint32_t* pay_len;
    const uint8_t* names; const uint32_t* name_off; pk_decision* out; uint32_t lo, hi;
} job_t;

static void* worker(void* arg)
{
    job_t* j = (job_t*)arg;
    for (uint32_t i = j->lo; i < j->hi; ++i)
        route_one(j->t, j->payload + j->pay_off[i], j->pay_len[i], j->names + j->name_off[i],
                  j->name_off[i + 1] - j->name_off[i], &j->out[i]);
    return NULL;
}

/* No argument validation: callers (bench.py, tests) pass batches already accepted by
 * pk_ref_route_batch.  Static partition by request count over nthreads. */
int pk_cpu_route_batch_mt(const pk_table* t, const uint8_t* payload, const uint64_t* pay_off,
                          const uint32_t* pay_len, const uint8_t* names, const uint32_t* name_off,
                          uint32_t n, pk_decision* out, uint32_t nthreads)
{
    if (n == 0) return PK_OK;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 1024) nthreads = 1024;
    if (nthreads > n) nthreads = n;
    if (nthreads == 1) {
        job_t one = {t, payload, pay_off, pay_len, names, name_off, out, 0, n};
        worker(&one);
        return PK_OK;
    }
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
    job_t* jobs = (job_t*)malloc(sizeof(job_t) * nthreads);
    if (!th || !jobs) { free(th); free(jobs); return PK_ERR_ARG; }
    for (uint32_t k = 0; k < nthreads; ++k) {
        jobs[k] = (job_t){t, payload, pay_off, pay_len, names, name_off, out,
                          (uint32_t)(((uint64_t)n * k) / nthreads), (uint32_t)(((uint64_t)n * (k + 1)) / nthreads)};
        pthread_create(&th[k], NULL, worker, &jobs[k]);
    }
    for (uint32_t k = 0; k < nthreads; ++k) pthread_join(th[k], NULL);
    free(th);
    free(jobs);
    return PK_OK;
}

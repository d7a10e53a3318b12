/*
 * pk_cpu_fast.c — an efficient CPU port of the same routing SPEC, used ONLY as the timed CPU
 * baseline (bench.py cpu_baseline / --impl reference).  TEST INFRASTRUCTURE, like the rest of
 * oracle/.  The checker stays pk_oracle.c (byte at a time, obviously correct); this file is
 * checked against it in tests/test_oracle_golden.py.
 *
 * It is meant to be a FAIR opponent for the GPU path — what a careful programmer would run on
 * the host cores:
 *   - ONE pass over the payload: the token classifier (SPEC §B) and the 32 hash stripes of a
 *     segment (SPEC §C) are fed from the same 64-byte load;
 *   - AVX-512 (F/BW/DQ/VL) when the CPU has it: four 16-byte blocks = four neighbouring stripes
 *     per zmm register, the eight registers of a segment's stripe state stay in registers across
 *     a 512-byte step, byte classes as 64-bit compare masks, stripe digests eight at a time
 *     (vprolvd, vpmullq); otherwise a scalar path (64-bit SIMD-in-word scan, block-major
 *     stripes) that gcc vectorises where it can.  Chosen at run time (__builtin_cpu_supports),
 *     so the prebuilt .so also runs on a host without AVX-512;
 *   - a PERSISTENT thread pool: workers are created once, spin briefly for the next batch
 *     before they sleep, and take requests in small chunks from a shared counter (dynamic
 *     balance); the calling thread works too.  No thread is created or joined per batch.
 * The tool switch is internal/service/mock.go:32-64.
 */
#define _GNU_SOURCE
#include <immintrin.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/pk_route.h"

#define P1 0x9E3779B1u
#define P2 0x85EBCA77u
#define P5 0x165667B1u
#define NAME_MUL 0xD6E8FEB86659FD93ull
#define FMIX_C1 0xFF51AFD7ED558CCDull
#define FMIX_C2 0xC4CEB9FE1A85EC53ull

static inline uint32_t rotl32(uint32_t x, unsigned r) { return (x << r) | (x >> (32u - r)); }
static inline uint64_t fmix64(uint64_t x)
{
    x ^= x >> 33; x *= FMIX_C1;
    x ^= x >> 33; x *= FMIX_C2;
    x ^= x >> 33;
    return x;
}

/* ---------------------------------------------------------------- scalar path -------- */
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

static inline void stripe_init(uint32_t a[4], uint32_t q)
{
    const uint32_t base = q * P5;
    a[0] = P1 + P2 + base; a[1] = P2 + base; a[2] = base; a[3] = base - P1;
}

static inline uint64_t stripe_digest(const uint32_t a[4])
{
    const uint32_t lo = rotl32(a[0], 1) + rotl32(a[1], 7) + rotl32(a[2], 12) + rotl32(a[3], 18);
    const uint32_t hi = rotl32(a[0], 24) + rotl32(a[1], 3) + rotl32(a[2], 19) + rotl32(a[3], 10);
    return fmix64(((uint64_t)hi << 32) | lo);
}

/* One block (16 bytes, already zero padded) into its stripe, and its tokens. */
static inline uint32_t block_scalar(uint32_t st[4], const uint8_t* blk)
{
    uint32_t w[4];
    uint64_t x[2];
    memcpy(w, blk, 16);
    memcpy(x, blk, 16);
    for (int j = 0; j < 4; ++j) st[j] = rotl32(st[j] + w[j] * P2, 13) * P1;
    return scan_word64(x[0]) + scan_word64(x[1]);
}

/* Blocks [r0, nb) of segment `seg` (nb <= 1024 blocks of this segment; `p` points at the
 * segment's first byte; the payload's last block may be partial: `tail` = its byte count, 0 if
 * whole), stripe states in a[32][4].  One pass: hash round + token scan per block. */
static uint32_t segment_tail_scalar(uint32_t a[32][4], const uint8_t* p, uint32_t r0, uint32_t nb, uint32_t tail, uint32_t smask)
{
    uint32_t tok = 0;
    const uint32_t whole = tail ? nb - 1u : nb;
    for (uint32_t r = r0; r < whole; ++r) tok += block_scalar(a[r & smask], p + (uint64_t)r * 16u);
    if (tail && nb > r0) {
        uint8_t last[16] = {0};
        memcpy(last, p + (uint64_t)(nb - 1u) * 16u, tail);
        tok += block_scalar(a[(nb - 1u) & smask], last);
    }
    return tok;
}

static void scan_hash_scalar(const uint8_t* p, uint32_t len, uint32_t* tok_out, uint64_t* sum_out)
{
    const uint32_t nblk = (len >> 4) + ((len & 15u) ? 1u : 0u);
    uint64_t sum = 0;
    uint32_t tok = 0;
    const uint32_t stripes = nblk <= 16u ? 8u : 32u;                  /* SPEC v3: short payloads have 8 stripes */
    for (uint32_t first = 0, seg = 0; first < nblk; first += 1024u, ++seg) {
        const uint32_t nb = nblk - first < 1024u ? nblk - first : 1024u;
        const uint32_t ns = nb < stripes ? nb : stripes;
        uint32_t a[32][4];
        for (uint32_t s = 0; s < ns; ++s) stripe_init(a[s], seg * 32u + s);
        const uint32_t tail = (first + nb == nblk) ? (len & 15u) : 0u;
        tok += segment_tail_scalar(a, p + (uint64_t)first * 16u, 0, nb, tail, stripes - 1u);
        for (uint32_t s = 0; s < ns; ++s) sum += stripe_digest(a[s]);
    }
    *tok_out = tok;
    *sum_out = sum;
}

/* ---------------------------------------------------------------- AVX-512 path ------- */
#define AVX512 __attribute__((target("avx512f,avx512bw,avx512dq,avx512vl,popcnt,bmi2")))

/* token starts among the 64 bytes of x (payload offset of byte 0 is a multiple of 8) */
AVX512 static inline uint32_t tokens64(__m512i x)
{
    const __m512i t = _mm512_and_si512(x, _mm512_set1_epi8(0x5F));
    const __mmask64 hi    = _mm512_movepi8_mask(x);                                  /* byte >= 0x80 */
    const __mmask64 lead  = _mm512_cmpge_epu8_mask(x, _mm512_set1_epi8((char)0xC0));
    const __mmask64 ge_a  = _mm512_cmpge_epu8_mask(t, _mm512_set1_epi8(0x41));
    const __mmask64 gt_z  = _mm512_cmpge_epu8_mask(t, _mm512_set1_epi8(0x5B));
    const __mmask64 nz    = _mm512_test_epi8_mask(t, t);                             /* not NUL / space (ASCII) */
    const uint64_t lm   = (uint64_t)ge_a & ~(uint64_t)gt_z & ~(uint64_t)hi;          /* ASCII letter */
    const uint64_t cont = lm & (lm << 1) & ~0x0101010101010101ull;                   /* continues a run; runs cut at j % 8 == 0 */
    const uint64_t m    = ((uint64_t)nz & ~(uint64_t)hi & ~cont) | (uint64_t)lead;
    return (uint32_t)_mm_popcnt_u64(m);
}

AVX512 static inline __m512i round512(__m512i a, __m512i w)
{
    a = _mm512_add_epi32(a, _mm512_mullo_epi32(w, _mm512_set1_epi32((int)P2)));
    return _mm512_mullo_epi32(_mm512_rol_epi32(a, 13), _mm512_set1_epi32((int)P1));
}

/* digests of the four stripes held in one register (lane 4b + j = word j of stripe b), summed */
AVX512 static inline __m512i digest512(__m512i a)
{
    const __m512i rl = _mm512_setr_epi32(1, 7, 12, 18, 1, 7, 12, 18, 1, 7, 12, 18, 1, 7, 12, 18);
    const __m512i rh = _mm512_setr_epi32(24, 3, 19, 10, 24, 3, 19, 10, 24, 3, 19, 10, 24, 3, 19, 10);
    __m512i lo = _mm512_rolv_epi32(a, rl), hi = _mm512_rolv_epi32(a, rh);
    /* horizontal sums inside each 128-bit group (all four lanes end up holding the sum) */
    lo = _mm512_add_epi32(lo, _mm512_shuffle_epi32(lo, (_MM_PERM_ENUM)0xB1));
    hi = _mm512_add_epi32(hi, _mm512_shuffle_epi32(hi, (_MM_PERM_ENUM)0xB1));
    lo = _mm512_add_epi32(lo, _mm512_shuffle_epi32(lo, (_MM_PERM_ENUM)0x4E));
    hi = _mm512_add_epi32(hi, _mm512_shuffle_epi32(hi, (_MM_PERM_ENUM)0x4E));
    /* qword 0 of each group = hi:lo; qword 1 is made zero-contributing below */
    __m512i x = _mm512_mask_blend_epi32((__mmask16)0xAAAA, lo, hi);
    x = _mm512_xor_si512(x, _mm512_srli_epi64(x, 33)); x = _mm512_mullo_epi64(x, _mm512_set1_epi64((long long)FMIX_C1));
    x = _mm512_xor_si512(x, _mm512_srli_epi64(x, 33)); x = _mm512_mullo_epi64(x, _mm512_set1_epi64((long long)FMIX_C2));
    x = _mm512_xor_si512(x, _mm512_srli_epi64(x, 33));
    return _mm512_maskz_mov_epi64((__mmask8)0x55, x);        /* keep one digest per stripe */
}

AVX512 static void scan_hash_avx512(const uint8_t* p, uint32_t len, uint32_t* tok_out, uint64_t* sum_out)
{
    const uint32_t nblk = (len >> 4) + ((len & 15u) ? 1u : 0u);
    __m512i vsum = _mm512_setzero_si512();
    uint64_t sum = 0;
    uint32_t tok = 0;
    const __m512i iv = _mm512_setr_epi32((int)(P1 + P2), (int)P2, 0, (int)(0u - P1), (int)(P1 + P2), (int)P2, 0, (int)(0u - P1),
                                         (int)(P1 + P2), (int)P2, 0, (int)(0u - P1), (int)(P1 + P2), (int)P2, 0, (int)(0u - P1));
    const __m512i q4 = _mm512_setr_epi32(0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3);
    const uint32_t stripes = nblk <= 16u ? 8u : 32u;                  /* SPEC v3: short payloads have 8 stripes = 2 registers */
    for (uint32_t first = 0, seg = 0; first < nblk; first += 1024u, ++seg) {
        const uint32_t nb = nblk - first < 1024u ? nblk - first : 1024u;
        const uint32_t tail = (first + nb == nblk) ? (len & 15u) : 0u;
        const uint32_t whole = tail ? nb - 1u : nb;                   /* blocks wholly inside the payload */
        const uint8_t* sp = p + (uint64_t)first * 16u;
        __m512i st[8];
        for (int v = 0; v < 8; ++v) {
            const __m512i q = _mm512_add_epi32(q4, _mm512_set1_epi32((int)(seg * 32u + 4u * (uint32_t)v)));
            st[v] = _mm512_add_epi32(iv, _mm512_mullo_epi32(q, _mm512_set1_epi32((int)P5)));
        }
        uint32_t r = 0;
        for (; r + 32u <= whole; r += 32u) {                          /* 512-byte steps: state stays in registers */
            const uint8_t* bp = sp + (uint64_t)r * 16u;
            for (int v = 0; v < 8; ++v) {
                const __m512i x = _mm512_loadu_si512((const void*)(bp + 64 * v));
                st[v] = round512(st[v], x);
                tok += tokens64(x);
            }
        }
        for (; r + 4u <= whole; r += 4u) {                            /* 64-byte steps */
            const __m512i x = _mm512_loadu_si512((const void*)(sp + (uint64_t)r * 16u));
            const int v = (int)((r & (stripes - 1u)) >> 2);
            st[v] = round512(st[v], x);
            tok += tokens64(x);
        }
        if (r < nb) {                                                 /* up to 3 whole blocks + the partial one */
            const uint32_t left = nb - r;                             /* 1..4 blocks, all in vector (r & 31) / 4 */
            const uint32_t bytes = (left - (tail ? 1u : 0u)) * 16u + tail;
            const __mmask64 km = bytes >= 64u ? ~0ull : ((1ull << bytes) - 1ull);
            const __m512i x = _mm512_maskz_loadu_epi8(km, (const void*)(sp + (uint64_t)r * 16u));
            const int v = (int)((r & (stripes - 1u)) >> 2);
            const __mmask16 blocks = (__mmask16)((1u << (4u * left)) - 1u);   /* only stripes that own a block advance */
            st[v] = _mm512_mask_mov_epi32(st[v], blocks, round512(st[v], x));
            tok += tokens64(x);
        }
        /* digests of the stripes that saw a block */
        const uint32_t ns = nb < stripes ? nb : stripes;
        for (uint32_t v = 0; v * 4u < ns; ++v) {
            __m512i d = digest512(st[v]);
            const uint32_t have = ns - v * 4u;                        /* stripes of this register in use */
            if (have < 4u) d = _mm512_maskz_mov_epi64((__mmask8)((1u << (2u * have)) - 1u), d);
            vsum = _mm512_add_epi64(vsum, d);
        }
    }
    sum += (uint64_t)_mm512_reduce_add_epi64(vsum);
    *tok_out = tok;
    *sum_out = sum;
}

static void (*g_scan_hash)(const uint8_t*, uint32_t, uint32_t*, uint64_t*) = NULL;

static void pick_impl(void)
{
    const char* force = getenv("PK_CPU_FAST_IMPL");                  /* "scalar" | "avx512": tests exercise both */
    int want512 = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
                  __builtin_cpu_supports("avx512dq") && __builtin_cpu_supports("avx512vl");
    if (force && strcmp(force, "scalar") == 0) want512 = 0;
    g_scan_hash = want512 ? scan_hash_avx512 : scan_hash_scalar;
}

/* 0 = scalar, 1 = AVX-512: which inner loop this process uses (reported by bench.py) */
int pk_cpu_fast_impl(void)
{
    if (!g_scan_hash) pick_impl();
    return g_scan_hash == scan_hash_avx512;
}

/* re-reads PK_CPU_FAST_IMPL (tests switch implementations inside one process) */
void pk_cpu_fast_reselect(void) { pick_impl(); }

static void route_one(const pk_table* t, const uint8_t* p, uint32_t len, const uint8_t* name,
                      uint32_t nlen, pk_decision* d)
{
    int unknown = 1;
    uint32_t cls = t->default_class;
    for (uint32_t k = 0; k < t->n_tools; ++k)                         /* mock.go:32-64 */
        if (t->tools[k].name_len == nlen && memcmp(t->tools[k].name, name, nlen) == 0) {
            cls = t->tools[k].class_id; unknown = 0; break;
        }
    uint32_t tok;
    uint64_t s;
    g_scan_hash(p, len, &tok, &s);
    for (uint32_t j = 0; 4u * j < nlen; ++j) {                        /* name as 32-bit words, zero padded */
        uint32_t w = 0;
        const uint32_t left = nlen - 4u * j;
        memcpy(&w, name + 4u * j, left < 4u ? left : 4u);
        uint64_t x = ((uint64_t)(j + 1u) << 32) | w;
        x *= NAME_MUL; x ^= x >> 32; x *= FMIX_C1;
        s += x;
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
    d->flags = (uint16_t)((best == PK_PROVIDER_NONE ? PK_FLAG_NO_PROVIDER : 0) |
                          (len == 0 ? PK_FLAG_EMPTY_PAYLOAD : 0) | (unknown ? PK_FLAG_UNKNOWN_TOOL : 0));
    d->status = 200;                                                  /* mock.go:25-28 */
}

/* ---------------------------------------------------------------- thread pool -------- */
/* One batch at a time (callers are serialised).  The shared cursor `next` is the whole protocol:
 * between batches it holds CLOSED (any fetch_add on it yields an index past every batch), the
 * caller writes the job, then opens the batch with next = 0 (release); a worker that obtains an
 * index below job.n from its (acquire) fetch_add therefore sees that batch's job completely,
 * however late it woke up.  The batch is over when `done` reaches n; the caller closes the cursor
 * before it returns. */
typedef struct {
    const pk_table* t; const uint8_t* payload; const uint64_t* pay_off; const uint32_t* pay_len;
    const uint8_t* names; const uint32_t* name_off; pk_decision* out;
    _Atomic uint32_t n, chunk;
} job_t;

#define POOL_MAX 1024
#define CLOSED 0x80000000u
static pthread_mutex_t g_call_mu = PTHREAD_MUTEX_INITIALIZER;      /* serialises callers, guards thread creation */
static pthread_mutex_t g_sleep_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t  g_sleep_cv = PTHREAD_COND_INITIALIZER;
static pthread_t       g_threads[POOL_MAX];
static uint32_t        g_n_workers = 0;                            /* threads created so far (they never exit) */
static job_t           g_job;
static _Atomic uint64_t g_gen = 0;                                 /* bumped once per batch */
static _Atomic uint32_t g_active = 0;                              /* workers with id < active join the batch */
static _Atomic uint32_t g_next = CLOSED;
static _Atomic uint32_t g_done = 0;
static _Atomic uint32_t g_sleepers = 0;

static inline uint64_t now_ns(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}

static void work(void)
{
    const uint32_t chunk = atomic_load_explicit(&g_job.chunk, memory_order_relaxed);
    if (chunk == 0) return;
    for (;;) {
        const uint32_t lo = atomic_fetch_add_explicit(&g_next, chunk, memory_order_acq_rel);
        const uint32_t n = atomic_load_explicit(&g_job.n, memory_order_relaxed);
        if (lo >= n) return;
        const uint32_t hi = lo + chunk < n ? lo + chunk : n;
        const job_t* j = &g_job;
        for (uint32_t i = lo; i < hi; ++i)
            route_one(j->t, j->payload + j->pay_off[i], j->pay_len[i], j->names + j->name_off[i],
                      j->name_off[i + 1] - j->name_off[i], &j->out[i]);
        atomic_fetch_add_explicit(&g_done, hi - lo, memory_order_release);
    }
}

static void* worker(void* arg)
{
    const uint32_t id = (uint32_t)(uintptr_t)arg;
    uint64_t seen = 0;
    for (;;) {
        /* the next batch usually follows within microseconds: poll for it, then sleep */
        const uint64_t spin_until = now_ns() + 200000ull;
        uint32_t polls = 0;
        while (atomic_load_explicit(&g_gen, memory_order_acquire) == seen) {
            _mm_pause();
            if ((++polls & 63u) == 0 && now_ns() >= spin_until) {
                pthread_mutex_lock(&g_sleep_mu);
                atomic_fetch_add(&g_sleepers, 1);                  /* seq_cst, then the predicate: see the caller */
                while (atomic_load(&g_gen) == seen) pthread_cond_wait(&g_sleep_cv, &g_sleep_mu);
                atomic_fetch_sub(&g_sleepers, 1);
                pthread_mutex_unlock(&g_sleep_mu);
            }
        }
        seen = atomic_load_explicit(&g_gen, memory_order_acquire);
        if (id < atomic_load_explicit(&g_active, memory_order_relaxed)) work();
    }
    return NULL;
}

/* Number of pool threads created so far (bench.py reports it). */
uint32_t pk_cpu_fast_pool_threads(void) { return g_n_workers; }

/* No argument validation: callers (bench.py, tests) pass batches already accepted by
 * pk_ref_route_batch.  nthreads = threads that work on the batch, the caller included. */
int pk_cpu_route_batch_mt(const pk_table* t, const uint8_t* payload, const uint64_t* pay_off,
                          const uint32_t* pay_len, const uint8_t* names, const uint32_t* name_off,
                          uint32_t n, pk_decision* out, uint32_t nthreads)
{
    if (n == 0) return PK_OK;
    if (n >= CLOSED) return PK_ERR_CAPACITY;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > POOL_MAX) nthreads = POOL_MAX;
    if (nthreads > n) nthreads = n;
    pthread_mutex_lock(&g_call_mu);
    if (!g_scan_hash) pick_impl();
    if (nthreads == 1) {
        for (uint32_t i = 0; i < n; ++i)
            route_one(t, payload + pay_off[i], pay_len[i], names + name_off[i], name_off[i + 1] - name_off[i], &out[i]);
        pthread_mutex_unlock(&g_call_mu);
        return PK_OK;
    }
    while (g_n_workers < nthreads - 1u) {
        if (pthread_create(&g_threads[g_n_workers], NULL, worker, (void*)(uintptr_t)g_n_workers) != 0) break;
        ++g_n_workers;
    }
    const uint32_t helpers = g_n_workers < nthreads - 1u ? g_n_workers : nthreads - 1u;
    uint32_t chunk = n / ((helpers + 1u) * 8u);
    chunk = chunk < 1u ? 1u : chunk > 64u ? 64u : chunk;

    g_job.t = t; g_job.payload = payload; g_job.pay_off = pay_off; g_job.pay_len = pay_len;
    g_job.names = names; g_job.name_off = name_off; g_job.out = out;
    atomic_store_explicit(&g_job.n, n, memory_order_relaxed);
    atomic_store_explicit(&g_job.chunk, chunk, memory_order_relaxed);
    atomic_store_explicit(&g_done, 0, memory_order_relaxed);
    atomic_store_explicit(&g_active, helpers, memory_order_relaxed);
    atomic_store_explicit(&g_next, 0, memory_order_release);       /* open */
    atomic_fetch_add(&g_gen, 1);                                    /* seq_cst, then the sleepers check */
    if (atomic_load(&g_sleepers) != 0) {
        pthread_mutex_lock(&g_sleep_mu);
        pthread_cond_broadcast(&g_sleep_cv);
        pthread_mutex_unlock(&g_sleep_mu);
    }
    work();
    while (atomic_load_explicit(&g_done, memory_order_acquire) != n) _mm_pause();
    atomic_store_explicit(&g_next, CLOSED, memory_order_release);   /* close: stragglers find nothing */
    pthread_mutex_unlock(&g_call_mu);
    return PK_OK;
}

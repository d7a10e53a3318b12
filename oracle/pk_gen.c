/*
 * pk_gen.c — seeded synthetic request streams.  TEST INFRASTRUCTURE
 * ONLY (same rule as pk_oracle.c: tests/, smoke() and bench.py may load it; the
 * product never does).
 *
 * Generator: SplitMix64.  Request i of a stream with seed S is drawn from its own
 * state splitmix(S ^ i*GOLDEN), so a batch is identical however many threads fill it
 * and whichever slice of the stream a rank generates.
 *
 * Payload modes
 *   0  english-like text: 85 % lowercase words (geometric length, mean ~5, 1 in 12
 *      capitalised) followed by a space, 5 % punctuation/digit runs, 5 % newlines,
 *      5 % 2- or 3-byte UTF-8 code points; always valid UTF-8
 *   1  all ASCII spaces                         (adversarial: zero tokens)
 *   2  uniformly random VALID UTF-8 code points of 1-4 bytes
 *   3  uniformly random bytes                   (opaque-bytes stress; not valid UTF-8)
 *   4  base64-like unbroken alphanumeric run    (adversarial: letter-run breaking)
 * The first 8 bytes of every payload of 8+ bytes in modes 0/1/4 are left as
 * generated — payload identity comes from the per-request seed.
 *
 * Tool names (the reference's literals are internal/service/mock.go:33,37,52):
 *   25 % each "example_tool" / "struct_tool" / "file_tool", 15 % random unknown names
 *   of 4-32 bytes, 10 % near misses (case flip, trailing space, prefix, suffix).
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define GOLDEN 0x9E3779B97F4A7C15ull

static inline uint64_t sm64(uint64_t* s)
{
    uint64_t z = (*s += GOLDEN);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static inline uint64_t req_seed(uint64_t seed, uint64_t i)
{
    uint64_t s = seed ^ (i * GOLDEN);
    sm64(&s);
    return s;
}

static unsigned put_utf8(uint8_t* d, uint32_t cp)
{
    if (cp < 0x80) { d[0] = (uint8_t)cp; return 1; }
    if (cp < 0x800) { d[0] = 0xC0 | (cp >> 6); d[1] = 0x80 | (cp & 0x3F); return 2; }
    if (cp < 0x10000) {
        d[0] = 0xE0 | (cp >> 12); d[1] = 0x80 | ((cp >> 6) & 0x3F); d[2] = 0x80 | (cp & 0x3F);
        return 3;
    }
    d[0] = 0xF0 | (cp >> 18); d[1] = 0x80 | ((cp >> 12) & 0x3F);
    d[2] = 0x80 | ((cp >> 6) & 0x3F); d[3] = 0x80 | (cp & 0x3F);
    return 4;
}

void pk_gen_payload(uint8_t* dst, uint64_t len, uint64_t seed, int mode)
{
    uint64_t s = seed;
    uint64_t j = 0;
    static const char punct[] = ".,;:!?\"'()[]{}-_/\\@#$%&*+=<>|~^`";
    static const char b64[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789";
    switch (mode) {
    case 1:
        memset(dst, ' ', len);
        return;
    case 3:
        while (j + 8 <= len) { uint64_t r = sm64(&s); memcpy(dst + j, &r, 8); j += 8; }
        if (j < len) { uint64_t r = sm64(&s); memcpy(dst + j, &r, len - j); }
        return;
    case 4:
        while (j < len) {
            uint64_t r = sm64(&s);
            for (int k = 0; k < 10 && j < len; ++k, r >>= 6) dst[j++] = (uint8_t)b64[(r & 63) % 62];
        }
        return;
    case 2:
        while (j < len) {
            uint64_t r = sm64(&s);
            uint32_t kind = (uint32_t)(r & 3);
            uint32_t v = (uint32_t)(r >> 8);
            uint32_t cp;
            if (kind == 0) cp = v % 0x80;
            else if (kind == 1) cp = 0x80 + v % (0x800 - 0x80);
            else if (kind == 2) { cp = 0x800 + v % (0x10000 - 0x800); if (cp >= 0xD800 && cp < 0xE000) cp = 0x4E00 + (cp & 0x3FF); }
            else cp = 0x10000 + v % (0x110000 - 0x10000);
            uint8_t tmp[4];
            unsigned n = put_utf8(tmp, cp);
            if (j + n > len) { while (j < len) dst[j++] = ' '; break; }
            memcpy(dst + j, tmp, n);
            j += n;
        }
        return;
    default:
        break;
    }
    /* mode 0: english-like */
    while (j < len) {
        uint64_t r = sm64(&s);
        uint32_t pick = (uint32_t)(r % 100);
        r >>= 8;
        if (pick < 85) {
            unsigned wl = 1;
            uint64_t g = sm64(&s);
            while (wl < 24 && (g & 7) < 6) { g >>= 3; ++wl; if ((wl % 20) == 0) g = sm64(&s); }  /* mean ~4.9 */
            int cap = (r % 12) == 0;
            r >>= 4;
            uint64_t l = sm64(&s);
            for (unsigned k = 0; k < wl && j < len; ++k) {
                uint8_t c = (uint8_t)('a' + (l & 0xFF) % 26);
                l >>= 8;
                if (k % 8 == 7) l = sm64(&s);
                if (cap && k == 0) c = (uint8_t)(c - 32);
                dst[j++] = c;
            }
            if (j < len) dst[j++] = ' ';
        } else if (pick < 90) {
            unsigned n = 1 + (unsigned)(r & 3);
            r >>= 2;
            int digits = (int)(r & 1);
            r >>= 1;
            for (unsigned k = 0; k < n && j < len; ++k) {
                dst[j++] = digits ? (uint8_t)('0' + (r & 0xF) % 10)
                                  : (uint8_t)punct[(r & 0xFF) % (sizeof punct - 1)];
                r >>= 8;
            }
        } else if (pick < 95) {
            dst[j++] = '\n';
        } else {
            uint32_t cp = (r & 1) ? 0x00A1 + (uint32_t)((r >> 1) % 0x500)       /* 2-byte */
                                  : 0x4E00 + (uint32_t)((r >> 1) % 0x5000);     /* 3-byte */
            uint8_t tmp[4];
            unsigned n = put_utf8(tmp, cp);
            if (j + n > len) { while (j < len) dst[j++] = ' '; break; }
            memcpy(dst + j, tmp, n);
            j += n;
        }
    }
}

/* lens[i] uniform in [len_min, len_max] */
void pk_gen_lengths(uint64_t seed, uint64_t first_id, uint32_t n, uint32_t len_min,
                    uint32_t len_max, uint32_t* lens)
{
    for (uint32_t i = 0; i < n; ++i) {
        uint64_t s = req_seed(seed ^ 0x4C454E53ull, first_id + i);
        uint64_t span = (uint64_t)len_max - len_min + 1;
        lens[i] = len_min + (uint32_t)(sm64(&s) % span);
    }
}

/* Writes request first_id+i's tool name at names + name_off[i]; fills name_off[0..n].
 * Returns total bytes, or 0 if cap is too small. */
uint64_t pk_gen_names(uint64_t seed, uint64_t first_id, uint32_t n, uint8_t* names, uint64_t cap,
                      uint32_t* name_off)
{
    static const char* known[3] = {"example_tool", "struct_tool", "file_tool"};
    uint64_t at = 0;
    for (uint32_t i = 0; i < n; ++i) {
        uint64_t s = req_seed(seed ^ 0x4E414D45ull, first_id + i);
        uint64_t r = sm64(&s);
        uint32_t pick = (uint32_t)(r % 100);
        r >>= 8;
        uint8_t buf[80];
        uint32_t l = 0;
        if (pick < 75) {
            const char* k = known[pick / 25];
            l = (uint32_t)strlen(k);
            memcpy(buf, k, l);
        } else if (pick < 90) {
            l = 4 + (uint32_t)(r % 29);
            r >>= 8;
            for (uint32_t k = 0; k < l; ++k) {
                if (k % 8 == 0) r = sm64(&s);
                uint32_t v = (uint32_t)(r & 0xFF) % 37;
                r >>= 8;
                buf[k] = v < 26 ? (uint8_t)('a' + v) : v < 36 ? (uint8_t)('0' + v - 26) : '_';
            }
        } else {
            const char* k = known[r % 3];
            r >>= 2;
            l = (uint32_t)strlen(k);
            memcpy(buf, k, l);
            switch (r & 3) {
            case 0: buf[(r >> 2) % l] ^= 0x20; break;                 /* case flip (or '_'→0x7F) */
            case 1: buf[l++] = ' '; break;                            /* trailing space */
            case 2: l -= 1 + (uint32_t)((r >> 2) % 3); break;         /* prefix */
            default: buf[l++] = 's'; break;                           /* suffix */
            }
        }
        if (at + l > cap) return 0;
        name_off[i] = (uint32_t)at;
        memcpy(names + at, buf, l);
        at += l;
    }
    name_off[n] = (uint32_t)at;
    return at;
}

typedef struct {
    uint64_t seed, first_id; uint32_t lo, hi; uint8_t* payload; const uint64_t* pay_off;
    const uint32_t* pay_len; int mode;
} fill_job;

static void* fill_worker(void* arg)
{
    fill_job* j = (fill_job*)arg;
    for (uint32_t i = j->lo; i < j->hi; ++i) {
        uint8_t* d = j->payload + j->pay_off[i];
        pk_gen_payload(d, j->pay_len[i], req_seed(j->seed, j->first_id + i), j->mode);
        /* deterministic padding so whole arenas can be compared byte for byte */
        uint64_t end = j->pay_off[i] + j->pay_len[i];
        if (j->pay_off[i + 1] > end) memset(j->payload + end, 0xA5, j->pay_off[i + 1] - end);
    }
    return NULL;
}

/* mode < 0: per-request mode = request id % 5 (a mixed stream). Padding between
 * requests is filled with 0xA5 — deliberately NOT zero, so a kernel that fails to
 * mask the tail block is caught. */
void pk_gen_fill(uint64_t seed, uint64_t first_id, uint32_t n, uint8_t* payload,
                 const uint64_t* pay_off, const uint32_t* pay_len, int mode, uint32_t nthreads)
{
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    if (nthreads > n) nthreads = n ? n : 1;
    if (mode < 0) {                       /* mixed: single thread, mode varies per request */
        for (uint32_t i = 0; i < n; ++i) {
            fill_job one = {seed, first_id, i, i + 1, payload, pay_off, pay_len,
                            (int)((first_id + i) % 5)};
            fill_worker(&one);
        }
        return;
    }
    pthread_t th[256];
    fill_job jobs[256];
    for (uint32_t k = 0; k < nthreads; ++k) {
        uint32_t lo = (uint32_t)(((uint64_t)n * k) / nthreads);
        uint32_t hi = (uint32_t)(((uint64_t)n * (k + 1)) / nthreads);
        jobs[k] = (fill_job){seed, first_id, lo, hi, payload, pay_off, pay_len, mode};
        pthread_create(&th[k], NULL, fill_worker, &jobs[k]);
    }
    for (uint32_t k = 0; k < nthreads; ++k) pthread_join(th[k], NULL);
}

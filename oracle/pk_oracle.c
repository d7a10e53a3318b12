/*
 * pk_oracle.c — CPU ORACLE for the batched routing path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (libpkroute.so) never links,
 * loads or calls anything in oracle/.
 *
 * What is pinned by the reference and what is not
 * ------------------------------------------------
 *  PINNED by the reference's literals (there are no tests or golden vectors in the
 *  reference — SURVEY.md §4 — so the literals are the only pins):
 *    - tool_class: Go string `switch toolName` over "example_tool", "struct_tool",
 *      "file_tool" and `default`  — internal/service/mock.go:32-64.  Whole-name,
 *      case-sensitive byte equality, no trimming.
 *    - status is always 200 / "Tool executed successfully" and an unknown tool is
 *      NOT an error — mock.go:24-29, 60-66.
 *    - the response bodies built per class — mock.go:33-63 (pk_ref_execute_tool).
 *  PARITY UNPINNED (build-defined; the reference has no such code — SURVEY.md §0.3,
 *  §8 a6): the byte scan / token estimate (SPEC §B), the route-key hash (SPEC §C)
 *  and the provider selection (SPEC §D).  For those this file IS the specification;
 *  oracle/pk_spec_py.py is a second, independent restatement of the same written
 *  spec (DESIGN.md "SPEC") used to cross-check this one.
 *
 *  The reference itself (Go + un-vendored spounge-proto v1.2.0, go.mod:6) cannot be
 *  built in this environment (no Go toolchain, no module cache), so there is no
 *  oracle/_ref.
 *
 * Written as plain scalar C on purpose: one byte at a time, no SWAR, no SIMD, so it
 * shares no trick with the CUDA kernels it checks.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/pk_route.h"

/* ---- SPEC constants (duplicated on purpose from the product's pk_spec.cuh) ---- */
#define SPEC_BLOCK        16u
#define SPEC_STRIPES      32u
#define SPEC_TINY_BLOCKS  16u      /* payloads of at most this many blocks ... */
#define SPEC_TINY_STRIPES 8u       /* ... have this many stripes */
#define SPEC_SEG_BLOCKS   1024u                 /* 16 KiB segments                 */
#define SPEC_P1           0x9E3779B1u
#define SPEC_P2           0x85EBCA77u
#define SPEC_P5           0x165667B1u
#define SPEC_NAME_MUL     0xD6E8FEB86659FD93ull
#define SPEC_FMIX_C1      0xFF51AFD7ED558CCDull
#define SPEC_FMIX_C2      0xC4CEB9FE1A85EC53ull

static inline uint32_t rotl32(uint32_t x, unsigned r) { return (x << r) | (x >> (32u - r)); }

static inline uint64_t fmix64(uint64_t x)
{
    x ^= x >> 33; x *= SPEC_FMIX_C1;
    x ^= x >> 33; x *= SPEC_FMIX_C2;
    x ^= x >> 33;
    return x;
}

/* ------------------------------------------------------------------ SPEC §A ---- */
/* internal/service/mock.go:32-64 — the string switch.  Returns the class and sets
 * *unknown when the `default:` branch (mock.go:60) is taken. */
uint32_t pk_ref_tool_class(const pk_table* t, const uint8_t* name, uint32_t len, int* unknown)
{
    for (uint32_t k = 0; k < t->n_tools; ++k) {
        const pk_tool* e = &t->tools[k];
        if (e->name_len == len && memcmp(e->name, name, len) == 0) {
            if (unknown) *unknown = 0;
            return e->class_id;
        }
    }
    if (unknown) *unknown = 1;
    return t->default_class;
}

/* ------------------------------------------------------------------ SPEC §B ---- */
enum { C_SEP, C_LETTER, C_LEAD, C_CONT, C_OTHER };

static int byte_class(uint8_t b)
{
    if (b == 0x00 || b == 0x20) return C_SEP;        /* NUL and space; zero padding counts nothing */
    if ((b >= 'A' && b <= 'Z') || (b >= 'a' && b <= 'z')) return C_LETTER;
    if (b >= 0xC0) return C_LEAD;
    if (b >= 0x80) return C_CONT;
    return C_OTHER;                       /* any other ASCII byte: digits, punctuation, control, DEL */
}

uint32_t pk_ref_est_tokens(const uint8_t* p, uint64_t len, int* nonascii)
{
    uint32_t tokens = 0;
    int hi = 0;
    for (uint64_t j = 0; j < len; ++j) {
        int c = byte_class(p[j]);
        if (p[j] >= 0x80) hi = 1;
        if (c == C_OTHER || c == C_LEAD) {
            ++tokens;
        } else if (c == C_LETTER) {
            int continues = (j > 0) && byte_class(p[j - 1]) == C_LETTER && (j % 8u) != 0;
            if (!continues) ++tokens;
        }
    }
    if (nonascii) *nonascii = hi;
    return tokens;
}

/* ------------------------------------------------------------------ SPEC §C ---- */
static inline uint32_t load_le32_padded(const uint8_t* p, uint64_t len, uint64_t at)
{
    uint32_t w = 0;
    for (unsigned i = 0; i < 4; ++i) {
        uint64_t j = at + i;
        if (j < len) w |= (uint32_t)p[j] << (8u * i);
    }
    return w;
}

/* Sum over all non-empty stripes of the 64-bit stripe digests. */
uint64_t pk_ref_payload_sum(const uint8_t* p, uint64_t len)
{
    uint64_t nblk = (len + SPEC_BLOCK - 1) / SPEC_BLOCK;
    uint64_t nseg = (nblk + SPEC_SEG_BLOCKS - 1) / SPEC_SEG_BLOCKS;
    uint64_t sum = 0;
    /* SPEC v3: a payload of at most 16 blocks (256 bytes) has 8 stripes, not 32 — blocks k and
     * k + 8 share one — so that a short payload pays half as many stripe digests */
    const uint32_t stripes = nblk <= SPEC_TINY_BLOCKS ? SPEC_TINY_STRIPES : SPEC_STRIPES;
    for (uint64_t seg = 0; seg < nseg; ++seg) {
        uint64_t first = seg * SPEC_SEG_BLOCKS;
        uint64_t in_seg = nblk - first;
        if (in_seg > SPEC_SEG_BLOCKS) in_seg = SPEC_SEG_BLOCKS;
        for (uint32_t s = 0; s < stripes && s < in_seg; ++s) {
            uint32_t q = (uint32_t)(seg * SPEC_STRIPES + s);
            uint32_t a[4];
            a[0] = SPEC_P1 + SPEC_P2 + q * SPEC_P5;
            a[1] = SPEC_P2 + q * SPEC_P5;
            a[2] = 0u + q * SPEC_P5;
            a[3] = 0u - SPEC_P1 + q * SPEC_P5;
            for (uint64_t r = s; r < in_seg; r += stripes) {
                uint64_t at = (first + r) * SPEC_BLOCK;
                for (unsigned j = 0; j < 4; ++j) {
                    uint32_t w = load_le32_padded(p, len, at + 4u * j);
                    a[j] = rotl32(a[j] + w * SPEC_P2, 13) * SPEC_P1;
                }
            }
            uint32_t lo = rotl32(a[0], 1) + rotl32(a[1], 7) + rotl32(a[2], 12) + rotl32(a[3], 18);
            uint32_t hi = rotl32(a[0], 24) + rotl32(a[1], 3) + rotl32(a[2], 19) + rotl32(a[3], 10);
            sum += fmix64(((uint64_t)hi << 32) | lo);
        }
    }
    return sum;
}

/* The tool name's share: the name as little-endian 32-bit words n_j (the last zero padded), word j
 * contributing mixn((j + 1) << 32 | n_j); mixn(x): x *= NAME_MUL; x ^= x >> 32; x *= FMIX_C1. */
uint64_t pk_ref_name_sum(const uint8_t* name, uint32_t len)
{
    uint64_t sum = 0;
    for (uint32_t j = 0; 4u * j < len; ++j) {
        uint32_t w = 0;
        for (uint32_t k = 0; k < 4u && 4u * j + k < len; ++k) w |= (uint32_t)name[4u * j + k] << (8u * k);
        uint64_t x = ((uint64_t)(j + 1u) << 32) | w;
        x *= SPEC_NAME_MUL;
        x ^= x >> 32;
        x *= SPEC_FMIX_C1;
        sum += x;
    }
    return sum;
}

uint64_t pk_ref_route_hash(const uint8_t* p, uint64_t len, const uint8_t* name, uint32_t name_len)
{
    uint64_t s = pk_ref_payload_sum(p, len) + pk_ref_name_sum(name, name_len);
    s += ((uint64_t)name_len << 32) + (uint64_t)(uint32_t)len;
    return fmix64(s);
}

/* ------------------------------------------------------------------ SPEC §D ---- */
/* Among the eligible providers choose one with probability weight / total weight, the random
 * number being the top 32 bits of the route hash: x = (hi32(hash) * W) >> 32 lies in [0, W);
 * the winner is the first eligible provider whose running weight total exceeds x. */
static int provider_eligible(const pk_provider* e, uint32_t est_tokens, uint32_t tool_class)
{
    return e->weight != 0 && ((e->tool_mask >> tool_class) & 1u) != 0 && est_tokens <= e->max_tokens;
}

uint32_t pk_ref_select_provider(const pk_table* t, uint64_t route_hash, uint32_t est_tokens,
                                uint32_t tool_class)
{
    uint64_t total = 0;
    for (uint32_t p = 0; p < t->n_providers; ++p)
        if (provider_eligible(&t->providers[p], est_tokens, tool_class)) total += t->providers[p].weight;
    if (total == 0) return PK_PROVIDER_NONE;
    uint64_t x = ((route_hash >> 32) * total) >> 32;
    uint64_t running = 0;
    for (uint32_t p = 0; p < t->n_providers; ++p) {
        if (!provider_eligible(&t->providers[p], est_tokens, tool_class)) continue;
        running += t->providers[p].weight;
        if (running > x) return p;
    }
    return PK_PROVIDER_NONE;   /* unreachable */
}

/* ------------------------------------------------------------ one request ------ */
void pk_ref_route_one(const pk_table* t, const uint8_t* p, uint32_t len, const uint8_t* name,
                      uint32_t name_len, pk_decision* d)
{
    int unknown = 0, nonascii = 0;
    memset(d, 0, sizeof *d);
    d->byte_len = len;
    d->tool_class = (uint8_t)pk_ref_tool_class(t, name, name_len, &unknown);
    d->est_tokens = pk_ref_est_tokens(p, len, &nonascii);
    d->route_hash = pk_ref_route_hash(p, len, name, name_len);
    uint32_t prov = pk_ref_select_provider(t, d->route_hash, d->est_tokens, d->tool_class);
    d->provider = (uint8_t)prov;
    (void)nonascii;
    d->flags = (uint16_t)((prov == PK_PROVIDER_NONE ? PK_FLAG_NO_PROVIDER : 0) |
                          (len == 0 ? PK_FLAG_EMPTY_PAYLOAD : 0) |
                          (unknown ? PK_FLAG_UNKNOWN_TOOL : 0));
    d->status = 200;                       /* mock.go:25-28 */
}

static int validate_table(const pk_table* t)
{
    if (!t || t->n_tools > PK_MAX_TOOLS || t->n_providers > PK_MAX_PROVIDERS) return PK_ERR_ARG;
    if (t->default_class >= 32) return PK_ERR_ARG;
    for (uint32_t k = 0; k < t->n_tools; ++k)
        if (t->tools[k].name_len > PK_MAX_TOOL_NAME || t->tools[k].class_id >= 32) return PK_ERR_ARG;
    for (uint32_t p = 0; p < t->n_providers; ++p)
        if (t->providers[p].weight > PK_MAX_WEIGHT) return PK_ERR_ARG;
    return PK_OK;
}

static int validate_batch(const uint64_t* pay_off, const uint32_t* pay_len,
                          const uint32_t* name_off, uint32_t n)
{
    for (uint32_t i = 0; i < n; ++i) {
        if (pay_off[i] % PK_ALIGN) return PK_ERR_ALIGN;
        if (pay_off[i + 1] < pay_off[i]) return PK_ERR_OFFSETS;
        if ((uint64_t)pay_len[i] > pay_off[i + 1] - pay_off[i]) return PK_ERR_OFFSETS;
        if (name_off[i + 1] < name_off[i]) return PK_ERR_OFFSETS;
        if (name_off[i + 1] - name_off[i] > PK_MAX_NAME_LEN) return PK_ERR_ARG;
    }
    return PK_OK;
}

/* Same argument list and error behaviour as pk_route_batch, minus the handle. */
int pk_ref_route_batch(const pk_table* t, const uint8_t* payload, const uint64_t* pay_off,
                       const uint32_t* pay_len, const uint8_t* names, const uint32_t* name_off,
                       uint32_t n, pk_decision* out)
{
    int rc = validate_table(t);
    if (rc) return rc;
    if (n == 0) return PK_OK;
    if (!pay_off || !pay_len || !name_off || !out) return PK_ERR_ARG;
    rc = validate_batch(pay_off, pay_len, name_off, n);
    if (rc) return rc;
    for (uint32_t i = 0; i < n; ++i)
        pk_ref_route_one(t, payload + pay_off[i], pay_len[i], names + name_off[i],
                         name_off[i + 1] - name_off[i], &out[i]);
    return PK_OK;
}

/* ---- multi-threaded form for the CPU baseline (static partition, pthreads) ---- */
typedef struct {
    const pk_table* t; const uint8_t* payload; const uint64_t* pay_off; const uint32_t* pay_len;
    const uint8_t* names; const uint32_t* name_off; pk_decision* out; uint32_t lo, hi;
    int switch_only;
} mt_job;

static void* mt_worker(void* arg)
{
    mt_job* j = (mt_job*)arg;
    for (uint32_t i = j->lo; i < j->hi; ++i) {
        if (j->switch_only) {
            int unknown = 0;
            memset(&j->out[i], 0, sizeof(pk_decision));
            j->out[i].tool_class = (uint8_t)pk_ref_tool_class(
                j->t, j->names + j->name_off[i], j->name_off[i + 1] - j->name_off[i], &unknown);
            j->out[i].flags = unknown ? PK_FLAG_UNKNOWN_TOOL : 0;
            j->out[i].status = 200;
        } else {
            pk_ref_route_one(j->t, j->payload + j->pay_off[i], j->pay_len[i],
                             j->names + j->name_off[i], j->name_off[i + 1] - j->name_off[i],
                             &j->out[i]);
        }
    }
    return NULL;
}

/* switch_only != 0 runs CPU-A (the reference-faithful work: mock.go's switch and
 * nothing else); 0 runs CPU-B (the full work-equivalent oracle). */
int pk_ref_route_batch_mt(const pk_table* t, const uint8_t* payload, const uint64_t* pay_off,
                          const uint32_t* pay_len, const uint8_t* names, const uint32_t* name_off,
                          uint32_t n, pk_decision* out, uint32_t nthreads, int switch_only)
{
    int rc = validate_table(t);
    if (rc) return rc;
    if (n == 0) return PK_OK;
    rc = validate_batch(pay_off, pay_len, name_off, n);
    if (rc) return rc;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 1024) nthreads = 1024;
    if (nthreads > n) nthreads = n;
    if (nthreads == 1) {                       /* no thread spawn for the single-core arm */
        mt_job one = {t, payload, pay_off, pay_len, names, name_off, out, 0, n, switch_only};
        mt_worker(&one);
        return PK_OK;
    }
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
    mt_job* jobs = (mt_job*)malloc(sizeof(mt_job) * nthreads);
    if (!th || !jobs) { free(th); free(jobs); return PK_ERR_ARG; }
    /* partition by payload bytes so ragged batches stay balanced */
    uint64_t total = pay_off[n] - pay_off[0];
    uint32_t cur = 0;
    for (uint32_t k = 0; k < nthreads; ++k) {
        uint32_t lo = cur;
        uint64_t target = pay_off[0] + (total / nthreads) * (k + 1);
        uint32_t want = (uint32_t)(((uint64_t)n * (k + 1)) / nthreads);
        if (total >= (uint64_t)nthreads * 4096u) {
            while (cur < n && pay_off[cur + 1] <= target) ++cur;
        } else {
            cur = want;
        }
        if (k == nthreads - 1) cur = n;
        jobs[k] = (mt_job){t, payload, pay_off, pay_len, names, name_off, out, lo, cur, switch_only};
    }
    for (uint32_t k = 0; k < nthreads; ++k) pthread_create(&th[k], NULL, mt_worker, &jobs[k]);
    for (uint32_t k = 0; k < nthreads; ++k) pthread_join(th[k], NULL);
    free(th);
    free(jobs);
    return PK_OK;
}

/* ---------------------------------------------------------------------------------
 * pk_ref_execute_tool — restatement of MockService.ExecuteTool's response
 * construction (internal/service/mock.go:22-67), for the host mirror's tests.
 *   kind 0: string output  (mock.go:33-36 "Mock execution of %s at %s", and the
 *           default branch mock.go:60-63 "Unknown tool: %s")
 *   kind 1: struct output  (mock.go:37-51) — rendered here as canonical JSON text
 *   kind 2: file output    (mock.go:52-59)
 * `now_unix` replaces time.Now() so outputs are reproducible; the RFC3339 stamp is
 * rendered in UTC ("Z"), where Go would use the local zone — tests mask it.
 * Returns the status code (always 200) and never fails: an unknown tool is not an
 * error (mock.go:60-66).
 * --------------------------------------------------------------------------------- */
typedef struct pk_ref_response {
    uint32_t status_code;
    char     status_message[64];
    uint32_t kind;
    char     string_output[512];
    char     file_name[32];
    char     mime_type[32];
    char     file_content[64];
} pk_ref_response;

uint32_t pk_ref_execute_tool(const uint8_t* name, uint32_t len, int64_t now_unix,
                             pk_ref_response* r)
{
    memset(r, 0, sizeof *r);
    r->status_code = 200;                                           /* mock.go:26 */
    snprintf(r->status_message, sizeof r->status_message, "Tool executed successfully"); /* :27 */
    char nm[256];
    uint32_t nl = len < sizeof nm - 1 ? len : (uint32_t)sizeof nm - 1;
    memcpy(nm, name, nl);
    nm[nl] = 0;
    if (len == 12 && memcmp(name, "example_tool", 12) == 0) {        /* mock.go:33 */
        time_t tt = (time_t)now_unix;
        struct tm tmv;
        gmtime_r(&tt, &tmv);
        char stamp[40];
        strftime(stamp, sizeof stamp, "%Y-%m-%dT%H:%M:%SZ", &tmv);
        r->kind = 0;
        snprintf(r->string_output, sizeof r->string_output, "Mock execution of %s at %s", nm, stamp);
    } else if (len == 11 && memcmp(name, "struct_tool", 11) == 0) {  /* mock.go:37 */
        r->kind = 1;
        snprintf(r->string_output, sizeof r->string_output,
                 "{\"data\":{\"count\":42,\"processed\":true},\"result\":\"success\",\"timestamp\":%lld}",
                 (long long)now_unix);
    } else if (len == 9 && memcmp(name, "file_tool", 9) == 0) {      /* mock.go:52 */
        r->kind = 2;
        snprintf(r->file_name, sizeof r->file_name, "example.txt");
        snprintf(r->mime_type, sizeof r->mime_type, "text/plain");
        snprintf(r->file_content, sizeof r->file_content, "This is mock file content");
    } else {                                                         /* mock.go:60 */
        r->kind = 0;
        snprintf(r->string_output, sizeof r->string_output, "Unknown tool: %s", nm);
    }
    return r->status_code;
}

/* The reference table: the three case labels of mock.go:33,37,52 + default (:60),
 * and the 8 build-defined providers used by every BASELINE config. */
int pk_ref_default_table(pk_table* t)
{
    static const char* names[3] = {"example_tool", "struct_tool", "file_tool"};
    static const uint32_t max_tokens[8] = {4096, 8192, 8192, 32768, 32768, 131072, 131072, 1048576};
    static const uint32_t weight[8] = {1, 2, 3, 4, 5, 6, 7, 8};
    if (!t) return PK_ERR_ARG;
    memset(t, 0, sizeof *t);
    t->n_tools = 3;
    t->default_class = 3;
    for (uint32_t k = 0; k < 3; ++k) {
        t->tools[k].name_len = (uint32_t)strlen(names[k]);
        memcpy(t->tools[k].name, names[k], t->tools[k].name_len);
        t->tools[k].class_id = k;
    }
    t->n_providers = 8;
    for (uint32_t p = 0; p < 8; ++p) {
        t->providers[p].max_tokens = max_tokens[p];
        t->providers[p].weight = weight[p];
        t->providers[p].tool_mask = (p % 4 == 3) ? 0x7u : 0xFu;  /* every 4th: known tools only */
        t->providers[p].reserved = 0;
    }
    return PK_OK;
}

size_t pk_ref_sizeof_response(void) { return sizeof(pk_ref_response); }

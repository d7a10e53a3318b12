// security.cpp — see security.hpp.
#include "security.hpp"

#include <sys/random.h>

#include <algorithm>
#include <cstring>
#include <memory>
#include <mutex>

namespace polykey::security {

namespace {
inline uint64_t round16(uint64_t x) { return (x + 15u) & ~uint64_t(15); }
}

Error validateKey(const Bytes& key)
{
    if (key.size() != 32)                               // cipher.go:16-17
        return "key length must be 32 bytes, got " + std::to_string(key.size()) + " bytes";
    return std::nullopt;
}

GpuCipher::GpuCipher(const Bytes& key, int device, uint32_t max_messages, uint64_t max_bytes)
{
    init_error_ = validateKey(key);
    if (init_error_) return;
    const int rc = pk_gcm_init(device, key.data(), (uint32_t)key.size(), max_messages, max_bytes, &h_);
    if (rc != PK_GCM_OK) {
        h_ = nullptr;
        init_error_ = std::string("failed to create AES cipher: ") + pk_gcm_strerror(rc);   // cipher.go:37
    }
}

GpuCipher::~GpuCipher()
{
    if (!h_) return;
    for (uint8_t* a : arena_) pk_gcm_host_free(h_, a);
    pk_gcm_destroy(h_);
}

uint8_t* GpuCipher::arena(int which, size_t bytes)
{
    if (bytes > arena_bytes_[which]) {
        pk_gcm_host_free(h_, arena_[which]);
        const size_t want = bytes + bytes / 2;             // head-room so a slowly growing batch does not reallocate every call
        arena_[which] = static_cast<uint8_t*>(pk_gcm_host_alloc(h_, want));
        arena_bytes_[which] = arena_[which] ? want : 0;
    }
    return arena_[which];
}

BatchResult GpuCipher::BatchEncrypt(const std::vector<Bytes>& plaintexts, const std::vector<Bytes>* nonces)
{
    if (init_error_) return {{}, init_error_};
    const uint32_t n = (uint32_t)plaintexts.size();
    if (n == 0) return {{}, std::nullopt};
    uint64_t* poff = reinterpret_cast<uint64_t*>(arena(2, (size_t)n * 8));
    uint64_t* soff = reinterpret_cast<uint64_t*>(arena(3, (size_t)n * 8));
    uint32_t* plen = reinterpret_cast<uint32_t*>(arena(4, (size_t)n * 4));
    uint8_t* nonce_arr = arena(5, (size_t)n * PK_GCM_NONCE_BYTES);
    if (!poff || !soff || !plen || !nonce_arr) return {{}, std::string("failed to encrypt plaintext: ") + pk_gcm_strerror(PK_GCM_ERR_CUDA)};
    const size_t nonce_bytes = (size_t)n * PK_GCM_NONCE_BYTES;
    uint64_t pat = 0, sat = 0;
    for (uint32_t i = 0; i < n; ++i) {
        plen[i] = (uint32_t)plaintexts[i].size();
        poff[i] = pat; soff[i] = sat;
        pat += round16(plen[i]);
        sat += round16((uint64_t)plen[i] + PK_GCM_OVERHEAD);
    }
    uint8_t* plain = arena(0, pat + 16);
    uint8_t* sealed = arena(1, sat + 16);
    if (!plain || !sealed) return {{}, std::string("failed to encrypt plaintext: ") + pk_gcm_strerror(PK_GCM_ERR_CUDA)};
    for (uint32_t i = 0; i < n; ++i)
        if (plen[i]) std::memcpy(plain + poff[i], plaintexts[i].data(), plen[i]);
    if (nonces) {
        for (uint32_t i = 0; i < n; ++i) {
            if ((*nonces)[i].size() != PK_GCM_NONCE_BYTES) return {{}, std::string("failed to encrypt plaintext: bad nonce length")};
            std::memcpy(nonce_arr + (size_t)i * PK_GCM_NONCE_BYTES, (*nonces)[i].data(), PK_GCM_NONCE_BYTES);
        }
    } else {
        size_t got = 0;                                  // generateNonce, cipher.go:25-31
        while (got < nonce_bytes) {
            const ssize_t r = getrandom(nonce_arr + got, nonce_bytes - got, 0);
            if (r <= 0) return {{}, std::string("failed to encrypt plaintext: failed to generate nonce")};
            got += (size_t)r;
        }
    }
    const int rc = pk_gcm_seal_batch(h_, plain, poff, plen, nonce_arr, n, sealed, soff);
    if (rc != PK_GCM_OK) return {{}, std::string("failed to encrypt plaintext: ") + pk_gcm_strerror(rc)};   // cipher.go:119
    BatchResult out;
    out.values.reserve(n);
    for (uint32_t i = 0; i < n; ++i)
        out.values.emplace_back(sealed + soff[i], sealed + soff[i] + plen[i] + PK_GCM_OVERHEAD);
    return out;
}

BatchResult GpuCipher::BatchDecrypt(const std::vector<Bytes>& ciphertexts)
{
    if (init_error_) return {{}, init_error_};
    const uint32_t n = (uint32_t)ciphertexts.size();
    if (n == 0) return {{}, std::nullopt};
    uint64_t* soff = reinterpret_cast<uint64_t*>(arena(2, (size_t)n * 8));
    uint64_t* poff = reinterpret_cast<uint64_t*>(arena(3, (size_t)n * 8));
    uint32_t* slen = reinterpret_cast<uint32_t*>(arena(4, (size_t)n * 4));
    uint32_t* status = reinterpret_cast<uint32_t*>(arena(6, (size_t)n * 4));
    if (!soff || !poff || !slen || !status) return {{}, std::string("failed to decrypt ciphertext: ") + pk_gcm_strerror(PK_GCM_ERR_CUDA)};
    std::memset(status, 0, (size_t)n * 4);
    uint64_t sat = 0, pat = 0;
    for (uint32_t i = 0; i < n; ++i) {
        slen[i] = (uint32_t)ciphertexts[i].size();
        soff[i] = sat; poff[i] = pat;
        sat += round16(slen[i]);
        pat += round16(slen[i] >= PK_GCM_OVERHEAD ? slen[i] - PK_GCM_OVERHEAD : 0);
    }
    uint8_t* sealed = arena(0, sat + 16);
    uint8_t* plain = arena(1, pat + 16);
    if (!sealed || !plain) return {{}, std::string("failed to decrypt ciphertext: ") + pk_gcm_strerror(PK_GCM_ERR_CUDA)};
    for (uint32_t i = 0; i < n; ++i)
        if (slen[i]) std::memcpy(sealed + soff[i], ciphertexts[i].data(), slen[i]);
    const int rc = pk_gcm_open_batch(h_, sealed, soff, slen, n, plain, poff, status);
    if (rc == PK_GCM_ERR_AUTH) {
        for (uint32_t i = 0; i < n; ++i) {               // the first failing item decides the message, cipher.go:133-137
            if (status[i] == 2)
                return {{}, "failed to decrypt ciphertext: ciphertext too short: " + std::to_string(slen[i]) +
                                " bytes, expected at least " + std::to_string(PK_GCM_NONCE_BYTES) + " bytes"};   // cipher.go:70
            if (status[i] == 1)
                return {{}, std::string("failed to decrypt ciphertext: failed to decrypt: cipher: message authentication failed")};   // cipher.go:80
        }
    }
    if (rc != PK_GCM_OK) return {{}, std::string("failed to decrypt ciphertext: ") + pk_gcm_strerror(rc)};
    BatchResult out;
    out.values.reserve(n);
    for (uint32_t i = 0; i < n; ++i)
        out.values.emplace_back(plain + poff[i], plain + poff[i] + slen[i] - PK_GCM_OVERHEAD);
    return out;
}

// ---- the reference's four exported functions (cipher.go:92-141) ----
// They take the key on every call.  Building a GPU context (device buffers, pinned arenas, the
// key's tables) costs a millisecond or two — more than sealing 65,536 secrets — so the last few
// contexts are kept, keyed by the key bytes, and rebuilt only when a call needs more room.
namespace {
struct Cached {
    Bytes key;
    uint32_t max_messages = 0;
    uint64_t max_bytes = 0;
    std::unique_ptr<GpuCipher> cipher;
    uint64_t last_use = 0;
};
std::mutex g_cache_mu;                                    // also serialises calls: a context is one-caller-at-a-time
std::vector<Cached> g_cache;
uint64_t g_cache_clock = 0;
constexpr size_t kCachedContexts = 4;

// Caller holds g_cache_mu.
GpuCipher& context_for(const Bytes& key, uint32_t messages, uint64_t bytes)
{
    Cached* slot = nullptr;
    for (Cached& c : g_cache)
        if (c.key.size() == key.size() && std::memcmp(c.key.data(), key.data(), key.size()) == 0) slot = &c;
    if (!slot) {
        if (g_cache.size() < kCachedContexts) {
            g_cache.emplace_back();
            slot = &g_cache.back();
        } else {
            slot = &g_cache[0];
            for (Cached& c : g_cache) if (c.last_use < slot->last_use) slot = &c;
            std::fill(slot->key.begin(), slot->key.end(), (uint8_t)0);
            slot->cipher.reset();
            slot->max_messages = 0;
            slot->max_bytes = 0;
        }
        slot->key = key;
    }
    if (!slot->cipher || slot->cipher->init_error() || messages > slot->max_messages || bytes > slot->max_bytes) {
        slot->cipher.reset();                             // free the old buffers before allocating larger ones
        slot->max_messages = std::max<uint32_t>(messages + messages / 2, 1024u);
        slot->max_bytes = std::max<uint64_t>(bytes + bytes / 2, 1u << 20);
        slot->cipher.reset(new GpuCipher(key, 0, slot->max_messages, slot->max_bytes));
    }
    slot->last_use = ++g_cache_clock;
    return *slot->cipher;
}
}  // namespace

void DropCachedContexts()
{
    std::lock_guard<std::mutex> g(g_cache_mu);
    for (Cached& c : g_cache) std::fill(c.key.begin(), c.key.end(), (uint8_t)0);
    g_cache.clear();
}

BatchResult BatchEncrypt(const Bytes& key, const std::vector<Bytes>& plaintexts)
{
    if (Error e = validateKey(key)) return {{}, e};       // cipher.go:111-113
    uint64_t bytes = 64;
    for (const Bytes& p : plaintexts) bytes += round16(p.size() + PK_GCM_OVERHEAD);
    std::lock_guard<std::mutex> g(g_cache_mu);
    return context_for(key, (uint32_t)plaintexts.size() + 1, bytes).BatchEncrypt(plaintexts);
}

BatchResult BatchDecrypt(const Bytes& key, const std::vector<Bytes>& ciphertexts)
{
    if (Error e = validateKey(key)) return {{}, e};       // cipher.go:127-129
    uint64_t bytes = 64;
    for (const Bytes& c : ciphertexts) bytes += round16(c.size());
    std::lock_guard<std::mutex> g(g_cache_mu);
    return context_for(key, (uint32_t)ciphertexts.size() + 1, bytes).BatchDecrypt(ciphertexts);
}

namespace {
// a single Encrypt / Decrypt returns encryptKey's / decryptKey's error without the batch wrapper
std::string unwrap(const std::string& msg, const std::string& prefix)
{
    return msg.rfind(prefix, 0) == 0 ? msg.substr(prefix.size()) : msg;
}
}  // namespace

BytesResult Encrypt(const Bytes& key, const Bytes& plaintext)
{
    if (Error e = validateKey(key)) return {{}, e};       // cipher.go:93-95
    BatchResult r = BatchEncrypt(key, {plaintext});
    if (r.error) return {{}, unwrap(*r.error, "failed to encrypt plaintext: ")};
    return {std::move(r.values[0]), std::nullopt};
}

BytesResult Decrypt(const Bytes& key, const Bytes& ciphertext)
{
    if (Error e = validateKey(key)) return {{}, e};       // cipher.go:100-102
    BatchResult r = BatchDecrypt(key, {ciphertext});
    if (r.error) return {{}, unwrap(*r.error, "failed to decrypt ciphertext: ")};   // cipher.go:103, 69-81
    return {std::move(r.values[0]), std::nullopt};
}

}  // namespace polykey::security

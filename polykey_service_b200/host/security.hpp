// security.hpp — C++ mirror of internal/adapters/security/cipher.go over the batched GPU stage.
//
//   Go (reference)                                   here
//   -----------------------------------------------  -------------------------------------------
//   Encrypt(key, plaintext)        cipher.go:92-97    polykey::security::Encrypt
//   Decrypt(key, ciphertext)       cipher.go:99-104   polykey::security::Decrypt
//   BatchEncrypt(key, plaintexts)  cipher.go:110-124  polykey::security::BatchEncrypt
//   BatchDecrypt(key, ciphertexts) cipher.go:126-141  polykey::security::BatchDecrypt
//   validateKey                    cipher.go:15-19    same message: "key length must be 32 bytes, got N bytes"
//   generateNonce (crypto/rand)    cipher.go:25-31    getrandom(2)
//
// Same results and error behaviour: output nonce || ciphertext || tag; a batch fails as a whole
// on the first bad item, with the reference's wrapped messages.  The free functions take the
// key on every call like the reference's; the last few GPU contexts are cached by key (a context
// costs more to build than a batch of 65,536 secrets to seal), DropCachedContexts() releases them.
#pragma once
#include <cstdint>
#include <optional>
#include <string>
#include <vector>

#include "../../include/pk_gcm.h"

namespace polykey::security {

using Bytes = std::vector<uint8_t>;
using Error = std::optional<std::string>;              // Go's error; nullopt = nil
struct BytesResult { Bytes value; Error error; };
struct BatchResult { std::vector<Bytes> values; Error error; };

Error validateKey(const Bytes& key);                    // cipher.go:15-19

// One key on one GPU (key schedule and GHASH table built once — the reference rebuilds them per item).
class GpuCipher {
public:
    GpuCipher(const Bytes& key, int device = 0, uint32_t max_messages = 65536, uint64_t max_bytes = 64u << 20);
    ~GpuCipher();
    GpuCipher(const GpuCipher&) = delete;
    GpuCipher& operator=(const GpuCipher&) = delete;
    const Error& init_error() const { return init_error_; }
    // nonces == nullptr: drawn from the OS like generateNonce; otherwise 12 bytes per plaintext
    BatchResult BatchEncrypt(const std::vector<Bytes>& plaintexts, const std::vector<Bytes>* nonces = nullptr);
    BatchResult BatchDecrypt(const std::vector<Bytes>& ciphertexts);
private:
    // grow-only page-locked arenas (pk_gcm_host_alloc): 0 = input, 1 = output, 2 / 3 = their slot
    // offsets, 4 = lengths, 5 = nonces, 6 = status — every array of a call page-locked, so that
    // small batches of long messages are worked on in place (pk_gcm_set_direct_bytes)
    uint8_t* arena(int which, size_t bytes);
    pk_gcm* h_ = nullptr;
    Error init_error_;
    uint8_t* arena_[7] = {};
    size_t arena_bytes_[7] = {};
};

BytesResult Encrypt(const Bytes& key, const Bytes& plaintext);
BytesResult Decrypt(const Bytes& key, const Bytes& ciphertext);
BatchResult BatchEncrypt(const Bytes& key, const std::vector<Bytes>& plaintexts);
BatchResult BatchDecrypt(const Bytes& key, const std::vector<Bytes>& ciphertexts);
void DropCachedContexts();                              // zeroes and frees the cached per-key GPU contexts

}  // namespace polykey::security

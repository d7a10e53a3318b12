// host_security_test.cpp — the C++ mirror of internal/adapters/security/cipher.go
// (polykey_service_b200/host/security.{hpp,cpp}).  `cpu` runs what needs no GPU (key validation,
// error texts); `cuda` runs round trips and failure behaviour on the device.  Built and run by
// tests/test_host_service.py.  Test-only code.
#include <cstdio>
#include <cstring>

#include "../polykey_service_b200/host/security.hpp"

using namespace polykey::security;
static int g_fail = 0;
#define CHECK(c) do { if (!(c)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); ++g_fail; } } while (0)

static Bytes B(const char* s) { return Bytes(s, s + std::strlen(s)); }

static void test_key_validation()
{
    for (size_t n : {0u, 16u, 31u, 33u}) {
        const Bytes key(n, 7);
        BytesResult e = Encrypt(key, B("secret"));                                  // cipher.go:15-19
        CHECK(e.error && *e.error == "key length must be 32 bytes, got " + std::to_string(n) + " bytes");
        CHECK(Decrypt(key, B("x")).error.has_value());
        CHECK(BatchEncrypt(key, {B("a")}).error.has_value());
        CHECK(BatchDecrypt(key, {B("a")}).error.has_value());
    }
}

static void test_round_trip()
{
    const Bytes key(32, 0x42);
    std::vector<Bytes> pts = {B(""), B("sk-live-1234567890abcdef"), Bytes(1000, 0xAB), B("x")};
    BatchResult enc = BatchEncrypt(key, pts);
    CHECK(!enc.error && enc.values.size() == pts.size());
    for (size_t i = 0; i < pts.size(); ++i) CHECK(enc.values[i].size() == pts[i].size() + 28);   // nonce + tag
    BatchResult enc2 = BatchEncrypt(key, pts);
    CHECK(!enc2.error && enc2.values[1] != enc.values[1]);                          // fresh random nonces
    BatchResult dec = BatchDecrypt(key, enc.values);
    CHECK(!dec.error && dec.values == pts);
    BytesResult one = Encrypt(key, pts[1]);
    CHECK(!one.error);
    BytesResult back = Decrypt(key, one.value);
    CHECK(!back.error && back.value == pts[1]);
    // failures: the batch fails as a whole with the reference's wrapped messages
    std::vector<Bytes> bad = enc.values;
    bad[2][40] ^= 1;
    BatchResult f = BatchDecrypt(key, bad);
    CHECK(f.error && *f.error == "failed to decrypt ciphertext: failed to decrypt: cipher: message authentication failed" && f.values.empty());
    BytesResult s = Decrypt(key, B("short"));                                       // cipher.go:69-71
    CHECK(s.error && *s.error == "ciphertext too short: 5 bytes, expected at least 12 bytes");
    BytesResult w = Decrypt(Bytes(32, 0x43), one.value);                            // wrong key
    CHECK(w.error && *w.error == "failed to decrypt: cipher: message authentication failed");
    // a held cipher amortises the key set-up and accepts caller nonces (deterministic output)
    GpuCipher c(key, 0, 16, 1 << 16);
    CHECK(!c.init_error());
    std::vector<Bytes> nonces(pts.size(), Bytes(12, 9));
    BatchResult d1 = c.BatchEncrypt(pts, &nonces), d2 = c.BatchEncrypt(pts, &nonces);
    CHECK(!d1.error && d1.values == d2.values);
    // the free functions cache contexts by key: more keys than cache slots, growing batches, and
    // every result still opens under its own key only
    std::vector<Bytes> keys;
    std::vector<BatchResult> sealed;
    for (int k = 0; k < 7; ++k) {
        keys.emplace_back(32, (uint8_t)(0x50 + k));
        std::vector<Bytes> batch((size_t)(3 + 700 * k), Bytes((size_t)(10 + 37 * k), (uint8_t)k));
        sealed.push_back(BatchEncrypt(keys.back(), batch));
        CHECK(!sealed.back().error && sealed.back().values.size() == batch.size());
    }
    for (int k = 6; k >= 0; --k) {
        BatchResult d = BatchDecrypt(keys[(size_t)k], sealed[(size_t)k].values);
        CHECK(!d.error && d.values.size() == (size_t)(3 + 700 * k) && d.values[0] == Bytes((size_t)(10 + 37 * k), (uint8_t)k));
        BatchResult wrong = BatchDecrypt(keys[(size_t)((k + 1) % 7)], sealed[(size_t)k].values);
        CHECK(wrong.error.has_value());
    }
    DropCachedContexts();
    BatchResult again = BatchDecrypt(keys[2], sealed[2].values);                     // rebuilt on demand
    CHECK(!again.error && again.values.size() == sealed[2].values.size());
    DropCachedContexts();
}

int main(int argc, char** argv)
{
    test_key_validation();
    if (argc > 1 && std::strcmp(argv[1], "cuda") == 0) test_round_trip();
    std::printf(g_fail ? "SECURITY TESTS FAILED (%d)\n" : "SECURITY TESTS OK\n", g_fail);
    return g_fail ? 1 : 0;
}

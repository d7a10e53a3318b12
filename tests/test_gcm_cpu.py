"""CPU: the AES-256-GCM arithmetic the GPU kernels run (csrc/pk_gcm.cuh, compiled here with g++)
against the published known-answer tests and OpenSSL (through Python `cryptography`).

Reference: internal/adapters/security/cipher.go:92-141 — AES-256-GCM, 12-byte nonce, no AAD, output
nonce || ciphertext || tag.  The reference has no tests; the algorithm is NIST SP 800-38D, so parity
is pinned by the standard's test cases and an independent implementation."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
from cryptography.hazmat.primitives.ciphers import Cipher, algorithms, modes
from cryptography.hazmat.primitives.ciphers.aead import AESGCM

HERE = os.path.dirname(os.path.abspath(__file__))

# McGrew & Viega, "The Galois/Counter Mode of Operation (GCM)", AES-256 cases with a 96-bit IV and
# no additional data: (key, iv, plaintext, ciphertext, tag)
KATS = [
    ("00" * 32, "00" * 12, "", "", "530f8afbc74536b9a963b4f1c4cb738b"),                       # test case 13
    ("00" * 32, "00" * 12, "00" * 16, "cea7403d4d606b6e074ec5d3baf39d18",
     "d0d1c8a799996bf0265b98b5d48ab919"),                                                     # test case 14
    ("feffe9928665731c6d6a8f9467308308feffe9928665731c6d6a8f9467308308", "cafebabefacedbaddecaf888",
     "d9313225f88406e5a55909c5aff5269a86a7a9531534f7da2e4c303d8a318a72"
     "1c3c0c95956809532fcf0e2449a6b525b16aedf5aa0de657ba637b391aafd255",
     "522dc1f099567d07f47f37a32a84427d643a8cdcbfe5c0c97598a2bd2555d1aa"
     "8cb08e48590dbb3da7b08b1056828838c5f61e6393ba7a0abcc9f662898015ad",
     "b094dac5d93471bdec1a502270e3cc6c"),                                                     # test case 15
]


@pytest.fixture(scope="module")
def gcmh(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("gcmh") / "libgcmhost.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so,
                           os.path.join(HERE, "gcm_host_check.cpp")])
    L = C.CDLL(so)
    L.gcmh_seal.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint32, C.c_char_p]
    L.gcmh_open.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, C.c_char_p]
    L.gcmh_seal_lanes.argtypes = L.gcmh_seal.argtypes
    L.gcmh_open_lanes.argtypes = L.gcmh_open.argtypes
    L.gcmh_group.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint32, C.c_int, C.c_char_p, C.c_char_p]
    L.gcmh_encrypt_block.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p]
    L.gcmh_check_ghash.argtypes = [C.c_char_p, C.c_uint64, C.c_int]
    L.gcmh_check_table4.argtypes = [C.c_char_p, C.c_uint64, C.c_int]
    return L


def seal(gcmh, key, nonce, pt):
    out = C.create_string_buffer(len(pt) + 28)
    gcmh.gcmh_seal(key, nonce, pt, len(pt), out)
    return out.raw


def test_known_answer_tests(gcmh):
    for key, iv, pt, ct, tag in KATS:
        got = seal(gcmh, bytes.fromhex(key), bytes.fromhex(iv), bytes.fromhex(pt))
        assert got.hex() == iv + ct + tag


def test_aes_block_fips197_c3(gcmh):
    """FIPS-197 Appendix C.3 (AES-256)."""
    out = C.create_string_buffer(16)
    gcmh.gcmh_encrypt_block(bytes(range(32)), bytes.fromhex("00112233445566778899aabbccddeeff"), out)
    assert out.raw.hex() == "8ea2b7ca516745bfeafc49904b496089"
    rng = np.random.default_rng(1)
    for _ in range(50):
        key, blk = rng.bytes(32), rng.bytes(16)
        gcmh.gcmh_encrypt_block(key, blk, out)
        enc = Cipher(algorithms.AES(key), modes.ECB()).encryptor()
        assert out.raw == enc.update(blk) + enc.finalize()


def test_ghash_table_method_equals_bitwise(gcmh):
    rng = np.random.default_rng(2)
    for _ in range(20):
        assert gcmh.gcmh_check_ghash(rng.bytes(32), int(rng.integers(1, 2**62)), 200) == 0


def test_ghash_4bit_tables_of_the_closing_powers_equal_bitwise(gcmh):
    """pw4[e-1] (16 entries of H^e's Shoup table) walked nibble by nibble = X · H^e, e = 1..32."""
    rng = np.random.default_rng(12)
    for _ in range(3):
        assert gcmh.gcmh_check_table4(rng.bytes(32), int(rng.integers(1, 2**62)), 20) == 0


def test_seal_matches_openssl_all_lengths(gcmh):
    rng = np.random.default_rng(3)
    for length in list(range(0, 70)) + [127, 128, 129, 255, 256, 1000, 4096, 4099]:
        key, nonce, pt = rng.bytes(32), rng.bytes(12), rng.bytes(length)
        assert seal(gcmh, key, nonce, pt) == nonce + AESGCM(key).encrypt(nonce, pt, None), length


def test_open_round_trip_and_tamper(gcmh):
    rng = np.random.default_rng(4)
    for length in (0, 1, 15, 16, 17, 100, 1000):
        key, nonce, pt = rng.bytes(32), rng.bytes(12), rng.bytes(length)
        sealed = nonce + AESGCM(key).encrypt(nonce, pt, None)            # produced by OpenSSL
        out = C.create_string_buffer(max(1, length))
        assert gcmh.gcmh_open(key, sealed, len(sealed), out) == 0 and out.raw[:length] == pt
        for pos in {0, 11, 12, len(sealed) - 1, len(sealed) // 2}:        # nonce, body, tag
            bad = bytearray(sealed)
            bad[pos] ^= 0x01
            assert gcmh.gcmh_open(key, bytes(bad), len(bad), out) == 1
        assert gcmh.gcmh_open(rng.bytes(32), sealed, len(sealed), out) == 1      # wrong key
    assert gcmh.gcmh_open(bytes(32), b"short", 5, C.create_string_buffer(1)) == 2  # cipher.go:69-71
    assert gcmh.gcmh_open(bytes(32), bytes(20), 20, C.create_string_buffer(1)) == 1


def test_warp_per_message_form_matches_openssl(gcmh):
    """The 32-lane split (blocks strided over lanes, Horner with H^32, closing power per lane) is
    the same function as the serial GHASH: every length around the 32-block period, and the tag
    check of the opening direction."""
    rng = np.random.default_rng(5)
    lengths = list(range(0, 40)) + [495, 496, 497, 511, 512, 513, 527, 528, 529, 1007, 1008, 1024,
                                    1039, 1040, 1041, 2048, 4096, 4099, 16384 + 5]
    for length in lengths:
        key, nonce, pt = rng.bytes(32), rng.bytes(12), rng.bytes(length)
        want = nonce + AESGCM(key).encrypt(nonce, pt, None)
        out = C.create_string_buffer(length + 28)
        gcmh.gcmh_seal_lanes(key, nonce, pt, length, out)
        assert out.raw == want, length
        back = C.create_string_buffer(max(1, length))
        assert gcmh.gcmh_open_lanes(key, want, len(want), back) == 0 and back.raw[:length] == pt
        bad = bytearray(want)
        bad[len(bad) // 2] ^= 0x20
        assert gcmh.gcmh_open_lanes(key, bytes(bad), len(bad), back) == 1


def test_lane_group_form_matches_openssl(gcmh):
    """The 4-lane split the short-message kernel uses (blocks strided over 4 lanes, Horner with H^4,
    closing powers H^1..H^4 by table, E_K(J0) on the lane that owns the lengths block) is the same
    function as the serial form: every length from 0 to 300 bytes, both directions."""
    rng = np.random.default_rng(6)
    for length in list(range(0, 301)) + [511, 512, 513, 1024, 2047]:
        key, nonce, pt = rng.bytes(32), rng.bytes(12), rng.bytes(length)
        want = AESGCM(key).encrypt(nonce, pt, None)
        ct, tag = C.create_string_buffer(max(1, length)), C.create_string_buffer(16)
        assert gcmh.gcmh_group(key, nonce, pt, length, 0, ct, tag) == 1          # exactly one lane made E_K(J0)
        assert ct.raw[:length] + tag.raw == want, length
        back, tag2 = C.create_string_buffer(max(1, length)), C.create_string_buffer(16)
        gcmh.gcmh_group(key, nonce, want[:length], length, 1, back, tag2)
        assert back.raw[:length] == pt and tag2.raw == want[length:], length

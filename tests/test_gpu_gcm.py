"""GPU: batched AES-256-GCM through the C ABI (include/pk_gcm.h) against OpenSSL (Python
`cryptography`) and the standard's known-answer tests — bit-exact.

Reference: BatchEncrypt / BatchDecrypt, internal/adapters/security/cipher.go:110-141; output format
nonce || ciphertext || tag (cipher.go:51-52); key must be 32 bytes (cipher.go:15-19); a ciphertext
shorter than a nonce and any authentication failure are errors (cipher.go:69-71, 78-81)."""
import numpy as np
import pytest
from cryptography.hazmat.primitives.ciphers.aead import AESGCM

from test_gcm_cpu import KATS

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G(pk):
    from polykey_service_b200 import gcm
    return gcm


def test_known_answer_tests_on_gpu(G):
    for key, iv, pt, ct, tag in KATS:
        with G.Gcm(bytes.fromhex(key), max_messages=4, max_bytes=4096) as g:
            out = g.batch_encrypt([bytes.fromhex(pt)], [bytes.fromhex(iv)])
            assert out[0].hex() == iv + ct + tag
            assert g.batch_decrypt(out) == [bytes.fromhex(pt)]


def test_batch_matches_openssl(G):
    rng = np.random.default_rng(11)
    key = rng.bytes(32)
    lengths = list(range(0, 80)) + [127, 128, 129, 255, 256, 257, 1000, 4096, 4099, 20000] + \
        rng.integers(0, 600, 3000).tolist()
    pts = [rng.bytes(int(l)) for l in lengths]
    nonces = [rng.bytes(12) for _ in lengths]
    ref = AESGCM(key)
    with G.Gcm(key, max_messages=8192, max_bytes=8 << 20) as g:
        g.set_chunks(1)
        before = g.launch_count
        sealed = g.batch_encrypt(pts, nonces)
        assert g.launch_count == before + 2                        # thread-per-message + warp-per-message kernels
        for s, p, nce in zip(sealed, pts, nonces):
            assert s == nce + ref.encrypt(nce, p, None)
        assert g.batch_decrypt(sealed) == pts                      # round trip on the GPU
        # and the GPU opens what OpenSSL sealed
        theirs = [nce + ref.encrypt(nce, p, None) for p, nce in zip(pts[:200], nonces[:200])]
        assert g.batch_decrypt(theirs) == pts[:200]


def test_tampering_and_short_ciphertexts(G):
    rng = np.random.default_rng(12)
    key = rng.bytes(32)
    pts = [rng.bytes(50) for _ in range(64)]
    nonces = [rng.bytes(12) for _ in range(64)]
    with G.Gcm(key, max_messages=128, max_bytes=1 << 20) as g:
        sealed = g.batch_encrypt(pts, nonces)
        bad = list(sealed)
        bad[3] = bytes([bad[3][0] ^ 1]) + bad[3][1:]                # nonce bit
        bad[17] = bad[17][:20] + bytes([bad[17][20] ^ 0x80]) + bad[17][21:]     # ciphertext bit
        bad[40] = bad[40][:-1] + bytes([bad[40][-1] ^ 2])           # tag bit
        bad[50] = b"short"                                          # cipher.go:69-71
        bad[51] = bytes(20)                                         # nonce but no room for a tag
        with pytest.raises(G.GcmError) as e:                        # BatchDecrypt fails as a whole
            g.batch_decrypt(bad)
        assert e.value.code == -10
        st = e.value.status
        assert [int(i) for i in np.nonzero(st)[0]] == [3, 17, 40, 50, 51]
        assert st[50] == 2 and st[51] == 1 and st[3] == 1
        assert g.batch_decrypt(sealed) == pts                       # the handle is still healthy
    with G.Gcm(rng.bytes(32), max_messages=128, max_bytes=1 << 20) as other:   # wrong key
        with pytest.raises(G.GcmError):
            other.batch_decrypt(sealed[:4])


def test_key_length_and_argument_errors(G):
    for bad_len in (0, 16, 24, 31, 33):
        with pytest.raises(G.GcmError) as e:                        # validateKey, cipher.go:15-19
            G.Gcm(bytes(bad_len) or b"", max_messages=4, max_bytes=4096)
        assert e.value.code in (-9, -1)
    with G.Gcm(bytes(32), max_messages=4, max_bytes=256) as g:
        assert g.batch_encrypt([], []) == []
        with pytest.raises(G.GcmError) as e:
            g.batch_encrypt([b"x"] * 5, [bytes(12)] * 5)
        assert e.value.code == -4
        with pytest.raises(G.GcmError) as e:
            g.batch_encrypt([bytes(400)], [bytes(12)])
        assert e.value.code == -4
        plen = np.array([4, 4], dtype=np.uint32)
        off = np.array([0, 8], dtype=np.uint64)                    # not 16-byte aligned
        with pytest.raises(G.GcmError) as e:
            g.seal_arena(np.zeros(64, np.uint8), off, plen, np.zeros(24, np.uint8), np.zeros(256, np.uint8),
                         np.array([0, 48], dtype=np.uint64))
        assert e.value.code == -2
        # an offset near 2^64: offset + length wraps to a small number — must be refused, not indexed
        huge = np.array([0, (1 << 64) - 16], dtype=np.uint64)
        for in_off, out_off in ((huge, np.array([0, 48], dtype=np.uint64)), (np.array([0, 16], dtype=np.uint64), huge)):
            with pytest.raises(G.GcmError) as e:
                g.seal_arena(np.zeros(64, np.uint8), in_off, np.array([4, 32], dtype=np.uint32), np.zeros(24, np.uint8),
                             np.zeros(256, np.uint8), out_off)
            assert e.value.code == -4


def test_large_batch_of_secrets(G):
    """65,536 secrets of 32-256 bytes: the shape this helper exists for."""
    rng = np.random.default_rng(13)
    key = rng.bytes(32)
    n = 65536
    lens = rng.integers(32, 257, n)
    blob = rng.bytes(int(lens.sum()))
    cuts = np.concatenate([[0], np.cumsum(lens)])
    pts = [blob[int(cuts[i]):int(cuts[i + 1])] for i in range(n)]
    nonce_blob = rng.bytes(12 * n)
    nonces = [nonce_blob[12 * i:12 * i + 12] for i in range(n)]
    with G.Gcm(key, max_messages=n, max_bytes=32 << 20) as g:
        sealed = g.batch_encrypt(pts, nonces)
        ref = AESGCM(key)
        for i in rng.integers(0, n, 2000).tolist() + [0, n - 1]:
            assert sealed[i] == nonces[i] + ref.encrypt(nonces[i], pts[i], None)
        assert g.batch_decrypt(sealed) == pts


def test_long_messages_take_the_warp_kernel(G):
    """Only long messages: one launch (the warp-per-message kernel), every length around the
    32-block period, sealed bytes equal OpenSSL's, tampered long messages fail and release nothing."""
    rng = np.random.default_rng(14)
    key = rng.bytes(32)
    lengths = [256, 257, 495, 496, 497, 511, 512, 513, 527, 528, 529, 1007, 1008, 1024, 1039, 1040,
               1041, 2048, 4095, 4096, 4097, 16384 + 5, 65536, 100001] + rng.integers(256, 9000, 300).tolist()
    pts = [rng.bytes(int(l)) for l in lengths]
    nonces = [rng.bytes(12) for _ in lengths]
    ref = AESGCM(key)
    with G.Gcm(key, max_messages=512, max_bytes=8 << 20) as g:
        g.set_chunks(1)
        before = g.launch_count
        sealed = g.batch_encrypt(pts, nonces)
        assert g.launch_count == before + 1
        for s, p, nce in zip(sealed, pts, nonces):
            assert s == nce + ref.encrypt(nce, p, None), len(p)
        assert g.batch_decrypt(sealed) == pts
        bad = list(sealed)
        for k in (0, 5, 20, 23):
            b = bytearray(bad[k])
            b[len(b) // 2] ^= 0x04
            bad[k] = bytes(b)
        with pytest.raises(G.GcmError) as e:
            g.batch_decrypt(bad)
        assert e.value.code == -10
        assert [int(i) for i in np.nonzero(e.value.status)[0]] == [0, 5, 20, 23]


def test_large_batch_with_a_long_tail(G):
    """A batch big enough for the higher thresholds (long_threshold in pk_gcm.cu): mostly short
    secrets plus a tail of multi-KB ones, both kernels in one call."""
    rng = np.random.default_rng(15)
    key = rng.bytes(32)
    n = 12000
    lens = rng.integers(16, 200, n)
    lens[rng.integers(0, n, 400)] = rng.integers(900, 6000, 400)
    blob = rng.bytes(int(lens.sum()))
    cuts = np.concatenate([[0], np.cumsum(lens)])
    pts = [blob[int(cuts[i]):int(cuts[i + 1])] for i in range(n)]
    nonce_blob = rng.bytes(12 * n)
    nonces = [nonce_blob[12 * i:12 * i + 12] for i in range(n)]
    ref = AESGCM(key)
    with G.Gcm(key, max_messages=n, max_bytes=16 << 20) as g:
        sealed = g.batch_encrypt(pts, nonces)
        for i in np.nonzero(lens >= 900)[0].tolist() + rng.integers(0, n, 500).tolist():
            assert sealed[i] == nonces[i] + ref.encrypt(nonces[i], pts[i], None)
        assert g.batch_decrypt(sealed) == pts


def test_pipelined_chunks_give_the_same_bytes(G):
    """A batch cut into 2..8 chunks (copy in / kernels / copy out overlapped) equals the unchunked
    call, in both directions, with long messages straddling the cuts and a failing message whose
    status lands in the right place."""
    rng = np.random.default_rng(16)
    key = rng.bytes(32)
    n = 3000
    lens = rng.integers(0, 700, n)
    lens[rng.integers(0, n, 60)] = rng.integers(2000, 60000, 60)
    pts = [rng.bytes(int(l)) for l in lens]
    nonces = [rng.bytes(12) for _ in range(n)]
    with G.Gcm(key, max_messages=n, max_bytes=16 << 20) as g:
        g.set_chunks(1)
        want = g.batch_encrypt(pts, nonces)
        ref = AESGCM(key)
        for i in rng.integers(0, n, 300).tolist():
            assert want[i] == nonces[i] + ref.encrypt(nonces[i], pts[i], None)
        for chunks in (0, 2, 3, 8):
            g.set_chunks(chunks)
            assert g.batch_encrypt(pts, nonces) == want, chunks
            assert g.batch_decrypt(want) == pts, chunks
            bad = list(want)
            for k in (1, n // 2, n - 1):
                b = bytearray(bad[k])
                b[-1] ^= 1
                bad[k] = bytes(b)
            with pytest.raises(G.GcmError) as e:
                g.batch_decrypt(bad)
            assert [int(i) for i in np.nonzero(e.value.status)[0]] == [1, n // 2, n - 1], chunks
        with pytest.raises(G.GcmError):
            g.set_chunks(9)


def test_small_batches_in_place(G):
    """Every array of the call page-locked and the batch small: sealed and opened in place (no
    copies issued) — same bytes as OpenSSL, tampering still detected, pageable arrays or a zero
    threshold take the copying path."""
    rng = np.random.default_rng(17)
    key = rng.bytes(32)
    ref = AESGCM(key)
    lens = np.array([256, 257, 271, 272, 273, 495, 496, 511, 512, 513, 1000, 4096, 5000] + rng.integers(256, 3000, 200).tolist(),
                    dtype=np.uint32)                                              # all long: the warp-per-message kernel's
    n = lens.size
    with G.Gcm(key, max_messages=512, max_bytes=4 << 20) as g:
        poff, ptotal = G.slots(lens)
        soff, stotal = G.slots(lens.astype(np.uint64) + G.OVERHEAD)
        pin = lambda a: (lambda d: (d.__setitem__(slice(0, a.size), a), d[:a.size])[1])(g.pinned(max(a.nbytes, 16), a.dtype))
        plain = pin(rng.integers(0, 256, max(ptotal, 16), dtype=np.uint8))
        nonces = pin(rng.integers(0, 256, 12 * n, dtype=np.uint8))
        p_lens, p_poff, p_soff = pin(lens), pin(poff), pin(soff)
        sealed = pin(np.zeros(max(stotal, 16), dtype=np.uint8))
        before = g.direct_count
        g.seal_arena(plain, p_poff, p_lens, nonces, sealed, p_soff)
        assert g.direct_count == before + 1
        for i in range(n):
            nce = bytes(nonces[12 * i:12 * i + 12])
            pt = bytes(plain[int(poff[i]):int(poff[i]) + int(lens[i])])
            got = bytes(sealed[int(soff[i]):int(soff[i]) + int(lens[i]) + G.OVERHEAD])
            assert got == nce + ref.encrypt(nce, pt, None), int(lens[i])
        # open in place, then with one ciphertext bit flipped
        slens = pin((lens + G.OVERHEAD).astype(np.uint32))
        back = pin(np.zeros(max(ptotal, 16), dtype=np.uint8))
        rc = g._L.pk_gcm_open_batch(g._h, sealed.ctypes.data, p_soff.ctypes.data, slens.ctypes.data, n, back.ctypes.data,
                                    p_poff.ctypes.data, pin(np.zeros(n, dtype=np.uint32)).ctypes.data)
        assert rc == 0 and g.direct_count == before + 2
        for i in range(n):
            assert bytes(back[int(poff[i]):int(poff[i]) + int(lens[i])]) == bytes(plain[int(poff[i]):int(poff[i]) + int(lens[i])])
        sealed[int(soff[10]) + 20] ^= 1
        status = pin(np.zeros(n, dtype=np.uint32))
        rc = g._L.pk_gcm_open_batch(g._h, sealed.ctypes.data, p_soff.ctypes.data, slens.ctypes.data, n, back.ctypes.data,
                                    p_poff.ctypes.data, status.ctypes.data)
        assert rc == -10 and [int(i) for i in np.nonzero(status)[0]] == [10] and g.direct_count == before + 3
        assert not back[int(poff[10]):int(poff[10]) + int(lens[10])].any()        # nothing released for the failed one
        sealed[int(soff[10]) + 20] ^= 1
        # pageable lengths array -> copies; one short message in the batch -> copies; threshold 0 -> copies
        g.seal_arena(plain, p_poff, lens, nonces, sealed, p_soff)
        p_lens[3] = 40
        g.seal_arena(plain, p_poff, p_lens, nonces, sealed, p_soff)
        nce, pt = bytes(nonces[36:48]), bytes(plain[int(poff[3]):int(poff[3]) + 40])
        assert bytes(sealed[int(soff[3]):int(soff[3]) + 68]) == nce + ref.encrypt(nce, pt, None)
        p_lens[3] = lens[3]
        g.set_direct_bytes(0)
        g.seal_arena(plain, p_poff, p_lens, nonces, sealed, p_soff)
        assert g.direct_count == before + 3


def test_in_place_with_a_nonce_array_at_an_odd_address(G):
    """The kernels read a nonce as three words only where it is 4-byte aligned (sealed slots always
    are; a caller's nonce array need not be): the same batch sealed in place with the nonce array
    at offsets 0..3 of a page-locked block gives OpenSSL's bytes every time."""
    rng = np.random.default_rng(19)
    key = rng.bytes(32)
    ref = AESGCM(key)
    lens = np.array([256, 300, 511, 512, 777, 1024, 2049] + rng.integers(256, 2000, 40).tolist(), dtype=np.uint32)
    n = lens.size
    with G.Gcm(key, max_messages=64, max_bytes=1 << 20) as g:
        poff, ptotal = G.slots(lens)
        soff, stotal = G.slots(lens.astype(np.uint64) + G.OVERHEAD)
        pin = lambda a: (lambda d: (d.__setitem__(slice(0, a.size), a), d[:a.size])[1])(g.pinned(max(a.nbytes, 16), a.dtype))
        plain = pin(rng.integers(0, 256, max(ptotal, 16), dtype=np.uint8))
        p_lens, p_poff, p_soff = pin(lens), pin(poff), pin(soff)
        sealed = pin(np.zeros(max(stotal, 16), dtype=np.uint8))
        nonce_bytes = rng.integers(0, 256, 12 * n, dtype=np.uint8)
        block = g.pinned(12 * n + 16, np.uint8)
        for shift in range(4):
            nonces = block[shift:shift + 12 * n]
            nonces[:] = nonce_bytes
            assert nonces.ctypes.data % 4 == (block.ctypes.data + shift) % 4
            sealed[:] = 0
            before = g.direct_count
            g.seal_arena(plain, p_poff, p_lens, nonces, sealed, p_soff)
            assert g.direct_count == before + 1
            for i in range(n):
                nce = bytes(nonce_bytes[12 * i:12 * i + 12])
                pt = bytes(plain[int(poff[i]):int(poff[i]) + int(lens[i])])
                got = bytes(sealed[int(soff[i]):int(soff[i]) + int(lens[i]) + G.OVERHEAD])
                assert got == nce + ref.encrypt(nce, pt, None), (shift, int(lens[i]))

"""CPU: the build-defined parts of the SPEC behave as advertised — checked on the
oracle, which the GPU matches bit for bit.  The reference has no counterpart for any of this
(SURVEY.md §8 a6); these are properties, not parity."""
import numpy as np


def hashes(oracle, payloads, name=b"example_tool"):
    b = oracle.pack_batch([(name, p) for p in payloads])
    rc, out = oracle.route_batch(oracle.default_table(), b, threads=4, fast=True)
    assert rc == 0
    return out


def popcount64(x):
    return bin(int(x)).count("1")


def test_route_hash_avalanche(oracle):
    """Flipping one payload bit flips about half of the 64 hash bits, wherever the bit sits
    (first block, a later stripe, the last partial block, a second segment)."""
    rng = np.random.default_rng(1)
    for length in (1, 16, 17, 300, 2048, 16384 + 40):
        base = bytearray(rng.bytes(length))
        positions = sorted(set([0, length - 1, length // 2] + rng.integers(0, length, 24).tolist()))
        variants = [bytes(base)]
        for pos in positions:
            v = bytearray(base)
            v[pos] ^= 1 << int(rng.integers(0, 8))
            variants.append(bytes(v))
        h = hashes(oracle, variants)["route_hash"]
        flips = [popcount64(h[0] ^ x) for x in h[1:]]
        assert min(flips) >= 12 and 24 <= np.mean(flips) <= 40, (length, flips)


def test_same_flip_in_different_blocks_gives_different_hashes(oracle):
    """Stripe digests are ADDED, so a stripe's digest must not be an affine function of its block:
    otherwise the same bit flipped at the same offset of two different blocks moves the sum by the
    same amount and the two payloads collide.  (A digest without the per-stripe finaliser does
    exactly that — tried in round 2 to save its 64-bit multiplications on tiny payloads, and
    caught by this test: every single-bit variant of a payload must hash differently.)"""
    rng = np.random.default_rng(5)
    for length in (16, 48, 256, 2048):
        base = bytearray(rng.bytes(length))
        variants = [bytes(base)]
        for pos in range(min(length, 64)):
            for bit in range(8):
                v = bytearray(base)
                v[pos] ^= 1 << bit
                variants.append(bytes(v))
        h = hashes(oracle, variants)["route_hash"]
        assert len(set(h.tolist())) == len(variants), length
        flips = [popcount64(h[0] ^ x) for x in h[1:]]
        assert min(flips) >= 12, (length, min(flips))


def test_route_hash_depends_on_name_length_and_order(oracle):
    rng = np.random.default_rng(2)
    p = rng.bytes(500)
    seen = set()
    for name in (b"example_tool", b"example_too", b"example_tool ", b"struct_tool", b"", b"loot_elpmaxe"):
        seen.add(int(hashes(oracle, [p], name)["route_hash"][0]))
    assert len(seen) == 6
    # zero padding is not confusable with real zero bytes, and block order matters
    a, b = bytes(16) + b"x" * 16, b"x" * 16 + bytes(16)
    hs = hashes(oracle, [b"abc", b"abc\0", b"abc\0\0", a, b])["route_hash"]
    assert len(set(hs.tolist())) == 5


def test_no_collisions_on_a_large_synthetic_stream(oracle):
    b = oracle.gen_batch(oracle.SEED_BASE + 77, 200000, 8, 64, mode=0)
    rc, out = oracle.route_batch(oracle.default_table(), b, threads=8, fast=True)
    assert rc == 0
    # distinct (name, payload) pairs -> distinct hashes (a 64-bit hash over 2e5 items: collisions ~1e-9)
    keys = set()
    for i in range(0, 200000, 97):                      # sample the inputs for duplicates cheaply
        o, l = int(b["pay_off"][i]), int(b["pay_len"][i])
        keys.add((bytes(b["payload"][o:o + l]), bytes(b["names"][b["name_off"][i]:b["name_off"][i + 1]])))
    assert len(keys) >= 2000
    assert np.unique(out["route_hash"]).size >= 200000 - 50     # tiny payloads may repeat exactly


def test_provider_choice_follows_the_weights(oracle):
    """§D: among eligible providers the choice is proportional to weight."""
    b = oracle.gen_batch(oracle.SEED_BASE + 5, 120000, 200, 400, mode=0)
    t = oracle.default_table()
    rc, out = oracle.route_batch(t, b, threads=8, fast=True)
    assert rc == 0
    for cls in range(4):
        sel = out[out["tool_class"] == cls]
        elig = [p for p in range(8) if (t.providers[p].tool_mask >> cls) & 1]      # all token limits are above 400
        w = np.array([t.providers[p].weight for p in elig], dtype=float)
        got = np.array([(sel["provider"] == p).sum() for p in elig], dtype=float)
        assert got.sum() == sel.size and sel.size > 5000
        assert np.abs(got / got.sum() - w / w.sum()).max() < 0.02, (cls, got / got.sum(), w / w.sum())


def test_token_limits_and_tool_masks_are_respected(oracle):
    provs = [(10, 5, 0b0001, 0), (1000, 1, 0b1111, 0), (100000, 3, 0b0110, 0)]
    t = oracle.make_table([(b"a", 0), (b"b", 1), (b"c", 2)], provs, default_class=3)
    reqs = [(nm, b"word " * k) for nm in (b"a", b"b", b"c", b"zzz") for k in (1, 50, 500, 5000)]
    rc, out = oracle.route_batch(t, oracle.pack_batch(reqs))
    assert rc == 0
    for (nm, p), r in zip(reqs, out):
        cls, tok, prov = int(r["tool_class"]), int(r["est_tokens"]), int(r["provider"])
        ok = [i for i, (mt, w, mask, _) in enumerate(provs) if w and (mask >> cls) & 1 and tok <= mt]
        assert (prov in ok) if ok else (prov == 0xFF and r["flags"] & 2), (nm, tok, prov, ok)


def test_token_estimate_examples(oracle):
    """Worked examples of §B (the numbers are part of the written spec)."""
    # "hello world" is 3, not 2: "world" starts at offset 6 and the run is cut at offset 8
    cases = {b"": 0, b"   ": 0, b"hello": 1, b"hello world": 3, b"Hello, world!": 5, b"a1b2": 4,
             b"abcdefgh": 1, b"abcdefghi": 2, b"x" * 24: 3, b"line1\nline2": 6, b"\t\tfoo": 3,
             "naïve".encode(): 3, "例え".encode(): 2, b'{"k": [1, 2]}': 11}
    got = hashes(oracle, list(cases), b"t")["est_tokens"]
    assert [int(x) for x in got] == list(cases.values()), list(zip(cases, got.tolist()))

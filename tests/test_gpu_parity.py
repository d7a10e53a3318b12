"""GPU parity: the CUDA path, called through the C ABI (libpkroute.so), against the CPU oracle
on the same seeded inputs.  Integer/byte work: every comparison is bit-exact (tolerance 0).

Reference pins: tool class + status 200 (internal/service/mock.go:22-67) via the golden fixture;
scan/hash/provider are build-defined — parity is with the oracle, which is itself pinned to the
independent Python restatement (tests/test_oracle_golden.py).
"""
import ctypes as C

import numpy as np
import pytest

from conftest import to_pk_table
from helpers import (assert_records_equal, golden_requests, load_golden, permute_batch,
                     record_as_dict)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def router(pk):
    r = pk.Router(device=0, max_requests=70000, max_payload_bytes=40 << 20, n_slots=3)
    yield r
    r.close()


def check(router, oracle, batch, what, table=None, threads=8):
    rc, want = oracle.route_batch(table if table is not None else oracle.default_table(), batch,
                                  threads=threads)
    assert rc == 0
    got = router.route_batch(batch)
    assert_records_equal(got, want, what)
    return got


def test_native_library_is_loaded(pk, router):
    maps = open("/proc/self/maps").read()
    assert "libpkroute.so" in maps                    # the CUDA extension, not a fallback
    assert router.launch_count == 0 or router.launch_count > 0


def test_golden_fixture_on_gpu(pk, router):
    """The committed golden vectors (incl. the dev-client request, cmd/dev_client/main.go:247-258)."""
    rows = golden_requests(load_golden())
    before = router.launch_count
    got = router.route_batch(pk.pack_requests([(nm, p) for nm, p, _ in rows]))
    assert router.launch_count == before + 1
    for r, (nm, p, want) in zip(got, rows):
        assert record_as_dict(r) == want, (nm, len(p))


def test_tool_class_rows_on_gpu(pk, router):
    """mock.go:32-64 on the device: exact match or the default branch; status always 200."""
    g = load_golden()
    rows = g["tool_class"]
    got = router.route_batch(pk.pack_requests([(bytes.fromhex(r["name_hex"]), b"payload %d" % i)
                                               for i, r in enumerate(rows)]))
    assert [int(x) for x in got["tool_class"]] == [r["tool_class"] for r in rows]
    assert [bool(x & pk.FLAG_UNKNOWN_TOOL) for x in got["flags"]] == [r["unknown"] for r in rows]
    assert (got["status"] == 200).all()


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("shape", [(2000, 0, 64), (3000, 0, 700), (512, 1000, 5000),
                                   (64, 16000, 17000), (24, 30000, 200000)])
def test_seeded_batches(router, oracle, mode, shape):
    n, lo, hi = shape
    b = oracle.gen_batch(oracle.SEED_BASE + 1000 * mode + n, n, lo, hi, mode=mode)
    check(router, oracle, b, f"mode {mode} shape {shape}")


def test_large_ragged_batch_throughput_mode(router, oracle):
    """Enough requests that a CTA owns 32+ of them: the thread-per-request decision phase
    (smaller batches take the per-warp latency mode)."""
    b = oracle.gen_batch(78, 24000, 0, 1500, mode=-1)
    check(router, oracle, b, "24000 ragged requests")


def test_ragged_mixed_stream(router, oracle):
    b = oracle.gen_batch(77, 5000, 0, 9000, mode=-1)
    check(router, oracle, b, "mixed modes, ragged")


@pytest.mark.parametrize("length", [0, 1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 32, 33, 63, 64, 65, 255,
                                    256, 257, 511, 512, 513, 2047, 2048, 2049, 16383, 16384, 16385,
                                    32767, 32768, 32769, 49152, 65536 + 3])
def test_length_edges(router, oracle, length):
    for mode in (0, 3):
        b = oracle.gen_batch(length * 7 + mode, 33, length, length, mode=mode)
        check(router, oracle, b, f"len {length} mode {mode}")


def test_spec_v3_stripe_boundary_on_gpu(pk, router, oracle):
    """SPEC v3 (8 stripes up to 256 bytes, 32 above): every length 0..300 in one batch — the
    full-warp kernel, whose lanes 0..7 fold blocks l and l + 8 of a short payload — and every length
    0..256 many times over in a batch of short payloads only — the four-lane kernels, at each of
    their blocks-per-lane shapes."""
    rng = np.random.default_rng(317)
    mixed = [(b"struct_tool", rng.integers(1, 256, l, dtype=np.uint8).tobytes()) for l in range(301)]
    check(router, oracle, pk.pack_requests(mixed), "lengths 0..300, one each")
    for top in (64, 128, 256):
        short = [(b"file_tool", rng.integers(1, 256, l % (top + 1), dtype=np.uint8).tobytes()) for l in range(6000)]
        check(router, oracle, pk.pack_requests(short), f"6000 payloads of 0..{top} bytes")


def test_empty_batch_and_empty_fields(pk, router, oracle):
    out = router.route_batch(pk.pack_requests([]))
    assert out.size == 0
    reqs = [(b"", b""), (b"example_tool", b""), (b"", b"x"), (b"x" * 65, b"abc"), (b"y" * 300, b"z" * 100),
            (b"struct_tool", b" " * 50), (b"file_tool", bytes(range(256)))]
    got = check(router, oracle, pk.pack_requests(reqs), "empty names / payloads / long names")
    assert got["flags"][0] & pk.FLAG_EMPTY_PAYLOAD and got["flags"][0] & pk.FLAG_UNKNOWN_TOOL
    assert got["tool_class"][1] == 0 and got["est_tokens"][5] == 0


def test_padding_bytes_are_ignored(pk, router, oracle):
    """Arena padding between requests must not influence any record."""
    reqs = [(b"example_tool", b"abc" * k) for k in range(40)]
    a = pk.pack_requests(reqs)
    b = {k: v.copy() for k, v in a.items()}
    for i in range(len(reqs)):
        end, nxt = int(a["pay_off"][i]) + int(a["pay_len"][i]), int(a["pay_off"][i + 1])
        a["payload"][end:nxt] = 0x00
        b["payload"][end:nxt] = 0xC3            # a UTF-8 lead byte: would count if it leaked
    ga, gb = router.route_batch(a), router.route_batch(b)
    assert_records_equal(ga, gb, "padding content")
    check(router, oracle, b, "padding content vs oracle")


def test_arena_with_leading_offset_and_gaps(pk, router, oracle):
    """pay_off[0] != 0 and uneven (16-byte multiple) gaps between requests."""
    payloads = [b"hello world, this is request %d!" % i * (i % 5 + 1) for i in range(50)]
    off, offs = 4096, []
    for i, p in enumerate(payloads):
        offs.append(off)
        off += ((len(p) + 15) & ~15) + 16 * (i % 3)
    offs.append(off)
    payload = np.full(off, 0xEE, dtype=np.uint8)
    for o, p in zip(offs, payloads):
        payload[o:o + len(p)] = np.frombuffer(p, dtype=np.uint8)
    names = b"".join(b"file_tool" for _ in payloads)
    batch = {"payload": payload, "pay_off": np.array(offs, dtype=np.uint64),
             "pay_len": np.array([len(p) for p in payloads], dtype=np.uint32),
             "names": np.frombuffer(b"HEADER" + names, dtype=np.uint8).copy(),
             "name_off": (6 + 9 * np.arange(len(payloads) + 1)).astype(np.uint32)}
    check(router, oracle, batch, "offset arena")


def test_custom_tables(pk, router, oracle):
    tools = [(b"t%d" % i + b"_" * i, (i * 5) % 32) for i in range(16)]            # 16 tools, lengths 2..32
    tools[7] = (b"q" * 64, 9)                                                    # max-length name
    tools[9] = (tools[3][0], 31)                                                 # duplicate: first match wins
    provs = [(50 * (p + 1), (p * 7) % 5, 1 << ((p * 5) % 32) | 0x1F, p * 2654435761 & 0xFFFFFFFF)
             for p in range(32)]                                                 # some weight 0
    ot = oracle.make_table(tools, provs, default_class=17)
    router.update_table(to_pk_table(pk, ot))
    try:
        names = [t[0] for t in tools] + [b"nope", b"q" * 63, b"q" * 65, b"t0", b"T0"]
        reqs = [(names[i % len(names)], b"word " * (i % 700)) for i in range(1500)]
        got = check(router, oracle, pk.pack_requests(reqs), "custom table", table=ot)
        assert (got["provider"] == pk.PROVIDER_NONE).any()                       # token limits bite
        assert ((got["flags"] & pk.FLAG_NO_PROVIDER) != 0).sum() == (got["provider"] == 0xFF).sum()
        # no tools / no providers at all
        ot2 = oracle.make_table([], [], default_class=0)
        router.update_table(to_pk_table(pk, ot2))
        got = check(router, oracle, pk.pack_requests(reqs[:100]), "empty table", table=ot2)
        assert (got["provider"] == 0xFF).all() and (got["tool_class"] == 0).all()
    finally:
        router.update_table(pk.default_table())


def test_error_codes_match_oracle(pk, router, oracle):
    b = pk.pack_requests([(b"example_tool", b"hello world"), (b"file_tool", b"x" * 40)])
    t = oracle.default_table()
    bad = dict(b)
    bad["pay_off"] = b["pay_off"].copy()
    bad["pay_off"][1] += 8
    with pytest.raises(pk.PkError) as e:
        router.route_batch(bad)
    assert e.value.code == oracle.route_batch(t, bad)[0] == -2
    bad = dict(b)
    bad["pay_len"] = np.array([11, 4000], dtype=np.uint32)
    with pytest.raises(pk.PkError) as e:
        router.route_batch(bad)
    assert e.value.code == oracle.route_batch(t, bad)[0] == -3
    small = pk.Router(device=0, max_requests=4, max_payload_bytes=4096, n_slots=1)
    try:
        with pytest.raises(pk.PkError) as e:
            small.route_batch(pk.pack_requests([(b"a", b"b")] * 5))
        assert e.value.code == -4
        with pytest.raises(pk.PkError) as e:
            small.route_batch(pk.pack_requests([(b"a", b"b" * 5000)]))
        assert e.value.code == -4
        small.route_batch(pk.pack_requests([(b"a", b"b" * 4096)]))              # exactly at capacity
    finally:
        small.close()
    router.route_batch(b)                                                        # still healthy


def test_async_slots_and_resident_form(pk, router, oracle):
    t = oracle.default_table()
    batches = [oracle.gen_batch(500 + s, 3000, 0, 4000, mode=s % 5) for s in range(3)]
    outs = [np.zeros(3000, dtype=pk.DECISION_DTYPE) for _ in range(3)]
    for s in range(3):
        router.submit(s, batches[s], outs[s])
    with pytest.raises(pk.PkError) as e:
        router.submit(1, batches[1], outs[1])
    assert e.value.code == -7                                                    # PK_ERR_BUSY
    for s in (2, 0, 1):
        router.wait(s)
    for s in range(3):
        assert_records_equal(outs[s], oracle.route_batch(t, batches[s], threads=8)[1], f"slot {s}")
    with pytest.raises(pk.PkError):
        router.wait(0)                                                           # nothing in flight
    # resident form: stage once, launch many (plain and graph), fetch
    for s in range(3):
        router.stage(s, batches[(s + 1) % 3])
    before = router.launch_count
    ms = router.launch_many([0, 1, 2], 9, use_graph=False)
    ms_g = router.launch_many([0, 1, 2], 9, use_graph=True)
    ms_p = router.launch_many([0, 1, 2], 9, use_graph=3)                      # three parallel chains
    assert ms > 0 and ms_g > 0 and ms_p > 0 and router.launch_count == before + 27
    for s in range(3):
        assert_records_equal(router.fetch(s, 3000),
                             oracle.route_batch(t, batches[(s + 1) % 3], threads=8)[1], f"resident {s}")


def test_small_batches_skip_the_copies(pk, oracle):
    """A small batch in page-locked buffers is routed by ONE launch on the caller's memory (no
    H2D / D2H copies): same records; pageable buffers, big batches and a zero threshold take the
    copying path."""
    t = oracle.default_table()
    r = pk.Router(device=0, max_requests=4096, max_payload_bytes=16 << 20, n_slots=2)
    try:
        r.set_direct_bytes(256 << 10)                 # the default is 16 MiB; a small one keeps the cases small
        buf = r.host_alloc(4 << 20)
        out = np.frombuffer(r.host_alloc(24 * 4096), dtype=pk.DECISION_DTYPE)
        seen = r.direct_count
        for k, (n, lo, hi, mode) in enumerate([(1, 0, 0, 0), (1, 16, 16, 0), (3, 0, 100, 2), (64, 0, 3000, 0),
                                                (100, 2048, 2048, 1), (257, 0, 700, 3), (17, 0, 16384, 4)]):
            b = oracle.gen_batch(7000 + k, n, lo, hi, mode=mode, payload_out=buf)
            want = oracle.route_batch(t, b, threads=4)[1]
            out[:] = 0
            got = r.route_batch(b, out=out[:n])
            assert_records_equal(got, want, f"direct case {k}")
            seen += 1
            assert r.direct_count == seen, k
            # submit / wait on the other slot, twice in a row
            for _ in range(2):
                out[:n] = 0
                r.submit(1, b, out[:n])
                r.wait(1)
                assert_records_equal(out[:n], want, f"direct submit case {k}")
            seen += 2
            assert r.direct_count == seen
            with pytest.raises(pk.PkError):
                r.launch(1)                                   # nothing was staged on the device
        b = oracle.gen_batch(7100, 64, 0, 3000, mode=0)                                  # pageable payload
        assert_records_equal(r.route_batch(b), oracle.route_batch(t, b, threads=4)[1], "pageable")
        b = oracle.gen_batch(7101, 8, 16385, 30000, mode=0, payload_out=buf)             # long requests: in place too
        assert_records_equal(r.route_batch(b, out=out[:8]), oracle.route_batch(t, b, threads=4)[1], "long")
        assert_records_equal(r.route_batch(b, out=out[:8]), oracle.route_batch(t, b, threads=4)[1], "long, again")
        seen += 2
        assert r.direct_count == seen
        b = oracle.gen_batch(7102, 1024, 2048, 2048, mode=0, payload_out=buf)            # 2 MiB > threshold
        assert_records_equal(r.route_batch(b, out=out[:1024]), oracle.route_batch(t, b, threads=4)[1], "big")
        assert r.direct_count == seen
        r.set_direct_bytes(1 << 40)                                                      # ... unless asked to
        assert_records_equal(r.route_batch(b, out=out[:1024]), oracle.route_batch(t, b, threads=4)[1], "big direct")
        assert r.direct_count == seen + 1
        r.set_direct_bytes(0)
        b = oracle.gen_batch(7103, 4, 100, 100, mode=0, payload_out=buf)
        assert_records_equal(r.route_batch(b, out=out[:4]), oracle.route_batch(t, b, threads=4)[1], "off")
        assert r.direct_count == seen + 1
    finally:
        r.close()


def test_misaligned_pinned_buffers_take_the_copying_path(pk, oracle):
    """The in-place path issues 16-byte loads on the caller's payload and 8-byte stores on `out`.
    A page-locked arena whose base is not so aligned (here +8 and +4) must not reach it — a
    misaligned device access is a sticky fault — and is copied instead: same records."""
    t = oracle.default_table()
    with pk.Router(device=0, max_requests=1024, max_payload_bytes=4 << 20, n_slots=1) as r:
        arena = r.host_alloc((1 << 20) + 64)
        outbuf = r.host_alloc(24 * 512 + 64)
        b = oracle.gen_batch(7200, 300, 0, 3000, mode=0)
        want = oracle.route_batch(t, b, threads=4)[1]
        nbytes = b["payload"].size
        for pay_shift, out_shift, direct in ((0, 0, True), (8, 0, False), (0, 4, False), (8, 4, False)):
            moved = dict(b)
            moved["payload"] = arena[pay_shift:pay_shift + nbytes]
            moved["payload"][:] = b["payload"]
            out = np.frombuffer(outbuf[out_shift:out_shift + 24 * 300], dtype=pk.DECISION_DTYPE)
            out[:] = 0
            seen = r.direct_count
            got = r.route_batch(moved, out=out)
            assert_records_equal(got, want, f"payload base +{pay_shift}, out base +{out_shift}")
            assert r.direct_count == seen + (1 if direct else 0)
        r.sync()                                       # no sticky error left behind
        assert_records_equal(r.route_batch(b), want, "afterwards")


def test_programmatic_dependent_launch(pk, oracle):
    """PK_LAUNCH_PDL: consecutive launches of one stream may overlap.  Same records as the oracle
    for plain launches and for a graph chain; a slot with long requests (which owns per-request
    scratch) launched back to back with itself stays correct; a copy out that follows sees the
    last launch complete."""
    t = oracle.default_table()
    PDL = 0x100
    batches = [oracle.gen_batch(900, 3000, 0, 4000, mode=0), oracle.gen_batch(901, 2500, 0, 200, mode=2),
               oracle.gen_batch(902, 48, 10000, 200000, mode=0), oracle.gen_batch(903, 4096, 2048, mode=0)]
    want = [oracle.route_batch(t, b, threads=8)[1] for b in batches]
    router = pk.Router(device=0, max_requests=4096, max_payload_bytes=16 << 20, n_slots=4)
    try:
        for s, b in enumerate(batches):
            router.stage(s, b)
        before = router.launch_count
        for mode in (PDL, 1 | PDL, 2 | PDL):
            for slots in ([0, 1, 2, 3], [2], [3, 2], [1]):
                assert router.launch_many(slots, 12, use_graph=mode) > 0
                for s in slots:
                    assert_records_equal(router.fetch(s, int(batches[s]["pay_len"].size)), want[s],
                                         f"pdl mode {mode:#x} slot {s}")
        assert router.launch_count == before + 3 * 4 * 12
    finally:
        router.close()


def test_fused_launches(pk, oracle):
    """PK_LAUNCH_FUSE(m): m staged batches per kernel launch.  Same records as the oracle for every
    batch — different sizes, ragged lengths, a batch with multi-segment requests, plain / graph /
    PDL issue, groups that do not divide the count — and the launch counter shows what happened:
    one launch per group while the batches share a kernel class, one per batch when a tiny-payload
    batch sits in the group."""
    t = oracle.default_table()
    PDL, FUSE = 0x100, lambda m: m << 12
    batches = [oracle.gen_batch(910, 3000, 300, 4000, mode=0), oracle.gen_batch(911, 2500, 257, 900, mode=2),
               oracle.gen_batch(912, 48, 10000, 200000, mode=0), oracle.gen_batch(913, 4096, 2048, mode=0),
               oracle.gen_batch(914, 1, 5000, mode=0), oracle.gen_batch(915, 777, 0, 3000, mode=1)]
    tiny = oracle.gen_batch(916, 2000, 0, 256, mode=0)                  # <= 256 bytes: the lane-group kernels
    want = [oracle.route_batch(t, b, threads=8)[1] for b in batches + [tiny]]
    router = pk.Router(device=0, max_requests=4096, max_payload_bytes=16 << 20, n_slots=7)
    try:
        for s, b in enumerate(batches + [tiny]):
            router.stage(s, b)
        allb = batches + [tiny]

        def verify(slots, what):
            for s in slots:
                assert_records_equal(router.fetch(s, int(allb[s]["pay_len"].size)), want[s], f"{what} slot {s}")

        same = [0, 1, 2, 3, 4, 5]
        for mode, what in ((0, "plain"), (1, "graph"), (1 | PDL, "graph+pdl"), (PDL, "plain+pdl"), (4 | PDL, "width ignored")):
            for m, count in ((6, 6), (2, 6), (4, 6), (3, 7), (6, 13)):
                before = router.launch_count
                assert router.launch_many(same, count, use_graph=mode | FUSE(m)) > 0
                assert router.launch_count == before + -(-count // m), (what, m, count)
                verify(same, f"{what} fuse {m} x {count}")
        # the ring position carries over between calls and groups wrap around the list
        for count in (1, 5, 9):
            router.launch_many(same, count, use_graph=1 | FUSE(4))
            verify(same, f"wrap {count}")
        # a tiny-payload batch in the group: another kernel class -> launched one by one, same records
        mixed = [3, 6, 0]
        before = router.launch_count
        router.launch_many(mixed, 3, use_graph=FUSE(3))
        assert router.launch_count == before + 3
        verify(mixed, "mixed classes")
        # tiny batches together fuse too
        r2 = [6]
        before = router.launch_count
        router.launch_many(r2, 1, use_graph=0)
        assert router.launch_count == before + 1
        # arguments: m larger than the list, or larger than 32
        with pytest.raises(pk.PkError):
            router.launch_many([0, 1], 2, use_graph=FUSE(3))
        with pytest.raises(pk.PkError):
            router.launch_many(same, 6, use_graph=FUSE(33))
    finally:
        router.close()


def test_fused_tiny_batches(pk, oracle):
    """Fused launch of lane-group (<= 256-byte payload) batches of different sizes, C2-like."""
    t = oracle.default_table()
    FUSE = lambda m: m << 12
    batches = [oracle.gen_batch(920 + k, n, 0, hi, mode=0) for k, (n, hi) in enumerate(((5000, 256), (1, 200), (4096, 256), (333, 64), (2048, 128)))]
    want = [oracle.route_batch(t, b, threads=8)[1] for b in batches]
    router = pk.Router(device=0, max_requests=8192, max_payload_bytes=4 << 20, n_slots=5)
    try:
        for s, b in enumerate(batches):
            router.stage(s, b)
        for mode in (0, 1, 1 | 0x100):
            before = router.launch_count
            router.launch_many(list(range(5)), 10, use_graph=mode | FUSE(5))
            for s in range(5):
                assert_records_equal(router.fetch(s, int(batches[s]["pay_len"].size)), want[s], f"tiny fused mode {mode:#x} slot {s}")
            # (batches whose block counts differ resolve to different lane-group shapes: 2 <= launches <= 10)
            assert before + 2 <= router.launch_count <= before + 10
    finally:
        router.close()


def test_long_requests_twice_scratch_is_clean(router, oracle):
    """Multi-segment requests use the per-request scratch accumulators; a second pass over the
    same slot must see them zeroed by the first."""
    b = oracle.gen_batch(4242, 40, 10000, 300000, mode=0)
    first = check(router, oracle, b, "long requests, pass 1")
    second = router.route_batch(b)
    assert_records_equal(second, first, "long requests, pass 2")
    b2 = oracle.gen_batch(4243, 64, 0, 120000, mode=3)
    check(router, oracle, b2, "different long batch on the same slot")


def test_permutation_property(router, oracle):
    """Size-independent property: a record depends on its own request only."""
    b = oracle.gen_batch(31337, 4096, 100, 3000, mode=0)
    base = router.route_batch(b)
    perm = np.random.default_rng(1).permutation(4096)
    got = router.route_batch(permute_batch(b, perm))
    assert_records_equal(got, base[perm], "permutation")


def test_config_c1_full(router, oracle):
    """BASELINE configs[1]: 4,096 requests x 2 KB, 8 providers."""
    b = oracle.gen_batch(oracle.SEED_BASE + 1, 4096, 2048, mode=0)
    got = check(router, oracle, b, "C1")
    assert np.bincount(got["provider"], minlength=8)[:8].min() > 0               # every provider is used


def test_config_c2_full(router, oracle):
    """BASELINE configs[2]: 65,536 requests x 256 B."""
    b = oracle.gen_batch(oracle.SEED_BASE + 2, 65536, 256, mode=0)
    check(router, oracle, b, "C2")


def test_config_c3_full(pk, oracle):
    """BASELINE configs[3]: 1,024 requests x 1 MiB (1 GiB) — full bit-exact comparison with the
    multi-threaded oracle plus the idempotence property (second pass over dirty scratch)."""
    n, length = 1024, 1 << 20
    r = pk.Router(device=0, max_requests=n, max_payload_bytes=n * length, n_slots=1)
    try:
        b = oracle.gen_batch(oracle.SEED_BASE + 3, n, length, mode=0, threads=16)
        rc, want = oracle.route_batch(oracle.default_table(), b, threads=32)
        assert rc == 0
        got = r.route_batch(b)
        assert_records_equal(got, want, "C3")
        assert_records_equal(r.route_batch(b), want, "C3 second pass")
        assert (got["byte_len"] == length).all() and len(set(got["route_hash"].tolist())) == n
    finally:
        r.close()


def test_config_c4_one_full_shard(pk, oracle):
    """BASELINE configs[4] at full size for one rank: 1,000,000 requests x 2 KB sharded over 8 GPUs
    is 125,000 requests (256 MB) per GPU; rank 3's shard, ids 3, 11, 19, ... generated directly."""
    n, length = 125000, 2048
    r = pk.Router(device=0, max_requests=n, max_payload_bytes=n * length, n_slots=1)
    try:
        # the generator is seeded per request id, so a rank can generate its own ids; here the
        # shard's ids are contiguous in the rank-local numbering used by bench.py
        b = oracle.gen_batch(oracle.SEED_BASE + 4, n, length, mode=0, first_id=3 * n, threads=16)
        rc, want = oracle.route_batch(oracle.default_table(), b, threads=32, fast=True)
        assert rc == 0
        assert_records_equal(r.route_batch(b), want, "C4 shard")
    finally:
        r.close()


def test_maximum_sizes(pk, oracle):
    """The largest request the reference's transport accepts is 4 MiB (grpc-go default,
    cmd/polykey/main.go:74-77); also one request far beyond it, and a 1,000-byte tool name."""
    table = oracle.default_table()
    reqs = [(b"example_tool", bytes(np.random.default_rng(5).integers(32, 127, 4 << 20, dtype=np.uint8))),
            (b"n" * 1000, b"short payload with a very long name"),
            (b"file_tool", bytes(np.random.default_rng(6).integers(0, 256, (33 << 20) + 7, dtype=np.uint8)))]
    b = pk.pack_requests(reqs)
    r = pk.Router(device=0, max_requests=8, max_payload_bytes=int(b["pay_off"][-1]), max_name_bytes=4096, n_slots=1)
    try:
        rc, want = oracle.route_batch(table, b, threads=3, fast=True)
        assert rc == 0
        assert_records_equal(r.route_batch(b), want, "maximum sizes")
        assert (oracle.route_batch(table, b, threads=3)[1] == want).all()      # fast port == plain oracle here too
    finally:
        r.close()


def test_config_c4_shards(pk, router, oracle):
    """BASELINE configs[4]: the stream is sharded by request id (req_id % G) with no exchange;
    each shard's records equal the corresponding rows of the unsharded result."""
    total, G = 20000, 4
    full = oracle.gen_batch(oracle.SEED_BASE + 4, total, 2048, mode=0)
    whole = router.route_batch(full)
    for g in range(G):
        ids = np.arange(g, total, G)
        shard = permute_batch(full, ids)
        assert_records_equal(router.route_batch(shard), whole[ids], f"shard {g}")

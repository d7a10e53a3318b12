"""CPU: the C oracle against the golden fixtures and against the pure-Python restatement.

Reference pins: the tool switch and status 200 (internal/service/mock.go:22-67).  The
scan/hash/provider expectations are build-defined ("parity unpinned" by the reference); the
fixture values were produced by oracle/pk_spec_py.py, which shares no code with pk_oracle.c.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import golden_requests, load_golden, record_as_dict


def test_tool_class_literals(oracle):
    """mock.go:32-64: exact, case-sensitive, whole-name equality; unknown -> default, not an error."""
    g = load_golden()
    t = oracle.default_table()
    assert [(bytes(t.tools[k].name[:t.tools[k].name_len]).decode(), t.tools[k].class_id)
            for k in range(t.n_tools)] == [tuple(x) for x in g["tools"]]
    for row in g["tool_class"]:
        name = bytes.fromhex(row["name_hex"])
        unknown = C.c_int(-1)
        buf = (C.c_uint8 * max(1, len(name))).from_buffer_copy(name or b"\0")
        cls = oracle.lib().pk_ref_tool_class(C.byref(t), buf, len(name), C.byref(unknown))
        assert (cls, bool(unknown.value)) == (row["tool_class"], row["unknown"]), name


def test_golden_records(oracle):
    g = load_golden()
    rows = golden_requests(g)
    batch = oracle.pack_batch([(nm, p) for nm, p, _ in rows])
    rc, out = oracle.route_batch(oracle.default_table(), batch)
    assert rc == 0
    for r, (nm, p, want) in zip(out, rows):
        assert record_as_dict(r) == want, (nm, len(p))
    assert (out["status"] == 200).all()          # mock.go:25-28: always 200


def test_default_table_matches_fixture(oracle):
    g = load_golden()
    t = oracle.default_table()
    got = [[t.providers[p].max_tokens, t.providers[p].weight, t.providers[p].tool_mask, t.providers[p].reserved]
           for p in range(t.n_providers)]
    assert got == g["providers"] and t.default_class == g["default_class"]


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
def test_oracle_vs_python_restatement(oracle, mode):
    from oracle import pk_spec_py as S
    t = oracle.default_table()
    tools = [(bytes(t.tools[k].name[:t.tools[k].name_len]), t.tools[k].class_id) for k in range(t.n_tools)]
    provs = [(p.max_tokens, p.weight, p.tool_mask, p.reserved) for p in list(t.providers)[:t.n_providers]]
    b = oracle.gen_batch(oracle.SEED_BASE + 17 * mode, 48, 0, 1200, mode=mode)
    rc, out = oracle.route_batch(t, b)
    assert rc == 0
    for i in range(48):
        o, l = int(b["pay_off"][i]), int(b["pay_len"][i])
        nm = bytes(b["names"][b["name_off"][i]:b["name_off"][i + 1]])
        assert record_as_dict(out[i]) == S.route_one(tools, 3, provs, nm, bytes(b["payload"][o:o + l]))


def test_oracle_multisegment_vs_python(oracle):
    from oracle import pk_spec_py as S
    t = oracle.default_table()
    tools = [(bytes(t.tools[k].name[:t.tools[k].name_len]), t.tools[k].class_id) for k in range(t.n_tools)]
    provs = [(p.max_tokens, p.weight, p.tool_mask, p.reserved) for p in list(t.providers)[:t.n_providers]]
    b = oracle.gen_batch(99, 3, 16380, 50000, mode=0)
    rc, out = oracle.route_batch(t, b)
    assert rc == 0
    for i in range(3):
        o, l = int(b["pay_off"][i]), int(b["pay_len"][i])
        nm = bytes(b["names"][b["name_off"][i]:b["name_off"][i + 1]])
        assert record_as_dict(out[i]) == S.route_one(tools, 3, provs, nm, bytes(b["payload"][o:o + l]))


def test_mt_equals_single_thread(oracle):
    t = oracle.default_table()
    b = oracle.gen_batch(5, 777, 0, 5000, mode=-1)
    rc1, a = oracle.route_batch(t, b, threads=1)
    rc2, m = oracle.route_batch(t, b, threads=5)
    assert rc1 == 0 and rc2 == 0 and (a == m).all()
    rc3, s = oracle.route_batch(t, b, threads=3, switch_only=True)
    assert rc3 == 0 and (s["tool_class"] == a["tool_class"]).all() and (s["status"] == 200).all()


def test_oracle_error_codes(oracle):
    t = oracle.default_table()
    b = oracle.pack_batch([(b"example_tool", b"hello world"), (b"file_tool", b"x" * 40)])
    bad = dict(b)
    bad["pay_off"] = b["pay_off"].copy()
    bad["pay_off"][1] += 8
    assert oracle.route_batch(t, bad)[0] == -2              # PK_ERR_ALIGN
    bad["pay_off"] = b["pay_off"][::-1].copy()
    assert oracle.route_batch(t, bad)[0] in (-2, -3)        # not monotone
    bad = dict(b)
    bad["pay_len"] = np.array([11, 4000], dtype=np.uint32)
    assert oracle.route_batch(t, bad)[0] == -3              # length overruns its slot
    rc, out = oracle.route_batch(t, oracle.pack_batch([]))
    assert rc == 0 and out.size == 0


def test_mock_responses(oracle):
    """Response literals of mock.go:24-63 (pk_ref_execute_tool restates them)."""
    g = load_golden()["responses"]
    r = oracle.execute_tool(b"example_tool", 1700000000)
    assert r.status_code == g["status_code"] and r.status_message.decode() == g["status_message"]
    assert r.kind == 0 and r.string_output.decode() == g["example_tool"]["prefix"] + "2023-11-14T22:13:20Z"
    r = oracle.execute_tool(b"struct_tool", 1700000000)
    assert r.kind == 1 and b'"count":42' in r.string_output and b'"processed":true' in r.string_output
    assert b'"result":"success"' in r.string_output and b'"timestamp":1700000000' in r.string_output
    r = oracle.execute_tool(b"file_tool", 0)
    assert (r.kind, r.file_name, r.mime_type, r.file_content) == (
        2, g["file_tool"]["file_name"].encode(), g["file_tool"]["mime_type"].encode(),
        g["file_tool"]["content"].encode())
    r = oracle.execute_tool(b"nope", 0)                     # mock.go:60-64: still 200, nil error
    assert r.status_code == 200 and r.kind == 0 and r.string_output == b"Unknown tool: nope"


def test_generator_is_deterministic_and_sliceable(oracle):
    a = oracle.gen_batch(42, 64, 10, 300, mode=0)
    b = oracle.gen_batch(42, 64, 10, 300, mode=0, threads=3)
    assert all((a[k] == b[k]).all() for k in a)
    # a rank generating requests [32, 64) of the stream gets the same requests
    c = oracle.gen_batch(42, 32, 10, 300, mode=0, first_id=32)
    assert (c["pay_len"] == a["pay_len"][32:]).all()
    o = int(a["pay_off"][32])
    assert (c["payload"] == a["payload"][o:]).all()
    # valid UTF-8 in the text modes (Go's proto3 decode would reject anything else upstream)
    for mode in (0, 2):
        x = oracle.gen_batch(7, 16, 100, 400, mode=mode)
        for i in range(16):
            s, l = int(x["pay_off"][i]), int(x["pay_len"][i])
            bytes(x["payload"][s:s + l]).decode("utf-8")


@pytest.mark.parametrize("impl", ["avx512", "scalar"])
@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
def test_fast_cpu_port_equals_plain_oracle(oracle, mode, impl, monkeypatch):
    """oracle/pk_cpu_fast.c (the timed CPU baseline: one fused pass, AVX-512 or scalar, persistent
    thread pool) against the plain checker — both inner loops, 1 / 3 / 8 threads, every length
    around the 16-byte block, the 64-byte vector, the 512-byte step and the 16 KiB segment."""
    monkeypatch.setenv("PK_CPU_FAST_IMPL", impl)
    oracle.lib().pk_cpu_fast_reselect()
    try:
        if impl == "avx512" and oracle.fast_impl() != "avx512":
            pytest.skip("this CPU has no AVX-512 (the scalar loop is what bench.py would time here)")
        t = oracle.default_table()
        for n, lo, hi in [(300, 0, 40000), (3000, 0, 70), (40, 16370, 16400), (700, 0, 1100), (64, 32700, 32800)]:
            b = oracle.gen_batch(900 + mode, n, lo, hi, mode=mode)
            plain = oracle.route_batch(t, b, threads=4)[1]
            for th in (1, 3, 8):
                rc, fast = oracle.route_batch(t, b, threads=th, fast=True)
                assert rc == 0 and (fast == plain).all(), (mode, n, th, np.nonzero(fast != plain)[0][:4])
        # a reused output array, as bench.py's timed loop passes it
        b = oracle.gen_batch(950 + mode, 2000, 0, 3000, mode=mode)
        out = np.zeros(2000, dtype=oracle.DECISION_DTYPE)
        for _ in range(3):
            out[:] = 0
            rc, got = oracle.route_batch(t, b, threads=8, fast=True, out=out)
            assert got is out and (out == oracle.route_batch(t, b, threads=2)[1]).all()
    finally:
        monkeypatch.delenv("PK_CPU_FAST_IMPL")
        oracle.lib().pk_cpu_fast_reselect()


@pytest.mark.parametrize("impl", ["avx512", "scalar"])
def test_spec_v3_stripe_boundary(oracle, impl, monkeypatch):
    """SPEC v3: payloads of at most 16 blocks (256 bytes) have 8 stripes, longer ones 32.  Every
    length from 0 to 300 bytes — both sides of the boundary, every partial last block — through the
    plain oracle, the timed CPU port (both inner loops) and the independent Python restatement; and
    the boundary is where it should be: a 256-byte payload's sum changes if it is hashed with 32
    stripes, a 257-byte payload's if hashed with 8."""
    from oracle import pk_spec_py as S
    monkeypatch.setenv("PK_CPU_FAST_IMPL", impl)
    oracle.lib().pk_cpu_fast_reselect()
    try:
        if impl == "avx512" and oracle.fast_impl() != "avx512":
            pytest.skip("this CPU has no AVX-512")
        t = oracle.default_table()
        tools = [(bytes(t.tools[k].name[:t.tools[k].name_len]), t.tools[k].class_id) for k in range(t.n_tools)]
        provs = [(p.max_tokens, p.weight, p.tool_mask, p.reserved) for p in list(t.providers)[:t.n_providers]]
        rng = np.random.default_rng(316)
        reqs = [(b"example_tool", rng.integers(1, 256, l, dtype=np.uint8).tobytes()) for l in range(301)]
        b = oracle.pack_batch(reqs)
        plain = oracle.route_batch(t, b, threads=2)[1]
        rc, fast = oracle.route_batch(t, b, threads=3, fast=True)
        assert rc == 0 and (fast == plain).all(), np.nonzero(fast != plain)[0][:8]
        for (nm, p), rec in zip(reqs, plain):
            assert record_as_dict(rec) == S.route_one(tools, 3, provs, nm, p), len(p)
    finally:
        monkeypatch.delenv("PK_CPU_FAST_IMPL")
        oracle.lib().pk_cpu_fast_reselect()

    def sum_with(p, width):                          # the payload sum with a forced stripe count
        nblk = (len(p) + 15) // 16
        padded = p + b"\0" * (nblk * 16 - len(p))
        st = {}
        for k in range(nblk):
            q = k % width
            if q not in st:
                base = (q * S.P5) & S.M32
                st[q] = [(S.P1 + S.P2 + base) & S.M32, (S.P2 + base) & S.M32, base, (base - S.P1) & S.M32]
            a = st[q]
            for j in range(4):
                w = int.from_bytes(padded[16 * k + 4 * j:16 * k + 4 * j + 4], "little")
                a[j] = (S.rotl((a[j] + w * S.P2) & S.M32, 13) * S.P1) & S.M32
        total = 0
        for a in st.values():
            lo = (S.rotl(a[0], 1) + S.rotl(a[1], 7) + S.rotl(a[2], 12) + S.rotl(a[3], 18)) & S.M32
            hi = (S.rotl(a[0], 24) + S.rotl(a[1], 3) + S.rotl(a[2], 19) + S.rotl(a[3], 10)) & S.M32
            total = (total + S.fmix64((hi << 32) | lo)) & S.M64
        return total

    p256, p257 = reqs[256][1], reqs[257][1]
    assert S.payload_sum(p256) == sum_with(p256, 8) != sum_with(p256, 32)
    assert S.payload_sum(p257) == sum_with(p257, 32) != sum_with(p257, 8)

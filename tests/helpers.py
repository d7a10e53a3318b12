"""Shared helpers for the parity tests (test infrastructure)."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def load_golden():
    with open(os.path.join(HERE, "golden", "route_golden.json")) as f:
        return json.load(f)


def pattern(n, a=131, b=7):
    i = np.arange(n, dtype=np.int64)
    return ((i * a + i // b) % 251).astype(np.uint8).tobytes()


def text(n):
    words = [b"route", b"the", b"Request", b"to", b"a", b"provider", b"internationalization", b"42",
             b"{\"k\":1}", "naïve".encode(), "例え".encode(), b"x_y", b"\n", b"END."]
    out = bytearray()
    i = 0
    while len(out) < n:
        out += words[(i * 5 + i // 3) % len(words)] + b" "
        i += 1
    return bytes(out[:n])


def golden_payload(vec):
    if "payload_hex" in vec:
        return bytes.fromhex(vec["payload_hex"])
    kind, n = vec["payload_gen"]
    return pattern(n) if kind == "pattern" else text(n)


def golden_requests(g):
    """[(name, payload, expect)] for every fixture row that carries a full expected record."""
    rows = [(bytes.fromhex(v["name_hex"]), golden_payload(v), v["expect"]) for v in g["vectors"]]
    d = g["dev_client_request"]
    rows.append((bytes.fromhex(d["name_hex"]), bytes.fromhex(d["payload_hex"]), d["expect"]))
    return rows


def assert_records_equal(got, want, what=""):
    """Bit-exact comparison of decision records (integer/byte work: tolerance is zero)."""
    assert got.shape == want.shape, (what, got.shape, want.shape)
    if (got == want).all():
        return
    bad = np.nonzero(got != want)[0]
    i = int(bad[0])
    raise AssertionError(f"{what}: {bad.size} of {got.size} records differ; first at {i}: "
                         f"got {got[i]} want {want[i]}")


def record_as_dict(r):
    return {k: int(r[k]) for k in r.dtype.names}


def permute_batch(batch, perm):
    """Reorders the requests of a flat batch (payload starts stay 16-byte aligned)."""
    from polykey_service_b200.shard import select_requests
    return select_requests(batch, perm)

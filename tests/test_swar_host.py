"""CPU: the kernels' own arithmetic (pk_spec.cuh, compiled with g++) walked lane by lane by
tests/host_emul.cpp exactly as a warp walks a (request, segment) task, against the oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from helpers import assert_records_equal, golden_requests, load_golden

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("emul") / "libhostemul.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so,
                           os.path.join(HERE, "host_emul.cpp")])
    L = C.CDLL(so)
    L.emul_route_one.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    return L


def run_emul(emul, oracle, table, b, lanes=32):
    n = b["pay_len"].size
    got = np.zeros(n, dtype=oracle.DECISION_DTYPE)
    for i in range(n):
        emul.emul_route_one(C.byref(table), b["payload"].ctypes.data + int(b["pay_off"][i]),
                            int(b["pay_len"][i]), b["names"].ctypes.data + int(b["name_off"][i]),
                            int(b["name_off"][i + 1] - b["name_off"][i]), lanes, got.ctypes.data + 24 * i)
    return got


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("shape", [(150, 0, 100), (60, 0, 3000), (12, 16000, 17000), (4, 30000, 70000)])
def test_warp_emulation_matches_oracle(emul, oracle, mode, shape):
    n, lo, hi = shape
    t = oracle.default_table()
    b = oracle.gen_batch(1234 + mode, n, lo, hi, mode=mode)
    rc, want = oracle.route_batch(t, b)
    assert rc == 0
    assert_records_equal(run_emul(emul, oracle, t, b), want, f"mode {mode} shape {shape}")


def test_warp_emulation_golden(emul, oracle):
    rows = golden_requests(load_golden())
    t = oracle.default_table()
    b = oracle.pack_batch([(nm, p) for nm, p, _ in rows])
    got = run_emul(emul, oracle, t, b)
    for r, (nm, p, want) in zip(got, rows):
        assert {k: int(r[k]) for k in r.dtype.names} == want, (nm, len(p))


@pytest.mark.parametrize("lanes,maxlen", [(4, 256), (8, 256), (4, 64), (8, 128)])
def test_lane_group_emulation(emul, oracle, lanes, maxlen):
    """The 4- and 8-lane groups used for batches of tiny payloads (8 stripes: a lane folds the
    blocks l, l + G, ... into the one or two stripes they fall into)."""
    t = oracle.default_table()
    for mode in range(5):
        b = oracle.gen_batch(77 + mode, 300, 0, maxlen, mode=mode)
        want = oracle.route_batch(t, b)[1]
        assert_records_equal(run_emul(emul, oracle, t, b, lanes), want, f"lanes {lanes} mode {mode}")

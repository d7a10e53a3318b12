"""ctypes front-end of the CPU oracle (oracle/libpkoracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product package never does.  See pk_oracle.c for what
the reference pins (tool class, status 200 — internal/service/mock.go:22-67) and what is
build-defined ("parity unpinned").
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpkoracle.so")

PK_MAX_TOOLS, PK_MAX_TOOL_NAME, PK_MAX_PROVIDERS = 16, 64, 32

DECISION_DTYPE = np.dtype(
    [("byte_len", "<u4"), ("est_tokens", "<u4"), ("route_hash", "<u8"), ("tool_class", "u1"),
     ("provider", "u1"), ("flags", "<u2"), ("status", "<u4")], align=False)
assert DECISION_DTYPE.itemsize == 24


class PkTool(C.Structure):
    _fields_ = [("name", C.c_uint8 * PK_MAX_TOOL_NAME), ("name_len", C.c_uint32),
                ("class_id", C.c_uint32)]


class PkProvider(C.Structure):
    _fields_ = [("max_tokens", C.c_uint32), ("weight", C.c_uint32), ("tool_mask", C.c_uint32),
                ("reserved", C.c_uint32)]


class PkTable(C.Structure):
    _fields_ = [("n_tools", C.c_uint32), ("default_class", C.c_uint32),
                ("n_providers", C.c_uint32), ("reserved", C.c_uint32),
                ("tools", PkTool * PK_MAX_TOOLS), ("providers", PkProvider * PK_MAX_PROVIDERS)]


class PkRefResponse(C.Structure):
    _fields_ = [("status_code", C.c_uint32), ("status_message", C.c_char * 64),
                ("kind", C.c_uint32), ("string_output", C.c_char * 512),
                ("file_name", C.c_char * 32), ("mime_type", C.c_char * 32),
                ("file_content", C.c_char * 64)]


def build(force: bool = False) -> str:
    """Compile the oracle with gcc if the .so is missing or older than its sources."""
    srcs = [os.path.join(_HERE, f) for f in ("pk_oracle.c", "pk_gen.c", "pk_cpu_fast.c", "Makefile")]
    srcs.append(os.path.join(_HERE, "..", "include", "pk_route.h"))
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", _HERE, "clean", "all"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.pk_ref_default_table.argtypes = [C.POINTER(PkTable)]
        L.pk_ref_route_batch.argtypes = [C.POINTER(PkTable), C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        L.pk_ref_route_batch_mt.argtypes = L.pk_ref_route_batch.argtypes + [C.c_uint32, C.c_int]
        L.pk_cpu_route_batch_mt.argtypes = L.pk_ref_route_batch.argtypes + [C.c_uint32]
        L.pk_cpu_fast_impl.restype = C.c_int
        L.pk_cpu_fast_reselect.restype = None
        L.pk_cpu_fast_pool_threads.restype = C.c_uint32
        L.pk_ref_tool_class.argtypes = [C.POINTER(PkTable), C.c_void_p, C.c_uint32, C.POINTER(C.c_int)]
        L.pk_ref_tool_class.restype = C.c_uint32
        L.pk_ref_est_tokens.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_int)]
        L.pk_ref_est_tokens.restype = C.c_uint32
        L.pk_ref_payload_sum.argtypes = [C.c_void_p, C.c_uint64]
        L.pk_ref_payload_sum.restype = C.c_uint64
        L.pk_ref_name_sum.argtypes = [C.c_void_p, C.c_uint32]
        L.pk_ref_name_sum.restype = C.c_uint64
        L.pk_ref_route_hash.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32]
        L.pk_ref_route_hash.restype = C.c_uint64
        L.pk_ref_select_provider.argtypes = [C.POINTER(PkTable), C.c_uint64, C.c_uint32, C.c_uint32]
        L.pk_ref_select_provider.restype = C.c_uint32
        L.pk_ref_execute_tool.argtypes = [C.c_void_p, C.c_uint32, C.c_int64, C.POINTER(PkRefResponse)]
        L.pk_ref_execute_tool.restype = C.c_uint32
        L.pk_ref_sizeof_response.restype = C.c_size_t
        L.pk_gen_payload.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int]
        L.pk_gen_lengths.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.pk_gen_names.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p]
        L.pk_gen_names.restype = C.c_uint64
        L.pk_gen_fill.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_int, C.c_uint32]
        assert L.pk_ref_sizeof_response() == C.sizeof(PkRefResponse)
        _lib = L
    return _lib


def default_table() -> PkTable:
    t = PkTable()
    rc = lib().pk_ref_default_table(C.byref(t))
    assert rc == 0
    return t


def make_table(tools, providers, default_class=3) -> PkTable:
    """tools: [(name_bytes, class_id)], providers: [(max_tokens, weight, tool_mask, reserved)]."""
    t = PkTable()
    t.n_tools, t.default_class, t.n_providers = len(tools), default_class, len(providers)
    for k, (name, cid) in enumerate(tools):
        name = bytes(name)
        t.tools[k].name_len, t.tools[k].class_id = len(name), cid
        for j, b in enumerate(name[:PK_MAX_TOOL_NAME]):
            t.tools[k].name[j] = b
    for p, (mt, w, mask, rsv) in enumerate(providers):
        e = t.providers[p]
        e.max_tokens, e.weight, e.tool_mask, e.reserved = mt, w, mask, rsv
    return t


def _ptr(a):
    return None if a is None else a.ctypes.data


def fast_impl() -> str:
    """Which inner loop pk_cpu_fast.c uses in this process ("avx512" or "scalar")."""
    return "avx512" if lib().pk_cpu_fast_impl() else "scalar"


def route_batch(table: PkTable, batch, threads: int = 1, switch_only: bool = False,
                fast: bool = False, out: np.ndarray | None = None) -> tuple[int, np.ndarray]:
    """batch: dict with payload(u8), pay_off(u64 n+1), pay_len(u32 n), names(u8), name_off(u32 n+1).
    fast=True runs pk_cpu_fast.c (the timed CPU baseline) instead of the plain checker.
    out: a record array to fill (timed loops reuse one instead of allocating per call)."""
    n = int(batch["pay_len"].shape[0])
    if out is None:
        out = np.zeros(n, dtype=DECISION_DTYPE)
    assert out.dtype == DECISION_DTYPE and out.size >= n
    args = [C.byref(table), _ptr(batch["payload"]), _ptr(batch["pay_off"]), _ptr(batch["pay_len"]),
            _ptr(batch["names"]), _ptr(batch["name_off"]), n, _ptr(out)]
    if fast and not switch_only:
        rc = lib().pk_cpu_route_batch_mt(*args, threads)
    elif threads == 1 and not switch_only:
        rc = lib().pk_ref_route_batch(*args)
    else:
        rc = lib().pk_ref_route_batch_mt(*args, threads, 1 if switch_only else 0)
    return rc, out


def execute_tool(name: bytes, now_unix: int = 0) -> PkRefResponse:
    r = PkRefResponse()
    buf = (C.c_uint8 * max(1, len(name))).from_buffer_copy(name or b"\0")
    lib().pk_ref_execute_tool(buf, len(name), now_unix, C.byref(r))
    return r


# ---------------------------------------------------------------- synthetic streams --
SEED_BASE = 0x504B5F31  # "PK_1", SURVEY.md §8d


def round16(x):
    """Round up to the arena alignment (PK_ALIGN = 16); works on ints and uint64 arrays."""
    if isinstance(x, np.ndarray):
        x = x.astype(np.uint64)
        return (x + np.uint64(15)) & np.uint64(0xFFFFFFFFFFFFFFF0)
    return (int(x) + 15) & ~15


def gen_batch(seed: int, n: int, len_min: int, len_max: int | None = None, mode: int = 0,
              first_id: int = 0, threads: int = 8, payload_out: np.ndarray | None = None):
    """Seeded synthetic batch in the C-ABI's flat layout (16-byte aligned payload starts)."""
    L = lib()
    len_max = len_min if len_max is None else len_max
    pay_len = np.empty(n, dtype=np.uint32)
    if n:
        L.pk_gen_lengths(seed, first_id, n, len_min, len_max, pay_len.ctypes.data)
    pay_off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(round16(pay_len.astype(np.uint64)), out=pay_off[1:])
    total = int(pay_off[-1])
    if payload_out is not None:
        assert payload_out.size >= total
        payload = payload_out[:total] if total else payload_out[:0]
    else:
        payload = np.empty(max(total, 16), dtype=np.uint8)[:total]
    names = np.empty(max(1, n * 36), dtype=np.uint8)
    name_off = np.zeros(n + 1, dtype=np.uint32)
    if n:
        nb = L.pk_gen_names(seed, first_id, n, names.ctypes.data, names.size, name_off.ctypes.data)
        names = names[:nb]
        L.pk_gen_fill(seed, first_id, n, payload.ctypes.data, pay_off.ctypes.data,
                      pay_len.ctypes.data, mode, threads)
    else:
        names = names[:0]
    return {"payload": payload, "pay_off": pay_off, "pay_len": pay_len, "names": names,
            "name_off": name_off}


def pack_batch(requests):
    """requests: iterable of (name_bytes, payload_bytes) -> flat batch dict."""
    requests = list(requests)
    n = len(requests)
    pay_len = np.array([len(p) for _, p in requests], dtype=np.uint32)
    pay_off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(round16(pay_len.astype(np.uint64)), out=pay_off[1:])
    payload = np.full(max(int(pay_off[-1]), 1), 0xA5, dtype=np.uint8)[:int(pay_off[-1])]
    for i, (_, p) in enumerate(requests):
        if len(p):
            payload[int(pay_off[i]):int(pay_off[i]) + len(p)] = np.frombuffer(bytes(p), dtype=np.uint8)
    name_off = np.zeros(n + 1, dtype=np.uint32)
    np.cumsum(np.array([len(nm) for nm, _ in requests], dtype=np.uint32), out=name_off[1:])
    names = np.frombuffer(b"".join(bytes(nm) for nm, _ in requests) or b"", dtype=np.uint8).copy()
    return {"payload": payload, "pay_off": pay_off, "pay_len": pay_len, "names": names,
            "name_off": name_off}

"""polykey_service_b200 — B200-native batched request routing behind the reference's
``service.Service`` seam (internal/service/service.go:13-15).

This package is a thin ctypes binding of the C ABI in ``include/pk_route.h``
(``libpkroute.so``: hand-written sm_100a kernels + the host shim).  It exists for the
tests and the benchmark; the production caller is the Go batcher shown in INTEGRATION.md.

There is NO CPU fallback: if ``libpkroute.so`` is missing, importing the binding raises,
and if no CUDA device is present ``Router(...)`` raises ``PkError(PK_ERR_NO_DEVICE)``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

__all__ = ["Router", "PkError", "PkTable", "PkTool", "PkProvider", "DECISION_DTYPE", "default_table",
           "make_table", "pack_requests", "round16", "lib", "LIB_PATH", "STATUS", "FLAG_NO_PROVIDER", "FLAG_EMPTY_PAYLOAD", "FLAG_UNKNOWN_TOOL", "PROVIDER_NONE"]

# POLYKEY_PKROUTE_LIB: developer aid — load a tuning variant built with build.build(defines=..., out=...)
LIB_PATH = os.environ.get("POLYKEY_PKROUTE_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libpkroute.so")

PK_MAX_TOOLS, PK_MAX_TOOL_NAME, PK_MAX_PROVIDERS = 16, 64, 32
FLAG_NO_PROVIDER, FLAG_EMPTY_PAYLOAD, FLAG_UNKNOWN_TOOL = 2, 4, 8
PROVIDER_NONE = 0xFF
STATUS = {0: "PK_OK", -1: "PK_ERR_ARG", -2: "PK_ERR_ALIGN", -3: "PK_ERR_OFFSETS", -4: "PK_ERR_CAPACITY",
          -5: "PK_ERR_CUDA", -6: "PK_ERR_NO_DEVICE", -7: "PK_ERR_BUSY", -8: "PK_ERR_STATE"}

# the 24-byte pk_decision record
DECISION_DTYPE = np.dtype(
    [("byte_len", "<u4"), ("est_tokens", "<u4"), ("route_hash", "<u8"), ("tool_class", "u1"),
     ("provider", "u1"), ("flags", "<u2"), ("status", "<u4")], align=False)
assert DECISION_DTYPE.itemsize == 24


class PkTool(C.Structure):
    _fields_ = [("name", C.c_uint8 * PK_MAX_TOOL_NAME), ("name_len", C.c_uint32), ("class_id", C.c_uint32)]


class PkProvider(C.Structure):
    _fields_ = [("max_tokens", C.c_uint32), ("weight", C.c_uint32), ("tool_mask", C.c_uint32),
                ("reserved", C.c_uint32)]


class PkTable(C.Structure):
    _fields_ = [("n_tools", C.c_uint32), ("default_class", C.c_uint32), ("n_providers", C.c_uint32),
                ("reserved", C.c_uint32), ("tools", PkTool * PK_MAX_TOOLS),
                ("providers", PkProvider * PK_MAX_PROVIDERS)]


class PkConfig(C.Structure):
    _fields_ = [("max_requests", C.c_uint32), ("n_slots", C.c_uint32),
                ("max_payload_bytes", C.c_uint64), ("max_name_bytes", C.c_uint64)]


class PkError(RuntimeError):
    def __init__(self, code: int, where: str, cuda: int = 0):
        self.code, self.cuda = code, cuda
        msg = f"{where}: {STATUS.get(code, code)}"
        try:
            msg += f" ({lib().pk_strerror(code).decode()})"
        except Exception:
            pass
        if cuda:
            msg += f" [cudaError {cuda}]"
        super().__init__(msg)


_lib = None


def lib() -> C.CDLL:
    """Loads libpkroute.so; raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m polykey_service_b200.build` "
                "(nvcc, sm_100a). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        vp, u32 = C.c_void_p, C.c_uint32
        batch = [vp, vp, vp, vp, vp, u32]           # payload, pay_off, pay_len, names, name_off, n
        L.pk_abi_version.restype = u32
        L.pk_strerror.restype, L.pk_strerror.argtypes = C.c_char_p, [C.c_int]
        L.pk_default_table.argtypes = [C.POINTER(PkTable)]
        L.pk_init.argtypes = [C.c_int, C.POINTER(PkTable), C.POINTER(PkConfig), C.POINTER(vp)]
        L.pk_destroy.argtypes, L.pk_destroy.restype = [vp], None
        L.pk_update_table.argtypes = [vp, C.POINTER(PkTable)]
        L.pk_route_batch.argtypes = [vp] + batch + [vp]
        L.pk_submit.argtypes = [vp, u32] + batch + [vp]
        L.pk_wait.argtypes = [vp, u32]
        L.pk_stage.argtypes = [vp, u32] + batch
        L.pk_launch.argtypes = [vp, u32]
        L.pk_set_direct_bytes.argtypes = [vp, C.c_uint64]
        L.pk_direct_count.argtypes, L.pk_direct_count.restype = [vp], C.c_uint64
        L.pk_launch_many.argtypes = [vp, vp, u32, u32, C.c_int, C.POINTER(C.c_float)]
        L.pk_fetch.argtypes = [vp, u32, vp]
        L.pk_sync.argtypes = [vp]
        L.pk_slot_stream.argtypes, L.pk_slot_stream.restype = [vp, u32], vp
        L.pk_launch_count.argtypes, L.pk_launch_count.restype = [vp], C.c_uint64
        L.pk_host_alloc.argtypes, L.pk_host_alloc.restype = [vp, C.c_size_t], vp
        L.pk_host_free.argtypes, L.pk_host_free.restype = [vp, vp], None
        L.pk_probe_host_read.argtypes = [vp, vp, C.c_size_t, u32, C.POINTER(C.c_float)]
        L.pk_last_cuda_error.argtypes = [vp]
        L.pk_route_device_noalloc.argtypes = [C.POINTER(PkTable), vp, vp, vp, vp, vp, vp, u32, u32, u32, vp, vp]
        _lib = L
    return _lib


def default_table() -> PkTable:
    """example_tool→0, struct_tool→1, file_tool→2, default→3 (mock.go:33,37,52,60) + 8 providers."""
    t = PkTable()
    rc = lib().pk_default_table(C.byref(t))
    if rc:
        raise PkError(rc, "pk_default_table")
    return t


def make_table(tools, providers, default_class: int = 3) -> PkTable:
    """tools: [(name_bytes, class_id)]; providers: [(max_tokens, weight, tool_mask, reserved)]."""
    t = PkTable()
    t.n_tools, t.default_class, t.n_providers = len(tools), default_class, len(providers)
    for k, (name, cid) in enumerate(tools[:PK_MAX_TOOLS]):
        name = bytes(name)
        t.tools[k].name_len, t.tools[k].class_id = len(name), cid
        for j, b in enumerate(name[:PK_MAX_TOOL_NAME]):
            t.tools[k].name[j] = b
    for p, (mt, w, mask, rsv) in enumerate(providers[:PK_MAX_PROVIDERS]):
        e = t.providers[p]
        e.max_tokens, e.weight, e.tool_mask, e.reserved = mt, w, mask, rsv
    return t


def round16(x: int) -> int:
    return (int(x) + 15) & ~15


def pack_requests(requests):
    """[(tool_name_bytes, payload_bytes)] -> the flat batch layout of include/pk_route.h:
    payload arena with 16-byte aligned starts, pay_off (n+1), pay_len (n), names, name_off (n+1).
    This is what the Go batcher does per flush."""
    requests = list(requests)
    n = len(requests)
    pay_len = np.array([len(p) for _, p in requests], dtype=np.uint32)
    pay_off = np.zeros(n + 1, dtype=np.uint64)
    if n:
        np.cumsum((pay_len.astype(np.uint64) + np.uint64(15)) & np.uint64(0xFFFFFFFFFFFFFFF0), out=pay_off[1:])
    payload = np.zeros(max(int(pay_off[-1]), 16), dtype=np.uint8)[:int(pay_off[-1])]
    for i, (_, p) in enumerate(requests):
        if len(p):
            payload[int(pay_off[i]):int(pay_off[i]) + len(p)] = np.frombuffer(bytes(p), dtype=np.uint8)
    name_off = np.zeros(n + 1, dtype=np.uint32)
    if n:
        np.cumsum(np.array([len(nm) for nm, _ in requests], dtype=np.uint32), out=name_off[1:])
    names = np.frombuffer(b"".join(bytes(nm) for nm, _ in requests), dtype=np.uint8).copy()
    return {"payload": payload, "pay_off": pay_off, "pay_len": pay_len, "names": names, "name_off": name_off}


def _batch_args(batch):
    n = int(batch["pay_len"].shape[0])
    for key, dt in (("payload", np.uint8), ("pay_off", np.uint64), ("pay_len", np.uint32),
                    ("names", np.uint8), ("name_off", np.uint32)):
        a = batch[key]
        if a.dtype != dt or not a.flags["C_CONTIGUOUS"]:
            raise TypeError(f"batch[{key!r}] must be a C-contiguous {np.dtype(dt).name} array")
    return [batch["payload"].ctypes.data, batch["pay_off"].ctypes.data, batch["pay_len"].ctypes.data,
            batch["names"].ctypes.data, batch["name_off"].ctypes.data, n]


class Router:
    """One handle = one GPU.  Mirrors the C ABI one to one; see include/pk_route.h."""

    def __init__(self, device: int = 0, table: PkTable | None = None, max_requests: int = 65536,
                 max_payload_bytes: int = 32 << 20, max_name_bytes: int | None = None, n_slots: int = 1):
        self._L = lib()
        self._h = C.c_void_p()
        table = table if table is not None else default_table()
        cfg = PkConfig(max_requests, n_slots, max_payload_bytes,
                       max_name_bytes if max_name_bytes is not None else 64 * max_requests)
        rc = self._L.pk_init(device, C.byref(table), C.byref(cfg), C.byref(self._h))
        if rc:
            raise PkError(rc, "pk_init")
        self.device, self.n_slots, self.max_requests = device, n_slots, max_requests
        self._pinned = []

    # -- lifetime -------------------------------------------------------------------
    def close(self):
        if self._h:
            for p in self._pinned:
                self._L.pk_host_free(self._h, p)
            self._pinned = []
            self._L.pk_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, where: str):
        if rc:
            raise PkError(rc, where, self._L.pk_last_cuda_error(self._h))

    # -- the hot call ---------------------------------------------------------------
    def route_batch(self, batch, out: np.ndarray | None = None) -> np.ndarray:
        args = _batch_args(batch)
        if out is None:
            out = np.empty(args[-1], dtype=DECISION_DTYPE)
        self._check(self._L.pk_route_batch(self._h, *args, out.ctypes.data), "pk_route_batch")
        return out

    def submit(self, slot: int, batch, out: np.ndarray):
        self._check(self._L.pk_submit(self._h, slot, *_batch_args(batch), out.ctypes.data), "pk_submit")

    def wait(self, slot: int):
        self._check(self._L.pk_wait(self._h, slot), "pk_wait")

    # -- device-resident form ---------------------------------------------------------
    def stage(self, slot: int, batch):
        self._check(self._L.pk_stage(self._h, slot, *_batch_args(batch)), "pk_stage")

    def launch(self, slot: int):
        self._check(self._L.pk_launch(self._h, slot), "pk_launch")

    def launch_many(self, slots, count: int, use_graph=False) -> float:
        """use_graph: False = plain launches, True / 1 = one graph (a chain), k >= 2 = a graph of k parallel chains."""
        arr = np.asarray(slots, dtype=np.uint32)
        ms = C.c_float(0)
        self._check(self._L.pk_launch_many(self._h, arr.ctypes.data, arr.size, count, int(use_graph),
                                           C.byref(ms)), "pk_launch_many")
        return float(ms.value)

    def fetch(self, slot: int, n: int) -> np.ndarray:
        out = np.empty(n, dtype=DECISION_DTYPE)
        self._check(self._L.pk_fetch(self._h, slot, out.ctypes.data), "pk_fetch")
        return out

    def sync(self):
        self._check(self._L.pk_sync(self._h), "pk_sync")

    def update_table(self, table: PkTable):
        self._check(self._L.pk_update_table(self._h, C.byref(table)), "pk_update_table")

    def slot_stream(self, slot: int) -> int:
        return int(self._L.pk_slot_stream(self._h, slot) or 0)

    @property
    def launch_count(self) -> int:
        return int(self._L.pk_launch_count(self._h))

    @property
    def direct_count(self) -> int:
        """Batches routed straight out of page-locked host memory (no copies)."""
        return int(self._L.pk_direct_count(self._h))

    def set_direct_bytes(self, nbytes: int):
        """Largest batch (payload + metadata bytes) that takes the no-copy path; 0 = never."""
        self._check(self._L.pk_set_direct_bytes(self._h, int(nbytes)), "pk_set_direct_bytes")

    def probe_host_read(self, host: np.ndarray, reps: int = 20) -> float:
        """GB/s at which a plain read kernel pulls the page-locked array `host` over PCIe."""
        ms = C.c_float(0)
        self._check(self._L.pk_probe_host_read(self._h, host.ctypes.data, host.nbytes, reps, C.byref(ms)),
                    "pk_probe_host_read")
        return host.nbytes * reps / (ms.value / 1e3) / 1e9

    def host_alloc(self, nbytes: int, dtype=np.uint8) -> np.ndarray:
        """Pinned host array owned by the C side (freed on close)."""
        p = self._L.pk_host_alloc(self._h, nbytes)
        if not p:
            raise PkError(-5, "pk_host_alloc", self._L.pk_last_cuda_error(self._h))
        self._pinned.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p)
        return np.frombuffer(buf, dtype=dtype)

"""ctypes binding of the batched AES-256-GCM ABI (include/pk_gcm.h), for tests and benchmarks.

The batched form of BatchEncrypt / BatchDecrypt (internal/adapters/security/cipher.go:110-141).
No CPU fallback: without a CUDA device the constructor raises GcmError(PK_GCM_ERR_NO_DEVICE).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib as _route_lib

NONCE, TAG, OVERHEAD = 12, 16, 28
STATUS = {0: "PK_GCM_OK", -1: "PK_GCM_ERR_ARG", -2: "PK_GCM_ERR_ALIGN", -3: "PK_GCM_ERR_OFFSETS",
          -4: "PK_GCM_ERR_CAPACITY", -5: "PK_GCM_ERR_CUDA", -6: "PK_GCM_ERR_NO_DEVICE",
          -9: "PK_GCM_ERR_KEYLEN", -10: "PK_GCM_ERR_AUTH"}


class GcmError(RuntimeError):
    def __init__(self, code: int, where: str, status=None):
        self.code, self.status = code, status
        super().__init__(f"{where}: {STATUS.get(code, code)} ({lib().pk_gcm_strerror(code).decode()})")


_bound = False


def lib() -> C.CDLL:
    global _bound
    L = _route_lib()
    if not _bound:
        vp, u32 = C.c_void_p, C.c_uint32
        L.pk_gcm_init.argtypes = [C.c_int, C.c_char_p, u32, u32, C.c_uint64, C.POINTER(vp)]
        L.pk_gcm_destroy.argtypes, L.pk_gcm_destroy.restype = [vp], None
        L.pk_gcm_seal_batch.argtypes = [vp, vp, vp, vp, vp, u32, vp, vp]
        L.pk_gcm_open_batch.argtypes = [vp, vp, vp, vp, u32, vp, vp, vp]
        L.pk_gcm_strerror.argtypes, L.pk_gcm_strerror.restype = [C.c_int], C.c_char_p
        L.pk_gcm_launch_count.argtypes, L.pk_gcm_launch_count.restype = [vp], C.c_uint64
        L.pk_gcm_last_cuda_error.argtypes = [vp]
        L.pk_gcm_last_kernel_ms.argtypes, L.pk_gcm_last_kernel_ms.restype = [vp], C.c_float
        L.pk_gcm_set_chunks.argtypes = [vp, u32]
        L.pk_gcm_set_direct_bytes.argtypes = [vp, C.c_uint64]
        L.pk_gcm_direct_count.argtypes, L.pk_gcm_direct_count.restype = [vp], C.c_uint64
        L.pk_gcm_host_alloc.argtypes, L.pk_gcm_host_alloc.restype = [vp, C.c_size_t], vp
        L.pk_gcm_host_free.argtypes, L.pk_gcm_host_free.restype = [vp, vp], None
        _bound = True
    return L


def slots(lengths):
    """16-byte aligned slot offsets for messages of the given lengths -> (offsets u64[n], total)."""
    lengths = np.asarray(lengths, dtype=np.uint64)
    sizes = (lengths + np.uint64(15)) & np.uint64(0xFFFFFFFFFFFFFFF0)
    off = np.zeros(lengths.size, dtype=np.uint64)
    if lengths.size > 1:
        np.cumsum(sizes[:-1], out=off[1:])
    return off, int(off[-1] + sizes[-1]) if lengths.size else 0


class Gcm:
    """One key on one GPU."""

    def __init__(self, key: bytes, device: int = 0, max_messages: int = 65536, max_bytes: int = 64 << 20):
        self._L = lib()
        self._h = C.c_void_p()
        rc = self._L.pk_gcm_init(device, bytes(key), len(key), max_messages, max_bytes, C.byref(self._h))
        if rc:
            raise GcmError(rc, "pk_gcm_init")
        self._pinned = []

    def pinned(self, nbytes: int, dtype=np.uint8) -> np.ndarray:
        """A page-locked array (pk_gcm_host_alloc) of nbytes bytes viewed as dtype; freed by close()."""
        nbytes = max(int(nbytes), 16)
        p = self._L.pk_gcm_host_alloc(self._h, nbytes)
        if not p:
            raise GcmError(-5, "pk_gcm_host_alloc")
        self._pinned.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p)
        return np.frombuffer(buf, dtype=np.uint8).view(dtype)

    def set_chunks(self, chunks: int):
        """0 = pipeline large batches (default), 1 = never, 2..8 = that many chunks."""
        rc = self._L.pk_gcm_set_chunks(self._h, chunks)
        if rc:
            raise GcmError(rc, "pk_gcm_set_chunks")

    def set_direct_bytes(self, nbytes: int):
        """Largest batch (arena bytes in + out) worked on in place, no copies; 0 = never."""
        rc = self._L.pk_gcm_set_direct_bytes(self._h, int(nbytes))
        if rc:
            raise GcmError(rc, "pk_gcm_set_direct_bytes")

    @property
    def direct_count(self) -> int:
        return int(self._L.pk_gcm_direct_count(self._h))

    def close(self):
        if self._h:
            for p in self._pinned:
                self._L.pk_gcm_host_free(self._h, p)
            self._pinned = []
            self._L.pk_gcm_destroy(self._h)
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

    @property
    def launch_count(self) -> int:
        return int(self._L.pk_gcm_launch_count(self._h))

    @property
    def last_kernel_ms(self) -> float:
        return float(self._L.pk_gcm_last_kernel_ms(self._h))

    # flat-arena forms (what the ABI takes) ------------------------------------------------
    def seal_arena(self, plain, plain_off, plain_len, nonces, sealed, sealed_off):
        rc = self._L.pk_gcm_seal_batch(self._h, plain.ctypes.data, plain_off.ctypes.data, plain_len.ctypes.data,
                                       nonces.ctypes.data, plain_len.size, sealed.ctypes.data, sealed_off.ctypes.data)
        if rc:
            raise GcmError(rc, "pk_gcm_seal_batch")

    def open_arena(self, sealed, sealed_off, sealed_len, plain, plain_off):
        status = np.zeros(sealed_len.size, dtype=np.uint32)
        rc = self._L.pk_gcm_open_batch(self._h, sealed.ctypes.data, sealed_off.ctypes.data, sealed_len.ctypes.data,
                                       sealed_len.size, plain.ctypes.data, plain_off.ctypes.data, status.ctypes.data)
        if rc not in (0, -10):
            raise GcmError(rc, "pk_gcm_open_batch")
        return rc, status

    # list-of-bytes forms (BatchEncrypt / BatchDecrypt signatures) ---------------------------
    def batch_encrypt(self, plaintexts, nonces):
        n = len(plaintexts)
        if n == 0:
            return []
        plen = np.array([len(p) for p in plaintexts], dtype=np.uint32)
        poff, ptotal = slots(plen)
        plain = np.zeros(max(ptotal, 16), dtype=np.uint8)
        for p, o in zip(plaintexts, poff):
            plain[int(o):int(o) + len(p)] = np.frombuffer(bytes(p), dtype=np.uint8)
        soff, stotal = slots(plen.astype(np.uint64) + OVERHEAD)
        sealed = np.zeros(max(stotal, 16), dtype=np.uint8)
        nonce_arr = np.frombuffer(b"".join(bytes(x) for x in nonces), dtype=np.uint8).copy()
        assert nonce_arr.size == NONCE * n
        self.seal_arena(plain, poff, plen, nonce_arr, sealed, soff)
        return [bytes(sealed[int(o):int(o) + int(l) + OVERHEAD]) for o, l in zip(soff, plen)]

    def batch_decrypt(self, sealed_msgs):
        """Returns (plaintexts, status); raises GcmError(PK_GCM_ERR_AUTH) like BatchDecrypt's error
        if any message fails (cipher.go:135-137) — the exception carries the per-message status."""
        n = len(sealed_msgs)
        if n == 0:
            return []
        slen = np.array([len(s) for s in sealed_msgs], dtype=np.uint32)
        soff, stotal = slots(slen)
        sealed = np.zeros(max(stotal, 16), dtype=np.uint8)
        for s, o in zip(sealed_msgs, soff):
            sealed[int(o):int(o) + len(s)] = np.frombuffer(bytes(s), dtype=np.uint8)
        plen = np.maximum(slen.astype(np.int64) - OVERHEAD, 0).astype(np.uint64)
        poff, ptotal = slots(plen)
        plain = np.zeros(max(ptotal, 16), dtype=np.uint8)
        rc, status = self.open_arena(sealed, soff, slen, plain, poff)
        if rc:
            raise GcmError(rc, "pk_gcm_open_batch", status)
        return [bytes(plain[int(o):int(o) + int(l)]) for o, l in zip(poff, plen)]

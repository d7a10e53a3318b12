#!/usr/bin/env python
"""Batched AES-256-GCM (the cipher.go widening step): GPU vs OpenSSL on the same box.

For each workload: seal a batch on the GPU through the C ABI (host buffers in and out), check it
bit-exactly against OpenSSL, and time
  gpu_e2e_ms     one pk_gcm_seal_batch call (copies + kernel + copies), median, pageable arenas
  gpu_e2e_pinned_ms  the same with page-locked arenas (pk_gcm_host_alloc), library defaults;
                     _serial: copies, chunking off;  _copies: copies, chunked;  gpu_e2e_inplace_ms: no copies
  gpu_kernel_ms  the kernel alone (CUDA events on its stream)
  cpu_ref_ms     OpenSSL with a fresh cipher per message — what BatchEncrypt does
                 (internal/adapters/security/cipher.go:116-122) — all cores / 1 core
  cpu_reuse_ms   OpenSSL with the key set once per thread, all cores
Test/measurement tooling; prints one JSON object.
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from polykey_service_b200 import gcm  # noqa: E402

WORKLOADS = [("65536 x 64 B", 65536, 64), ("65536 x 256 B", 65536, 256), ("4096 x 4 KiB", 4096, 4096),
             ("1024 x 64 B", 1024, 64), ("256 x 4 KiB", 256, 4096), ("32768 x 4 KiB", 32768, 4096)]
# kernel-only sweep of the thread-per-message / warp-per-message split (PK_GCM_LONG_BYTES)
SWEEP = [(65536, 512), (65536, 1024), (65536, 2048), (32768, 4096), (16384, 1024), (16384, 4096), (4096, 512),
         (4096, 1024), (4096, 4096), (1024, 256), (1024, 1024), (256, 256), (256, 4096), (16, 16384)]


def med(f, reps):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    ts.sort()
    return ts[len(ts) // 2] * 1e3


def main():
    ref = C.CDLL(os.path.join(ROOT, "oracle", "libpkaesref.so"))
    ref.pk_aes_ref_seal_batch.argtypes = [C.c_char_p] + [C.c_void_p] * 4 + [C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int]
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(7)
    key = rng.bytes(32)
    rows = []
    for name, n, length in WORKLOADS:
        plen = np.full(n, length, dtype=np.uint32)
        poff, ptotal = gcm.slots(plen)
        soff, stotal = gcm.slots(plen.astype(np.uint64) + gcm.OVERHEAD)
        plain = rng.integers(0, 256, ptotal, dtype=np.uint8)
        nonces = rng.integers(0, 256, 12 * n, dtype=np.uint8)
        sealed = np.zeros(stotal, dtype=np.uint8)
        want = np.zeros(stotal, dtype=np.uint8)

        def cpu(threads, per_item):
            ref.pk_aes_ref_seal_batch(key, plain.ctypes.data, poff.ctypes.data, plen.ctypes.data, nonces.ctypes.data,
                                      n, want.ctypes.data, soff.ctypes.data, threads, per_item)
        with gcm.Gcm(key, max_messages=n, max_bytes=stotal + 64) as g:
            g.seal_arena(plain, poff, plen, nonces, sealed, soff)
            cpu(cores, 1)
            for i in range(n):                                   # compare message bytes, not slot padding
                o, l = int(soff[i]), length + gcm.OVERHEAD
                if not (sealed[o:o + l] == want[o:o + l]).all():
                    raise SystemExit(f"GPU and OpenSSL differ at message {i} of {name}")
            t_end = time.perf_counter() + 0.3
            while time.perf_counter() < t_end:
                g.seal_arena(plain, poff, plen, nonces, sealed, soff)
            e2e = med(lambda: g.seal_arena(plain, poff, plen, nonces, sealed, soff), 30)
            g.set_chunks(1)                                      # kernels alone: no copy overlapping them
            ks = []
            for _ in range(10):
                g.seal_arena(plain, poff, plen, nonces, sealed, soff)
                ks.append(g.last_kernel_ms)
            kernel = sorted(ks)[len(ks) // 2]
            # the same call with every arena page-locked (pk_gcm_host_alloc), as the host mirror packs them
            pin = [g.pinned(a.nbytes, a.dtype) for a in (plain, poff, plen, nonces, sealed, soff)]
            for dst, src in zip(pin, (plain, poff, plen, nonces, sealed, soff)):
                dst[:src.size] = src
            p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff = (d[:s_.size] for d, s_ in zip(pin, (plain, poff, plen, nonces, sealed, soff)))
            g.set_direct_bytes(0)                                # copies, one chunk
            p_sealed[:] = 0
            g.seal_arena(p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff)
            if not (np.asarray(p_sealed) == sealed).all():
                raise SystemExit(f"pinned and pageable results differ for {name}")
            e2e_pinned_serial = med(lambda: g.seal_arena(p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff), 30)
            g.set_chunks(0)                                      # default: large batches pipelined in chunks
            p_sealed[:] = 0
            g.seal_arena(p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff)
            if not (np.asarray(p_sealed) == sealed).all():
                raise SystemExit(f"pipelined result differs for {name}")
            e2e_pinned_copies = med(lambda: g.seal_arena(p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff), 30)
            g.set_direct_bytes(1 << 40)                          # in place whatever the size
            p_sealed[:] = 0
            g.seal_arena(p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff)
            if not (np.asarray(p_sealed) == sealed).all():
                raise SystemExit(f"in-place result differs for {name}")
            e2e_inplace = med(lambda: g.seal_arena(p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff), 30)
            g.set_direct_bytes(4 << 20)                          # the library's default
            direct_before = g.direct_count
            e2e_pinned = med(lambda: g.seal_arena(p_plain, p_poff, p_plen, p_nonces, p_sealed, p_soff), 30)
            in_place = g.direct_count > direct_before
        row = {"workload": name, "messages": n, "bytes": length, "payload_mb": n * length / 1e6,
               "gpu_e2e_ms": e2e, "gpu_e2e_pinned_ms": e2e_pinned, "gpu_e2e_pinned_serial_ms": e2e_pinned_serial, "gpu_e2e_pinned_copies_ms": e2e_pinned_copies,
               "gpu_e2e_inplace_ms": e2e_inplace, "default_is_in_place": in_place, "gpu_kernel_ms": kernel,
               "cpu_ref_all_ms": med(lambda: cpu(cores, 1), 10), "cpu_ref_1core_ms": med(lambda: cpu(1, 1), 3),
               "cpu_reuse_all_ms": med(lambda: cpu(cores, 0), 10)}
        row["gpu_kernel_gbs"] = n * length / row["gpu_kernel_ms"] / 1e6
        row["gpu_e2e_msgs_per_s"] = n / row["gpu_e2e_ms"] * 1e3
        row["cpu_ref_all_msgs_per_s"] = n / row["cpu_ref_all_ms"] * 1e3
        rows.append(row)
        print(row, file=sys.stderr, flush=True)
    sweep = []
    for n, length in SWEEP:
        plen = np.full(n, length, dtype=np.uint32)
        poff, ptotal = gcm.slots(plen)
        soff, stotal = gcm.slots(plen.astype(np.uint64) + gcm.OVERHEAD)
        plain = rng.integers(0, 256, ptotal, dtype=np.uint8)
        nonces = rng.integers(0, 256, 12 * n, dtype=np.uint8)
        sealed = np.zeros(stotal, dtype=np.uint8)
        row = {"messages": n, "bytes": length}
        with gcm.Gcm(key, max_messages=n, max_bytes=stotal + 64) as g:
            g.set_chunks(1)
            for label, env in (("thread_ms", str(1 << 30)), ("warp_ms", "64"), ("default_ms", "")):
                os.environ["PK_GCM_LONG_BYTES"] = env
                ks = []
                for _ in range(8):
                    g.seal_arena(plain, poff, plen, nonces, sealed, soff)
                    ks.append(g.last_kernel_ms)
                row[label] = sorted(ks[2:])[len(ks[2:]) // 2]
        os.environ["PK_GCM_LONG_BYTES"] = ""
        sweep.append(row)
        print(row, file=sys.stderr, flush=True)
    print(json.dumps({"cores": cores, "rows": rows, "sweep": sweep}))


if __name__ == "__main__":
    main()

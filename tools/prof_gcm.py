#!/usr/bin/env python
"""Seal one fixed batch a few times — a deterministic launch sequence for ncu captures of the
AES-GCM kernels:  python tools/prof_gcm.py MESSAGES BYTES [REPS]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from polykey_service_b200 import gcm  # noqa: E402

n, length = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 6
rng = np.random.default_rng(7)
plen = np.full(n, length, dtype=np.uint32)
poff, ptotal = gcm.slots(plen)
soff, stotal = gcm.slots(plen.astype(np.uint64) + gcm.OVERHEAD)
plain = rng.integers(0, 256, ptotal, dtype=np.uint8)
nonces = rng.integers(0, 256, 12 * n, dtype=np.uint8)
sealed = np.zeros(stotal, dtype=np.uint8)
with gcm.Gcm(rng.bytes(32), max_messages=n, max_bytes=stotal + 64) as g:
    g.set_chunks(1)
    for _ in range(reps):
        g.seal_arena(plain, poff, plen, nonces, sealed, soff)
        print("kernel_ms", g.last_kernel_ms)

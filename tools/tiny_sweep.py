#!/usr/bin/env python
"""Tiny payloads, few requests: serial-chain µs per batch with the packed lane-group path and with
a warp per request (PK_TINY_WARP_MAX forces the latter up to that batch size).  Run once per
setting: PK_TINY_WARP_MAX=0 / 1000000 python tools/tiny_sweep.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import polykey_service_b200 as pk
from oracle import pk_oracle as O

out = {}
for plen in (64, 256):
    for n in (1, 64, 512, 1024, 2048, 4096, 8192, 16384):
        ring = 9
        r = pk.Router(device=0, max_requests=n, max_payload_bytes=max(64, n * plen), n_slots=ring)
        for j in range(ring):
            r.stage(j, O.gen_batch(O.SEED_BASE + 1, n, plen, mode=0, first_id=j * n))
        slots = list(range(ring))
        r.launch_many(slots, 500, use_graph=False)
        r.launch_many(slots, 200, use_graph=1)
        out[f"{n} x {plen} B"] = round(r.launch_many(slots, 2000, use_graph=1) / 2000 * 1e3, 2)
        r.close()
print(json.dumps({"PK_TINY_WARP_MAX": os.environ.get("PK_TINY_WARP_MAX"), "serial_us": out}))

#!/usr/bin/env python
"""How much of a C1 step is not the kernel's own work: per-step time of a graph of K launches for
batches that are almost empty (1 request of 16 B; 4,096 requests of 16 B) next to C1 itself."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import polykey_service_b200 as pk
from oracle import pk_oracle as O

out = {}
for name, n, plen in (("1 x 16 B", 1, 16), ("4096 x 16 B", 4096, 16), ("4096 x 512 B", 4096, 512), ("4096 x 2048 B (C1)", 4096, 2048)):
    ring = 33
    r = pk.Router(device=0, max_requests=n, max_payload_bytes=max(64, n * plen), n_slots=ring)
    for j in range(ring):
        r.stage(j, O.gen_batch(O.SEED_BASE + 1, n, plen, mode=0, first_id=j * n))
    slots = list(range(ring))
    r.launch_many(slots, 2000, use_graph=False)
    r.launch_many(slots, 200, use_graph=True)
    out[name] = {"graph_us": r.launch_many(slots, 4000, use_graph=True) / 4000 * 1e3,
                 "plain_us": r.launch_many(slots, 4000, use_graph=False) / 4000 * 1e3,
                 "graph4_us": r.launch_many(slots, 4000, use_graph=4) / 4000 * 1e3}
    r.close()
print(json.dumps(out, indent=1))

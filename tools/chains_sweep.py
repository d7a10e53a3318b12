import sys, os
sys.path.insert(0, os.getcwd())
import polykey_service_b200 as pk
from oracle import pk_oracle as O
O.build()
n, plen = 4096, 2048
ring = 33
r = pk.Router(device=0, max_requests=n, max_payload_bytes=n*plen, n_slots=ring)
for j in range(ring): r.stage(j, O.gen_batch(O.SEED_BASE+1, n, plen, mode=0, first_id=j*n, threads=16))
slots = list(range(ring)); PDL = 0x100
import time
t=time.time()
while time.time()-t < 0.4: r.launch_many(slots, 512, use_graph=False)
for w in (1,2,3,4,6,8,11,16):
    r.launch_many(slots, 100, use_graph=w|PDL)
    print("chains", w, "K=4000: %.3f us   K=20: %.3f us" % (r.launch_many(slots, 4000, use_graph=w|PDL)/4000*1e3, min(r.launch_many(slots, 20, use_graph=w|PDL) for _ in range(5))/20*1e3))

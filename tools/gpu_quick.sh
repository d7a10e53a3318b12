#!/bin/bash
# quick kernel check: parity tests, C1/C2 lean bench lines, ncu capture of C2
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
for w in c1 c2; do python bench.py --workload $w --steps 2000 --warmup 10 --no-cpu --lean 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$w', 'pdl %.3f us frac %.3f serial %.3f us e2e %.4g'%(d['ms_per_step']*1e3, d['roofline']['frac'], d['serial']['ms_per_step']*1e3, d['e2e']['value']))"; done
python bench.py --workload c2 --steps 6 --warmup 3 --no-cpu --no-prewarm --lean > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 8 -c 2 -o gpurun_out/prof_c2 -f python bench.py --workload c2 --steps 6 --warmup 3 --no-cpu --no-prewarm --lean > gpurun_out/ncu_full_c2.log 2>&1; echo "ncu rc=$?"
for k in 1 2; do PK_SHARDS_PER_GPU=$k timeout 120 ./build/latency_bench open 20000000 2 4096 200 2048 12 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('shards/gpu',d['shards_per_gpu'],'gens',d['generators'],'completed %.3g mean_batch %.1f'%(d['completed_per_s'],d['mean_batch']))"; done

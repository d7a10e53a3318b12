#!/bin/bash
# tests + front-end A/B + the driver-shaped bench line
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
bash tools/gpu_fe_ab.sh 2>&1 | tee gpurun_out/fe_ab.txt
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_k20.json 2> gpurun_out/bench_k20.err; echo "bench k20 rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_k20.json'))
print('value %.4g frac %.3f steady %s serial %.3f e2e %.4g frac_pcie %s'%(d['value'],d['roofline']['frac'],d.get('steady_state',{}).get('frac'),d['serial']['frac'],d['e2e']['value'],d['e2e'].get('frac_of_pcie')))
for k,v in d.get('workloads',{}).items(): print(k,'frac %.3f e2e %.4g'%(v['frac'],v['e2e']['value']))
print('latency',[(r['offered_per_s'],r.get('completed_per_s'),r.get('p50_us'),r.get('p99_us')) for r in d['latency']['rates']])
PY

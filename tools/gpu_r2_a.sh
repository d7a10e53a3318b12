#!/bin/bash
# Round-2 first GPU pass: tests, smoke, the driver-shaped bench line, the reference arm, and ncu
# captures of route_kernel on C1 / C2 (instruction counts per request before the C2 work).
set -u
mkdir -p gpurun_out
{ nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv; nproc; lscpu | grep -E "Model name|Socket|Core|Thread|NUMA"; } > gpurun_out/box.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_k20.json 2> gpurun_out/bench_k20.err; echo "bench k20 rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
for w in c1 c2; do
  timeout 600 python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm --lean > gpurun_out/plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 8 -c 2 -o gpurun_out/prof_$w \
      python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm --lean > gpurun_out/ncu_full_$w.log 2>&1; echo "ncu full $w rc=$?"
done
cat gpurun_out/bench_k20.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value %.4g frac %.3f steady %s serial %.3f e2e %.4g frac_pcie %s'%(d['value'],d['roofline']['frac'],d.get('steady_state',{}).get('frac'),d['serial']['frac'],d['e2e']['value'],d['e2e'].get('frac_of_pcie')))
for k,v in d.get('workloads',{}).items(): print(k,'frac %.3f e2e %.4g'%(v['frac'],v['e2e']['value']))
print('c4',d.get('c4',{}).get('value'),d.get('c4',{}).get('e2e'))
print('latency',json.dumps(d.get('latency'))[:1500])
print('cpu',d['cpu_baseline'])
"

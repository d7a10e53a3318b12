#!/bin/bash
# Round-2 measurement pass on one GPU: tests, smoke, bench lines, reference arm, ncu launch list and
# full captures (each after its plain run exited 0), front-end sweep.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
{ nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv; nproc; lscpu | grep -E "Model name|Socket|Core|Thread|NUMA|L3"; go version 2>&1 | head -1; } > gpurun_out/box.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_k20.json 2> gpurun_out/bench_k20.err; echo "bench k20 rc=$?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_k20.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"
# ncu: launch list of the bench command, then full captures (each after its plain run exited 0)
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu --no-prewarm --lean > gpurun_out/plain_c1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c1.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu --no-prewarm --lean > gpurun_out/ncu_launches_c1.log 2>&1; echo "ncu launches rc=$?"
# (default issue = fused: the 2nd route_fused_kernel launch is the first timed one — all 20 batches of c1,
#  all 17 of c2's ring, 2 of c3's; profiles/ncu_traffic.json divides its DRAM bytes by that count)
for cfg in "c1 20" "c2 17" "c3 6"; do
  set -- $cfg; w=$1; k=$2
  timeout 600 python bench.py --workload $w --steps $k --warmup 3 --no-cpu --no-prewarm --lean > gpurun_out/plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:route_fused -s 1 -c 1 -f -o gpurun_out/prof_$w \
      python bench.py --workload $w --steps $k --warmup 3 --no-cpu --no-prewarm --lean > gpurun_out/ncu_full_$w.log 2>&1; echo "ncu full $w rc=$?"
done
# front-end: open loop, several rates
: > gpurun_out/frontend.jsonl
for cfg in "100000 4" "1000000 8" "5000000 12" "10000000 12" "20000000 12" "20000000 14"; do
  set -- $cfg
  timeout 120 ./build/latency_bench open $1 2 4096 200 2048 $2 >> gpurun_out/frontend.jsonl 2>> gpurun_out/frontend.err || echo "fail $cfg"
done
for cfg in "1 2 4096 200 2048" "16 2 4096 200 2048" "128 2 4096 200 2048" "1024 2 4096 200 2048"; do
  timeout 120 ./build/latency_bench $cfg >> gpurun_out/frontend.jsonl 2>> gpurun_out/frontend.err || echo "fail closed $cfg"
done
python - <<'PY'
import json
for f in ("bench_k20","bench_default"):
    d=json.load(open("gpurun_out/%s.json"%f))
    print(f,'value %.4g frac %.3f steady %.3f serial %.3f e2e %.4g frac_pcie %.3f cpu %.4g'%(d['value'],d['roofline']['frac'],d['steady_state']['frac'],d['serial']['frac'],d['e2e']['value'],d['e2e']['frac_of_pcie'],d['cpu_baseline']['value']))
    for k,v in d['workloads'].items(): print('  ',k,'frac %.3f e2e %.4g'%(v['frac'],v['e2e']['value']))
    print('   c4 %.4g e2e %.4g'%(d['c4']['value'],d['c4']['e2e']['value']), 'gcm', {k:(round(v['kernel_ms'],3),round(v['call_ms'],3),round(v['openssl_key_once_ms'],3)) for k,v in d['gcm'].items() if isinstance(v,dict)})
    print('   latency',[(r['offered_per_s'],round(r.get('completed_per_s',0)),r.get('p50_us'),r.get('p99_us')) for r in d['latency']['rates']])
for l in open('gpurun_out/frontend.jsonl'):
    d=json.loads(l)
    if d.get('mode')=='open-loop': print('open gens %d offered %.3g completed %.3g p50 %.1f p99 %.1f mean_batch %.1f'%(d['generators'],d['offered_per_s'],d['completed_per_s'],d['latency_us']['p50'],d['latency_us']['p99'],d['mean_batch']))
    else: print('closed callers %d req/s %.3g p50 %.1f p99 %.1f mean_batch %.1f'%(d['callers'],d['requests_per_s'],d['latency_us']['p50'],d['latency_us']['p99'],d['mean_batch']))
PY

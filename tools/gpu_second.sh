#!/bin/bash
# Second GPU call: kernel v2 — tests, bench lines, requests-per-CTA sweep, ncu.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
for w in c1 c2 c3; do
  steps=4000; [ $w = c3 ] && steps=60; [ $w = c2 ] && steps=2000
  timeout 900 python bench.py --workload $w --steps $steps --warmup 10 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
done
for r in 8 16 32 64; do
  PK_REQS_PER_CTA=$r timeout 600 python bench.py --workload c1 --steps 4000 --warmup 10 --no-cpu > gpurun_out/sweep_c1_R$r.json 2>/dev/null; echo "sweep c1 R=$r rc=$?"
done
for r in 16 32 64 128; do
  PK_REQS_PER_CTA=$r timeout 600 python bench.py --workload c2 --steps 2000 --warmup 10 --no-cpu > gpurun_out/sweep_c2_R$r.json 2>/dev/null; echo "sweep c2 R=$r rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_c*.json')+glob.glob('gpurun_out/sweep_*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print(f, 'us/step=%.2f frac=%.3f value=%.4g e2e=%.4g alt_us=%.2f'%(r['avg_launch_us'], r['frac'], d['value'], d['e2e']['value'], d['alt_issue_ms_per_step']*1e3))
    except Exception as e: print(f, 'ERR', e)
PY
for w in c3 c2 c1; do
  timeout 600 python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 8 -c 2 -o gpurun_out/prof2_$w \
      python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/ncu_full_$w.log 2>&1; echo "ncu full $w rc=$?"
done

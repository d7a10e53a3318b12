#!/bin/bash
# Full measurement round: tests, smoke, bench lines (all workloads), reference arm, break-even
# sweep, ncu launch list + full captures.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
{ nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv; nproc; lscpu | grep -E "Model name|Socket|Core|Thread"; go version 2>&1 | head -1; } > gpurun_out/box.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 (defaults) rc=$?"
timeout 900 python bench.py --impl reference --steps 400 --warmup 10 > gpurun_out/bench_ref_c1.json 2> gpurun_out/bench_ref_c1.err; echo "bench ref rc=$?"
timeout 900 python bench.py --workload c2 --steps 2000 --warmup 10 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
timeout 900 python bench.py --workload c3 --steps 300 --warmup 60 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"
timeout 900 python tools/break_even.py 2048 > gpurun_out/break_even_2k.json 2> gpurun_out/break_even_2k.err; echo "break-even 2k rc=$?"
timeout 900 python tools/break_even.py 256 > gpurun_out/break_even_256.json 2> gpurun_out/break_even_256.err; echo "break-even 256 rc=$?"
# ncu: launch list of the bench command, then full captures (each after its plain run exited 0)
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu --no-prewarm > gpurun_out/plain_c1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c1.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu --no-prewarm > gpurun_out/ncu_launches_c1.log 2>&1; echo "ncu launches rc=$?"
for w in c1 c2 c3; do
  timeout 600 python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 8 -c 2 -o gpurun_out/prof_$w \
      python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/ncu_full_$w.log 2>&1; echo "ncu full $w rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_c?.json')+glob.glob('gpurun_out/bench_ref*.json')):
    try:
        d=json.load(open(f)); r=d.get('roofline') or {}
        print('%-28s value=%.4g ms/step=%.5f frac=%s e2e=%.4g cpu=%s serial=%s clocks=%s'%(f.split('/')[-1], d['value'], d['ms_per_step'], r.get('frac'), d['e2e']['value'], (d.get('cpu_baseline') or {}).get('value'), (d.get('serial') or {}).get('ms_per_step'), d.get('clocks')))
    except Exception as e: print(f, 'ERR', e)
PY

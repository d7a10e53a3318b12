#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for w in c1 c2 c3; do
  steps=4000; [ $w = c3 ] && steps=60; [ $w = c2 ] && steps=2000
  timeout 900 python bench.py --workload $w --steps $steps --warmup 10 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
done
PK_LATENCY_MODE=0 timeout 600 python bench.py --workload c1 --steps 4000 --warmup 10 --no-cpu > gpurun_out/bench_c1_thr.json 2>/dev/null
PK_LATENCY_MODE=1 PK_REQS_PER_CTA=16 timeout 600 python bench.py --workload c1 --steps 4000 --warmup 10 --no-cpu > gpurun_out/bench_c1_lat16.json 2>/dev/null
for r in 32 128; do PK_REQS_PER_CTA=$r timeout 600 python bench.py --workload c2 --steps 2000 --warmup 10 --no-cpu > gpurun_out/bench_c2_R$r.json 2>/dev/null; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_c*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print('%-36s us/step=%8.2f frac=%.3f alt_us=%8.2f value=%.4g e2e=%.4g lat_p50=%.0f'%(f.split('/')[-1], r['avg_launch_us'], r['frac'], d['alt_issue_ms_per_step']*1e3, d['value'], d['e2e']['value'], d['e2e']['sync_call_latency_us']['p50']))
    except Exception as e: print(f, 'ERR', e)
PY

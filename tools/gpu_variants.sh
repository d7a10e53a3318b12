#!/bin/bash
# A/B of build-time tuning variants (built on the CPU box into build/, git-ignored).
set -u
mkdir -p gpurun_out/var; rm -f gpurun_out/var/*.json
for v in c5u4 c6u4 c8u4 c6u2 c8u2 c5u2 c6u3; do
  for w in c3 c1 c2; do
    steps=4000; [ $w = c3 ] && steps=60; [ $w = c2 ] && steps=2000
    POLYKEY_PKROUTE_LIB=$PWD/build/libpkroute_$v.so timeout 600 python bench.py --workload $w --steps $steps --warmup 10 --no-cpu > gpurun_out/var/${v}_$w.json 2>/dev/null || echo "fail $v $w"
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/var/*.json')):
    try:
        d=json.load(open(f)); r=d['roofline']
        print('%-36s us/step=%8.2f frac=%.3f alt_us=%8.2f e2e=%.4g lat_p50=%.0f'%(f.split('/')[-1], r['avg_launch_us'], r['frac'], d['alt_issue_ms_per_step']*1e3, d['e2e']['value'], d['e2e']['sync_call_latency_us']['p50']))
    except Exception as e: print(f, 'ERR', e)
PY
export POLYKEY_PKROUTE_LIB=$PWD/build/libpkroute_c5u4.so
for w in c1 c2; do
timeout 600 python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/plain_$w.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 8 -c 2 -o gpurun_out/prof3_$w \
    python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/ncu_full_$w.log 2>&1; echo "ncu full $w rc=$?"
done

#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python tools/break_even.py 2048 > gpurun_out/break_even_2k.json 2> gpurun_out/break_even_2k.err; echo "break-even 2k rc=$?"
timeout 900 python tools/break_even.py 256 > gpurun_out/break_even_256.json 2> gpurun_out/break_even_256.err; echo "break-even 256 rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/break_even_*.json')):
    d=json.load(open(f)); print(f, d['break_even_n'])
    for r in d['rows']: print('  %6d sync=%8.1f pipe=%8.1f cpuB_best=%8.1f cpuA=%8.2f'%(r['n'],r['gpu_sync_us'],r['gpu_pipe_us'],r['cpu_b_best_us'],r['cpu_a_1core_us']))
PY

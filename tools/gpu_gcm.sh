#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gcm.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python tools/bench_gcm.py > gpurun_out/bench_gcm.json 2> gpurun_out/bench_gcm.err; echo "bench gcm rc=$?"; tail -3 gpurun_out/bench_gcm.err
timeout 600 python tools/bench_gcm.py > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gcm_ -s 40 -c 3 -o gpurun_out/prof_gcm python tools/bench_gcm.py > gpurun_out/ncu_gcm.log 2>&1; echo "ncu gcm rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_gcm.json'))
for r in d['rows']: print('%-16s e2e=%.3f ms (pinned %.3f, unchunked %.3f) kernel=%.3f ms (%.1f GB/s) | cpu ref all=%.3f 1core=%.3f reuse all=%.3f'%(r['workload'],r['gpu_e2e_ms'],r['gpu_e2e_pinned_ms'],r['gpu_e2e_pinned_serial_ms'],r['gpu_kernel_ms'],r['gpu_kernel_gbs'],r['cpu_ref_all_ms'],r['cpu_ref_1core_ms'],r['cpu_reuse_all_ms']))
"
python -c "
import json
for r in json.load(open('gpurun_out/bench_gcm.json'))['sweep']: print('%6d x %6d B  thread %.3f  warp %.3f  default %.3f ms'%(r['messages'],r['bytes'],r['thread_ms'],r['warp_ms'],r['default_ms']))
"

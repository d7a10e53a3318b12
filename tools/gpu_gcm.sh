#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gcm.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python tools/bench_gcm.py > gpurun_out/bench_gcm.json 2> gpurun_out/bench_gcm.err; echo "bench gcm rc=$?"; tail -3 gpurun_out/bench_gcm.err
true &&
echo "ncu: use tools/prof_gcm.py for captures"
python -c "
import json
d=json.load(open('gpurun_out/bench_gcm.json'))
for r in d['rows']: print('%-16s e2e=%.3f ms (pinned: default %.3f%s, copies %.3f, unchunked %.3f, in place %.3f) kernel=%.3f ms (%.1f GB/s) | cpu ref all=%.3f 1core=%.3f reuse all=%.3f'%(r['workload'],r['gpu_e2e_ms'],r['gpu_e2e_pinned_ms'],'*' if r['default_is_in_place'] else '',r['gpu_e2e_pinned_copies_ms'],r['gpu_e2e_pinned_serial_ms'],r['gpu_e2e_inplace_ms'],r['gpu_kernel_ms'],r['gpu_kernel_gbs'],r['cpu_ref_all_ms'],r['cpu_ref_1core_ms'],r['cpu_reuse_all_ms']))
"
python -c "
import json
for r in json.load(open('gpurun_out/bench_gcm.json'))['sweep']: print('%6d x %6d B  thread %.3f  warp %.3f  default %.3f ms'%(r['messages'],r['bytes'],r['thread_ms'],r['warp_ms'],r['default_ms']))
"

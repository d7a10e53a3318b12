#!/bin/bash
# AES-GCM: GPU tests, then the bench with the lane-group kernel (default) and with one thread per message
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gcm.py tests/test_host_service.py -m gpu -x -q 2>&1 | tail -3
for g in 1 0; do
  PK_GCM_GROUP=$g timeout 900 python tools/bench_gcm.py > gpurun_out/bench_gcm_g$g.json 2> gpurun_out/bench_gcm_g$g.err; echo "bench gcm group=$g rc=$?"
  python -c "
import json
d=json.load(open('gpurun_out/bench_gcm_g$g.json'))
for r in d['rows']: print('group=$g %-16s e2e pinned %.3f ms  kernel=%.3f ms (%.1f GB/s) | cpu ref all=%.3f reuse all=%.3f'%(r['workload'],r['gpu_e2e_pinned_ms'],r['gpu_kernel_ms'],r['gpu_kernel_gbs'],r['cpu_ref_all_ms'],r['cpu_reuse_all_ms']))
"
done

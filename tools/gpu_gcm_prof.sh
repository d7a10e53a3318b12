#!/bin/bash
# AES-GCM kernels: parity tests, kernel times at a few shapes, one ncu capture each of the
# lane-group kernel (65,536 x 256 B) and the warp-per-message kernel (32,768 x 4 KiB).
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gcm.py tests/test_host_service.py -m gpu -x -q 2>&1 | tail -2
for shape in "65536 256 8" "65536 64 4" "4096 1024 4" "1024 64 4" "32768 4096 4"; do
  echo "== $shape"; python tools/prof_gcm.py $shape | sort | head -3
done
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gcm_group_kernel -s 3 -c 1 \
  -o gpurun_out/prof_gcm_group -f python tools/prof_gcm.py 65536 256 6 > gpurun_out/prof_gcm_group.log 2>&1; echo "ncu group rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gcm_long_kernel -s 2 -c 1 \
  -o gpurun_out/prof_gcm_long -f python tools/prof_gcm.py 32768 4096 4 > gpurun_out/prof_gcm_long.log 2>&1; echo "ncu long rc=$?"

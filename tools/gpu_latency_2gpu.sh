#!/bin/bash
# The batching front-end sharding requests over 2 GPUs (run under gpurun --gpus 2): closed loop
# with two backends, open loop with PK_BACKENDS=2.  Output: gpurun_out/latency_2gpu.jsonl
set -u
mkdir -p gpurun_out build
H=polykey_service_b200/host
g++ -std=c++20 -O2 -pthread -o build/latency_bench $H/latency_bench.cpp $H/batching_service.cpp $H/service.cpp \
    -Lpolykey_service_b200 -lpkroute -Wl,-rpath,$PWD/polykey_service_b200 || exit 1
: > gpurun_out/latency_2gpu.jsonl
for cfg in "16 3 4096 200 2048 2" "128 3 4096 200 2048 2" "512 3 4096 200 2048 2"; do
  timeout 120 ./build/latency_bench $cfg >> gpurun_out/latency_2gpu.jsonl 2>> gpurun_out/latency.err || echo "fail: $cfg"
done
for cfg in "100000 3 4096 200 2048 1" "1000000 3 4096 200 2048 2" "2000000 3 4096 200 2048 4" "3000000 3 4096 200 2048 4" "4000000 3 4096 200 2048 6"; do
  PK_BACKENDS=2 timeout 120 ./build/latency_bench open $cfg >> gpurun_out/latency_2gpu.jsonl 2>> gpurun_out/latency.err || echo "fail: open $cfg"
done
cat gpurun_out/latency_2gpu.jsonl

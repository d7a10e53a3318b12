#!/bin/bash
# Per-request latency through the C++ batching front-end on one GPU: closed loop (K blocking
# callers) and open loop (Poisson arrivals at a fixed rate), work-conserving flush (the default)
# and the hold-the-batch-open policy (trailing 0).  Output: gpurun_out/latency{,_open}.jsonl
set -u
mkdir -p gpurun_out build
H=polykey_service_b200/host
g++ -std=c++20 -O2 -pthread -o build/latency_bench $H/latency_bench.cpp $H/batching_service.cpp $H/service.cpp \
    -Lpolykey_service_b200 -lpkroute -Wl,-rpath,$PWD/polykey_service_b200 || exit 1
: > gpurun_out/latency.jsonl
# callers seconds max_batch max_wait_us payload_bytes [backends idle_flush_us eager]
for cfg in "1 2 4096 200 2048" "16 3 4096 200 2048" "128 3 4096 200 2048" "512 3 4096 200 2048" "1024 3 4096 200 256" \
           "1 2 4096 200 2048 1 20 0" "16 3 4096 200 2048 1 20 0" "128 3 4096 200 2048 1 20 0" "512 3 4096 200 2048 1 20 0"; do
  timeout 120 ./build/latency_bench $cfg >> gpurun_out/latency.jsonl 2>> gpurun_out/latency.err || echo "fail: $cfg"
done
: > gpurun_out/latency_open.jsonl
# requests_per_s seconds max_batch max_wait_us payload_bytes [generators idle_flush_us eager]
for cfg in "1000 2 4096 200 2048 1" "10000 2 4096 200 2048 1" "100000 3 4096 200 2048 1" "300000 3 4096 200 2048 1" \
           "1000000 3 4096 200 2048 2" "2000000 3 4096 200 2048 4" "3000000 3 4096 200 2048 4" "4000000 3 4096 200 256 4" \
           "10000 2 4096 200 2048 1 20 0" "100000 3 4096 200 2048 1 20 0" "1000000 3 4096 200 2048 2 20 0" "2000000 3 4096 200 2048 4 20 0"; do
  timeout 120 ./build/latency_bench open $cfg >> gpurun_out/latency_open.jsonl 2>> gpurun_out/latency.err || echo "fail: open $cfg"
done
cat gpurun_out/latency.jsonl gpurun_out/latency_open.jsonl

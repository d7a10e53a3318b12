#!/bin/bash
set -u
mkdir -p gpurun_out
: > gpurun_out/latency.jsonl
for cfg in "1 2 4096 200 2048" "16 3 4096 200 2048" "128 3 4096 200 2048" "512 3 4096 200 2048" "512 3 4096 50 2048" "512 3 256 200 2048" "1024 3 4096 200 256"; do
  timeout 120 ./build/latency_bench $cfg >> gpurun_out/latency.jsonl 2>> gpurun_out/latency.err || echo "fail: $cfg"
done
cat gpurun_out/latency.jsonl

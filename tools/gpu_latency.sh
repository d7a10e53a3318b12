#!/bin/bash
set -u
mkdir -p gpurun_out build
# the front-end binary: C++ host mirror + libpkroute.so
H=polykey_service_b200/host
g++ -std=c++20 -O2 -pthread -o build/latency_bench $H/latency_bench.cpp $H/batching_service.cpp $H/service.cpp \
    -Lpolykey_service_b200 -lpkroute -Wl,-rpath,$PWD/polykey_service_b200 || exit 1
: > gpurun_out/latency.jsonl
for cfg in "1 2 4096 200 2048" "16 3 4096 200 2048" "128 3 4096 200 2048" "512 3 4096 200 2048" "512 3 4096 50 2048" "512 3 256 200 2048" "1024 3 4096 200 256"; do
  timeout 120 ./build/latency_bench $cfg >> gpurun_out/latency.jsonl 2>> gpurun_out/latency.err || echo "fail: $cfg"
done
cat gpurun_out/latency.jsonl

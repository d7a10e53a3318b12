#!/bin/bash
# The batching front-end on the GPU box: host tests through CudaBackend, then open-loop sweeps
# (rate x shards per GPU x generators).  Output: gpurun_out/frontend.jsonl
set -u
mkdir -p gpurun_out build
python -c "from polykey_service_b200 import build; build.build_latency_bench(force=True)" || exit 1
timeout 600 python -m pytest tests/test_host_service.py -m gpu -x -q 2>&1 | tail -3
: > gpurun_out/frontend.jsonl
for k in 1 2 4; do
  for cfg in "1000000 12" "5000000 12" "10000000 12" "20000000 12" "20000000 8"; do
    set -- $cfg
    PK_SHARDS_PER_GPU=$k timeout 120 ./build/latency_bench open $1 2 4096 200 2048 $2 >> gpurun_out/frontend.jsonl 2>> gpurun_out/frontend.err || echo "fail $k $cfg"
  done
done
# closed loop for comparison with round 1
for cfg in "1 2 4096 200 2048" "16 2 4096 200 2048" "128 2 4096 200 2048" "1024 2 4096 200 2048"; do
  timeout 120 ./build/latency_bench $cfg >> gpurun_out/frontend.jsonl 2>> gpurun_out/frontend.err || echo "fail closed $cfg"
done
python - <<'PY'
import json
for l in open('gpurun_out/frontend.jsonl'):
    d=json.loads(l)
    if d.get('mode')=='open-loop':
        print('open shards/gpu %d gens %d offered %.3g completed %.3g p50 %.1f p99 %.1f mean_batch %.1f dropped %d'%(d['shards_per_gpu'],d['generators'],d['offered_per_s'],d['completed_per_s'],d['latency_us']['p50'],d['latency_us']['p99'],d['mean_batch'],d['dropped']))
    else:
        print('closed callers %d req/s %.3g p50 %.1f p99 %.1f mean_batch %.1f'%(d['callers'],d['requests_per_s'],d['latency_us']['p50'],d['latency_us']['p99'],d['mean_batch']))
PY

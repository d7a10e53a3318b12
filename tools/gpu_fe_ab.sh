set -u
python -c "from polykey_service_b200 import build; build.build_latency_bench()" || exit 1
nproc; uptime
for cfg in "5000000 12" "10000000 12" "20000000 12" "20000000 8" "20000000 14" "20000000 4"; do
  set -- $cfg
  timeout 120 ./build/latency_bench open $1 2 4096 200 2048 $2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('gens',d['generators'],'offered %.3g completed %.3g p50 %.1f p99 %.1f mean_batch %.1f'%(d['offered_per_s'],d['completed_per_s'],d['latency_us']['p50'],d['latency_us']['p99'],d['mean_batch']))"
done
echo "--- with an idle CUDA process alongside"
python -c "
import torch,time; torch.zeros(1,device='cuda'); time.sleep(25)" &
sleep 8
timeout 120 ./build/latency_bench open 20000000 2 4096 200 2048 12 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('gens',d['generators'],'offered %.3g completed %.3g p50 %.1f mean_batch %.1f'%(d['offered_per_s'],d['completed_per_s'],d['latency_us']['p50'],d['mean_batch']))"
wait

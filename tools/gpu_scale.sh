#!/bin/bash
# bench.py at N ranks through torchrun (the driver's launch), summary of the scaling-relevant keys
N=${1:-8}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo rc=$?
nproc; lscpu | grep -E "NUMA|Socket|Core" 
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n$N.json"))
e=d["e2e"]; p=e["pcie_probe"]
print("N",d["n_gpus"],"value %.4g frac %.3f steady %.3f"%(d["value"],d["roofline"]["frac"],d["steady_state"]["frac"]))
print("e2e %.4g ms/step %.4f host GB/s per gpu %.1f frac_of_pcie %.3f other_path %.4g"%(e["value"],e["ms_per_step"],e["host_gbs_per_gpu"],e["frac_of_pcie"],e["other_path"]["value"]))
print("probe memcpy %.1f kernel %.1f duplex %.1f GB/s per gpu"%(p["memcpy_h2d_gbs_per_gpu"],p["kernel_read_gbs_per_gpu"],p["memcpy_duplex_gbs_per_gpu_each_way"]))
c=d["c4"]; print("c4 value %.4g pass_ms %.4f e2e %.4g"%(c["value"],c["pass_ms"],c["e2e"]["value"]))
l=d["latency"]; print("latency", [(r["offered_per_s"],round(r.get("completed_per_s",0)),r.get("p50_us"),r.get("p99_us")) for r in l["rates"]])
PY

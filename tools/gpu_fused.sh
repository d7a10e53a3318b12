#!/bin/bash
# fused (multi-batch) launches: parity tests, the bench line at K=20 with one launch for the 20
# batches, and the ncu launch list of that same command (a profiler's own duration for the launch)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or dependent_launch or c1_full" 2>&1 | tail -3
for i in 1 2; do
python bench.py --issue fused --steps 20 --warmup 3 --no-cpu --lean --no-frontend > gpurun_out/bench_fused_k20.json 2> gpurun_out/bench_fused_k20.err; echo "bench fused rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_fused_k20.json"))
print("fused K=20: %.3f us/step, frac %.3f, launches %d, avg launch %.1f us for %.1f MB" % (
    d["ms_per_step"] * 1e3, d["roofline"]["frac"], d["gpu_launches"], d["roofline"]["avg_launch_us"], d["roofline"]["algorithmic_bytes_per_launch"] / 1e6))
print({k: round(v * 1e3, 3) for k, v in sorted(d["issue_modes_ms_per_step"].items(), key=lambda kv: kv[1])})
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c1_fused.csv \
  python bench.py --issue fused --steps 20 --warmup 3 --no-cpu --lean --no-frontend --no-prewarm > gpurun_out/ncu_fused.log 2>&1; echo "ncu rc=$?"
grep -c route_fused gpurun_out/launches_c1_fused.csv; grep route_fused gpurun_out/launches_c1_fused.csv | awk -F'","' '{print $(NF-0)}' | tr -d '"' | sort -n | tail -5

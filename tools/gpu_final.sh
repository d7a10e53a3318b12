#!/bin/bash
# what the driver does at round end, in one call: GPU tests, smoke, the bench line at K = 20
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
s=$(date +%s); python bench.py --steps 20 --warmup 3 > gpurun_out/bench_k20.json 2> gpurun_out/bench_k20.err; echo "bench rc=$? in $(( $(date +%s) - s )) s"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_k20.json"))
r = d["roofline"]
print("value %.4g frac %.3f traffic %s (%s) launches %d steady %.3f" % (d["value"], r["frac"], r["traffic"], r["traffic_source"][:50], d["gpu_launches"], d["steady_state"]["frac"]))
print("parity_gate:", d["parity_gate"][:80]); print("clocks", d["clocks"])
PY

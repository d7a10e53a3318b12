#!/bin/bash
# First GPU call: tests, smoke, bench lines, ncu launch list + full capture.
set -u
mkdir -p gpurun_out
{ nvidia-smi; nproc; go version 2>&1 | head -1; lscpu | head -20; } > gpurun_out/box.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py --steps 4000 --warmup 20 > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"
timeout 600 python bench.py --impl reference --steps 200 --warmup 5 > gpurun_out/bench_ref_c1.json 2> gpurun_out/bench_ref_c1.err; echo "bench ref rc=$?"
timeout 600 python bench.py --workload c2 --steps 1000 --warmup 10 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
timeout 900 python bench.py --workload c3 --steps 40 --warmup 5 --no-cpu > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"
# ncu: launch list of the bench command (after it exited 0 plain), then one full capture per workload
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu --no-prewarm > gpurun_out/plain_c1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c1.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu --no-prewarm > gpurun_out/ncu_launches_c1.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 30 -c 3 -o gpurun_out/prof_c1 \
    python bench.py --steps 20 --warmup 3 --no-cpu --no-prewarm > gpurun_out/ncu_full_c1.log 2>&1; echo "ncu full c1 rc=$?"
timeout 600 python bench.py --workload c3 --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/plain_c3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:route_kernel -s 5 -c 2 -o gpurun_out/prof_c3 \
    python bench.py --workload c3 --steps 6 --warmup 3 --no-cpu --no-prewarm > gpurun_out/ncu_full_c3.log 2>&1; echo "ncu full c3 rc=$?"
tail -5 gpurun_out/pytest_gpu.log
cat gpurun_out/bench_c1.json

#!/bin/bash
# final-state run of the round: all GPU tests, smoke, headline line (C4, with cpu_baseline), C5, C2, C3
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 200 python bench.py > gpurun_out/bench_default.log 2>&1; echo "c4 rc=$?"; tail -1 gpurun_out/bench_default.log | cut -c1-130
for w in c5 c2 c3 c1; do
  timeout 120 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"; tail -1 gpurun_out/bench_$w.log | cut -c1-130
done

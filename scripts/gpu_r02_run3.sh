#!/bin/bash
# round 2, third GPU run: megakernel back on the r02a schedule with fetch chunk 16, hierarchy kernel with S-local normal,
# |1/d| folded into the FFMAs, S threshold 12
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for w in c4 c5 c2 c3 c1; do
  timeout 200 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"; tail -1 gpurun_out/bench_$w.log | cut -c1-140
done

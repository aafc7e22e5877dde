#!/bin/bash
# round 2, first GPU run: all GPU tests on the refactored library (per-handle scene, split TUs, slimmer megakernel,
# LDG.256 hierarchy walk), bench lines, the LDG.128 A/B variant on C5, ncu captures of both dominant kernels
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 200 python bench.py --no-cpu-baseline > gpurun_out/bench_c4.log 2>&1; echo "c4 rc=$?"; tail -1 gpurun_out/bench_c4.log | cut -c1-140
for w in c5 c2 c3 c1; do
  timeout 120 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"; tail -1 gpurun_out/bench_$w.log | cut -c1-140
done
W=c5 bash scripts/gpu_variants.sh
bash scripts/gpu_ncu.sh r02a_c5 --workload c5
bash scripts/gpu_ncu.sh r02a_c4

#!/bin/bash
# hierarchy path: parity tests, C5 bench (+ optional C4 regression)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bvh.py -m gpu -x -q > gpurun_out/pytest_gpu_bvh.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu_bvh.log
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1; echo "c5 rc=$?"; tail -1 gpurun_out/bench_c5.log | cut -c1-1400
[ -n "$C4" ] && { timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.log 2>&1; echo "c4 rc=$?"; tail -1 gpurun_out/bench_c4.log | cut -c1-400; }
exit 0

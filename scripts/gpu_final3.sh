#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --workload c5 --steps 3 --warmup 3 > gpurun_out/bench_c5.log 2>&1; echo "c5 rc=$?"; tail -1 gpurun_out/bench_c5.log | cut -c1-200
timeout 600 python bench.py > gpurun_out/bench_default.log 2>&1; echo "c4 rc=$?"; tail -1 gpurun_out/bench_default.log | cut -c1-200

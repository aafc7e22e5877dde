#!/bin/bash
# tests + default bench (no ncu)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 ${BENCH_ARGS:---no-cpu-baseline} > gpurun_out/bench_c4.log 2>&1; echo "c4 rc=$?"; tail -1 gpurun_out/bench_c4.log | cut -c1-300

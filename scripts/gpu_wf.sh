#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bvh.py -m gpu -x -q -k wavefront > gpurun_out/pytest_wf.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_wf.log
timeout 900 python bench.py --workload c5 --variant wavefront --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_wf.log 2>&1; echo "c5 wf rc=$?"; tail -1 gpurun_out/bench_c5_wf.log | cut -c1-1400
timeout 900 python bench.py --workload c2 --variant wavefront --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_wf.log 2>&1; echo "c2 wf rc=$?"; tail -1 gpurun_out/bench_c2_wf.log | cut -c1-700
timeout 900 python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_mega.log 2>&1; echo "c2 mega rc=$?"; tail -1 gpurun_out/bench_c2_mega.log | cut -c1-700

#!/bin/bash
# first contact with the GPU: tests, smoke, small + full bench.  Everything under timeout.
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia-smi.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 300 python bench.py --workload tiny --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tiny.log 2>&1; echo "tiny rc=$?"; tail -2 gpurun_out/bench_tiny.log
timeout 300 python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1; echo "c2 rc=$?"; tail -2 gpurun_out/bench_c2.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c4.log 2>&1; echo "c4 rc=$?"; tail -2 gpurun_out/bench_c4.log

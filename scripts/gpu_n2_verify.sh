#!/bin/bash
# 2-GPU re-entry check: the single-process multi-GPU ABI test, both bench arms under torchrun (N=2)
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "multi_gpu" > gpurun_out/pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 200 $TR bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/ref_n2.log 2>&1; echo "ref rc=$?"; grep '^{' gpurun_out/ref_n2.log | cut -c1-160; grep -o '"cpu_baseline".*' gpurun_out/ref_n2.log | cut -c1-300
timeout 200 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.log 2>&1; echo "b200 rc=$?"; grep '^{' gpurun_out/bench_n2.log | cut -c1-200

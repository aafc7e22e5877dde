#!/bin/bash
# round-end style check: all GPU tests, smoke, default bench (with CPU baseline), reference arm, C5/C1/C3 lines
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_default.log | cut -c1-3500
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/bench_reference.log | cut -c1-900
for w in c5 c3 c1 c2 c2p c4p; do timeout 600 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"; tail -1 gpurun_out/bench_$w.log | cut -c1-250; done

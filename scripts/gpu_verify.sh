#!/bin/bash
# re-entry verification of freshly rebuilt artefacts: GPU tests, smoke, default bench line
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 200 python bench.py > gpurun_out/bench_default.log 2>&1; echo "c4 rc=$?"; tail -1 gpurun_out/bench_default.log | cut -c1-300

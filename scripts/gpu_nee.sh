#!/bin/bash
# next-event estimation: GPU tests of the row, equal-time comparison, then the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nee.py -q > gpurun_out/pytest_nee.log 2>&1; echo "nee rc=$?"; tail -25 gpurun_out/pytest_nee.log
timeout 600 python scripts/nee_compare.py > gpurun_out/nee_compare.log 2>&1; echo "compare rc=$?"; tail -12 gpurun_out/nee_compare.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log

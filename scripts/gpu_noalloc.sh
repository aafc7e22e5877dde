#!/bin/bash
# The no-alloc entry point's test: alone in a fresh process, then behind the whole GPU suite.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_gpu_all.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/pytest_gpu_all.log

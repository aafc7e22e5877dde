#!/bin/bash
# PlaneObject extension: GPU tests of the new row, then the whole GPU suite, then the plane benches + the C4 regression check
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_planes.py -q > gpurun_out/pytest_planes.log 2>&1; echo "planes rc=$?"; tail -15 gpurun_out/pytest_planes.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for w in c2p c4p c2 c4; do
  timeout 600 python bench.py --steps 3 --warmup 3 --workload $w $( [ $w = c2p ] || echo --no-cpu-baseline ) > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"; tail -1 gpurun_out/bench_$w.log | cut -c1-400
done

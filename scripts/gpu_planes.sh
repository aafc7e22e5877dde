#!/bin/bash
# PlaneObject extension: GPU tests of the row, the plane benches, C5 regression check, optional ncu evidence (c2p)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_planes.py -q > gpurun_out/pytest_planes.log 2>&1; echo "planes rc=$?"; tail -5 gpurun_out/pytest_planes.log
for w in c2p c4p c5; do
  timeout 600 python bench.py --steps 3 --warmup 3 --workload $w $( [ $w = c2p ] || echo --no-cpu-baseline ) > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"; tail -1 gpurun_out/bench_$w.log | cut -c1-200
done
if [ "$1" = ncu ]; then bash scripts/gpu_ncu.sh planes_c2p --workload c2p; fi

#!/bin/bash
# rebuilds kernels_radiance.cu on the box with different tuning defines and runs the C5 bench
mkdir -p gpurun_out
for cfg in "$@"; do
  flags=$(echo "$cfg" | sed 's/,/ /g')
  touch raytracingmin_b200/csrc/kernels_radiance.cu
  make -C raytracingmin_b200/csrc EXTRA="$flags" > gpurun_out/build_sweep.log 2>&1 || { echo "build failed $cfg"; tail -5 gpurun_out/build_sweep.log; continue; }
  timeout 600 python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/sweep.log 2>&1
  echo "$cfg: $(tail -1 gpurun_out/sweep.log | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"],1), "Mrays/s", d["config"]["closest_hit"])')"
done

#!/bin/bash
# ncu evidence: launch list of the whole bench run + one full capture of the dominant kernel.
# Each ncu pass only after the same command exited 0 without ncu.  usage: gpu_ncu.sh <tag> [bench args]
TAG=${1:-run}; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline $*"
timeout 600 $CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
tail -1 gpurun_out/${TAG}_plain.log | cut -c1-200
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"radiance_(mega|bvh)" -s 3 -c 1 \
    -o gpurun_out/${TAG}_mega_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"

#!/bin/bash
# ncu evidence for the megakernel-vs-wavefront comparison and the tuned hierarchy kernel
mkdir -p gpurun_out
C2="python bench.py --workload c2 --variant wavefront --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $C2 > gpurun_out/wf_c2_plain.log 2>&1 || { echo plain failed; tail -3 gpurun_out/wf_c2_plain.log; }
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:wf_ -s 600 -c 2 -o gpurun_out/wf_c2_full $C2 > gpurun_out/wf_c2_ncu.log 2>&1; echo "wf c2 rc=$?"
C5W="python bench.py --workload c5 --variant wavefront --steps 1 --warmup 3 --no-cpu-baseline"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:wf_ -s 2000 -c 2 -o gpurun_out/wf_c5_full $C5W > gpurun_out/wf_c5_ncu.log 2>&1; echo "wf c5 rc=$?"
C2M="python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:radiance_mega -s 2 -c 1 -o gpurun_out/mega_c2_full $C2M > gpurun_out/mega_c2_ncu.log 2>&1; echo "mega c2 rc=$?"
C5="python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 $C5 > gpurun_out/bvh3_c5_plain.log 2>&1; tail -1 gpurun_out/bvh3_c5_plain.log | cut -c1-200
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:radiance_bvh -s 3 -c 1 -o gpurun_out/bvh3_c5_full $C5 > gpurun_out/bvh3_c5_ncu.log 2>&1; echo "bvh3 c5 rc=$?"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/bvh3_c5_launches.csv $C5 > gpurun_out/bvh3_c5_ncu_launches.log 2>&1; echo "launch list rc=$?"

#!/bin/bash
# round 2, second GPU run: compact 32-byte hierarchy nodes (A/B against the float nodes through --knob), guided
# chunking + RNG prep in the megakernel, the reference-object band tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 200 python bench.py --no-cpu-baseline > gpurun_out/bench_c4.log 2>&1; echo "c4 rc=$?"; tail -1 gpurun_out/bench_c4.log | cut -c1-140
timeout 120 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1; echo "c5 rc=$?"; tail -1 gpurun_out/bench_c5.log | cut -c1-140
timeout 120 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline --knob debug_bvh_walk=2 > gpurun_out/bench_c5_floatnodes.log 2>&1; echo "c5 float nodes rc=$?"; tail -1 gpurun_out/bench_c5_floatnodes.log | cut -c1-140
for w in c2 c3 c1; do
  timeout 120 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"; tail -1 gpurun_out/bench_$w.log | cut -c1-140
done
bash scripts/gpu_ncu.sh r02b_c5 --workload c5

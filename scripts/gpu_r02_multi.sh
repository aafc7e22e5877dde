#!/bin/bash
# multi-GPU evidence (run under `gpurun --gpus N`): the single-process multi-GPU ABI test (rtm_options.n_gpus), then the
# one-rank-per-GPU bench at the requested rank counts.  usage: gpu_r02_multi.sh "2 4 8" [workloads]
NS=${1:-2}; WL=${2:-c4}
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/multi_gpus.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -q -k "single_process_multi_gpu or subpixel_grid_slices or sample_slices" -rs > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -4 gpurun_out/pytest_multi.log
for w in $WL; do
  timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/scale_${w}_n1.log 2>&1; echo "$w n1 rc=$?"; tail -1 gpurun_out/scale_${w}_n1.log | cut -c1-150
  for n in $NS; do
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --workload $w --steps 5 --warmup 3 > gpurun_out/scale_${w}_n$n.log 2>&1
    echo "$w n$n rc=$?"; grep '^{' gpurun_out/scale_${w}_n$n.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('  value %.1f Grays/s  ms %.3f  e2e %.1f Grays/s  image_check %s' % (d['value']/1e3, d['ms_per_step'], d['e2e']['value']/1e3, d.get('image_check')))" 2>/dev/null || tail -3 gpurun_out/scale_${w}_n$n.log
  done
done

#!/bin/bash
# multi-GPU bench: torchrun, one rank per GPU.  usage: gpu_scale.sh N [N2 ...]
mkdir -p gpurun_out
for N in "$@"; do
  if [ "$N" = "1" ]; then
    timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$N.log 2>&1
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/scale_$N.log 2>&1
  fi
  echo "N=$N rc=$?"; tail -1 gpurun_out/scale_$N.log | cut -c1-330
  grep -o '"e2e": {[^}]*}' gpurun_out/scale_$N.log | tail -1
done

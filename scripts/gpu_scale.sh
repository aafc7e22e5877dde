#!/bin/bash
# multi-GPU bench: torchrun, one rank per GPU.  usage: gpu_scale.sh N [extra bench args]
mkdir -p gpurun_out
N=$1; shift
tag=$(echo "$*" | tr -c 'a-zA-Z0-9' '_')
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 3 --warmup 3 "$@" > gpurun_out/scale_${N}${tag}.log 2>&1
echo "N=$N $* rc=$?"; grep '^{' gpurun_out/scale_${N}${tag}.log | tail -1 | python3 -c 'import sys,json; d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ("value","ms_per_step","e2e","reduce_check_max_abs_diff_vs_nccl")}, d["config"]["partition"])' || tail -20 gpurun_out/scale_${N}${tag}.log

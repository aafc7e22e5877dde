#!/bin/bash
mkdir -p gpurun_out
for w in c4 c5; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 8 --steps 3 --warmup 3 --workload $w > gpurun_out/scale8_$w.log 2>&1
echo "$w N=8 rc=$?"; grep '^{' gpurun_out/scale8_$w.log | tail -1 | cut -c1-400
done

#!/bin/bash
# SAH constants sweep for the sphere hierarchy (C5): traversal cost x largest leaf
mkdir -p gpurun_out
IFS=";" read -ra CFGS <<< "${SWEEP:-1 4;0.5 4;0.25 4;1 2;1 1;0.5 1}"
for cfg in "${CFGS[@]}"; do
  set -- $cfg
  RTM_BVH_CTRAV_ENV=$1 RTM_BVH_LEAFMAX_ENV=$2 timeout 300 python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/sweep_bvh_$1_$2.log 2>&1
  python - "$1" "$2" <<'PY'
import json,sys
l=[x for x in open(f"gpurun_out/sweep_bvh_{sys.argv[1]}_{sys.argv[2]}.log") if x.startswith("{")]
if not l: print(sys.argv[1:], "failed"); sys.exit()
d=json.loads(l[-1]); print("ctrav", sys.argv[1], "leafmax", sys.argv[2], "Grays/s %.3f" % (d["value"]/1e3), "ms %.1f" % d["ms_per_step"], d["config"]["closest_hit"], "rps %.4f" % d["rays_per_sample"])
PY
done

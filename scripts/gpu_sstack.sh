#!/bin/bash
# shared-memory traversal stack (radiance_bvh_kernel<..., SST=true>) against the local-memory stack, same box
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_bvh.py -m gpu -q > gpurun_out/pytest_bvh_sstack.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_bvh_sstack.log
for off in 1 0 1 0; do
  RTM_BVH_SSTACK_OFF=$off timeout 120 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_sstack_off$off.log 2>&1
  python - $off <<'PY'
import json,sys
l=[x for x in open(f"gpurun_out/bench_c5_sstack_off{sys.argv[1]}.log") if x.startswith("{")]
if not l: print("off", sys.argv[1], "failed"); sys.exit()
d=json.loads(l[-1]); print("SSTACK_OFF", sys.argv[1], "Grays/s %.3f" % (d["value"]/1e3), "ms %.1f" % d["ms_per_step"], "rps %.4f" % d["rays_per_sample"])
PY
done

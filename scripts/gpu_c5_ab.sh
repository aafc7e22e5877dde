#!/bin/bash
# C5 A/B: the default library (with optional knobs), then the prebuilt variants_*.so (scripts/gpu_variants3.sh), then the default again
mkdir -p gpurun_out
run() { # name, args
  timeout 300 python bench.py --workload c5 --steps ${STEPS:-4} --warmup 3 --no-cpu-baseline $2 > gpurun_out/ab_$1.log 2>&1
  python - $1 <<'PY'
import json,sys
l=[x for x in open(f"gpurun_out/ab_{sys.argv[1]}.log") if x.startswith("{")]
if not l: print(sys.argv[1], "failed"); sys.exit()
d=json.loads(l[-1]); print(sys.argv[1], "Grays/s %.3f" % (d["value"]/1e3), "ms %.2f" % d["ms_per_step"], "rps %.4f" % d["rays_per_sample"], "clk", d["clocks"]["sm_mhz"])
PY
}
run default ""
W=c5 STEPS=4 bash scripts/gpu_variants3.sh
run default2 ""
for k in $KNOBS; do run knob$k "--knob debug_bvh_walk=$k"; done

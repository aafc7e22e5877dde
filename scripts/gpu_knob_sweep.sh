#!/bin/bash
# SAH constants of the hierarchy build through rtm_options.debug_* (bench.py --knob): node-visit cost x largest leaf, on C5
mkdir -p gpurun_out
for lm in 4 6 8; do for ct in 0.5 1 1.5 2 3; do
  timeout 120 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline --knob debug_bvh_leaf_max=$lm --knob debug_bvh_ctrav=$ct > gpurun_out/knob_lm${lm}_ct${ct}.log 2>&1
  grep '^{' gpurun_out/knob_lm${lm}_ct${ct}.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('leaf_max $lm ctrav $ct: %.3f Grays/s  %.2f ms  %s' % (d['value']/1e3, d['ms_per_step'], d['config']['closest_hit']))"
done; done

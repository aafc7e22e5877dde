#!/bin/bash
# A/B of prebuilt library variants (variants_<name>.so in the repo root) on several workloads: W="c4 c2 c3"
mkdir -p gpurun_out
cp raytracingmin_b200/librtm_b200.so /tmp/librtm_keep.so
for f in ${PAT:-variants_*.so}; do
  name=${f#variants_}; name=${name%.so}
  cp $f raytracingmin_b200/librtm_b200.so
  for w in ${W:-c4}; do
    timeout 300 python bench.py --workload $w --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline > gpurun_out/variant_${name}_$w.log 2>&1
    python - $name $w <<'PY'
import json,sys
l=[x for x in open(f"gpurun_out/variant_{sys.argv[1]}_{sys.argv[2]}.log") if x.startswith("{")]
if not l: print(sys.argv[1], sys.argv[2], "failed"); sys.exit()
d=json.loads(l[-1]); print(sys.argv[1], sys.argv[2], "Grays/s %.3f" % (d["value"]/1e3), "ms %.2f" % d["ms_per_step"], "rps %.4f" % d["rays_per_sample"], "clk", d["clocks"]["sm_mhz"])
PY
  done
done
cp /tmp/librtm_keep.so raytracingmin_b200/librtm_b200.so

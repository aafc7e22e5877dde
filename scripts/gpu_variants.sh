#!/bin/bash
# tuning sweep: prebuilt library variants (variants_<name>.so in the repo root) swapped in for the C5 bench
mkdir -p gpurun_out
cp raytracingmin_b200/librtm_b200.so /tmp/librtm_keep.so
for f in variants_*.so; do
  name=${f#variants_}; name=${name%.so}
  cp $f raytracingmin_b200/librtm_b200.so
  timeout 300 python bench.py --workload ${W:-c5} --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/variant_$name.log 2>&1
  python - $name <<'PY'
import json,sys
l=[x for x in open(f"gpurun_out/variant_{sys.argv[1]}.log") if x.startswith("{")]
if not l: print(sys.argv[1], "failed"); sys.exit()
d=json.loads(l[-1]); print(sys.argv[1], "Grays/s %.3f" % (d["value"]/1e3), "ms %.1f" % d["ms_per_step"], "rps %.4f" % d["rays_per_sample"])
PY
done
cp /tmp/librtm_keep.so raytracingmin_b200/librtm_b200.so

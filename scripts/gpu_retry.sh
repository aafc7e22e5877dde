#!/bin/bash
# megakernel roulette rounds: parity tests + all brute-force workloads
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for w in c1 c3 c2 c4 c2p; do
  timeout 600 python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"
  python - $w <<'PY'
import json,sys
l=[x for x in open(f"gpurun_out/bench_{sys.argv[1]}.log") if x.startswith("{")]
d=json.loads(l[-1]); print(sys.argv[1], "Grays/s %.2f" % (d["value"]/1e3), "ms %.3f" % d["ms_per_step"], "Gsamples/s %.2f" % (d["msamples_per_s"]/1e3), {k:round(v,3) for k,v in d["kernel_ms"].items()}, "frac %.3f" % d["roofline"]["frac"])
PY
done

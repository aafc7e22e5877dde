#!/bin/bash
# The reference's CLI AS IS (its own main.cpp / Render(): shared racing mt19937, per-row OpenMP fork/join, file writes)
# on the box's host cores, C1 and C2 at full size (SURVEY.md §8d "CPU path timed beside it", variant 1).
# ref_abc_cli / ref_head_cli are built from /root/reference/src by oracle/Makefile (prebuilt here; the box has no reference tree).
mkdir -p gpurun_out
ROOT=$(pwd)
echo "{\"host_cores\": $(nproc)}" > gpurun_out/cpu_cli_baseline.jsonl
for cfg in "simple1 abc 48384000" "cornellBox abc 193536000" "cornellBox head 193536000"; do
  set -- $cfg
  tmp=$(mktemp -d); cd $tmp
  t0=$(date +%s.%N)
  timeout 1200 $ROOT/oracle/_ref/ref_$2_cli -json $ROOT/scenes/$1.head.json > cli.log 2>&1; rc=$?
  t1=$(date +%s.%N)
  python3 - "$1" "$2" "$3" "$t0" "$t1" "$rc" <<'PY' >> $ROOT/gpurun_out/cpu_cli_baseline.jsonl
import sys, json, os
name, level, samples, t0, t1, rc = sys.argv[1:]
dt = float(t1) - float(t0)
print(json.dumps({"scene": name, "level": level, "samples": int(samples), "wall_s": round(dt, 2), "rc": int(rc),
                  "msamples_per_s": round(int(samples) / dt / 1e6, 3), "wrote_bmp": os.path.exists("result.bmp")}))
PY
  cd $ROOT; rm -rf $tmp
done
cat gpurun_out/cpu_cli_baseline.jsonl

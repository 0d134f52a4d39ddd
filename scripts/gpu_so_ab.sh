#!/bin/bash
# A/B of two builds (default .so vs LB200_SO=<alt>) on a list of workloads: gpu_so_ab.sh <tag> <alt.so> <workload>...
T=$1; ALT=$2; shift 2
mkdir -p gpurun_out
for SO in "" "$ALT"; do
  N=default; [ -n "$SO" ] && N=alt
  for W in "$@"; do
    LB200_SO=$SO timeout 600 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_${N}_$W.json 2> gpurun_out/${T}_${N}_$W.err
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_${N}_$W.json").read().strip().splitlines()[-1])
    print("$N $W", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"], 3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"], 3), "frac", round(d["roofline"]["frac"], 3), "crc", d.get("result_crc32"))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/${T}_${N}_$W.err").read()[-1500:])
PY
  done
done

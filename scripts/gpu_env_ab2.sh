#!/bin/bash
# env-switch A/B on one GPU: gpu_env_ab2.sh <tag> "<ENV=.. ENV=..>" <workload>...   (first a run with the defaults)
T=$1; E=$2; shift 2
mkdir -p gpurun_out
for V in default alt; do
  for W in "$@"; do
    if [ $V = alt ]; then X="env $E"; else X=""; fi
    $X timeout 300 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_${V}_$W.json 2> gpurun_out/${T}_${V}_$W.err
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_${V}_$W.json").read().strip().splitlines()[-1])
    print("$V $W", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"], 3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"], 3), "frac", round(d["roofline"]["frac"], 3), "crc", d.get("result_crc32"))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/${T}_${V}_$W.err").read()[-800:])
PY
  done
done

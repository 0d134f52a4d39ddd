#!/bin/bash
# seed rows A/B: gpu_seedrows.sh <tag> <workload:rows>...
T=$1; shift
mkdir -p gpurun_out
for X in "$@"; do
  W=${X%%:*}; R=${X##*:}
  LB200_SEED_ROWS=$R timeout 200 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_${W}_$R.json 2> gpurun_out/${T}_${W}_$R.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_${W}_$R.json").read().strip().splitlines()[-1])
    print("$W rows $R", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"], 3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"], 3), "crc", d.get("result_crc32"))
except Exception as e:
    print("parse fail", e)
PY
done

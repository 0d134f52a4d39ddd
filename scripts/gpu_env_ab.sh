#!/bin/bash
# same-box A/B of environment switches on one workload: gpu_env_ab.sh <tag> <workload> "<ENV=..>" ["<ENV=..>" ...]   ("-" = defaults)
T=$1; W=$2; shift 2
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  i=$((i+1))
  if [ "$envs" = "-" ]; then E=""; else E="$envs"; fi
  env $E timeout 900 python bench.py --workload $W --steps 10 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_${W}_$i.json 2> gpurun_out/${T}_${W}_$i.err || tail -3 gpurun_out/${T}_${W}_$i.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_${W}_$i.json").read().strip().splitlines()[-1])
print("[$envs]", round(d["value"]), "q/s ms/step", round(d["ms_per_step"],3), "frac", round(d["roofline"]["frac"],3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "crc", d.get("result_crc32"))
PY
done

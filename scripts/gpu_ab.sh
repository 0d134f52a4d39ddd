#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; tail -2 gpurun_out/tests.log
run() { # name, env..., workload
  local name=$1; shift; local w=$1; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_$name.json").read().strip().splitlines()[-1])
    print("$name", round(j["value"]), "q/s  scan ms", round(j["roofline"]["scan_ms_per_launch"],2), "GB/s", round(j["roofline"]["achieved"]), "frac", round(j["roofline"]["frac"],3))
except Exception as e:
    print("$name parse fail", e); print(open("gpurun_out/ab_$name.err").read()[-800:])
PY
}
run cfg4_L32 cfg4 X=1
run cfg4_L16 cfg4 LB200_SKEW_L=16
run cfg4_L8 cfg4 LB200_SKEW_L=8
run cfg1_L8 cfg1 X=1
run cfg1_L4 cfg1 LB200_SKEW_L=4
run cfg1_L16 cfg1 LB200_SKEW_L=16

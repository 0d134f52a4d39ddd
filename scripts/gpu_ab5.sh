#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; tail -3 gpurun_out/tests.log
run() { local name=$1; shift; local args=$1; shift
  env "$@" timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline $args > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_$name.json").read().strip().splitlines()[-1])
    print("$name", round(j["value"]), "q/s  ms/step", round(j["ms_per_step"],2), " scan ms", round(j["roofline"]["scan_ms_per_launch"],2), "GB/s", round(j["roofline"]["achieved"]), "frac", round(j["roofline"]["frac"],3))
except Exception as e:
    print("$name parse fail", e); print(open("gpurun_out/ab_$name.err").read()[-800:])
PY
}
run cfg2 "--workload cfg2" X=1
run s1 "--workload cfg4" X=1

run cfg1 "--workload cfg1" X=1



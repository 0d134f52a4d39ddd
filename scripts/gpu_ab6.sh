#!/bin/bash
set -u
mkdir -p gpurun_out
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
nvidia-smi --query-gpu=name,clocks.sm,clocks.mem,power.limit,temperature.gpu --format=csv
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; tail -3 gpurun_out/tests.log
run b_955c595 "--workload cfg4" LB200_SO=/root/repo/lance_b200/liblance_b200_955c595.so
run head "--workload cfg4" X=1
run head_s8 "--workload cfg4 --shard-of 8" X=1
run head_cfg1 "--workload cfg1" X=1
run head_cfg5 "--workload cfg5" X=1
run head_cfg3 "--workload cfg3" X=1

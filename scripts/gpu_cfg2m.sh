#!/bin/bash
# flat kNN column, all three metrics: tests + bench lines.  usage: gpu_cfg2m.sh <tag> [ENV=..]
T=${1:-r2ax}; shift
for kv in "$@"; do export "$kv"; done
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_column.py -x -q > gpurun_out/${T}_tests_column.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests_column.log; tail -3 gpurun_out/${T}_tests_column.log
for m in l2 cosine dot; do
  timeout 600 python bench.py --workload cfg2 --flat-metric $m --steps 20 --warmup 3 > gpurun_out/${T}_bench_cfg2_$m.json 2> gpurun_out/${T}_bench_cfg2_$m.err; tail -2 gpurun_out/${T}_bench_cfg2_$m.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg2_$m.json").read().strip().splitlines()[-1])
print("$m", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), d["config"]["filter"], "exact ms", round(d["config"]["exact_path_ms_per_step"],2), "parity", d.get("parity",{}).get("unexplained_mismatches"), "cpu", round(d.get("cpu_baseline",{}).get("value",0)))
PY
done

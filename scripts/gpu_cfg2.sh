#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2s}
timeout 600 python bench.py --workload cfg2 --steps 20 --warmup 3 > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err; tail -3 gpurun_out/${T}_bench_cfg2.err
python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg2.json").read().strip().splitlines()[-1])
print(round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), d["config"]["filter"], "exact ms", round(d["config"]["exact_path_ms_per_step"],2), "parity", d.get("parity",{}).get("unexplained_mismatches"), "cpu", d.get("cpu_baseline",{}).get("value"))
PY
timeout 600 python bench.py --workload cfg2 --steps 20 --warmup 3 --queries 1024 --no-cpu-baseline > gpurun_out/${T}_bench_cfg2_q1024.json 2>/dev/null
python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg2_q1024.json").read().strip().splitlines()[-1])
print("Q=1024:", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), d["config"]["filter"])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${T}_launches_cfg2.csv python bench.py --workload cfg2 --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1; echo "launch list exit $?"

#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; tail -3 gpurun_out/tests.log
for w in ${@:-cfg4}; do
timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/q_$w.json 2> gpurun_out/q_$w.err
python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/q_$w.json").read().strip().splitlines()[-1])
    print("$w", round(j["value"]), "q/s  scan ms", round(j["roofline"]["scan_ms_per_launch"],2), "GB/s", round(j["roofline"]["achieved"]), "frac", round(j["roofline"]["frac"],3), "ms/step", round(j["ms_per_step"],2))
except Exception as e:
    print("$w parse fail", e); print(open("gpurun_out/q_$w.err").read()[-800:])
PY
done

#!/bin/bash
T=${1:-r2u}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_column.py -x -q > gpurun_out/${T}_tests_column.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests_column.log; tail -5 gpurun_out/${T}_tests_column.log
bash scripts/gpu_cfg2.sh $T
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/${T}_launches_cfg2.csv")) if len(r) > 10 and r[0].isdigit()]
# last step: take the last flat_tc_filter launch and the launches around it
names = [(r[4].split('(')[0][-40:], float(r[-1])) for r in rows]
idx = max(i for i, (n, _) in enumerate(names) if 'flat_tc_filter' in n)
for n, v in names[max(0, idx - 8): idx + 8]: print(f"{n:42s} {v}")
PY

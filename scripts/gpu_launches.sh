#!/bin/bash
# launch list (ncu, per-kernel durations) of one bench workload: gpu_launches.sh <tag> <workload> [env assignments...]
T=$1; W=$2; shift 2
for kv in "$@"; do export "$kv"; done
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_$W.csv python bench.py --workload $W --steps 2 --warmup 3 --no-cpu-baseline --no-recall > /dev/null 2>&1; echo "launch list exit $?"
python - <<PY
import csv
rows = [r for r in csv.reader(open("gpurun_out/${T}_launches_$W.csv")) if len(r) > 10 and r[0].isdigit()]
names = [(r[4].split('(')[0][-44:], float(r[-1])) for r in rows]
idx = max(i for i, (n, _) in enumerate(names) if "scan3_kernel" in n or "scan4" in n or "scanflat" in n)
for n, v in names[max(0, idx - 12): idx + 6]: print(f"{n:46s} {v/1e6:8.3f} ms")
PY

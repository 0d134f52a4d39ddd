#!/bin/bash
set -u
mkdir -p gpurun_out
for w in cfg4 cfg5; do
timeout 600 python bench.py --workload $w --queries 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/q1_$w.json 2> gpurun_out/q1_$w.err
python -c "
import json; j=json.loads(open('gpurun_out/q1_$w.json').read().strip().splitlines()[-1]); print('$w Q=1', round(j['ms_per_step'],3), 'ms device;', round(j['e2e']['ms_per_step'],3), 'ms e2e; scan', round(j['roofline']['scan_ms_per_launch'],3))"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_q1_$w.csv python bench.py --workload $w --queries 1 --steps 3 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
done
timeout 600 python scripts/recall.py 2>&1 | head -2

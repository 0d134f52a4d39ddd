#!/bin/bash
# launch lists (ncu, gpu__time_duration) of the given workloads
set -u
mkdir -p gpurun_out
for w in "$@"; do
timeout 600 python bench.py --workload $w --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_$w.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_$w.csv \
    python bench.py --workload $w --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_ll_$w.log 2>&1
echo "$w ncu exit $?"
done

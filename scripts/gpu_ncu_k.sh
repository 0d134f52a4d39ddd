#!/bin/bash
# ncu --set full of one kernel (regex $2) of workload $1
set -u
mkdir -p gpurun_out
timeout 600 python bench.py --workload $1 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_$1.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$2 -s 1 -c 1 -f -o gpurun_out/prof_$1_$2 \
    python bench.py --workload $1 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_$1.log 2>&1
echo "ncu exit $?"

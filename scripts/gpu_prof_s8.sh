#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python bench.py --workload cfg4 --shard-of 8 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_s8.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:scan3 -s 1 -c 1 -f -o gpurun_out/prof_scan_s8 \
    python bench.py --workload cfg4 --shard-of 8 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_s8.log 2>&1
echo "ncu exit $?"

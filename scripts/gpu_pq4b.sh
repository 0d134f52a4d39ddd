#!/bin/bash
# 4-bit PQ: parity of both device paths, then bench lines for the fused and the distance-scratch path.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_pq4.py -m gpu -q -x > gpurun_out/tests_pq4.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests_pq4.log
tail -30 gpurun_out/tests_pq4.log
timeout 300 python bench.py --workload cfg1_pq4 --steps 5 --warmup 3 > gpurun_out/pq4_fused.json 2> gpurun_out/pq4_fused.err; tail -c 1500 gpurun_out/pq4_fused.json; tail -3 gpurun_out/pq4_fused.err
LB200_PQ4_SIMPLE=1 timeout 300 python bench.py --workload cfg1_pq4 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/pq4_simple.json 2> gpurun_out/pq4_simple.err; tail -c 900 gpurun_out/pq4_simple.json; tail -3 gpurun_out/pq4_simple.err

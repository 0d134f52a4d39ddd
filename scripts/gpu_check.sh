#!/bin/bash
# One GPU-box visit: parity tests, smoke, benches.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
free -g | head -2 >> gpurun_out/gpu.txt; nproc >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests.log
tail -30 gpurun_out/tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -5 gpurun_out/smoke.log
for v in 1 2; do
  timeout 600 python bench.py --workload cfg1 --steps 5 --warmup 3 --scan-variant $v --no-cpu-baseline > gpurun_out/bench_cfg1_v$v.json 2> gpurun_out/bench_cfg1_v$v.err
  echo "cfg1 v$v exit $?"; tail -c 1500 gpurun_out/bench_cfg1_v$v.json
done
for v in 1 2; do
  timeout 900 python bench.py --workload cfg4 --steps 5 --warmup 3 --scan-variant $v --no-cpu-baseline > gpurun_out/bench_cfg4_v$v.json 2> gpurun_out/bench_cfg4_v$v.err
  echo "cfg4 v$v exit $?"; tail -c 1500 gpurun_out/bench_cfg4_v$v.json; tail -3 gpurun_out/bench_cfg4_v$v.err
done

#!/bin/bash
T=${1:-r2ar}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -k "prepass or baseline_config or table" > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -3 gpurun_out/${T}_tests.log
bash scripts/gpu_env_ab.sh $T cfg3 -
bash scripts/gpu_env_ab.sh $T cfg5 -

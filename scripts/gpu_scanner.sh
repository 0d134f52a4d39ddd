#!/bin/bash
# One short GPU-box visit: the reader/scanner parity tests, then the whole GPU suite, smoke, one cfg4 bench line.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_scanner.py -m gpu -q > gpurun_out/tests_scanner.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests_scanner.log
tail -40 gpurun_out/tests_scanner.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests.log
tail -5 gpurun_out/tests.log
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/q_cfg4.json 2> gpurun_out/q_cfg4.err; tail -c 1200 gpurun_out/q_cfg4.json

#!/bin/bash
# Short GPU visit: 4-bit PQ parity + native index open, then the whole GPU suite.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_pq4.py tests/test_gpu_scanner.py -m gpu -q > gpurun_out/tests_pq4.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests_pq4.log
tail -40 gpurun_out/tests_pq4.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests.log
tail -5 gpurun_out/tests.log

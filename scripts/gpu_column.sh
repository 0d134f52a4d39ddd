#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_column.py -x -q > gpurun_out/r2r_tests_column.log 2>&1; echo "exit $?" >> gpurun_out/r2r_tests_column.log; tail -30 gpurun_out/r2r_tests_column.log

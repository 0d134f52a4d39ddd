#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_group.py -x -q > gpurun_out/r2c_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r2c_tests.log
tail -15 gpurun_out/r2c_tests.log

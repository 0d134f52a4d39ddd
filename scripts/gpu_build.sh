#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_build.py -x -q > gpurun_out/r2p_tests_build.log 2>&1; echo "exit $?" >> gpurun_out/r2p_tests_build.log; tail -25 gpurun_out/r2p_tests_build.log

#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2n_noalloc_alone.log 2>&1; echo "alone exit $?" >> gpurun_out/r2n_noalloc_alone.log; tail -4 gpurun_out/r2n_noalloc_alone.log
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/r2n_tests_all.log 2>&1; echo "all exit $?" >> gpurun_out/r2n_tests_all.log; tail -6 gpurun_out/r2n_tests_all.log

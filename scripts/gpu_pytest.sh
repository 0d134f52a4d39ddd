#!/bin/bash
# usage: gpu_pytest.sh <tag> <pytest args...>
T=$1; shift
mkdir -p gpurun_out
timeout 1500 python -m pytest "$@" -x -q > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -25 gpurun_out/${T}_tests.log

#!/bin/bash
set -u
mkdir -p gpurun_out
run() { # label, env...
  local lab=$1; shift
  env "$@" timeout 600 python bench.py --workload cfg1 --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/c1_$lab.err | python -c "
import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lab', round(j['value']), 'q/s scan', round(j['roofline']['scan_ms_per_launch'],3), 'ms/step', round(j['ms_per_step'],3), 'crc', j.get('result_crc32'))"
}
run base X=1
run boot LB200_BOOT=1
run L4 LB200_SKEW_L=4
run L16 LB200_SKEW_L=16
run L32 LB200_SKEW_L=32
run noqthr LB200_NO_QTHR=1
run nocold LB200_NO_COLDPACK=1

#!/bin/bash
set -u
mkdir -p gpurun_out
run() { local lab=$1; shift
  env "$@" timeout 600 python bench.py --workload cfg4 --shard-of 8 --steps 8 --warmup 3 --no-cpu-baseline 2> gpurun_out/ch_$lab.err | python -c "
import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lab', 'scan', round(j['roofline']['scan_ms_per_launch'],3), 'ms/step', round(j['ms_per_step'],3), 'crc', j.get('result_crc32'))"
}
run none LB200_BOOT=1
run c48k LB200_BOOT=1 LB200_CHUNK_ROWS=49152
run c32k LB200_BOOT=1 LB200_CHUNK_ROWS=32768
run c24k LB200_BOOT=1 LB200_CHUNK_ROWS=24576
run c16k LB200_BOOT=1 LB200_CHUNK_ROWS=16384
run c8k LB200_BOOT=1 LB200_CHUNK_ROWS=8192
env LB200_CHUNK_ROWS=32768 timeout 600 python bench.py --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n1 c32k scan', round(j['roofline']['scan_ms_per_launch'],3), 'ms/step', round(j['ms_per_step'],3))"

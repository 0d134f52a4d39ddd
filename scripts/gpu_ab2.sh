#!/bin/bash
set -u
mkdir -p gpurun_out
run() { local name=$1; shift; local args=$1; shift
  env "$@" timeout 600 python bench.py --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline $args > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_$name.json").read().strip().splitlines()[-1])
    print("$name", round(j["value"]), "q/s  ms/step", round(j["ms_per_step"],2), " scan ms", round(j["roofline"]["scan_ms_per_launch"],2), "GB/s", round(j["roofline"]["achieved"]), "bytes", j["roofline"]["algorithmic_bytes_per_launch"]/1e9)
except Exception as e:
    print("$name parse fail", e); print(open("gpurun_out/ab_$name.err").read()[-800:])
PY
}
run s8_boot "--shard-of 8" X=1
run s8_noboot "--shard-of 8" LB200_NO_BOOT=1
run s8_noqthr "--shard-of 8" LB200_NO_QTHR=1
run s1_noboot "" LB200_NO_BOOT=1
run s1_noqthr "" LB200_NO_QTHR=1
run s1_noqthr_noboot "" LB200_NO_QTHR=1 LB200_NO_BOOT=1

#!/bin/bash
# how much of the scan is cold-start cost?  the same batch repeated with the bounds of the previous run kept
T=${1:-r2bi}
mkdir -p gpurun_out
for so in 0 8; do
  for keep in 0 1; do
    if [ $keep = 1 ]; then export LB200_KEEP_QTHR=1; else unset LB200_KEEP_QTHR; fi
    EX=""; if [ $so != 0 ]; then EX="--shard-of $so"; fi
    timeout 600 python bench.py $EX --steps 10 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_so${so}_keep$keep.json 2>/dev/null
    python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_so${so}_keep$keep.json").read().strip().splitlines()[-1])
print("shard-of $so keep=$keep", round(d["value"]), "q/s scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "frac", round(d["roofline"]["frac"],3), "crc", d.get("result_crc32"))
PY
  done
done

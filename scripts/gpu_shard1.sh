#!/bin/bash
# rank 0's shard of an 8-way split on ONE GPU: BOOT vs no-BOOT scan kernel, then ncu of the BOOT kernel
T=${1:-r2ag}
mkdir -p gpurun_out
for b in 1 0; do
  LB200_BOOT=$b timeout 600 python bench.py --shard-of 8 --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_shard8_boot$b.json 2> gpurun_out/${T}_shard8_boot$b.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_shard8_boot$b.json").read().strip().splitlines()[-1])
print("boot=$b", round(d["value"]), "q/s ms/step", round(d["ms_per_step"],3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "frac", round(d["roofline"]["frac"],3), d["roofline"]["kernel"])
PY
done
LB200_BOOT=1 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:scan3_kernel -s 3 -c 1 -f -o gpurun_out/${T}_prof_scan_shard8 python bench.py --shard-of 8 --steps 2 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_ncu.log 2>&1; echo "ncu exit $?"

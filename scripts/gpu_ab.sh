#!/bin/bash
# same-box A/B of library builds under lance_b200/ab/ (LB200_SO) against the in-tree build: cfg4 (+ $2 workloads)
mkdir -p gpurun_out
T=${1:-ab}; shift
WL=${@:-cfg4}
for w in $WL; do
for so in "" $(ls lance_b200/ab/*.so); do
  name=$(basename "${so:-main}" .so)
  LB200_SO=$so timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_${name}_$w.json 2> gpurun_out/${T}_${name}_$w.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_${name}_$w.json").read().strip().splitlines()[-1])
    print("$w %-28s" % "$name", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "crc", d["result_crc32"])
except Exception as e:
    print("$w $name parse fail", e); print(open("gpurun_out/${T}_${name}_$w.err").read()[-800:])
PY
done
done

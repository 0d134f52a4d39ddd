#!/bin/bash
T=${1:-r2ac}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -k "prepass or baseline_config or cfg or table" > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -4 gpurun_out/${T}_tests.log
for mode in image noimage; do
  if [ $mode = noimage ]; then export LB200_NO_LUT_IMAGE=1; else unset LB200_NO_LUT_IMAGE; fi
  timeout 900 python bench.py --workload cfg3 --steps 10 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_bench_cfg3_$mode.json 2> gpurun_out/${T}_bench_cfg3_$mode.err || tail -3 gpurun_out/${T}_bench_cfg3_$mode.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg3_$mode.json").read().strip().splitlines()[-1])
print("$mode", round(d["value"]), "q/s ms/step", round(d["ms_per_step"],2), "frac", round(d["roofline"]["frac"],3), "scan_ms", d["roofline"].get("kernel_ms") or d["config"].get("scan_ms"), "crc", d["config"].get("result_crc32"), "parity", d.get("parity",{}).get("unexplained_mismatches"))
PY
done

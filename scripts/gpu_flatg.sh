#!/bin/bash
# IVF_FLAT grouped L2 scan: tests, then A/B against the pair-wise kernel on cfg1_flat
T=${1:-r2ak}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf_flat.py tests/test_gpu_parity.py tests/test_gpu_pq4.py tests/test_gpu_column.py -x -q -k "flat or pq4 or column or 4bit" > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -4 gpurun_out/${T}_tests.log
for mode in grouped pairwise; do
  if [ $mode = pairwise ]; then export LB200_FLAT_PAIRWISE=1; else unset LB200_FLAT_PAIRWISE; fi
  timeout 900 python bench.py --workload cfg1_flat --steps 10 --warmup 3 --no-recall > gpurun_out/${T}_bench_cfg1_flat_$mode.json 2> gpurun_out/${T}_bench_cfg1_flat_$mode.err || tail -3 gpurun_out/${T}_bench_cfg1_flat_$mode.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg1_flat_$mode.json").read().strip().splitlines()[-1])
print("$mode", round(d["value"]), "q/s ms/step", round(d["ms_per_step"],2), "frac", round(d["roofline"]["frac"],3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],2), "crc", d.get("result_crc32"), "parity", d.get("parity",{}).get("unexplained_mismatches"))
PY
done

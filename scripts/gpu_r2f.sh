#!/bin/bash
# parity suite + the four IVF_PQ configs after a scan-kernel change
mkdir -p gpurun_out
T=${1:-r2f}
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dropin.py tests/test_gpu_group.py tests/test_gpu_scanner.py -x -q > gpurun_out/${T}_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -4 gpurun_out/${T}_tests.log
for w in cfg4 cfg1 cfg5 cfg3; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_$w.json 2> gpurun_out/${T}_$w.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_$w.json").read().strip().splitlines()[-1])
    print("$w", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "frac", round(d["roofline"]["frac"],3), "crc", d["result_crc32"])
except Exception as e:
    print("$w parse fail", e); print(open("gpurun_out/${T}_$w.err").read()[-1500:])
PY
done

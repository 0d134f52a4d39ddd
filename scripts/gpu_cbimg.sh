#!/bin/bash
# lane-order codebook image for the in-kernel table build: parity tests, then cfg1 / cfg4 per prefetch depth (alt builds)
T=${1:-r2bu}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_group.py tests/test_gpu_dropin.py tests/test_gpu_f16_ivfpq.py -x -q > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -3 gpurun_out/${T}_tests.log
for D in 1 2 3; do
  SO=""; [ $D -gt 1 ] && SO="lance_b200/liblance_b200_d$D.so"
  [ -n "$SO" ] && [ ! -f "$SO" ] && continue
  for W in cfg1 cfg4; do
    LB200_SO=$SO timeout 600 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_d${D}_$W.json 2> gpurun_out/${T}_d${D}_$W.err
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_d${D}_$W.json").read().strip().splitlines()[-1])
    print("depth $D $W", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"], 3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"], 3), "frac", round(d["roofline"]["frac"], 3), "crc", d.get("result_crc32"))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/${T}_d${D}_$W.err").read()[-1500:])
PY
  done
done

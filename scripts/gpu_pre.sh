#!/bin/bash
# ADC-table pre-pass A/B: tests, then each workload with the pre-pass forced off / on / auto
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; tail -3 gpurun_out/tests.log
run() { # name env workload
  env $2 timeout 600 python bench.py --workload $3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/pre_$1.json 2> gpurun_out/pre_$1.err
  python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    j=json.loads(open(f"gpurun_out/pre_{n}.json").read().strip().splitlines()[-1])
    print(n, round(j["value"]), "q/s  scan ms", round(j["roofline"]["scan_ms_per_launch"],2), "frac", round(j["roofline"]["frac"],3), "ms/step", round(j["ms_per_step"],2), "crc", j.get("result_crc32"))
except Exception as e:
    print(n, "parse fail", e); print(open(f"gpurun_out/pre_{n}.err").read()[-800:])
PY
}
for w in "$@"; do
  run ${w}_off LB200_LUT_PREPASS=0 $w
  run ${w}_on LB200_LUT_PREPASS=1 $w
done

#!/bin/bash
# cfg1 (BASELINE configs[0]) under the library's diagnostic switches: which per-item cost dominates?
set -u
mkdir -p gpurun_out
run() { # name, env...
  local name=$1; shift
  env "$@" timeout 200 python bench.py --workload cfg1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/cfg1_$name.json 2> gpurun_out/cfg1_$name.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/cfg1_$name.json").read().strip().splitlines()[-1])
    print("$name", round(j["value"]), "q/s scan ms", round(j["roofline"]["scan_ms_per_launch"],3), "frac", round(j["roofline"]["frac"],3), "ms/step", round(j["ms_per_step"],3), "crc", j["result_crc32"], j["roofline"]["kernel"])
except Exception as e:
    print("$name parse fail", e); print(open("gpurun_out/cfg1_$name.err").read()[-600:])
PY
}
run default A=1
run prepass LB200_LUT_PREPASS=1
run L4 LB200_SKEW_L=4
run L16 LB200_SKEW_L=16
run boot1 LB200_BOOT=1
run prepassL4 LB200_LUT_PREPASS=1 LB200_SKEW_L=4

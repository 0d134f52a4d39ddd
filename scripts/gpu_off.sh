#!/bin/bash
# single-block scan rewrite (offsets_kernel, queue_offsets_kernel): tests that reach them + cfg1 / cfg4 lines + launch list
T=${1:-r2ch}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_group.py tests/test_gpu_dropin.py tests/test_gpu_pq4.py tests/test_gpu_ivf_flat.py tests/test_gpu_f16_ivfpq.py -x -q > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -3 gpurun_out/${T}_tests.log
for W in cfg1 cfg4; do
  timeout 300 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_bench_$W.json 2> gpurun_out/${T}_bench_$W.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_bench_$W.json").read().strip().splitlines()[-1])
    print("$W", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"], 3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"], 3), "frac", round(d["roofline"]["frac"], 3), "crc", d.get("result_crc32"))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/${T}_bench_$W.err").read()[-800:])
PY
done
bash scripts/gpu_launches.sh $T cfg1 | grep -i "offsets\|launch list"

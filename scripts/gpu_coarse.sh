#!/bin/bash
# packed-arithmetic coarse_dist_l2_slab_kernel: parity tests that cover it, cfg4 / cfg2 lines, launch list of cfg4
T=${1:-r2bs}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_column.py tests/test_gpu_build.py tests/test_gpu_ivf_flat.py -x -q > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -3 gpurun_out/${T}_tests.log
for W in cfg4 cfg2; do
  timeout 600 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_bench_$W.json 2> gpurun_out/${T}_bench_$W.err; echo "$W exit $?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_$W.json").read().strip().splitlines()[-1])
print("$W", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 3), "parity", d.get("parity"), "crc", d.get("result_crc32"))
PY
done
LB200_FLAT_NO_TC=1 timeout 600 python bench.py --workload cfg2 --steps 5 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_bench_cfg2_exact.json 2> gpurun_out/${T}_bench_cfg2_exact.err; tail -c 600 gpurun_out/${T}_bench_cfg2_exact.json
bash scripts/gpu_launches.sh $T cfg4

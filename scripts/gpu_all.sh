#!/bin/bash
# tests + smoke + bench on every workload (1 GPU)
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests.log
tail -4 gpurun_out/tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
for w in cfg4 cfg1 cfg5 cfg3; do
  timeout 900 python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  echo "$w exit $?"; python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/bench_$w.json").read().strip().splitlines()[-1])
    print("$w", round(j["value"]), "q/s e2e", round(j["e2e"]["value"]), "scan GB/s", round(j["roofline"]["achieved"]), "frac", round(j["roofline"]["frac"],3), "variant", j["config"]["scan_variant"], j["config"]["scan_queries_per_cta"], "cpu", j.get("cpu_baseline",{}).get("value"))
except Exception as e:
    print("$w parse fail", e); print(open("gpurun_out/bench_$w.err").read()[-1500:])
PY
done
timeout 600 python bench.py --workload cfg2 --steps 5 --warmup 3 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "cfg2 exit $?"; cut -c1-300 gpurun_out/bench_cfg2.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; cut -c1-400 gpurun_out/bench_ref.json

#!/bin/bash
mkdir -p gpurun_out
for w in ${@:-cfg1 cfg4 cfg5 cfg3}; do
  SECONDS=0; timeout 900 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/r2q_bench_$w.json 2> gpurun_out/r2q_bench_$w.err; echo "$w wall ${SECONDS}s"; tail -2 gpurun_out/r2q_bench_$w.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2q_bench_$w.json").read().strip().splitlines()[-1])
    r = d.get("recall", {}); print("$w", round(d["value"]), "recall", r.get("recall_at_10"), "qps_real", round(r.get("qps_on_this_index", 0)), r.get("parity_vs_oracle"), r.get("seconds"), r.get("index"))
except Exception as e:
    print("$w fail", e)
PY
done

#!/bin/bash
# cfg2 regression at N=1, cfg3 at nprobes 1 / 20 (cfg 3 is run at P = 1, 20, 102)
T=${1:-r2bb}
mkdir -p gpurun_out
timeout 600 python bench.py --workload cfg2 --steps 20 --warmup 3 > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err || tail -5 gpurun_out/${T}_bench_cfg2.err
python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg2.json").read().strip().splitlines()[-1])
print("cfg2", round(d["value"]), "q/s ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "parity", d.get("parity",{}).get("unexplained_mismatches"), d["clocks"])
PY
for P in 1 20; do
  timeout 900 python bench.py --workload cfg3 --nprobes $P --steps 10 --warmup 3 --no-recall > gpurun_out/${T}_bench_cfg3_p$P.json 2> gpurun_out/${T}_bench_cfg3_p$P.err || tail -5 gpurun_out/${T}_bench_cfg3_p$P.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg3_p$P.json").read().strip().splitlines()[-1])
print("cfg3 P=$P", round(d["value"]), "q/s ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"],3), "parity", d.get("parity",{}).get("unexplained_mismatches"), "cpu", round((d.get("cpu_baseline") or {}).get("value",0)))
PY
done

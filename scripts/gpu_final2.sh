#!/bin/bash
# Lean end-of-round evidence (GPU budget nearly spent): whole GPU suite, smoke(), the headline line in full, the other
# workloads with cpu_baseline + parity but no recall block, drop-in mode, launch list of the headline.  usage: gpu_final2.sh <tag>
T=${1:-final}
mkdir -p gpurun_out
SECONDS=0
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_gpu_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_gpu_tests.log; tail -3 gpurun_out/${T}_gpu_tests.log; echo "tests ${SECONDS}s"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for w in cfg4 cfg1 cfg5 cfg3 cfg2 cfg1_pq4 cfg1_flat; do
  X="--no-recall"; [ $w = cfg4 ] && X=""
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 $X > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err || { echo "$w bench failed"; tail -5 gpurun_out/${T}_bench_$w.err; continue; }
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_$w.json").read().strip().splitlines()[-1])
r = d["roofline"]; print("$w", round(d["value"]), "q/s e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"],3), "frac", round(r.get("frac") or 0,3), "parity", d.get("parity",{}).get("unexplained_mismatches"), "recall", d.get("recall_at_10"), "cpu", round((d.get("cpu_baseline") or {}).get("value") or 0), "clocks", d["clocks"].get("sm_mhz"), d["clocks"].get("reasons"))
PY
  echo "  ${SECONDS}s"
done
timeout 300 python bench.py --mode dropin --steps 3 --warmup 1 > gpurun_out/${T}_dropin_cfg4.json 2>/dev/null; tail -c 500 gpurun_out/${T}_dropin_cfg4.json; echo
bash scripts/gpu_launches.sh $T cfg4 | tail -6
echo "total ${SECONDS}s"

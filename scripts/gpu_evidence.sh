#!/bin/bash
# Evidence pass for the current build: bench line + ncu launch list + one ncu --set full capture of the scan kernel per
# workload (each ncu run only after the plain run exited 0).  usage: gpu_evidence.sh <tag> <workloads...>
T=$1; shift
mkdir -p gpurun_out
for w in "$@"; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err || { echo "$w bench failed"; tail -5 gpurun_out/${T}_bench_$w.err; continue; }
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_$w.json").read().strip().splitlines()[-1])
r = d["roofline"]; print("$w", round(d["value"]), "q/s e2e", round(d["e2e"]["value"]), "scan_ms", r.get("scan_ms_per_launch"), "frac", round(r["frac"],3), "parity", d.get("parity",{}).get("unexplained_mismatches"), "crc", d.get("result_crc32"))
PY
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_$w.csv \
      python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1; echo "$w launch list exit $?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:scan3 -s 3 -c 1 -f -o gpurun_out/${T}_prof_scan_$w \
      python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu_full_$w.log 2>&1; echo "$w ncu full exit $?"
done

#!/bin/bash
# N GPUs: cfg4 with the probe-load-weighted placement, then the cfg5 batch-size sweep (BASELINE configs[4])
N=${1:-8}; T=${2:-r2af}
mkdir -p gpurun_out
run() {
  name=$1; shift
  SECONDS=0
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-recall "$@" > gpurun_out/${T}_${name}_n$N.json 2> gpurun_out/${T}_${name}_n$N.err
  echo "$name exit $? (${SECONDS}s)"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_${name}_n$N.json").read().strip().splitlines()[-1])
    print("$name", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "crc", d.get("result_crc32"))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/${T}_${name}_n$N.err").read()[-1500:])
PY
}
run cfg4
for q in 1 256 4096 16384; do run cfg5_q$q --workload cfg5 --queries $q; done

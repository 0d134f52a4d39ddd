#!/bin/bash
# one cfg4 line at N ranks (final build).  usage: gpu_scale1.sh <N> <tag>
N=$1; T=$2
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-recall > gpurun_out/${T}_cfg4_n$N.json 2> gpurun_out/${T}_cfg4_n$N.err
python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_cfg4_n$N.json").read().strip().splitlines()[-1])
print("N=$N", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "frac", round(d["roofline"]["frac"],3), "crc", d.get("result_crc32"))
PY

#!/bin/bash
# N = 8: peer exchange vs NCCL exchange, cfg4 (+ cfg5), and N = 1 beside it on the same box
N=${1:-8}
mkdir -p gpurun_out
run() {
  name=$1; shift
  env "$@" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2e_${name}_n$N.json 2> gpurun_out/r2e_${name}_n$N.err
  echo "$name exit $?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2e_${name}_n$N.json").read().strip().splitlines()[-1])
    print("$name", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "crc", d["result_crc32"])
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/r2e_${name}_n$N.err").read()[-2500:])
PY
}
run peer_cfg4 LB200_EXCHANGE=peer
run nccl_cfg4 LB200_EXCHANGE=nccl
run peer_cfg5 LB200_EXCHANGE=peer LB200_WORKLOAD=cfg5
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2e_cfg4_n1.json 2>/dev/null; python - <<PY
import json
d = json.loads(open("gpurun_out/r2e_cfg4_n1.json").read().strip().splitlines()[-1])
print("n1", round(d["value"]), "ms/step", round(d["ms_per_step"],3), "crc", d["result_crc32"])
PY

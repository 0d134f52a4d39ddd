#!/bin/bash
# multi-GPU: peer-memory exchange (lb200_shard_*) vs the NCCL all_gather exchange, N = $1
N=${1:-2}
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2d_${name}_n$N.json 2> gpurun_out/r2d_${name}_n$N.err
  echo "$name exit $?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2d_${name}_n$N.json").read().strip().splitlines()[-1])
    print("$name", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "crc", d["result_crc32"], d["config"].get("exchange"))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/r2d_${name}_n$N.err").read()[-2500:])
PY
}
run peer_cfg4 LB200_EXCHANGE=peer
run nccl_cfg4 LB200_EXCHANGE=nccl
run peer_cfg5 LB200_EXCHANGE=peer LB200_WORKLOAD=cfg5
run nccl_cfg5 LB200_EXCHANGE=nccl LB200_WORKLOAD=cfg5
run peer_cfg3 LB200_EXCHANGE=peer LB200_WORKLOAD=cfg3
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2d_cfg4_n1.json 2>/dev/null; python - <<PY
import json
d = json.loads(open("gpurun_out/r2d_cfg4_n1.json").read().strip().splitlines()[-1])
print("n1", round(d["value"]), "ms/step", round(d["ms_per_step"],3), "crc", d["result_crc32"])
PY

#!/bin/bash
# N-GPU A/B of the shared pruning bounds (peer tables over NVLink): N = $1
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "shard" > gpurun_out/tests_peer.log 2>&1; tail -3 gpurun_out/tests_peer.log
show() { python - "$1" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "n=%d"%j["n_gpus"], round(j["value"]), "q/s e2e", round(j["e2e"]["value"]), "ms/step", round(j["ms_per_step"],3), "scan ms", round(j["roofline"]["scan_ms_per_launch"],3), "crc", j.get("result_crc32"))
except Exception as e:
    print("parse fail", sys.argv[1], e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/peer_n1.json 2> gpurun_out/peer_n1.err; show gpurun_out/peer_n1.json
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/peer_on_n$N.json 2> gpurun_out/peer_on_n$N.err; echo "on exit $?"; show gpurun_out/peer_on_n$N.json
LB200_NO_PEER_THR=1 timeout 600 $TR --master-port 29522 bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/peer_off_n$N.json 2> gpurun_out/peer_off_n$N.err; echo "off exit $?"; show gpurun_out/peer_off_n$N.json

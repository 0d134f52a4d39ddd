#!/bin/bash
# N-GPU validation: cfg4 and cfg5 at N=1 and N=$1, result CRCs must agree
set -u
N=${1:-2}
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "n=%d"%j["n_gpus"], round(j["value"]), "q/s e2e", round(j["e2e"]["value"]), "ms/step", round(j["ms_per_step"],3), "scan ms", round(j["roofline"]["scan_ms_per_launch"],3), "crc", j.get("result_crc32"))
except Exception as e:
    print("parse fail", sys.argv[1], e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29530
for w in cfg4 cfg5; do
  timeout 600 python bench.py --workload $w --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/m_${w}_n1.json 2> gpurun_out/m_${w}_n1.err; show gpurun_out/m_${w}_n1.json
  port=$((port+1))
  timeout 600 $TR --master-port $port bench.py --workload $w --gpus $N --steps 8 --warmup 3 > gpurun_out/m_${w}_n$N.json 2> gpurun_out/m_${w}_n$N.err; show gpurun_out/m_${w}_n$N.json
done
port=$((port+1))
timeout 600 $TR --master-port $port bench.py --impl reference --gpus $N --steps 1 --warmup 0 > gpurun_out/m_ref_n$N.json 2> gpurun_out/m_ref_n$N.err; echo "ref exit $?"; cut -c1-200 gpurun_out/m_ref_n$N.json

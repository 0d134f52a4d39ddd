#!/bin/bash
# N-GPU bench under torchrun (N = $1), plus the GPU tests on rank-0's device
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "shard or find_partitions" > gpurun_out/tests_multi.log 2>&1; tail -2 gpurun_out/tests_multi.log
for n in 1 $N; do
  if [ $n = 1 ]; then CMD="python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline";
  else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 5 --warmup 3"; fi
  timeout 900 $CMD > gpurun_out/bench_cfg4_n$n.json 2> gpurun_out/bench_cfg4_n$n.err; echo "n=$n exit $?"
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/bench_cfg4_n$n.json").read().strip().splitlines()[-1])
    print("n=$n", round(j["value"]), "q/s  e2e", round(j["e2e"]["value"]), " ms/step", round(j["ms_per_step"],2), "scan ms", round(j["roofline"]["scan_ms_per_launch"],2))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/bench_cfg4_n$n.err").read()[-2000:])
PY
done
if [ $N -gt 1 ]; then
LB200_REPLICATED_COARSE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_cfg4_n${N}_repl.json 2> gpurun_out/bench_cfg4_n${N}_repl.err; echo "repl exit $?"
python -c "
import json; j=json.loads(open('gpurun_out/bench_cfg4_n${N}_repl.json').read().strip().splitlines()[-1]); print('replicated coarse n=$N', round(j['value']), 'q/s')"
fi

#!/bin/bash
# flat kNN (cfg2) with rows sharded over N GPUs, and N = 1 beside it on the same box.  usage: gpu_flat_n.sh <N> <tag> [queries]
N=${1:-2}; T=${2:-r2bc}; Q=${3:-4096}
mkdir -p gpurun_out
run() {
  n=$1
  if [ $n -gt 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513"; else L=python; fi
  timeout 600 $L bench.py --gpus $n --workload cfg2 --queries $Q --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg2_q${Q}_n$n.json 2> gpurun_out/${T}_bench_cfg2_q${Q}_n$n.err || tail -8 gpurun_out/${T}_bench_cfg2_q${Q}_n$n.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${T}_bench_cfg2_q${Q}_n$n.json").read().strip().splitlines()[-1])
print("Q=$Q N=$n", round(d["value"]), "q/s ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), d["config"]["filter"])
PY
}
run 1
run $N

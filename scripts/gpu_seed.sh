#!/bin/bash
# seeding launch A/B (LB200_SEED=0|1) on cfg4 / cfg1 at N ranks, plus the parity suite at N = 1
N=${1:-1}; T=${2:-r2bk}
mkdir -p gpurun_out
if [ $N -eq 1 ]; then
  timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_group.py tests/test_gpu_dropin.py -x -q > gpurun_out/${T}_tests.log 2>&1; echo "exit $?" >> gpurun_out/${T}_tests.log; tail -3 gpurun_out/${T}_tests.log
fi
run() {
  name=$1; shift
  if [ $N -gt 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"; else L=python; fi
  env "$@" timeout 900 $L bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-recall $EXTRA > gpurun_out/${T}_${name}_n$N.json 2> gpurun_out/${T}_${name}_n$N.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_${name}_n$N.json").read().strip().splitlines()[-1])
    print("$name", round(d["value"]), "q/s  ms/step", round(d["ms_per_step"],3), "scan_ms", round(d["roofline"]["scan_ms_per_launch"],3), "frac", round(d["roofline"]["frac"],3), "crc", d.get("result_crc32"))
except Exception as e:
    print("parse fail", e); print(open("gpurun_out/${T}_${name}_n$N.err").read()[-1500:])
PY
}
EXTRA=""
run seed_cfg4 LB200_SEED=1
if [ $N -gt 1 ]; then run seedall_cfg4 LB200_SEED=1 LB200_SEED_ALL=1; fi
if [ $N -eq 1 ]; then run noseed_cfg4 LB200_SEED=0; fi
if [ $N -eq 1 ]; then
  EXTRA="--workload cfg1"; run seed_cfg1 LB200_SEED=1; run noseed_cfg1 LB200_SEED=0
fi

#!/bin/bash
# batch-size sweep (BASELINE configs[4]): queries/s and ms per batch for Q = 1 ... 16384, one GPU
set -u
W=${1:-cfg5}
mkdir -p gpurun_out
: > gpurun_out/sweep_$W.jsonl
for q in 1 2 4 8 16 32 64 128 256 512 1024 2048 4096 8192 16384; do
  timeout 300 python bench.py --workload $W --queries $q --steps 20 --warmup 5 --no-cpu-baseline >> gpurun_out/sweep_$W.jsonl 2> gpurun_out/sweep_$W.err || { echo "Q=$q failed"; tail -3 gpurun_out/sweep_$W.err; }
done
python - "$W" <<'PY'
import json,sys
for l in open(f"gpurun_out/sweep_{sys.argv[1]}.jsonl"):
    j=json.loads(l); q=j["config"]["queries_per_step"]
    print(f"Q={q:6d}  {j['value']:12.0f} q/s  {j['ms_per_step']:8.3f} ms/batch  e2e {j['e2e']['value']:12.0f} q/s  {j['e2e']['ms_per_step']:8.3f} ms  launches/step {j['gpu_launches']/j['steps']:.0f}")
PY

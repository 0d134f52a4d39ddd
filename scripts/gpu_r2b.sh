#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_parity.py tests/test_gpu_scanner.py -x -q > gpurun_out/r2b_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r2b_tests.log
tail -5 gpurun_out/r2b_tests.log
timeout 600 python bench.py --mode dropin > gpurun_out/r2b_dropin_cfg4.json 2> gpurun_out/r2b_dropin_cfg4.err; python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2b_dropin_cfg4.json').read().strip().splitlines()[-1])
print(d['value'], d['latency_ms'], d['concurrent_clients'], d['cpu_baseline']['value'], d['parity']['unexplained_mismatches'])
PY
tail -3 gpurun_out/r2b_dropin_cfg4.err
LB200_NO_GRAPHS=1 timeout 600 python bench.py --mode dropin > gpurun_out/r2b_dropin_cfg4_nographs.json 2>/dev/null; python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2b_dropin_cfg4_nographs.json').read().strip().splitlines()[-1])
print('nographs', d['value'], d['latency_ms'])
PY
for w in cfg1 cfg3 cfg5; do timeout 600 python bench.py --mode dropin --workload $w --queries 500 > gpurun_out/r2b_dropin_$w.json 2>gpurun_out/r2b_dropin_$w.err; python - <<PY
import json
d = json.loads(open('gpurun_out/r2b_dropin_$w.json').read().strip().splitlines()[-1])
print('$w', d['value'], d['latency_ms'], d['cpu_baseline']['value'], d['parity']['unexplained_mismatches'])
PY
done

#!/bin/bash
# One GPU-box visit: parity tests, bench (plain), then ncu launch list + full capture of the scan kernel.
set -u
mkdir -p gpurun_out
W=${1:-cfg4}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests.log
tail -5 gpurun_out/tests.log
timeout 900 python bench.py --workload $W --steps 5 --warmup 3 > gpurun_out/bench_$W.json 2> gpurun_out/bench_$W.err; echo "bench exit $?"
cat gpurun_out/bench_$W.json
timeout 600 python bench.py --workload $W --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$W.csv \
    python bench.py --workload $W --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:scan -s 1 -c 1 -f -o gpurun_out/prof_scan_$W \
    python bench.py --workload $W --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
tail -3 gpurun_out/ncu_full.log

#!/bin/bash
# ncu --set full capture of the scan kernel of one workload (after a plain run exits 0)
W=${1:-cfg4}; T=${2:-r2g}
mkdir -p gpurun_out
timeout 600 python bench.py --workload $W --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_plain_$W.json 2>gpurun_out/${T}_plain_$W.err || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:scan3 -s 3 -c 1 -f -o gpurun_out/${T}_prof_scan_$W \
    python bench.py --workload $W --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu_full_$W.log 2>&1
echo "ncu full exit $?"; tail -3 gpurun_out/${T}_ncu_full_$W.log

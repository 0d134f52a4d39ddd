#!/bin/bash
# ncu --set full of one kernel (regex $1) in a bench workload ($2), tag $3, extra bench args $4..
K=$1; W=$2; T=$3; shift 3
mkdir -p gpurun_out
timeout 600 python bench.py --workload $W --steps 2 --warmup 3 --no-cpu-baseline "$@" > /dev/null 2>gpurun_out/${T}_plain.err || { tail -5 gpurun_out/${T}_plain.err; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/${T}_prof_${K}_$W \
    python bench.py --workload $W --steps 2 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${T}_ncu_${K}_$W.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/${T}_ncu_${K}_$W.log

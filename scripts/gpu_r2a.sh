#!/bin/bash
# round 2, first GPU pass: test suite, headline bench with parity, drop-in mode, the other BASELINE configs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r2a_tests.log
tail -5 gpurun_out/r2a_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_cfg4.json 2> gpurun_out/r2a_cfg4.err; tail -c 600 gpurun_out/r2a_cfg4.json
timeout 600 python bench.py --mode dropin > gpurun_out/r2a_dropin_cfg4.json 2> gpurun_out/r2a_dropin_cfg4.err; tail -c 1500 gpurun_out/r2a_dropin_cfg4.json; tail -5 gpurun_out/r2a_dropin_cfg4.err
for w in cfg1 cfg5 cfg3 cfg2; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/r2a_$w.json 2> gpurun_out/r2a_$w.err; tail -c 400 gpurun_out/r2a_$w.json
done

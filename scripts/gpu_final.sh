#!/bin/bash
# Final round visit: whole GPU suite, smoke, headline bench with CPU baseline, reference arm, 4-bit bench.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/tests.log
tail -4 gpurun_out/tests.log
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/final_cfg4.json 2> gpurun_out/final_cfg4.err; tail -c 700 gpurun_out/final_cfg4.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_ref_cfg4.json 2> gpurun_out/final_ref_cfg4.err; tail -c 500 gpurun_out/final_ref_cfg4.json
timeout 300 python bench.py --workload cfg1_pq4 --steps 5 --warmup 3 > gpurun_out/final_cfg1_pq4.json 2> gpurun_out/final_cfg1_pq4.err; tail -c 400 gpurun_out/final_cfg1_pq4.json

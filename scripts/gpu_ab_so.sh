#!/bin/bash
# same-box A/B of library builds: $@ = list of .so suffixes (lance_b200/liblance_b200_<sfx>.so)
set -u
mkdir -p gpurun_out
for sfx in "$@"; do
 for w in cfg4 cfg1; do
  LB200_SO=$PWD/lance_b200/liblance_b200_$sfx.so timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/ab_${sfx}_$w.err | python -c "
import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$sfx $w', round(j['value']), 'q/s scan', round(j['roofline']['scan_ms_per_launch'],3), 'frac', round(j['roofline']['frac'],4), 'crc', j.get('result_crc32'))"
 done
done

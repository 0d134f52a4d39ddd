#!/bin/bash
set -u
mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0, "tests")
from conftest import lance_files_dir
print(lance_files_dir())
open("gpurun_out/fixture_dir.txt", "w").write(lance_files_dir())
PY
gcc -std=c11 -Iinclude examples/open_legacy_index.c -Llance_b200 -llance_b200 -Wl,-rpath,$PWD/lance_b200 -o /tmp/open_legacy_index && /tmp/open_legacy_index $(cat gpurun_out/fixture_dir.txt)/v0.8.14_f358a219_index.idx > gpurun_out/c_host_example.log 2>&1; echo "example exit $?"; cat gpurun_out/c_host_example.log
timeout 100 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/last_cfg4.json 2> gpurun_out/last_cfg4.err; tail -c 400 gpurun_out/last_cfg4.json

"""bench.py's `parity` block: the oracle outputs of the cpu_baseline leg are compared with the GPU
results of the same queries on the same index, at the benchmarked size."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _mk(nq=6, k=4):
    rd = np.sort(np.random.default_rng(0).uniform(1, 2, size=(nq, k)).astype(np.float32), axis=1)
    ri = np.arange(nq * k, dtype=np.uint64).reshape(nq, k)
    rc = np.full(nq, k, dtype=np.uint32)
    rp = np.tile(np.arange(3, dtype=np.uint32), (nq, 1))
    return rd, ri, rc, rp


def test_parity_block_identical():
    import bench
    rd, ri, rc, rp = _mk()
    out = bench.parity_block((rd, ri, rc, rp), (rd.copy(), ri.copy(), rc.copy()), 4, gpu_probes=rp.copy())
    assert out["queries_checked"] == 6 and out["id_mismatches"] == 0 and out["unexplained_mismatches"] == 0
    assert out["probe_set_mismatches"] == 0 and out["max_rel_err"] == 0.0 and out["distance_bits_identical"]


def test_parity_block_flags_ties_and_errors():
    import bench
    rd, ri, rc, rp = _mk()
    gd, gi, gc = rd.copy(), ri.copy(), rc.copy()
    gi[0, 1], gi[0, 2] = ri[0, 2], ri[0, 1]
    gd[0, 1] = gd[0, 2] = rd[0, 1] = rd[0, 2]            # a swapped exact tie: explained
    gi[1, 0] = 999                                        # a wrong row with a different distance: unexplained
    gd[1, 0] = rd[1, 0] * np.float32(1.01)
    gp = rp.copy()
    gp[2] = [0, 1, 5]                                     # another probed partition
    gp[3] = [1, 0, 2]                                     # same set, other order
    gc[4] = 3
    out = bench.parity_block((rd, ri, rc, rp), (gd, gi, gc), 4, gpu_probes=gp)
    assert out["id_mismatches"] == 3 and out["explained_ties"] == 2
    assert out["unexplained_mismatches"] == 1 + 1         # the wrong row + the count mismatch
    assert out["probe_set_mismatches"] == 1 and out["probe_order_mismatches"] == 2
    assert out["max_rel_err"] > 5e-3 and not out["distance_bits_identical"]


@pytest.mark.gpu
@pytest.mark.parametrize("wl", ["tiny", "cfg2"])
def test_bench_line_carries_green_parity(wl):
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", wl, "--steps", "2", "--warmup", "3"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    par = line["parity"]
    assert par["queries_checked"] >= 32
    assert par["unexplained_mismatches"] == 0 and par["max_rel_err"] <= 1e-5
    assert par.get("probe_set_mismatches", 0) == 0
    assert line["roofline"]["frac"] > 0 and line["cpu_baseline"]["value"] > 0

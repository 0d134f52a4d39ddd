"""examples/dropin_host.cpp (the host that drives an index the way the reference's exec nodes do) against the oracle's
own batched search: find_partitions + per-partition search_in_partition on a thread pool + merge must reproduce
VectorIndex::search exactly (ivf/v2.rs:368-401 is that very loop)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_dropin_host_reproduces_oracle_search(oracle, metric):
    import bench
    from lance_b200.synth import build_ivf_pq
    rng = np.random.default_rng(4)
    vecs = rng.normal(size=(6000, 32)).astype(np.float32)
    data = build_ivf_pq(vecs, metric, nlist=16, M=8, seed=1, ivf_iters=3, pq_iters=2, sample=2048)
    q = vecs[:40] + 0.1 * rng.normal(size=(40, 32)).astype(np.float32)
    o = oracle.OracleIndex(data)
    olib = oracle.lib()
    be = bench._DropinBackend(1, C.addressof(o.ix), C.cast(olib.lo_find_partitions, C.c_void_p),
                              C.cast(olib.lo_search_in_partition, C.c_void_p), {"l2": 0, "cosine": 1, "dot": 2}[metric],
                              32, 16, data.centroids.ctypes.data_as(C.c_void_p))
    lib = bench._dropin_lib()
    for threads, clients in ((1, 1), (6, 1), (4, 3)):
        od, oi, oc, lat, wall = bench._dropin_call(lib, be, q, 5, 10, threads, clients)
        rd, ri, rc, _ = o.search(q, k=10, nprobes=5)
        assert oc.tolist() == rc.tolist() and oi.tolist() == ri.tolist()
        assert od.view(np.uint32).tolist() == rd.view(np.uint32).tolist()
        assert wall > 0 and np.all(lat > 0)

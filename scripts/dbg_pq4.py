import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import oracle; oracle.build()
from test_gpu_pq4 import _index4
from lance_b200 import IvfPqIndex, Query
sizes = [0, 1, 15, 16, 17, 100, 333, 0, 1000, 47, 256, 31]
data = _index4("dot", 64, len(sizes), 8, sizes, seed=72, keep_raw=True)
q = np.random.default_rng(1).normal(size=(13, 64)).astype(np.float32)
ix, ox = IvfPqIndex(data), oracle.OracleIndex(data)
ids = ix.find_partitions(q, 3)
for i in range(13):
    r, _ = oracle.find_partitions(data.centroids, q[i], 3, "dot")
    if ids[i].tolist() != r.tolist(): print("PROBES differ q", i, ids[i], r)
bad = 0
for p in range(len(sizes)):
    parts = np.full((13, 1), p, np.uint32)
    od, oi, oc = ix.search_in_partition(parts, q, 5)
    for i in range(13):
        rd, rr = ox.search_in_partition(p, q[i], 5)
        n = int(oc[i, 0])
        if n != len(rr) or oi[i, 0, :n].tolist() != rr.tolist() or od[i, 0, :n].view(np.uint32).tolist() != rd.view(np.uint32).tolist():
            bad += 1
            if bad <= 4:
                print("PAIR differs q", i, "p", p, "n_p", sizes[p]); print("  gpu", od[i, 0, :n], oi[i, 0, :n]); print("  ref", rd, rr)
                t = oracle.build_distance_table4("dot", data.codebook, q[i]); qmin, qmax, qt = oracle.quantize_distance_table(t)
                print("  qmin", qmin, "qmax", qmax, "table range", t.min(), t.max())
print("bad pairs", bad)
gd, gi, gc = ix.search(Query(q, k=1, nprobes=1)); rd, ri, rc, pr = ox.search(q, 1, 1)
print("search k1 P1 gpu", gi[:, 0].tolist()); print("ref", ri[:, 0].tolist(), "probes", pr[:, 0].tolist())

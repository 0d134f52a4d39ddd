"""recall@10 of the GPU IVF_PQ path on REAL-BUILT indexes (trained IVF + PQ, SURVEY 8(d) data shapes),
ground truth = the GPU flat kNN (lb200_flat_topk) on the raw vectors.  The bench's throughput
workloads use uniform random codes (scan cost is value independent), so recall is measured here.
Writes one JSON object per case to stdout / gpurun_out/recall.json.  Needs a GPU."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lance_b200 import IvfPqIndex, Query  # noqa: E402
from lance_b200.index import flat_knn  # noqa: E402
from lance_b200.synth import build_ivf_pq, embedding_like, gaussian, sift_like  # noqa: E402

def lowrank_blobs(N, d, seed, rank=12):
    """Blobs whose within-cluster spread lives in a rank-12 subspace (+ small isotropic noise): the
    neighbourhood structure real descriptors have and isotropic 128-d Gaussian noise has not
    (there all neighbours are nearly equidistant, so PQ recall is low for ANY implementation)."""
    rng = np.random.default_rng(seed)
    centres = np.random.default_rng(1234).uniform(0, 128, size=(256, d)).astype(np.float32)
    basis = np.random.default_rng(99).normal(size=(rank, d)).astype(np.float32)
    which = rng.integers(0, 256, size=N)
    x = centres[which] + 12.0 * (rng.normal(size=(N, rank)).astype(np.float32) @ basis)
    return (x + rng.normal(0, 1.0, size=(N, d)).astype(np.float32)).astype(np.float32)


CASES = [
    ("low-intrinsic-dimension blobs 500kx128 L2 IVF256,PQ16", lowrank_blobs, 500_000, 128, "l2", 256, 16, (20,), (0, 5)),
    # name, generator, N, d, metric, nlist, M, nprobes list, refine list
    ("cfg1-shaped SIFT-like 1Mx128 L2 IVF256,PQ16", sift_like, 1_000_000, 128, "l2", 256, 16, (20,), (0, 5)),
    ("cfg3-shaped embedding-like 200kx256 cosine IVF256,PQ16", embedding_like, 200_000, 256, "cosine", 256, 16,
     (1, 20, 26), (0, 5)),
    ("cfg5-shaped gaussian 200kx128 dot IVF128,PQ16", gaussian, 200_000, 128, "dot", 128, 16, (32,), (0, 5)),
]


def main():
    out = []
    Q, k = 1000, 10
    for name, gen, N, d, metric, nlist, M, probes, refines in CASES:
        t0 = time.time()
        base = gen(N, d, 1234)
        queries = gen(Q, d, 4321)
        data = build_ivf_pq(base, metric, nlist, M, seed=7, keep_raw=True)
        t_build = time.time() - t0
        # ground truth on the raw vectors in index order; positions -> row ids
        gd, gpos, gc = flat_knn(queries, metric, data.raw_vectors, k)
        truth = data.row_ids[gpos.astype(np.int64)]
        ix = IvfPqIndex(data)
        for P in probes:
            for rf in refines:
                res = ix.search(Query(key=queries, k=k, nprobes=P, refine_factor=rf or None))
                ids = res[1]
                hit = sum(len(set(ids[i, :k].tolist()) & set(truth[i].tolist())) for i in range(Q))
                rec = {"case": name, "metric": metric, "N": N, "dim": d, "nlist": nlist, "M": M, "queries": Q, "k": k,
                       "nprobes": P, "refine_factor": rf, "recall_at_10": hit / (Q * k), "index_build_s": round(t_build, 1)}
                print(json.dumps(rec), flush=True)
                out.append(rec)
        ix.close()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/recall.json", "w"), indent=1)


if __name__ == "__main__":
    main()

"""Generates tests/golden/ivfpq_small.npz: a small seeded IVF_PQ index per metric and the CPU
oracle's answers for it.  The Rust reference cannot run in this image (no cargo), so these
vectors come from the oracle, which is itself pinned to the reference's known-answer tests
(tests/test_oracle_golden.py).  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from lance_b200.synth import build_ivf_pq  # noqa: E402

N, D, NLIST, M, Q = 2000, 64, 12, 8, 16
rng = np.random.default_rng(20261104)
vecs = (rng.normal(size=(N, D)) + rng.normal(size=(1, D))).astype(np.float32)
queries = (vecs[rng.integers(0, N, Q)] + 0.1 * rng.normal(size=(Q, D))).astype(np.float32)
out = {"dim": D, "nlist": NLIST, "M": M, "queries": queries}
for metric in ("l2", "cosine", "dot"):
    data = build_ivf_pq(vecs, metric, NLIST, M, seed=5, keep_raw=True, ivf_iters=4, pq_iters=3)
    o = oracle.OracleIndex(data)
    out[f"{metric}_centroids"] = data.centroids
    out[f"{metric}_codebook"] = data.codebook
    out[f"{metric}_part_offsets"] = data.part_offsets
    out[f"{metric}_codes"] = data.codes
    out[f"{metric}_row_ids"] = data.row_ids
    out[f"{metric}_raw"] = data.raw_vectors
    for refine in (0, 3):
        d, i, c, p = o.search(queries, k=10, nprobes=6, refine_factor=refine)
        out[f"{metric}_r{refine}_dist"] = d
        out[f"{metric}_r{refine}_ids"] = i
        out[f"{metric}_r{refine}_cnt"] = c
        out[f"{metric}_r{refine}_probes"] = p
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ivfpq_small.npz"), **out)
print("written", {k: getattr(v, "shape", v) for k, v in out.items() if k.startswith("l2")})

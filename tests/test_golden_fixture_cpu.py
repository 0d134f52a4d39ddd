"""CPU: the committed golden fixture still equals what the oracle computes today."""
import os

import numpy as np

from lance_b200.synth import IvfPqData


def test_oracle_reproduces_committed_fixture(oracle):
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ivfpq_small.npz"))
    for metric in ("l2", "cosine", "dot"):
        data = IvfPqData(metric, int(g["dim"]), int(g["nlist"]), int(g["M"]), g[f"{metric}_centroids"],
                         g[f"{metric}_codebook"], g[f"{metric}_part_offsets"], g[f"{metric}_codes"],
                         g[f"{metric}_row_ids"], g[f"{metric}_raw"])
        o = oracle.OracleIndex(data)
        for refine in (0, 3):
            d, i, c, p = o.search(g["queries"], k=10, nprobes=6, refine_factor=refine)
            assert d.view(np.uint32).tolist() == g[f"{metric}_r{refine}_dist"].view(np.uint32).tolist()
            assert i.tolist() == g[f"{metric}_r{refine}_ids"].tolist()
            assert c.tolist() == g[f"{metric}_r{refine}_cnt"].tolist()
            assert p.tolist() == g[f"{metric}_r{refine}_probes"].tolist()

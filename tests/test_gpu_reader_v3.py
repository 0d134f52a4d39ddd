"""A v3 IVF_PQ index (index.idx + auxiliary.idx) opened natively and searched on the GPU equals the oracle's search of
the same index (lb200_index_open_v3: parse + upload in one call)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from lance_v3_writer import write_v3  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("metric,transposed", [("l2", True), ("cosine", False), ("dot", True)])
def test_search_on_v3_index_files(oracle, metric, transposed, tmp_path):
    from lance_b200 import Query
    from lance_b200.reader import open_index_v3
    from lance_b200.synth import build_ivf_pq
    rng = np.random.default_rng(2)
    vecs = rng.normal(size=(9000, 64)).astype(np.float32)
    data = build_ivf_pq(vecs, metric, nlist=20, M=16, seed=3, keep_raw=True, ivf_iters=3, pq_iters=2, sample=4096)
    idx, aux = write_v3(data, transposed=transposed, rows_per_page=1000)
    (tmp_path / "index.idx").write_bytes(idx)
    (tmp_path / "auxiliary.idx").write_bytes(aux)
    ix = open_index_v3(str(tmp_path / "index.idx"), str(tmp_path / "auxiliary.idx"), raw_vectors=data.raw_vectors)
    q = vecs[:33] + 0.05 * rng.normal(size=(33, 64)).astype(np.float32)
    o = oracle.OracleIndex(data)
    for refine in (0, 3):
        gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=6, refine_factor=refine or None))
        rd, ri, rc, _ = o.search(q, k=10, nprobes=6, refine_factor=refine)
        assert gc.tolist() == np.minimum(rc, 10).tolist()
        assert gi.tolist() == ri[:, :10].tolist()
        assert gd.view(np.uint32).tolist() == rd[:, :10].view(np.uint32).tolist()
    ix.close()

"""Multi-GPU form behind the C ABI (lb200_shard_* / lb200_group_*) exercised on ONE device: G shard
handles of one process exchange probe lists and candidate lists through their exchange blocks (same-process peer
pointers) and meet through epoch flags, each on its own stream.  The merged result must be bit-identical to the oracle's
search of the unsharded index."""
import numpy as np
import pytest
import torch

from lance_b200.synth import build_ivf_pq

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _data(metric, refine, seed=5):
    rng = np.random.default_rng(seed)
    d, nlist, M = 64, 24, 16
    vecs = rng.normal(size=(12000, d)).astype(np.float32)
    data = build_ivf_pq(vecs, metric, nlist, M, seed=2, keep_raw=refine > 0, ivf_iters=3, pq_iters=2, sample=4096)
    q = vecs[rng.integers(0, len(vecs), 45)] + 0.05 * rng.normal(size=(45, d)).astype(np.float32)
    return data, q


@pytest.mark.parametrize("metric,refine,G", [("l2", 0, 2), ("l2", 4, 3), ("cosine", 3, 2), ("dot", 0, 4)])
def test_shards_exchange_through_peer_memory(oracle, metric, refine, G):
    from lance_b200 import IvfPqIndex
    from lance_b200.shard import shard_of
    data, q = _data(metric, refine)
    k, nprobes, Q = 10, 9, len(q)
    shards = [IvfPqIndex(shard_of(data, r, G)[0]) for r in range(G)]
    ptrs = [ix.shard_export(Q, nprobes, k * max(refine, 1))[1] for ix in shards]
    for r, ix in enumerate(shards):
        ix.shard_attach(G, r, ptrs=ptrs)
    dev = torch.device("cuda", 0)
    d_q = torch.from_numpy(q).to(dev)
    od = torch.empty((Q, k), dtype=torch.float32, device=dev)
    oi = torch.empty((Q, k), dtype=torch.int64, device=dev)
    oc = torch.empty((Q,), dtype=torch.int32, device=dev)
    streams = [torch.cuda.Stream(device=dev) for _ in range(G)]
    torch.cuda.synchronize()
    rd, ri, rc, _ = oracle.OracleIndex(data).search(q, k=k, nprobes=nprobes, refine_factor=refine)
    for batch in range(3):                      # several batches: epochs, parity of the double-buffered tables
        for r in reversed(range(G)):            # the root last: its peers' kernels wait for it on the device
            shards[r].shard_search_device(d_q.data_ptr(), Q, nprobes, k, refine, od.data_ptr(), oi.data_ptr(), oc.data_ptr(),
                                          stream=streams[r].cuda_stream)
        torch.cuda.synchronize()
        for ix in shards:
            ix.shard_check()
        gd, gi, gc = od.cpu().numpy(), oi.cpu().numpy().view(np.uint64), oc.cpu().numpy()
        assert gc.tolist() == np.minimum(rc, k).tolist()
        for qi in range(Q):
            n = int(gc[qi])
            assert gi[qi, :n].tolist() == ri[qi, :n].tolist(), (batch, qi)
            assert _bits(gd[qi, :n]).tolist() == _bits(rd[qi, :n]).tolist(), (batch, qi)
    for ix in shards:
        ix.shard_detach()
        ix.close()


@pytest.mark.parametrize("metric,refine,G", [("l2", 0, 3), ("cosine", 2, 2), ("dot", 0, 1)])
def test_group_search_equals_single_index(oracle, metric, refine, G):
    from lance_b200 import IvfPqGroup, Query
    data, q = _data(metric, refine, seed=8)
    grp = IvfPqGroup(data, [0] * G, max_queries=64, max_nprobes=12, max_kprime=40)
    rd, ri, rc, _ = oracle.OracleIndex(data).search(q, k=10, nprobes=7, refine_factor=refine)
    for nq in (len(q), 5, 1):
        gd, gi, gc = grp.search(Query(key=q[:nq], k=10, nprobes=7, refine_factor=refine or None))
        assert gc.tolist() == np.minimum(rc[:nq], 10).tolist()
        for qi in range(nq):
            n = int(gc[qi])
            assert gi[qi, :n].tolist() == ri[qi, :n].tolist()
            assert _bits(gd[qi, :n]).tolist() == _bits(rd[qi, :n]).tolist()
    from lance_b200 import LanceB200Error
    with pytest.raises(LanceB200Error, match="exceeds the exported capacity"):
        grp.search(Query(key=np.zeros((65, 64), np.float32), k=10, nprobes=7))
    grp.close()

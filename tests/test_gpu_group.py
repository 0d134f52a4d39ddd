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


@pytest.mark.parametrize("kind,metric,refine,G", [("pq4", "l2", 0, 2), ("pq4", "cosine", 2, 3), ("flat", "l2", 0, 3), ("flat", "dot", 2, 2)])
def test_group_search_on_4bit_and_ivf_flat_handles(oracle, kind, metric, refine, G):
    """The peer-memory exchange with 4-bit PQ and IVF_FLAT shards (shard contract of csrc/api_pq4.inl): one
    lb200_group_search call == the single-index search, bit for bit."""
    from lance_b200 import IvfPqGroup, Query
    from lance_b200.synth import IvfPqData
    rng = np.random.default_rng(17)
    d, sizes = 32, [300, 0, 77, 512, 1, 950, 64, 200]
    nlist, N = len(sizes), sum(sizes)
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    cents = rng.normal(size=(nlist, d)).astype(np.float32)
    raw = (cents[np.repeat(np.arange(nlist), sizes)] + 0.3 * rng.normal(size=(N, d))).astype(np.float32)
    rids = rng.permutation(N * 2)[:N].astype(np.uint64)
    if kind == "pq4":
        M = 8
        cb = rng.normal(size=(M, 16, d // M)).astype(np.float32)
        codes = np.concatenate([rng.integers(0, 256, size=(M // 2, s), dtype=np.uint8).ravel() for s in sizes])
        data = IvfPqData(metric, d, nlist, M, cents, cb, offs, codes, rids, raw, 4)
        want = lambda x, k, P, rf: oracle.OracleIndex(data).search(x[None, :], k, P, rf or 0)
    else:
        data = IvfPqData(metric, d, nlist, 1, cents, np.zeros((1, 256, d), np.float32), offs, np.zeros(0, np.uint8), rids, raw, 32)
        want = None
    q = np.concatenate([raw[::397], rng.normal(size=(3, d)).astype(np.float32)])
    k, P = 10, 5
    grp = IvfPqGroup(data, [0] * G, max_queries=32, max_nprobes=8, max_kprime=40)
    for batch in range(2):
        gd, gi, gc = grp.search(Query(key=q, k=k, nprobes=P, refine_factor=refine or None))
        for j in range(len(q)):
            if kind == "pq4":
                rd, ri, rc, _ = want(q[j], k, P, refine)
                rd, ri = rd[0, :min(int(rc[0]), k)], ri[0, :min(int(rc[0]), k)]
            else:
                rd, ri = oracle.ivf_flat_search(data, q[j], k, P, refine or None)
            n = int(gc[j])
            assert n == len(ri) and gi[j, :n].tolist() == ri.tolist(), (kind, metric, batch, j)
            assert _bits(gd[j, :n]).tolist() == _bits(rd).tolist()
    grp.close()

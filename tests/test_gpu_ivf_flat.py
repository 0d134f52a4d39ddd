"""GPU parity of the IVF_FLAT sub-index (IVFIndex<FlatIndex, FlatQuantizer>: rust/lance-index/src/vector/flat/
{index.rs:76-137, storage.rs:315-375}, quantizer.rs:46) behind the same entry points (`lb200_index_create(num_bits =
32)`): exact distances on the raw rows of the probed partitions, bit-identical to the oracle."""
import numpy as np
import pytest
import torch

from lance_b200.synth import IvfPqData

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _flat_index(metric, d, sizes, seed):
    rng = np.random.default_rng(seed)
    sizes = np.asarray(sizes, dtype=np.int64)
    nlist, N = len(sizes), int(sizes.sum())
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    cents = rng.normal(size=(nlist, d)).astype(np.float32)
    part = np.repeat(np.arange(nlist), sizes)
    raw = (cents[part] + 0.3 * rng.normal(size=(N, d))).astype(np.float32)
    raw[3] = raw[1]                                    # an exact tie inside a partition -> ordered by row id
    return IvfPqData(metric, d, nlist, 1, cents, np.zeros((1, 256, d), np.float32), offs, np.zeros(0, np.uint8),
                     rng.permutation(N * 2)[:N].astype(np.uint64), raw, 32)


@pytest.mark.parametrize("metric,d", [("l2", 128), ("l2", 20), ("cosine", 96), ("cosine", 1536), ("dot", 64), ("dot", 70)])
def test_ivf_flat_search_parity(oracle, metric, d):
    from lance_b200 import IvfPqIndex, Query
    sizes = [50, 0, 1, 300, 17, 1200, 64]
    data = _flat_index(metric, d, sizes, seed=d)
    rng = np.random.default_rng(1)
    q = np.concatenate([data.raw_vectors[::211], rng.normal(size=(5, d)).astype(np.float32)])
    ix = IvfPqIndex(data)
    dall = np.concatenate([oracle.ivf_flat_search(data, x, 50, len(sizes))[0] for x in q])
    lo, hi = float(np.quantile(dall, 0.2)), float(np.quantile(dall, 0.8))
    for k, nprobes, rf, lo_, hi_ in ((1, 1, None, None, None), (10, 3, None, None, None), (100, len(sizes), None, None, None),
                                     (10, 4, None, lo, hi), (5, 3, 4, None, None), (10, 9, 2, None, hi)):
        gd, gi, gc = ix.search(Query(q, k=k, nprobes=nprobes, refine_factor=rf, lower_bound=lo_, upper_bound=hi_))
        for j in range(len(q)):
            rd, rr = oracle.ivf_flat_search(data, q[j], k, nprobes, rf, lo_, hi_)
            n = int(gc[j])
            what = f"{metric} d={d} k={k} P={nprobes} rf={rf} q{j}"
            assert n == len(rr), what
            assert gi[j, :n].tolist() == rr.tolist(), what
            assert _bits(gd[j, :n]).tolist() == _bits(rd).tolist(), what
    assert ix.stats()["scan_variant"] == 32
    # search_in_partition: explicit partitions, key used as given
    key = np.stack([oracle.normalize(x) for x in q]) if metric == "cosine" else q
    parts = np.tile(np.array([[5, 0, 1]], dtype=np.uint32), (len(q), 1))
    od, oi, oc = ix.search_in_partition(parts, key, 7)
    for j in range(len(q)):
        for c, p in enumerate((5, 0, 1)):
            o, e = int(data.part_offsets[p]), int(data.part_offsets[p + 1])
            rd, rr = oracle.flat_knn_rows(metric, key[j], data.raw_vectors[o:e], data.row_ids[o:e], 7)
            n = int(oc[j, c])
            assert n == len(rr) and oi[j, c, :n].tolist() == rr.tolist()
            assert _bits(od[j, c, :n]).tolist() == _bits(rd).tolist()
    # an indexed row's own vector, its partition probed: that row first (L2: at distance exactly 0)
    gd, gi, gc = ix.search(Query(data.raw_vectors[10:20], k=1, nprobes=len(sizes)))
    if metric == "l2":
        assert gd[:, 0].tolist() == [0.0] * 10
    ix.close()


@pytest.mark.parametrize("metric,d", [("l2", 128), ("l2", 20), ("cosine", 96), ("dot", 70)])
def test_ivf_flat_prefilter_and_shards(oracle, metric, d):
    """FlatIndex::search with a prefilter (flat/index.rs:106-122) on IVF_FLAT handles, as host masks and as a
    device-resident Prefilter; then the shard contract (merged k' list + exact distances) against the whole index."""
    from lance_b200 import IvfPqIndex
    from lance_b200.index import merge_topk_device
    from lance_b200.shard import assign_partitions
    sizes = [50, 0, 1, 300, 17, 1200, 64]
    data = _flat_index(metric, d, sizes, seed=d + 1)
    rng = np.random.default_rng(2)
    q = np.concatenate([data.raw_vectors[::311], rng.normal(size=(3, d)).astype(np.float32)])
    key = np.stack([oracle.normalize(x) for x in q]) if metric == "cosine" else q
    ix = IvfPqIndex(data)
    masks = {0: rng.random(50) < 0.5, 3: rng.random(300) < 0.05, 5: rng.random(1200) < 0.7, 6: np.zeros(64, bool)}
    parts = np.tile(np.array([[5, 0, 3, 6, 4]], dtype=np.uint32), (len(q), 1))
    pf = ix.prefilter(masks)
    for how in (masks, pf):
        od, oi, oc = ix.search_in_partition(parts, key, 9, prefilter=how)
        for j in range(len(q)):
            for c, p in enumerate(parts[j].tolist()):
                o, e = int(data.part_offsets[p]), int(data.part_offsets[p + 1])
                m = masks.get(p, np.ones(e - o, bool))
                rd, rr = oracle.flat_knn_rows(metric, key[j], data.raw_vectors[o:e][m], data.row_ids[o:e][m], 9)
                n = int(oc[j, c])
                assert n == len(rr) and oi[j, c, :n].tolist() == rr.tolist(), (metric, j, p)
                assert _bits(od[j, c, :n]).tolist() == _bits(rd).tolist()
    pf.close()
    ix.close()
    G, k, nprobes, Q = 3, 10, 4, len(q)
    owner = assign_partitions(np.asarray(sizes, dtype=np.int64), G)
    dev = torch.device("cuda", 0)
    dq = torch.from_numpy(q).to(dev)
    for refine in (0, 2):
        kp = k * max(refine, 1)
        g_d = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
        g_i = torch.empty((G, Q, kp), dtype=torch.int64, device=dev)
        g_c = torch.empty((G, Q), dtype=torch.int32, device=dev)
        g_e = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
        shards = [IvfPqIndex(data.subset(np.nonzero(owner == g)[0])) for g in range(G)]
        for g, sh in enumerate(shards):
            sh.search_device(dq.data_ptr(), Q, nprobes, k, refine, g_d[g].data_ptr(), g_i[g].data_ptr(), g_c[g].data_ptr(),
                             d_out_exact=g_e[g].data_ptr())
        torch.cuda.synchronize()
        fd = torch.empty((Q, k), dtype=torch.float32, device=dev)
        fi = torch.empty((Q, k), dtype=torch.int64, device=dev)
        fc = torch.empty((Q,), dtype=torch.int32, device=dev)
        merge_topk_device(0, g_d.data_ptr(), g_i.data_ptr(), g_e.data_ptr() if refine else None, g_c.data_ptr(), G, Q, kp, kp,
                          k, fd.data_ptr(), fi.data_ptr(), fc.data_ptr())
        torch.cuda.synchronize()
        fd, fi, fc = fd.cpu().numpy(), fi.cpu().numpy().astype(np.uint64), fc.cpu().numpy()
        for j in range(Q):
            rd, rr = oracle.ivf_flat_search(data, q[j], k, nprobes, refine or None)
            n = int(fc[j])
            assert n == len(rr) and fi[j, :n].tolist() == rr.tolist(), (metric, refine, j)
            assert _bits(fd[j, :n]).tolist() == _bits(rd).tolist()
        for sh in shards:
            sh.close()


def test_ivf_flat_needs_raw_vectors():
    from lance_b200 import IvfPqIndex, LanceB200Error
    data = _flat_index("l2", 8, [4, 4], seed=0)
    data.raw_vectors = None
    with pytest.raises(LanceB200Error, match="raw vectors"):
        IvfPqIndex(data)

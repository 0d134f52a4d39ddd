"""GPU parity for the callers either side of the hot path (SURVEY 8(f)-2,3): indexes loaded from files the
REFERENCE wrote, the union over delta indices, the re-rank of caller-fetched candidates, and the whole
Scanner::nearest plan (fan-out, refine through take, knn_combined) — each against the oracle's restatement,
bit-identical distances and identical row ids.
"""
import os

import numpy as np
import pytest
import torch

from lance_b200.reader import open_legacy_ivf_pq
from lance_b200.synth import build_ivf_pq

pytestmark = pytest.mark.gpu
from conftest import lance_files_dir  # noqa: E402

G = lance_files_dir()


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from lance_b200 import _lib
    assert _lib.lib() is not None


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _raw(tag):
    z = np.load(os.path.join(G, f"{tag}_raw_vectors.npz"))
    lut = {int(r): i for i, r in enumerate(z["row_ids"].tolist())}
    vec = z["vectors"]
    return (lambda ids: vec[[lut[int(r)] for r in np.asarray(ids).reshape(-1).tolist()]]), z["row_ids"], vec


def _same_row(dg, ig, n, dr, ir, what):
    assert n == len(ir), f"{what}: count {n} != {len(ir)}"
    assert ig[:n].tolist() == ir.tolist(), f"{what}: row ids differ"
    assert _bits(dg[:n]).tolist() == _bits(dr).tolist(), f"{what}: distances differ"


# ---- indexes written by the reference ----------------------------------------------------------------------
@pytest.mark.parametrize("tag,u", [("v0.8.14", "d32dac97"), ("v0.8.14", "f358a219"), ("v0.21.0", "dc833a6e"),
                                   ("v0.21.0", "ca9b1111")])
def test_search_on_reference_written_index(oracle, tag, u):
    from lance_b200 import IvfPqIndex, Query
    take, _, vec = _raw(tag)
    li = open_legacy_ivf_pq(os.path.join(G, f"{tag}_{u}_index.idx"), raw_lookup=take)
    data = li.data
    ox = oracle.OracleIndex(data)
    rng = np.random.default_rng(5)
    # the reference test's own key (all zeros, dataset.rs:4326-4329), rows of the dataset, and noise around them
    near = vec[::67][:12]
    q = np.concatenate([np.zeros((1, data.dim), np.float32), vec[::61][:12],
                        (near + rng.normal(scale=0.05, size=near.shape)).astype(np.float32)])
    ix = IvfPqIndex(data)
    for k, nprobes, rf in ((1, 1, 0), (10, 1, 0), (10, 4, 0), (100, 4, 0), (10, 4, 5), (7, 2, 3)):
        for lo, hi in ((None, None), (60.0, 170.0) if data.dim == 128 else (0.5, 3.0)):
            gd, gi, gc = ix.search(Query(q, k=k, nprobes=nprobes, refine_factor=rf or None, lower_bound=lo,
                                         upper_bound=hi))
            rd, ri, rc, _ = ox.search(q, k, nprobes, rf, lo, hi)
            assert gc.tolist() == rc.tolist()
            for j in range(len(q)):
                _same_row(gd[j], gi[j], int(gc[j]), rd[j, :rc[j]], ri[j, :rc[j]], f"{tag}/{u} k={k} P={nprobes} rf={rf} q{j}")
    # domain property: an indexed row's own vector, everything probed, refined -> that row at distance 0
    pick = np.arange(0, data.N, max(1, data.N // 16))
    gd, gi, gc = ix.search(Query(data.raw_vectors[pick], k=1, nprobes=data.nlist, refine_factor=min(20, data.N)))
    assert gd[:, 0].tolist() == [0.0] * len(pick)
    ix.close()


# ---- lb200_rerank ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("metric,d", [("l2", 128), ("l2", 16), ("l2", 100), ("cosine", 96), ("cosine", 1536),
                                      ("dot", 768), ("dot", 20)])
def test_rerank_parity(oracle, metric, d):
    from lance_b200.index import rerank
    rng = np.random.default_rng(d)
    Q = 9
    for Cn, k in ((1, 1), (37, 10), (300, 25), (64, 100)):
        key = rng.normal(size=(Q, d)).astype(np.float32)
        cv = rng.normal(size=(Q, Cn, d)).astype(np.float32)
        cr = rng.integers(0, 1 << 40, size=(Q, Cn)).astype(np.uint64)
        cr[:, Cn // 2] = cr[:, 0]                      # a duplicated row id must not break the order
        cv[0, Cn - 1] = cv[0, 0]                       # an exact distance tie -> ordered by row id
        cc = rng.integers(0, Cn + 1, size=Q).astype(np.uint32)
        cc[0], cc[-1] = Cn, 0
        dall = np.concatenate([oracle.distance_batch(metric, key[i], cv[i]) for i in range(Q)])
        for lo, hi in ((None, None), (float(np.quantile(dall, 0.2)), float(np.quantile(dall, 0.8)))):
            gd, gi, gc = rerank(key, metric, cv, cr, cc, k, lo, hi)
            for i in range(Q):
                n = int(cc[i])
                rd, rr = oracle.flat_knn_rows(metric, key[i], cv[i, :n], cr[i, :n], k, lo, hi)
                _same_row(gd[i], gi[i], int(gc[i]), rd, rr, f"rerank {metric} d={d} C={Cn} k={k} q{i}")
    # empty candidate stride
    gd, gi, gc = rerank(key, metric, np.zeros((Q, 0, d), np.float32), np.zeros((Q, 0), np.uint64),
                        np.zeros(Q, np.uint32), 3)
    assert gc.tolist() == [0] * Q
    with pytest.raises(Exception, match="exceeds"):
        rerank(key, metric, cv, cr, np.full(Q, Cn + 1, np.uint32), 3)


def test_merge_topk_host(oracle):
    from lance_b200.index import merge_topk
    rng = np.random.default_rng(0)
    Gn, Q, stride = 3, 11, 20
    dist = np.sort(rng.normal(size=(Gn, Q, stride)).astype(np.float32), axis=2)
    dist[1, :, 3] = dist[0, :, 3]                      # ties across lists
    dist = np.sort(dist, axis=2)
    rid = rng.integers(0, 1 << 34, size=(Gn, Q, stride)).astype(np.uint64)
    cnt = rng.integers(0, stride + 1, size=(Gn, Q)).astype(np.uint32)
    cnt[:, 0] = 0
    for k in (1, 10, 20, 45):
        gd, gi, gc = merge_topk(dist, rid, cnt, k)
        for q in range(Q):
            d = np.concatenate([dist[g, q, :cnt[g, q]] for g in range(Gn)])
            r = np.concatenate([rid[g, q, :cnt[g, q]] for g in range(Gn)])
            o = np.lexsort((r, d))[:k]
            _same_row(gd[q], gi[q], int(gc[q]), d[o], r[o], f"merge k={k} q{q}")


# ---- Scanner::nearest ------------------------------------------------------------------------------------------
def _check_plan(oracle, ds, oxs, take, q, un_ids, un_vec, cases):
    from lance_b200 import Query
    for k, nprobes, rf, lo, hi, fast in cases:
        gd, gi, gc = ds.nearest(Query(q, k=k, nprobes=nprobes, refine_factor=rf, lower_bound=lo, upper_bound=hi),
                                fast_search=fast)
        for j in range(len(q)):
            rd, rr = oracle.nearest_plan(oxs, take, q[j], k, nprobes, rf, lo, hi, un_ids, un_vec, fast)
            _same_row(gd[j], gi[j], int(gc[j]), rd, rr, f"nearest k={k} P={nprobes} rf={rf} [{lo},{hi}) fast={fast} q{j}")


def test_nearest_over_reference_written_delta_indices(oracle):
    """The v0.21.0 dataset: `vector_idx` has two deltas (256 rows of fragment 0, 32 rows of fragment 1), both
    written by the reference; the plan fans out over both."""
    from lance_b200 import IvfPqIndex
    from lance_b200.scanner import VectorDataset
    take, ids, vec = _raw("v0.21.0")
    datas = [open_legacy_ivf_pq(os.path.join(G, f"v0.21.0_{u}_index.idx")).data for u in ("dc833a6e", "ca9b1111")]
    ixs = [IvfPqIndex(d) for d in datas]
    oxs = [oracle.OracleIndex(d) for d in datas]
    rng = np.random.default_rng(1)
    q = np.concatenate([vec[::23], rng.random(size=(8, 16)).astype(np.float32)])
    ds = VectorDataset(take, 16, ixs)
    _check_plan(oracle, ds, oxs, take, q, None, None,
                [(5, 1, None, None, None, False), (10, 1, 3, None, None, False), (300, 1, None, None, None, False),
                 (10, 1, 2, 0.2, 1.5, False), (1, 1, 30, None, None, True)])
    # refined self-query over the union of the deltas finds each row of either fragment
    from lance_b200 import Query
    gd, gi, gc = ds.nearest(Query(vec[::9], k=1, nprobes=1, refine_factor=288))
    assert gi[:, 0].tolist() == ids[::9].tolist() and gd[:, 0].tolist() == [0.0] * len(ids[::9])
    for ix in ixs:
        ix.close()


def test_nearest_knn_combined_on_reference_written_index(oracle):
    """v0.8.14: the first index covers fragments 0-9 (1000 rows); the 1000 rows of fragment 10 were appended later
    -> unindexed fragment -> knn_combined (scanner.rs:1755-1835)."""
    from lance_b200 import IvfPqIndex
    from lance_b200.scanner import VectorDataset
    take, ids, vec = _raw("v0.8.14")
    data = open_legacy_ivf_pq(os.path.join(G, "v0.8.14_d32dac97_index.idx")).data
    un = (ids >> np.uint64(32)) == np.uint64(10)
    assert un.sum() == 1000 and not np.isin(ids[un], data.row_ids).any()
    ix, ox = IvfPqIndex(data), oracle.OracleIndex(data)
    rng = np.random.default_rng(2)
    q = np.concatenate([np.zeros((1, 128), np.float32), vec[::173],
                        (vec[5::211] + rng.normal(scale=0.1, size=vec[5::211].shape)).astype(np.float32)])
    ds = VectorDataset(take, 128, [ix], ids[un], vec[un])
    _check_plan(oracle, ds, [ox], take, q, ids[un], vec[un],
                [(10, 4, None, None, None, False), (10, 2, 4, None, None, False), (3, 1, None, None, None, True),
                 (25, 4, 2, 1.0, 200.0, False), (10, 4, None, None, None, True)])
    ix.close()


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_nearest_plan_synthetic_three_deltas_and_unindexed(oracle, metric):
    from lance_b200 import IvfPqIndex
    from lance_b200.scanner import VectorDataset
    rng = np.random.default_rng(11)
    d, n = 32, 4000
    x = rng.normal(size=(n, d)).astype(np.float32)
    rid = (rng.permutation(1 << 22)[:n]).astype(np.uint64)
    lut = {int(r): i for i, r in enumerate(rid.tolist())}
    take = lambda ids: x[[lut[int(r)] for r in np.asarray(ids).reshape(-1).tolist()]]
    cuts = [0, 2000, 3000, 3600, n]
    datas = []
    for a, b in zip(cuts[:3], cuts[1:4]):
        datas.append(build_ivf_pq(x[a:b], metric, 8, 4, seed=a, row_ids=rid[a:b]))
    ixs = [IvfPqIndex(dd) for dd in datas]
    oxs = [oracle.OracleIndex(dd) for dd in datas]
    q = np.concatenate([x[::500], rng.normal(size=(6, d)).astype(np.float32)])
    ds = VectorDataset(take, d, ixs, rid[3600:], x[3600:])
    _check_plan(oracle, ds, oxs, take, q, rid[3600:], x[3600:],
                [(10, 3, None, None, None, False), (10, 3, 5, None, None, False), (10, 8, None, None, None, True),
                 (4, 2, 2, None, None, True)])
    # flat path: use_index = False
    from lance_b200 import Query
    gd, gi, gc = ds.nearest(Query(q, k=5, use_index=False, metric_type=metric))
    for j in range(len(q)):
        rd, rr = oracle.flat_knn_rows(metric, q[j], x[3600:], rid[3600:], 5)
        _same_row(gd[j], gi[j], int(gc[j]), rd, rr, f"flat q{j}")
    with pytest.raises(ValueError, match="Refine factor"):
        ds.nearest(Query(q, k=5, refine_factor=0))
    for ix in ixs:
        ix.close()


@pytest.mark.parametrize("tag,u", [("v0.8.14", "f358a219"), ("v0.21.0", "ca9b1111")])
def test_native_index_open_equals_python_reader_path(oracle, tag, u):
    """lb200_index_open (C++ parse + upload inside the library) against the oracle on the Python reader's arrays."""
    from lance_b200 import Query
    from lance_b200.reader import open_index
    take, _, vec = _raw(tag)
    path = os.path.join(G, f"{tag}_{u}_index.idx")
    data = open_legacy_ivf_pq(path, raw_lookup=take).data
    ox = oracle.OracleIndex(data)
    q = np.concatenate([np.zeros((1, data.dim), np.float32), vec[3::41][:20]])
    for raw in (None, data.raw_vectors):
        ix = open_index(path, raw_vectors=raw)
        assert (ix.metric_type(), ix.dim, ix.nlist, ix.M) == (data.metric, data.dim, data.nlist, data.M)
        for k, nprobes, rf in ((10, 2, 0), (5, data.nlist, 4 if raw is not None else 0)):
            gd, gi, gc = ix.search(Query(q, k=k, nprobes=nprobes, refine_factor=rf or None))
            rd, ri, rc, _ = ox.search(q, k, nprobes, rf)
            assert gc.tolist() == rc.tolist()
            for j in range(len(q)):
                _same_row(gd[j], gi[j], int(gc[j]), rd[j, :rc[j]], ri[j, :rc[j]], f"open {tag} k={k} rf={rf} q{j}")
        ix.close()


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_nearest_with_prefilter(oracle, metric):
    """A row-id allow-list (deletions / scalar filter -> RowIdMask, prefilter.rs:26-51) through the whole plan: masked
    index scan per probed partition, filtered flat scan of the unindexed rows, refine, knn_combined."""
    from lance_b200 import IvfPqIndex
    from lance_b200.scanner import VectorDataset
    rng = np.random.default_rng(23)
    d, n = 32, 3000
    x = rng.normal(size=(n, d)).astype(np.float32)
    rid = (rng.permutation(1 << 20)[:n]).astype(np.uint64)
    lut = {int(r): i for i, r in enumerate(rid.tolist())}
    take = lambda ids: x[[lut[int(r)] for r in np.asarray(ids).reshape(-1).tolist()]]
    datas = [build_ivf_pq(x[a:b], metric, 6, 4, seed=a, row_ids=rid[a:b]) for a, b in ((0, 1500), (1500, 2500))]
    ixs = [IvfPqIndex(dd) for dd in datas]
    oxs = [oracle.OracleIndex(dd) for dd in datas]
    q = np.concatenate([x[::400], rng.normal(size=(4, d)).astype(np.float32)])
    ds = VectorDataset(take, d, ixs, rid[2500:], x[2500:])
    from lance_b200 import Query
    for frac in (0.5, 0.02):
        allowed = rid[rng.random(n) < frac]
        for k, nprobes, rf, fast in ((10, 3, None, False), (5, 6, 3, False), (10, 2, None, True)):
            gd, gi, gc = ds.nearest(Query(q, k=k, nprobes=nprobes, refine_factor=rf), fast_search=fast, allowed_row_ids=allowed)
            for j in range(len(q)):
                rd, rr = oracle.nearest_plan(oxs, take, q[j], k, nprobes, rf, None, None, rid[2500:], x[2500:], fast,
                                             allowed=allowed)
                _same_row(gd[j], gi[j], int(gc[j]), rd, rr, f"prefilter {metric} frac={frac} k={k} P={nprobes} rf={rf} q{j}")
                assert np.isin(gi[j, :int(gc[j])], allowed).all()
    for ix in ixs:
        ix.close()

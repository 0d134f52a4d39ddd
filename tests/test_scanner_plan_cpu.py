"""Host logic of lance_b200/scanner.py on CPU: the plan plumbing (delta fan-out, union, take, refine, knn_combined,
prefilter masks) with every C-ABI call replaced by an oracle-backed stand-in, against oracle.nearest_plan -- two
independent restatements of rust/lance/src/dataset/scanner.rs:1670-1835.  The GPU tests (tests/test_gpu_scanner.py)
run the same plans through the real library; this file only makes sure the host-side bookkeeping (shapes, counts,
row-id <-> vector bookkeeping, mask construction) is exercised by the CPU suite too."""
import numpy as np
import pytest

from lance_b200.synth import build_ivf_pq

NONE = np.iinfo(np.uint64).max


@pytest.fixture()
def standins(oracle, monkeypatch):
    import lance_b200.scanner as S

    class FakeIx:
        def __init__(self, data):
            self.data, self.ox, self.metric = data, oracle.OracleIndex(data), data.metric
            self.has_raw, self.part_offsets, self.row_ids = False, data.part_offsets, data.row_ids

        def metric_type(self):
            return self.metric

        def search(self, q):
            od, oi, oc, _ = self.ox.search(np.atleast_2d(q.key), q.k, q.nprobes, 0, q.lower_bound, q.upper_bound)
            return od, oi, oc

        def find_partitions(self, key, nprobes):
            return np.stack([oracle.find_partitions(self.data.centroids, k, nprobes, self.metric)[0]
                             for k in np.atleast_2d(key)]).astype(np.uint32)

        def search_in_partition(self, parts, key, k, lo=None, hi=None, prefilter=None):
            Q, P = parts.shape
            od = np.full((Q, P, k), np.inf, np.float32)
            oi = np.full((Q, P, k), NONE, np.uint64)
            oc = np.zeros((Q, P), np.uint32)
            for i in range(Q):
                for j in range(P):
                    p, bits = int(parts[i, j]), None
                    if prefilter is not None:
                        m = np.asarray(prefilter[p], bool)
                        bits = np.zeros((len(m) + 63) // 64 + 1, np.uint64)
                        on = np.nonzero(m)[0]
                        np.bitwise_or.at(bits, on >> 6, np.uint64(1) << (on & 63).astype(np.uint64))
                    d, r = self.ox.search_in_partition(p, key[i], k, lo, hi, bitmask=bits)
                    oc[i, j] = len(r)
                    od[i, j, :len(r)], oi[i, j, :len(r)] = d, r
            return od, oi, oc

    def out(Q, k):
        return np.full((Q, k), np.inf, np.float32), np.full((Q, k), NONE, np.uint64), np.zeros(Q, np.uint32)

    def f_rerank(key, metric, cv, cr, cc, k, lo=None, hi=None, device_id=0):
        key = np.atleast_2d(key)
        od, oi, oc = out(len(key), k)
        for i in range(len(key)):
            n = int(cc[i])
            d, r = oracle.flat_knn_rows(metric, key[i], cv[i, :n], cr[i, :n], k, lo, hi)
            od[i, :len(d)], oi[i, :len(d)], oc[i] = d, r, len(d)
        return od, oi, oc

    def f_merge(dist, rid, cnt, k, device_id=0, **_):
        G, Q, _s = dist.shape
        od, oi, oc = out(Q, k)
        for q in range(Q):
            d = np.concatenate([dist[g, q, :cnt[g, q]] for g in range(G)])
            r = np.concatenate([rid[g, q, :cnt[g, q]] for g in range(G)])
            o = np.lexsort((r, d))[:k]
            od[q, :len(o)], oi[q, :len(o)], oc[q] = d[o], r[o], len(o)
        return od, oi, oc

    def f_flat(key, metric, vectors, k, validity=None, lower_bound=None, upper_bound=None, device_id=0):
        key = np.atleast_2d(key)
        od, oi, oc = out(len(key), k)
        for i in range(len(key)):
            d, r = oracle.flat_knn_rows(metric, key[i], vectors, np.arange(len(vectors), dtype=np.uint64), k,
                                        lower_bound, upper_bound)
            od[i, :len(d)], oi[i, :len(d)], oc[i] = d, r, len(d)
        return od, oi, oc

    monkeypatch.setattr(S, "rerank", f_rerank)
    monkeypatch.setattr(S, "merge_topk", f_merge)
    monkeypatch.setattr(S, "flat_knn", f_flat)
    monkeypatch.setattr(S, "normalize", lambda key, device_id=0: np.stack([oracle.normalize(k) for k in np.atleast_2d(key)]))
    return FakeIx


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_scanner_plan_host_logic(oracle, standins, metric):
    from lance_b200 import Query
    from lance_b200.scanner import VectorDataset
    rng = np.random.default_rng(5)
    d, n = 16, 1200
    x = rng.normal(size=(n, d)).astype(np.float32)
    rid = rng.permutation(1 << 16)[:n].astype(np.uint64)
    lut = {int(r): i for i, r in enumerate(rid.tolist())}
    take = lambda ids: x[[lut[int(r)] for r in np.asarray(ids).reshape(-1).tolist()]]
    datas = [build_ivf_pq(x[a:b], metric, 4, 4, seed=a, ivf_iters=3, pq_iters=2, row_ids=rid[a:b])
             for a, b in ((0, 500), (500, 900))]
    ixs = [standins(dd) for dd in datas]
    oxs = [ix.ox for ix in ixs]
    q = np.concatenate([x[::300], rng.normal(size=(2, d)).astype(np.float32)])
    ds = VectorDataset(take, d, ixs, rid[900:], x[900:])
    allowed = rid[rng.random(n) < 0.4]
    for k, nprobes, rf, fast, allow in ((5, 2, None, False, None), (5, 4, 3, False, None), (7, 2, None, True, None),
                                        (5, 3, 2, False, allowed), (300, 4, None, False, None)):
        gd, gi, gc = ds.nearest(Query(q, k=k, nprobes=nprobes, refine_factor=rf), fast_search=fast, allowed_row_ids=allow)
        for j in range(len(q)):
            rd, rr = oracle.nearest_plan(oxs, take, q[j], k, nprobes, rf, None, None, rid[900:], x[900:], fast, allowed=allow)
            c = int(gc[j])
            assert c == len(rr) and gi[j, :c].tolist() == rr.tolist(), (metric, k, nprobes, rf, fast, j)
            assert np.array_equal(gd[j, :c].view(np.uint32), np.asarray(rd, np.float32).view(np.uint32))
    # argument validation mirrors the scanner's errors
    with pytest.raises(ValueError, match="Refine factor"):
        ds.nearest(Query(q, k=5, refine_factor=0))
    with pytest.raises(ValueError, match="query dim"):
        ds.nearest(Query(np.zeros((1, d + 1), np.float32), k=5))
    with pytest.raises(ValueError, match="k must be positive"):
        ds.nearest(Query(q, k=0))
    with pytest.raises(ValueError, match="go together"):
        VectorDataset(take, d, ixs, rid[:3], None)

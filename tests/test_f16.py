"""Float16 vectors.

CPU part: the oracle's restatement of the reference's scalar f16 fallbacks against (a) the REFERENCE'S OWN f16 kernels,
compiled from rust/lance-linalg/src/simd/f16.c with build.rs's flags into oracle/_ref/libf16_ref.so (`make -C oracle
ref`; -ffast-math, so only a tolerance is meaningful), (b) an f64 evaluation, with the error bound the reference's own
tests use (dot.rs:333-341: 2 eps sum |x||y|), and (c) the reference's known answers that carry over to f16
(l2.rs:316-335 rows `i..i+8` vs `2..10` -> [32, 8, 0, 8]; l2.rs:434-439 the f16::MAX case stays finite in f32).
GPU part: lb200_flat_distance_f16 / lb200_flat_topk_f16 bit-identical to the oracle's restatement."""
import numpy as np
import pytest

METRICS = ("l2", "cosine", "dot")


def _f64(metric, q, v):
    q, v = q.astype(np.float64), v.astype(np.float64)
    if metric == "l2":
        return ((q - v) ** 2).sum(1)
    if metric == "dot":
        return 1.0 - v @ q
    return 1.0 - (v @ q) / np.linalg.norm(q) / np.linalg.norm(v, axis=1)


REL_TOL = 1e-5  # the north star's tolerance, taken relative to the magnitude sum (as dot.rs:333-341 does)


def _bound(metric, q, v):
    q, v = np.abs(q.astype(np.float64)), np.abs(v.astype(np.float64))
    if metric == "l2":
        return REL_TOL * ((q + v) ** 2).sum(1) + 1e-30
    if metric == "dot":
        return REL_TOL * (1.0 + v @ q)
    return REL_TOL * np.ones(len(v))


@pytest.mark.parametrize("metric", METRICS)
@pytest.mark.parametrize("d", [1, 8, 15, 16, 33, 100, 128, 1536])
def test_oracle_f16_against_reference_kernels_and_f64(oracle, metric, d):
    rng = np.random.default_rng(d)
    q = rng.normal(size=d).astype(np.float16)
    v = rng.normal(size=(64, d)).astype(np.float16)
    got = oracle.distance_batch_f16(metric, q, v).astype(np.float64)
    assert (np.abs(got - _f64(metric, q, v)) <= _bound(metric, q, v)).all()
    if oracle.ref_f16() is None:
        pytest.skip("oracle/_ref/libf16_ref.so not built (no reference tree on this box)")
    ref = oracle.ref_distance_batch_f16(metric, q, v).astype(np.float64)
    assert (np.abs(got - ref) <= _bound(metric, q, v)).all()


def test_oracle_f16_known_answers(oracle):
    rows = np.stack([np.arange(i, i + 8) for i in range(4)]).astype(np.float16)
    point = np.arange(2, 10).astype(np.float16)
    assert oracle.distance_batch_f16("l2", point, rows).tolist() == [32.0, 8.0, 0.0, 8.0]     # l2.rs:316-335
    x = np.full(4048, np.finfo(np.float16).max, dtype=np.float16)                              # l2.rs:434-439
    d = oracle.distance_batch_f16("l2", x, (-x)[None, :])[0]
    assert np.isfinite(d) and abs(d / (4048 * (2.0 * 65504.0) ** 2) - 1.0) < 1e-6
    if oracle.ref_f16() is not None:
        assert oracle.ref_distance_batch_f16("l2", point, rows).tolist() == [32.0, 8.0, 0.0, 8.0]


@pytest.mark.gpu
@pytest.mark.parametrize("metric", METRICS)
@pytest.mark.parametrize("d", [1, 8, 15, 33, 100, 128, 768])
def test_gpu_flat_f16_parity(oracle, metric, d):
    import torch
    assert torch.cuda.is_available()
    from lance_b200.index import compute_distance_f16, flat_knn_f16
    rng = np.random.default_rng(d + 7)
    n, Q = 1500, 7
    q = rng.normal(size=(Q, d)).astype(np.float16)
    v = rng.normal(size=(n, d)).astype(np.float16)
    v[5] = v[3]                                     # an exact tie -> ordered by row index
    valid = np.ones(n, bool)
    valid[::17] = False
    vb = np.packbits(valid, bitorder="little")
    got = compute_distance_f16(q, metric, v, validity=vb)
    for i in range(Q):
        want = oracle.distance_batch_f16(metric, q[i], v)
        assert np.isnan(got[i][~valid]).all()
        assert np.array_equal(got[i][valid].view(np.uint32), want[valid].view(np.uint32)), f"{metric} d={d} q{i}"
        if oracle.ref_f16() is not None:            # and within rounding of the reference's own C kernels
            ref = oracle.ref_distance_batch_f16(metric, q[i], v).astype(np.float64)
            assert (np.abs(got[i][valid] - ref[valid]) <= _bound(metric, q[i], v)[valid]).all()
    dall = got[np.isfinite(got)]
    for k, lo, hi in ((1, None, None), (10, None, None), (100, float(np.quantile(dall, 0.2)), float(np.quantile(dall, 0.6)))):
        od, oi, oc = flat_knn_f16(q, metric, v, k, validity=vb, lower_bound=lo, upper_bound=hi)
        for i in range(Q):
            want = oracle.distance_batch_f16(metric, q[i], v)
            keep = valid.copy()
            if lo is not None:
                keep &= (want >= np.float32(lo)) & (want < np.float32(hi))
            idx = np.nonzero(keep)[0]
            order = idx[np.lexsort((idx, want[idx]))][:k]
            c = int(oc[i])
            assert c == len(order) and oi[i, :c].tolist() == order.tolist(), f"{metric} d={d} k={k} q{i}"
            assert np.array_equal(od[i, :c].view(np.uint32), want[order].view(np.uint32))
    # a single key, no validity, empty input
    one = compute_distance_f16(q[0], metric, v)
    assert np.array_equal(one.view(np.uint32), oracle.distance_batch_f16(metric, q[0], v).view(np.uint32))
    od, oi, oc = flat_knn_f16(q, metric, np.zeros((0, d), np.float16), 3)
    assert oc.tolist() == [0] * Q


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_f16_ivf_pq_checker_is_the_generic_path_with_f16_pieces(oracle, metric):
    """lo_search_batch_f16 against a numpy walk through the same steps built from the f16 primitives that
    oracle/_ref (the reference's own f16.c) pins: find_partitions by distance_batch_f16, residual rounded to f16,
    table = distance_batch_f16 per sub-vector, f32 code sums in ascending m, (distance, row id) order."""
    from lance_b200.synth import IvfPqData
    rng = np.random.default_rng(31)
    d, nlist, M, n = 64, 5, 16, 700
    dsub = d // M
    cents = rng.normal(size=(nlist, d)).astype(np.float16)
    cb = rng.normal(size=(M, 256, dsub)).astype(np.float16)
    sizes = np.array([100, 0, 250, 50, 300])
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    codes = np.concatenate([rng.integers(0, 256, size=(M, s), dtype=np.uint8).ravel() for s in sizes])
    rids = rng.permutation(n * 3)[:n].astype(np.uint64)
    data = IvfPqData(metric, d, nlist, M, cents.astype(np.float32), cb.astype(np.float32), offs, codes, rids, None)
    q = rng.normal(size=(6, d)).astype(np.float16)
    k, nprobes = 7, 3
    od, oi, oc, pr = oracle.OracleIndexF16(data).search(q, k, nprobes)
    for j in range(len(q)):
        key = oracle.normalize_f16(q[j]) if metric == "cosine" else q[j]
        cd = oracle.distance_batch_f16("dot" if metric == "dot" else "l2", key, cents)
        probes = np.lexsort((np.arange(nlist), cd))[:nprobes]
        assert probes.tolist() == pr[j].tolist()
        ds, rs = [], []
        for p in probes.tolist():
            o, e = int(offs[p]), int(offs[p + 1])
            if e == o:
                continue
            r = key if metric == "dot" else (key.astype(np.float32) - cents[p].astype(np.float32)).astype(np.float16)
            table = np.stack([oracle.distance_batch_f16("dot" if metric == "dot" else "l2", r[m * dsub:(m + 1) * dsub], cb[m])
                              for m in range(M)])
            c = codes[o * M:e * M].reshape(M, e - o)
            acc = np.zeros(e - o, dtype=np.float32)
            for m in range(M):
                acc = (acc + table[m][c[m]]).astype(np.float32)
            ds.append(acc)
            rs.append(rids[o:e])
        ds, rs = np.concatenate(ds), np.concatenate(rs)
        order = np.lexsort((rs, ds))[:k]
        n_ = int(oc[j])
        assert n_ == len(order) and oi[j, :n_].tolist() == rs[order].tolist()
        assert np.array_equal(od[j, :n_].view(np.uint32), ds[order].view(np.uint32))

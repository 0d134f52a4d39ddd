"""Pins the CPU oracle against the reference's own known-answer tests.

Each test cites the reference test it transcribes (paths relative to /root/reference/rust/).
"""
import numpy as np
import pytest

from lance_b200.synth import IvfPqData


def _f64_l2(x, y):
    return float(((x.astype(np.float64) - y.astype(np.float64)) ** 2).sum())


# lance-linalg/src/distance/l2.rs:316-335 test_euclidean_distance
def test_l2_known_answers(oracle):
    mat = np.stack([np.arange(i, i + 8) for i in range(4)]).astype(np.float32)
    point = np.arange(2, 10, dtype=np.float32)
    assert oracle.distance_batch("l2", point, mat).tolist() == [32.0, 8.0, 0.0, 8.0]


# l2.rs:337-350 test_not_aligned (same answers from unaligned slices)
def test_l2_not_aligned(oracle):
    mat = np.concatenate([np.arange(0, 6)] + [np.arange(i, i + 8) for i in range(4)]).astype(np.float32)
    point = np.arange(0, 10, dtype=np.float32)
    out = oracle.distance_batch("l2", point[2:], mat[6:].reshape(4, 8))
    assert out.tolist() == [32.0, 8.0, 0.0, 8.0]


# l2.rs:352-359 test_odd_length_vector
def test_l2_odd_length(oracle):
    mat = np.arange(0, 5, dtype=np.float32)[None, :]
    point = np.arange(2, 7, dtype=np.float32)
    assert oracle.distance_batch("l2", point, mat).tolist() == [20.0]


# proptests of l2.rs / dot.rs / norm_l2.rs / cosine.rs: agreement with an f64 reference
@pytest.mark.parametrize("d", [1, 3, 7, 8, 15, 16, 17, 31, 32, 63, 64, 65, 100, 128, 200, 1536])
def test_distances_vs_f64(oracle, d):
    rng = np.random.default_rng(d)
    x = rng.normal(size=d).astype(np.float32)
    y = rng.normal(size=(5, d)).astype(np.float32)
    l2 = oracle.distance_batch("l2", x, y)
    dot = oracle.distance_batch("dot", x, y)
    cos = oracle.distance_batch("cosine", x, y)
    for i in range(5):
        assert l2[i] == pytest.approx(_f64_l2(x, y[i]), rel=1e-5)
        assert dot[i] == pytest.approx(1.0 - float(x.astype(np.float64) @ y[i].astype(np.float64)), rel=1e-4, abs=1e-5)
        c = 1.0 - float(x.astype(np.float64) @ y[i].astype(np.float64)) / (
            np.linalg.norm(x.astype(np.float64)) * np.linalg.norm(y[i].astype(np.float64)))
        assert cos[i] == pytest.approx(c, rel=1e-4, abs=1e-5)
    assert oracle.norm_l2(x) == pytest.approx(float(np.linalg.norm(x.astype(np.float64))), rel=1e-6)


# lance/src/index/vector/fixture_test.rs:165-231 test_ivf_residual_handling
def test_residual_and_cosine_normalisation(oracle):
    cents = np.array([[1, 1], [-1, -1], [-1, 1], [1, -1]], dtype=np.float32)

    def mk(metric):
        return oracle.OracleIndex(IvfPqData(metric, 2, 4, 1, cents, np.zeros((1, 256, 2), np.float32),
                                            np.zeros(5, np.uint64), np.zeros(0, np.uint8), np.zeros(0, np.uint64)))

    ix = mk("l2")
    q = np.array([1, 1], dtype=np.float32)
    part, _ = oracle.find_partitions(cents, q, 1, "l2")
    assert part.tolist() == [0]
    assert ix.preprocess_query(0, q).tolist() == [0.0, 0.0]
    ixc = mk("cosine")
    for q, want in [(np.array([1, 1], np.float32), np.float32(1.0) / np.sqrt(np.float32(2.0)) - np.float32(1.0)),
                    (np.array([2, 2], np.float32), np.float32(2.0) / np.sqrt(np.float32(8.0)) - np.float32(1.0))]:
        key = oracle.normalize(q)
        part, _ = oracle.find_partitions(cents, key, 1, "cosine")
        assert part.tolist() == [0]
        got = ixc.preprocess_query(0, key)
        assert got.tolist() == [float(want)] * 2


def _transpose(codes_rowmajor):
    return np.ascontiguousarray(codes_rowmajor.T)


# lance-index/src/vector/pq/distance.rs:281-308 test_compute_on_transposed_codes
def test_pq_transposed_equals_rowmajor(oracle):
    nv, M, dim = 100, 4, 16
    # the reference builds an (undersized) iota codebook; pad it to the full [M][256][dsub]
    cb = codebook_resize(np.arange(M * nv * dim, dtype=np.float32), M, dim)
    query = np.arange(dim, dtype=np.float32)
    table = oracle.build_distance_table("l2", cb, query)
    codes = (np.arange(nv * M) % 256).astype(np.uint8).reshape(nv, M)
    got = oracle.compute_pq_distance(table, _transpose(codes))
    want = np.array([sum_f32([table[m, codes[i, m]] for m in range(M)]) for i in range(nv)], dtype=np.float32)
    assert got.tolist() == want.tolist()


def codebook_resize(flat, M, dim):
    need = M * 256 * (dim // M)
    out = np.zeros(need, dtype=np.float32)
    out[:min(need, flat.size)] = flat[:need]
    return out.reshape(M, 256, dim // M)


def sum_f32(vals):
    s = np.float32(0.0)
    for v in vals:
        s = np.float32(s + np.float32(v))
    return s


# lance-index/src/vector/pq/storage.rs:741-751 test_distance_all (distance_all == [distance(i)])
def test_distance_all_equals_distance_one(oracle):
    DIM, M, TOTAL = 32, 16, 512
    cb = np.arange(256 * DIM, dtype=np.float32).reshape(M, 256, DIM // M)
    vecs = np.arange(TOTAL * DIM, dtype=np.float32).reshape(TOTAL, DIM)
    codes = oracle.pq_quantize("l2", cb, vecs)
    query = np.arange(DIM, dtype=np.float32)
    table = oracle.build_distance_table("l2", cb, query)
    ct = _transpose(codes)
    allv = oracle.compute_pq_distance(table, ct)
    one = [oracle.pq_distance_one(table, ct, i) for i in range(TOTAL)]
    assert allv.tolist() == one


# lance-index/src/vector/pq.rs:546-592 test_l2_distance (DIM 512, M 16, 66 rows incl. remainder)
def test_pq_l2_distance_definition(oracle):
    DIM, M, TOTAL = 512, 16, 66
    rng = np.random.default_rng(7)
    cb = rng.random((M, 256, DIM // M), dtype=np.float32)
    codes = (np.arange(16 * TOTAL) % 256).astype(np.uint8).reshape(TOTAL, 16)
    query = rng.random(DIM, dtype=np.float32)
    table = oracle.build_distance_table("l2", cb, query)
    got = oracle.compute_pq_distance(table, _transpose(codes))
    sub = DIM // M
    for i in range(TOTAL):
        want = sum(_f64_l2(query[m * sub:(m + 1) * sub], cb[m, codes[i, m]]) for m in range(M))
        assert abs(got[i] - want) <= 1e-4 * max(1.0, want)


# pq.rs:595-632 test_pq_transform: quantise = per-sub-space argmin
def test_pq_quantize_is_argmin(oracle):
    DIM, TOTAL, M = 16, 64, 4
    rng = np.random.default_rng(11)
    cb = rng.random((M, 256, DIM // M), dtype=np.float32)
    vecs = rng.random((TOTAL, DIM), dtype=np.float32)
    codes = oracle.pq_quantize("l2", cb, vecs)
    sub = DIM // M
    for i in range(TOTAL):
        for m in range(M):
            d = oracle.distance_batch("l2", vecs[i, m * sub:(m + 1) * sub], cb[m])
            assert codes[i, m] == int(np.argmin(d))


# lance-linalg/src/kmeans.rs:785-810: find_partitions(nprobes=1) == naive argmin L2
def test_find_partitions_is_argmin(oracle):
    rng = np.random.default_rng(3)
    cents = rng.random((128, 64), dtype=np.float32)
    for _ in range(20):
        q = rng.random(64, dtype=np.float32)
        ids, d = oracle.find_partitions(cents, q, 1, "l2")
        naive = np.array([_f64_l2(q, c) for c in cents])
        assert ids[0] == int(np.argmin(naive))
    ids, d = oracle.find_partitions(cents, q, 10, "l2")
    assert np.all(np.diff(d) >= 0) and len(set(ids.tolist())) == 10
    ids, _ = oracle.find_partitions(cents, q, 1000, "l2")
    assert len(ids) == 128  # nprobes > nlist returns every partition


def _test_vector_dataset():
    """TestVectorDataset (lance/src/dataset/scanner.rs:2410-2510): 5 copies of 80 vectors
    v_j = [32j .. 32j+31]; row i = copy*80 + j."""
    base = np.arange(32 * 80, dtype=np.float32).reshape(80, 32)
    return np.concatenate([base] * 5, axis=0)


# scanner.rs:2758-2816 test_knn_nodes without index (flat) and io/exec/knn.rs:667-747
def test_flat_knn_golden(oracle):
    data = _test_vector_dataset()
    key = np.arange(32, 64, dtype=np.float32)
    d, ids, cnt = oracle.flat_knn("l2", key[None, :], data, 5)
    assert cnt[0] == 5 and set(ids[0].tolist()) == {1, 81, 161, 241, 321}
    assert d[0].tolist() == [0.0] * 5
    # knn.rs:667-747: result == sort_to_indices(compute_distance(all), limit)
    rng = np.random.default_rng(5)
    vec = rng.random((400, 128), dtype=np.float32)
    q = vec[5]
    d, ids, cnt = oracle.flat_knn("l2", q[None, :], vec, 10)
    alld = oracle.distance_batch("l2", q, vec)
    order = np.lexsort((np.arange(400), alld))[:10]
    assert ids[0].tolist() == order.tolist() and d[0].tolist() == alld[order].tolist()


# scanner.rs:2758-2816 test_knn_nodes with ivf_pq(2, 8, 2, L2, 2) + refine(5) -> {1,81,161,241,321}
def test_ivf_pq_refine_golden(oracle):
    from lance_b200.synth import build_ivf_pq
    data = _test_vector_dataset()
    ix = build_ivf_pq(data, "l2", nlist=2, M=2, seed=0, keep_raw=True)
    o = oracle.OracleIndex(ix)
    key = np.arange(32, 64, dtype=np.float32)
    d, ids, cnt, _ = o.search(key[None, :], k=5, nprobes=2, refine_factor=5)
    assert cnt[0] == 5
    assert set(ids[0, :5].tolist()) == {1, 81, 161, 241, 321}
    assert d[0, :5].tolist() == [0.0] * 5


# ivf/v2.rs:1192-1302 test_distance_range: [lower, upper) and prefix consistency
@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_distance_range_semantics(oracle, metric):
    from lance_b200.synth import build_ivf_pq
    rng = np.random.default_rng(1)
    vecs = rng.random((1000, 32), dtype=np.float32)
    ix = build_ivf_pq(vecs, metric, nlist=4, M=16, seed=1)
    o = oracle.OracleIndex(ix)
    q = vecs[:3]
    d, ids, cnt, _ = o.search(q, k=10, nprobes=4)
    for qi in range(3):
        assert cnt[qi] == 10
        mid = float(d[qi, 5])
        lo_d, lo_i, lo_c, _ = o.search(q[qi:qi + 1], k=10, nprobes=4, upper=mid)
        hi_d, hi_i, hi_c, _ = o.search(q[qi:qi + 1], k=10, nprobes=4, lower=mid)
        assert np.all(lo_d[0, :lo_c[0]] < mid) and np.all(hi_d[0, :hi_c[0]] >= mid)
        n_lo = int(lo_c[0])
        assert lo_i[0, :n_lo].tolist() == ids[qi, :n_lo].tolist()
        assert hi_i[0, :10 - n_lo].tolist() == ids[qi, n_lo:10].tolist()


# ivf.rs:2470-2481,2581-2593: cosine PQ distances in [0,2); dot PQ distance = M - sum(dot)
def test_metric_sign_and_range(oracle):
    from lance_b200.synth import build_ivf_pq
    rng = np.random.default_rng(2)
    vecs = rng.random((2000, 32), dtype=np.float32)
    oc = oracle.OracleIndex(build_ivf_pq(vecs, "cosine", nlist=4, M=8, seed=2))
    d, _, cnt, _ = oc.search(vecs[:8], k=20, nprobes=4)
    assert np.all(d >= 0) and np.all(d < 2)
    ixd = build_ivf_pq(vecs, "dot", nlist=4, M=8, seed=2)
    od = oracle.OracleIndex(ixd)
    d, ids, cnt, _ = od.search(vecs[:4], k=5, nprobes=4)
    # reconstruct: M - sum_m dot(q_m, cb[m][code])
    pos = {int(r): i for i, r in enumerate(ixd.row_ids)}
    for qi in range(4):
        rid = int(ids[qi, 0])
        i = pos[rid]
        p = int(np.searchsorted(ixd.part_offsets, i, side="right") - 1)
        codes = ixd.part_codes(p)[:, i - int(ixd.part_offsets[p])]
        approx = ixd.M - sum(float(vecs[qi, m * 4:(m + 1) * 4] @ ixd.codebook[m, codes[m]]) for m in range(ixd.M))
        assert d[qi, 0] == pytest.approx(approx, rel=1e-5, abs=1e-5)


# ivf/v2.rs:852-916 recall floors on 1000x32 uniform data, nprobes = nlist
@pytest.mark.parametrize("metric,floor", [("l2", 0.9), ("cosine", 0.9), ("dot", 0.85)])
def test_recall_floor(oracle, metric, floor):
    from lance_b200.synth import build_ivf_pq
    rng = np.random.default_rng(9)
    vecs = rng.random((1000, 32), dtype=np.float32)
    if metric != "l2":
        vecs = vecs / np.linalg.norm(vecs, axis=1, keepdims=True)
    ix = build_ivf_pq(vecs, metric, nlist=4, M=16, seed=3, keep_raw=True, pq_iters=10)
    o = oracle.OracleIndex(ix)
    q = vecs[:20]
    d, ids, cnt, _ = o.search(q, k=100, nprobes=4)
    gt_d, gt_i, _ = oracle.flat_knn(metric, q, vecs, 100)
    rec = np.mean([len(set(ids[i].tolist()) & set(gt_i[i].tolist())) / 100 for i in range(20)])
    assert rec >= floor

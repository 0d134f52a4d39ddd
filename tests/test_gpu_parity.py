"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same inputs.

Bar (BASELINE.json north_star): identical probed partitions; row ids identical except for
distance ties within 1e-5 relative; distances within 1e-5 relative.  Because the kernels keep
the reference's operation order, these tests assert the stronger property: BIT-IDENTICAL
distances and identical row ids (ties ordered by row id on both sides); REL_TOL documents the
contract the bit-exact check implies.
"""
import numpy as np
import pytest
import torch

from lance_b200.synth import IvfPqData, build_ivf_pq, random_index, queries_near_centroids

pytestmark = pytest.mark.gpu
REL_TOL = 1e-5  # the stated tolerance; every check below is exact, which is stricter


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from lance_b200 import _lib
    assert _lib.lib() is not None  # the in-tree .so is what runs; no fallback exists


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _assert_same(d_gpu, i_gpu, c_gpu, d_ref, i_ref, c_ref, what=""):
    assert c_gpu.tolist() == c_ref.tolist(), f"{what}: counts differ"
    for q in range(len(c_ref)):
        n = int(c_ref[q])
        assert i_gpu[q, :n].tolist() == i_ref[q, :n].tolist(), f"{what}: row ids differ for query {q}"
        assert _bits(d_gpu[q, :n]).tolist() == _bits(d_ref[q, :n]).tolist(), f"{what}: distances differ for query {q}"


def _ragged_index(metric, d, nlist, M, sizes, seed, keep_raw=False):
    """Index with explicit partition sizes (incl. empty / tiny / non-multiple-of-16), random
    codes, unique row ids in random order."""
    rng = np.random.default_rng(seed)
    sizes = np.asarray(sizes, dtype=np.int64)
    N = int(sizes.sum())
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    cents = rng.normal(size=(nlist, d)).astype(np.float32)
    cb = rng.normal(size=(M, 256, d // M)).astype(np.float32)
    codes = rng.integers(0, 256, size=N * M, dtype=np.uint8)
    rid = rng.permutation(N * 3)[:N].astype(np.uint64)
    raw = rng.normal(size=(N, d)).astype(np.float32) if keep_raw else None
    return IvfPqData(metric, d, nlist, M, cents, cb, offs, codes, rid, raw)


# ---- A1/A2: find_partitions -------------------------------------------------------------
@pytest.mark.parametrize("metric,d,nlist", [("l2", 128, 256), ("l2", 8, 33), ("l2", 16, 64), ("l2", 100, 50),
                                            ("dot", 768, 128), ("dot", 70, 20), ("cosine", 1536, 64),
                                            ("l2", 256, 300), ("l2", 32, 5000), ("cosine", 64, 700),
                                            ("dot", 64, 300), ("dot", 128, 129), ("dot", 1536, 40)])
def test_find_partitions_parity(oracle, metric, d, nlist):
    from lance_b200 import IvfPqIndex
    rng = np.random.default_rng(d + nlist)
    data = _ragged_index(metric, d, nlist, 1 if d % 2 else 2, [3] * nlist, seed=1)
    q = rng.normal(size=(37, d)).astype(np.float32)
    if metric == "cosine":
        q = np.stack([oracle.normalize(x) for x in q])
    ix = IvfPqIndex(data)
    for nprobes in (1, 7, min(nlist, 256), min(nlist, 300), nlist, nlist + 5):
        ids, dist = ix.find_partitions(q, nprobes, return_distances=True)
        for i in range(q.shape[0]):
            rid, rd = oracle.find_partitions(data.centroids, q[i], nprobes, metric)
            assert ids[i].tolist() == rid.tolist()
            assert _bits(dist[i]).tolist() == _bits(rd).tolist()
    ix.close()


def test_find_partitions_ties_and_clustered_layout(oracle):
    """Select fast path: (a) all centroids identical -> all distances tie -> order by id;
    (b) the smallest distances all sit in ONE strided group (ids = 0 mod 256), the layout that
    overflows the candidate buffer and takes the exact-select fallback."""
    from lance_b200 import IvfPqIndex
    rng = np.random.default_rng(0)
    d, nlist = 64, 4096
    q = rng.normal(size=(9, d)).astype(np.float32)
    data = _ragged_index("l2", d, nlist, 2, [1] * nlist, seed=1)
    data.centroids[:] = data.centroids[0]
    ix = IvfPqIndex(data)
    ids, _ = ix.find_partitions(q, 32, return_distances=True)
    assert ids.tolist() == [list(range(32))] * 9
    ix.close()
    data = _ragged_index("l2", d, nlist, 2, [1] * nlist, seed=2)
    data.centroids[:] = rng.normal(size=(nlist, d)).astype(np.float32) * 0.01 + 5.0
    far = np.arange(nlist) % 256 != 0
    data.centroids[far] += 100.0 + np.arange(nlist)[far, None] % 7   # 16 near centroids, 4080 far ones
    data.centroids[~far] = q[0] + rng.normal(size=(16, d)).astype(np.float32) * 0.1
    ix = IvfPqIndex(data)
    for nprobes in (5, 16, 40, 256):
        ids, dist = ix.find_partitions(q, nprobes, return_distances=True)
        for i in range(len(q)):
            rid, rd = oracle.find_partitions(data.centroids, q[i], nprobes, "l2")
            assert ids[i].tolist() == rid.tolist()
            assert _bits(dist[i]).tolist() == _bits(rd).tolist()
    ix.close()


# ---- A3-A6: search_in_partition ------------------------------------------------------------
CASES = [
    # metric, d, M (dsub), variant
    ("l2", 128, 16, 1), ("l2", 128, 16, 2), ("cosine", 128, 16, 2), ("dot", 128, 16, 2),
    ("l2", 64, 4, 1), ("l2", 64, 4, 2),        # dsub 16
    ("dot", 96, 8, 2), ("l2", 96, 8, 2),       # dsub 12: generic table path
    ("l2", 32, 16, 2), ("l2", 24, 2, 1),       # dsub 2 / 12, M=2 only fits variant 1
    ("l2", 256, 32, 2), ("dot", 768, 48, 2),   # QL sweeps through VG=32/16/8/4
    # variant 3 (skew streams): D=8/QL=4 for M<=48, D=16/QL=2 for M in (48,96]
    ("l2", 128, 16, 3), ("cosine", 128, 16, 3), ("dot", 128, 16, 3), ("l2", 32, 16, 3),
    ("l2", 256, 32, 3), ("dot", 768, 48, 3), ("l2", 576, 48, 3), ("l2", 1024, 64, 3),
    ("cosine", 192, 96, 3), ("dot", 1536, 96, 3),
]


@pytest.mark.parametrize("metric,d,M,variant", CASES)
def test_search_in_partition_parity(oracle, metric, d, M, variant):
    from lance_b200 import IvfPqIndex
    sizes = [0, 1, 5, 15, 16, 17, 100, 1000, 33, 0, 2049, 64, 7, 300, 4100, 9]
    nlist = len(sizes)
    data = _ragged_index(metric, d, nlist, M, sizes, seed=d + M)
    o = oracle.OracleIndex(data)
    ix = IvfPqIndex(data)
    ix.set_scan_variant(variant)
    rng = np.random.default_rng(5)
    Q, P = 19, nlist
    q = rng.normal(size=(Q, d)).astype(np.float32)
    if metric == "cosine":
        q = np.stack([oracle.normalize(x) for x in q])
    parts = np.stack([rng.permutation(nlist) for _ in range(Q)]).astype(np.uint32)
    for k in (1, 10, 50):
        od, oi, oc = ix.search_in_partition(parts, q, k)
        assert ix.stats()["scan_variant"] == variant
        for qi in range(Q):
            for pi in range(P):
                rd, ri = o.search_in_partition(int(parts[qi, pi]), q[qi], k)
                n = len(rd)
                assert oc[qi, pi] == n == min(k, sizes[parts[qi, pi]])
                assert oi[qi, pi, :n].tolist() == ri.tolist()
                assert _bits(od[qi, pi, :n]).tolist() == _bits(rd).tolist()
    ix.close()


@pytest.mark.parametrize("variant", [1, 2, 3])
def test_bounds_and_prefilter(oracle, variant):
    from lance_b200 import IvfPqIndex
    sizes = [500, 0, 77, 1300]
    data = _ragged_index("l2", 64, 4, 16, sizes, seed=3)
    o = oracle.OracleIndex(data)
    ix = IvfPqIndex(data)
    ix.set_scan_variant(variant)
    rng = np.random.default_rng(8)
    q = rng.normal(size=(6, 64)).astype(np.float32)
    parts = np.tile(np.arange(4, dtype=np.uint32), (6, 1))
    base_d, _, _ = ix.search_in_partition(parts, q, 20)
    lo, hi = float(np.median(base_d[np.isfinite(base_d)])), float(np.max(base_d[np.isfinite(base_d)]))
    masks = {p: rng.random(sizes[p]) < 0.3 for p in (0, 3)}
    for lower, upper, pf in [(lo, None, None), (None, hi, None), (lo, hi, None), (None, None, masks), (lo, hi, masks)]:
        od, oi, oc = ix.search_in_partition(parts, q, 20, lower_bound=lower, upper_bound=upper, prefilter=pf)
        for qi in range(6):
            for p in range(4):
                bm = None
                if pf is not None and p in pf:
                    bm = np.zeros((sizes[p] + 63) // 64 + 1, dtype=np.uint64)
                    packed = np.packbits(pf[p], bitorder="little")
                    bm.view(np.uint8)[:packed.size] = packed
                rd, ri = o.search_in_partition(p, q[qi], 20, lower=lower, upper=upper, bitmask=bm)
                n = len(rd)
                assert oc[qi, p] == n
                assert oi[qi, p, :n].tolist() == ri.tolist()
                assert _bits(od[qi, p, :n]).tolist() == _bits(rd).tolist()
    ix.close()


def test_distance_ties_ordered_by_row_id(oracle):
    """All rows share one code -> every distance ties; both sides must order by row id."""
    from lance_b200 import IvfPqIndex
    data = _ragged_index("l2", 32, 2, 4, [700, 45], seed=4)
    data.codes[:] = 7
    o = oracle.OracleIndex(data)
    q = np.random.default_rng(1).normal(size=(3, 32)).astype(np.float32)
    for variant in (1, 2):  # M=4: variant 3 needs M % 16 == 0, see the M=16 tie test below
        ix = IvfPqIndex(data)
        ix.set_scan_variant(variant)
        od, oi, oc = ix.search_in_partition(np.tile(np.arange(2, dtype=np.uint32), (3, 1)), q, 10)
        for qi in range(3):
            for p in range(2):
                rd, ri = o.search_in_partition(p, q[qi], 10)
                assert oi[qi, p, :len(ri)].tolist() == ri.tolist()
        ix.close()


def test_variant3_ties_and_nonfinite(oracle):
    """Variant 3: all-equal distances are ordered by row id (whole search and per partition), and
    a table with non-finite / huge entries takes the select-based switch path with results
    still bit-identical."""
    from lance_b200 import IvfPqIndex, Query
    data = _ragged_index("l2", 64, 2, 16, [2100, 300], seed=4)
    data.codes[:] = 7
    o = oracle.OracleIndex(data)
    q = np.random.default_rng(1).normal(size=(5, 64)).astype(np.float32)
    ix = IvfPqIndex(data)
    ix.set_scan_variant(3)
    od, oi, oc = ix.search_in_partition(np.tile(np.arange(2, dtype=np.uint32), (5, 1)), q, 10)
    for qi in range(5):
        for p in range(2):
            rd, ri = o.search_in_partition(p, q[qi], 10)
            assert oi[qi, p, :len(ri)].tolist() == ri.tolist()
    gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=2))
    rd, ri, rc, _ = o.search(q, k=10, nprobes=2)
    _assert_same(gd, gi, gc, rd, ri, rc, "ties")
    ix.close()
    # huge codebook entries: partial sums overflow to +inf for some rows
    data = _ragged_index("l2", 64, 3, 16, [1500, 40, 700], seed=9)
    data.codebook[3, 5, :] = 3e19   # (3e19)^2*4 ~ 3.6e39 -> +inf table entry
    data.codebook[7, 9, :] = 1e18
    o = oracle.OracleIndex(data)
    ix = IvfPqIndex(data)
    ix.set_scan_variant(3)
    parts = np.tile(np.arange(3, dtype=np.uint32), (5, 1))
    od, oi, oc = ix.search_in_partition(parts, q, 20)
    for qi in range(5):
        for p in range(3):
            rd, ri = o.search_in_partition(p, q[qi], 20)
            n = len(rd)
            assert oc[qi, p] == n
            assert oi[qi, p, :n].tolist() == ri.tolist()
            assert _bits(od[qi, p, :n]).tolist() == _bits(rd).tolist()
    ix.close()


@pytest.mark.parametrize("L", [1, 4, 32])
def test_variant3_superblock_depths(oracle, L, monkeypatch):
    from lance_b200 import IvfPqIndex, Query
    monkeypatch.setenv("LB200_SKEW_L", str(L))
    data = random_index(30000, 64, 8, 16, "l2", seed=12)
    q = queries_near_centroids(data, 50, seed=5)
    o = oracle.OracleIndex(data)
    ix = IvfPqIndex(data)
    ix.set_scan_variant(3)
    for Q in (1, 3, 50):   # small batches chunk the partitions across CTAs
        rd, ri, rc, _ = o.search(q[:Q], k=10, nprobes=4)
        gd, gi, gc = ix.search(Query(key=q[:Q], k=10, nprobes=4))
        _assert_same(gd, gi, gc, rd, ri, rc, f"L={L} Q={Q}")
    ix.close()


# ---- whole path: VectorIndex::search (+ refine) ---------------------------------------------
@pytest.mark.parametrize("metric,d,nlist,M,refine", [("l2", 128, 64, 16, 0), ("l2", 128, 64, 16, 5),
                                                     ("cosine", 96, 32, 6, 5), ("dot", 64, 16, 8, 2),
                                                     ("cosine", 64, 16, 16, 0)])
def test_search_parity_trained_index(oracle, metric, d, nlist, M, refine):
    from lance_b200 import IvfPqIndex, Query
    rng = np.random.default_rng(nlist)
    vecs = rng.normal(size=(20000, d)).astype(np.float32) + rng.normal(size=(1, d)).astype(np.float32)
    data = build_ivf_pq(vecs, metric, nlist, M, seed=1, keep_raw=True, ivf_iters=4, pq_iters=3, sample=8192)
    o = oracle.OracleIndex(data)
    ix = IvfPqIndex(data)
    q = vecs[rng.integers(0, len(vecs), 64)] + 0.05 * rng.normal(size=(64, d)).astype(np.float32)
    for nprobes in (1, 8, nlist):
        rd, ri, rc, rp = o.search(q, k=10, nprobes=nprobes, refine_factor=refine)
        probes = ix.find_partitions(np.stack([oracle.normalize(x) for x in q]) if metric == "cosine" else q, nprobes)
        assert probes.tolist() == rp.tolist(), "probed partitions differ"
        gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=nprobes, refine_factor=refine or None))
        kout = gd.shape[1]
        _assert_same(gd, gi, gc, rd[:, :kout], ri[:, :kout], rc, f"{metric} nprobes={nprobes}")
    # batched == one at a time (the reference has no inter-query batching)
    gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=8, refine_factor=refine or None))
    for i in (0, 17, 63):
        d1, i1, c1 = ix.search(Query(key=q[i], k=10, nprobes=8, refine_factor=refine or None))
        assert i1[0].tolist() == gi[i].tolist() and _bits(d1[0]).tolist() == _bits(gd[i]).tolist()
    ix.close()


def test_variants_agree_and_bounds_in_search(oracle):
    from lance_b200 import IvfPqIndex, Query
    data = random_index(60000, 128, 32, 16, "l2", seed=2)
    q = queries_near_centroids(data, 200, seed=3)
    o = oracle.OracleIndex(data)
    rd, ri, rc, _ = o.search(q, k=10, nprobes=8)
    mid = float(np.median(rd[:, 5]))
    rdb, rib, rcb, _ = o.search(q, k=10, nprobes=8, lower=mid * 0.9, upper=mid * 1.1)
    for variant in (1, 2, 3):
        ix = IvfPqIndex(data)
        ix.set_scan_variant(variant)
        gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=8))
        _assert_same(gd, gi, gc, rd, ri, rc, f"variant {variant}")
        gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=8, lower_bound=mid * 0.9, upper_bound=mid * 1.1))
        _assert_same(gd, gi, gc, rdb, rib, rcb, f"variant {variant} bounds")
        ix.close()


def test_golden_test_vector_dataset(oracle):
    """scanner.rs:2758-2816 test_knn_nodes: ivf_pq(2,8,2,L2,2) + refine(5) -> {1,81,161,241,321}."""
    from lance_b200 import IvfPqIndex, Query, flat_knn
    base = np.arange(32 * 80, dtype=np.float32).reshape(80, 32)
    rows = np.concatenate([base] * 5, axis=0)
    key = np.arange(32, 64, dtype=np.float32)
    data = build_ivf_pq(rows, "l2", nlist=2, M=2, seed=0, keep_raw=True)
    ix = IvfPqIndex(data)
    d, ids, cnt = ix.search(Query(key=key, k=5, nprobes=2, refine_factor=5))
    assert cnt[0] == 5 and set(ids[0].tolist()) == {1, 81, 161, 241, 321} and d[0].tolist() == [0.0] * 5
    d, ids, cnt = flat_knn(key, "l2", rows, 5)
    assert cnt[0] == 5 and set(ids[0].tolist()) == {1, 81, 161, 241, 321}
    ix.close()


def test_committed_golden_fixture():
    """tests/golden/ivfpq_small.npz was produced by tests/golden/make_golden.py (oracle output at
    commit time); the GPU must reproduce it bit for bit."""
    import os
    from lance_b200 import IvfPqIndex, Query
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ivfpq_small.npz"))
    for metric in ("l2", "cosine", "dot"):
        data = IvfPqData(metric, int(g["dim"]), int(g["nlist"]), int(g["M"]), g[f"{metric}_centroids"],
                         g[f"{metric}_codebook"], g[f"{metric}_part_offsets"], g[f"{metric}_codes"],
                         g[f"{metric}_row_ids"], g[f"{metric}_raw"])
        ix = IvfPqIndex(data)
        for refine in (0, 3):
            gd, gi, gc = ix.search(Query(key=g["queries"], k=10, nprobes=6, refine_factor=refine or None))
            _assert_same(gd, gi, gc, g[f"{metric}_r{refine}_dist"][:, :gd.shape[1]],
                         g[f"{metric}_r{refine}_ids"][:, :gd.shape[1]], g[f"{metric}_r{refine}_cnt"], metric)
        ix.close()


# ---- flat kNN (KNNVectorDistanceExec) -----------------------------------------------------------
@pytest.mark.parametrize("metric,d", [("l2", 128), ("cosine", 128), ("dot", 128), ("l2", 8), ("cosine", 16),
                                      ("l2", 100), ("cosine", 1536), ("dot", 70)])
def test_flat_parity(oracle, metric, d):
    from lance_b200 import compute_distance, flat_knn
    rng = np.random.default_rng(d)
    n = 3001
    vecs = rng.normal(size=(n, d)).astype(np.float32)
    q = rng.normal(size=(9, d)).astype(np.float32)
    valid = rng.random(n) < 0.9
    vbits = np.packbits(valid, bitorder="little")
    dist = compute_distance(q, metric, vecs, validity=vbits)
    for i in range(9):
        ref = oracle.distance_batch(metric, q[i], vecs)
        assert _bits(dist[i][valid]).tolist() == _bits(ref[valid]).tolist()
        assert np.all(np.isnan(dist[i][~valid]))
    for k in (1, 10, 100):
        gd, gi, gc = flat_knn(q, metric, vecs, k, validity=vbits)
        rd, ri, rc = oracle.flat_knn(metric, q, vecs, k, validity=vbits)
        _assert_same(gd, gi, gc, rd, ri, rc, f"flat {metric} k={k}")


def test_flat_empty_batch():
    """n = 0 (e.g. the allow-list removed every unindexed row): empty results, not an error."""
    from lance_b200 import flat_knn
    q = np.ones((3, 16), dtype=np.float32)
    gd, gi, gc = flat_knn(q, "l2", np.zeros((0, 16), dtype=np.float32), 5)
    assert gc.tolist() == [0, 0, 0] and np.all(np.isinf(gd))


def test_flat_segmented_select(oracle):
    """n large enough that each distance row is cut into several select segments + merge; many
    duplicated rows so that ties straddle segment borders."""
    from lance_b200 import flat_knn
    rng = np.random.default_rng(11)
    n, d = 150_001, 32
    base = rng.normal(size=(500, d)).astype(np.float32)
    vecs = base[rng.integers(0, 500, n)]          # every row has ~300 exact duplicates
    q = base[:12] + 0.01 * rng.normal(size=(12, d)).astype(np.float32)
    valid = rng.random(n) < 0.95
    vbits = np.packbits(valid, bitorder="little")
    for k in (1, 10, 400):
        gd, gi, gc = flat_knn(q, "l2", vecs, k, validity=vbits)
        rd, ri, rc = oracle.flat_knn("l2", q, vecs, k, validity=vbits)
        _assert_same(gd, gi, gc, rd, ri, rc, f"segmented flat k={k}")


def test_flat_edge_cases(oracle):
    from lance_b200 import flat_knn
    rng = np.random.default_rng(0)
    vecs = rng.normal(size=(40, 32)).astype(np.float32)
    vecs[10:30] = vecs[5]  # 21 identical rows: ties resolved by row position
    q = vecs[5:6]
    gd, gi, gc = flat_knn(q, "l2", vecs, 8)
    rd, ri, rc = oracle.flat_knn("l2", q, vecs, 8)
    _assert_same(gd, gi, gc, rd, ri, rc, "ties")
    gd, gi, gc = flat_knn(q, "l2", vecs, 64)  # k > n
    assert gc[0] == 40 and sorted(gi[0, :40].tolist()) == list(range(40))
    mid = float(rd[0, 7]) + 1.0
    gd, gi, gc = flat_knn(q, "l2", vecs, 8, lower_bound=0.5, upper_bound=mid)
    assert np.all(gd[0, :gc[0]] >= 0.5) and np.all(gd[0, :gc[0]] < mid)


# ---- reference-style error behaviour ---------------------------------------------------------------
def test_error_behaviour():
    from lance_b200 import IvfPqIndex, LanceB200Error, Query
    data = _ragged_index("l2", 32, 4, 4, [10, 20, 30, 40], seed=1)
    ix = IvfPqIndex(data)
    with pytest.raises(ValueError):  # scanner.rs:642: dim mismatch
        ix.search(Query(key=np.zeros(16, np.float32), k=5))
    with pytest.raises(ValueError):  # k must be positive
        ix.search(Query(key=np.zeros(32, np.float32), k=0))
    with pytest.raises(LanceB200Error, match="does not exist"):  # ivf/v2.rs:288-293
        ix.search_in_partition(np.array([[9]], np.uint32), np.zeros(32, np.float32), 5)
    with pytest.raises(LanceB200Error, match="raw vectors"):
        ix.search(Query(key=np.zeros(32, np.float32), k=5, nprobes=2, refine_factor=2))
    # k' larger than the partition: short result, like the reference's short batches
    od, oi, oc = ix.search_in_partition(np.array([[0]], np.uint32), np.zeros(32, np.float32), 50)
    assert oc[0, 0] == 10
    ix.close()


# ---- full-size properties (BASELINE cfg 1 scale; oracle on a sample) ---------------------------------
def test_full_size_properties(oracle):
    from lance_b200 import IvfPqIndex, Query
    data = random_index(1_000_000, 128, 256, 16, "l2", seed=11)
    q = queries_near_centroids(data, 2000, seed=12)
    ix = IvfPqIndex(data)
    gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=20))
    assert np.all(gc == 10)
    assert np.all(np.diff(gd, axis=1) >= 0), "distances must be sorted ascending"
    assert all(len(set(r.tolist())) == 10 for r in gi), "row ids must be unique per query"
    # idempotence / batch-composition independence
    gd2, gi2, _ = ix.search(Query(key=q[::-1].copy(), k=10, nprobes=20))
    assert gi2[::-1].tolist() == gi.tolist() and _bits(gd2[::-1]).tolist() == _bits(gd).tolist()
    # nprobes monotonicity: more probes can only improve the k-th distance
    gd3, _, _ = ix.search(Query(key=q[:200], k=10, nprobes=40))
    assert np.all(gd3[:, -1] <= gd[:200, -1])
    # oracle on a sample
    o = oracle.OracleIndex(data)
    sel = np.arange(0, 2000, 40)
    rd, ri, rc, _ = o.search(q[sel], k=10, nprobes=20)
    _assert_same(gd[sel], gi[sel], gc[sel], rd, ri, rc, "cfg1-size sample")
    ix.close()


# ---- multi-GPU protocol on one device: G shard handles + lb200_merge_topk_device --------------
@pytest.mark.parametrize("probed", [False, True])
@pytest.mark.parametrize("metric,refine,G", [("l2", 0, 2), ("l2", 4, 3), ("cosine", 3, 2), ("dot", 0, 4)])
def test_sharded_search_merge_equals_single_index(oracle, metric, refine, G, probed):
    """every shard returns its k' candidates (+ exact distances), the gathered lists
    are merged on the device; the result must be bit-identical to the unsharded oracle search."""
    from lance_b200 import IvfPqIndex
    from lance_b200.index import merge_topk_device
    from lance_b200.shard import assign_partitions
    rng = np.random.default_rng(G)
    d, nlist, M, k, nprobes, Q = 64, 24, 16, 10, 9, 40
    vecs = rng.normal(size=(12000, d)).astype(np.float32)
    data = build_ivf_pq(vecs, metric, nlist, M, seed=2, keep_raw=True, ivf_iters=3, pq_iters=2, sample=4096)
    q = vecs[rng.integers(0, len(vecs), Q)] + 0.05 * rng.normal(size=(Q, d)).astype(np.float32)
    rd, ri, rc, _ = oracle.OracleIndex(data).search(q, k=k, nprobes=nprobes, refine_factor=refine)
    owner = assign_partitions(np.diff(data.part_offsets.astype(np.int64)), G)
    kp = k * max(refine, 1)
    dev = torch.device("cuda", 0)
    dq = torch.from_numpy(q).to(dev)
    g_d = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
    g_i = torch.empty((G, Q, kp), dtype=torch.int64, device=dev)
    g_c = torch.empty((G, Q), dtype=torch.int32, device=dev)
    g_e = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
    shards = [IvfPqIndex(data.subset(np.nonzero(owner == g)[0])) for g in range(G)]
    probes = torch.zeros((Q, nprobes), dtype=torch.int32, device=dev)
    if probed:
        # coarse step sharded by queries: "rank" g selects the probes of its slice of the batch
        qs = (Q + G - 1) // G
        for g in range(G):
            lo, n = g * qs, min(qs, Q - g * qs)
            shards[g].find_partitions_device(dq[lo:].data_ptr(), n, nprobes, probes[lo:].data_ptr(), normalize=True)
    for g, ix in enumerate(shards):
        if probed:
            ix.search_probed_device(dq.data_ptr(), Q, probes.data_ptr(), nprobes, k, refine, g_d[g].data_ptr(),
                                    g_i[g].data_ptr(), g_c[g].data_ptr(), d_out_exact=g_e[g].data_ptr())
        else:
            ix.search_device(dq.data_ptr(), Q, nprobes, k, refine, g_d[g].data_ptr(), g_i[g].data_ptr(),
                             g_c[g].data_ptr(), d_out_exact=g_e[g].data_ptr())
    torch.cuda.synchronize()
    fd = torch.empty((Q, k), dtype=torch.float32, device=dev)
    fi = torch.empty((Q, k), dtype=torch.int64, device=dev)
    fc = torch.empty((Q,), dtype=torch.int32, device=dev)
    merge_topk_device(0, g_d.data_ptr(), g_i.data_ptr(), g_e.data_ptr() if refine else None, g_c.data_ptr(), G, Q, kp,
                      kp, k, fd.data_ptr(), fi.data_ptr(), fc.data_ptr())
    torch.cuda.synchronize()
    _assert_same(fd.cpu().numpy(), fi.cpu().numpy().astype(np.uint64), fc.cpu().numpy(), rd[:, :k], ri[:, :k], rc,
                 f"sharded {metric} G={G}")
    for ix in shards:
        ix.close()


@pytest.mark.parametrize("metric,refine,G", [("l2", 0, 2), ("l2", 3, 4), ("cosine", 2, 3), ("dot", 0, 2)])
def test_sharded_search_with_shared_bounds(oracle, metric, refine, G):
    """lb200_peer_thresholds_*: the shards publish their per-query pruning bounds into each other's
    tables (here: G shard handles of one process on one device, attached by address; across
    processes the same tables are opened through CUDA IPC and written over NVLink).  The merged
    result must stay bit-identical to the unsharded oracle search for consecutive batches with
    different queries (tables are double-buffered by batch parity), and shards searched later must
    have been pruned by the bounds of the earlier ones."""
    from lance_b200 import IvfPqIndex
    from lance_b200.index import merge_topk_device
    from lance_b200.shard import assign_partitions
    rng = np.random.default_rng(10 + G)
    d, nlist, M, k, nprobes, Q = 64, 24, 16, 10, 12, 48
    vecs = rng.normal(size=(16000, d)).astype(np.float32)
    data = build_ivf_pq(vecs, metric, nlist, M, seed=2, keep_raw=True, ivf_iters=3, pq_iters=2, sample=4096)
    owner = assign_partitions(np.diff(data.part_offsets.astype(np.int64)), G)
    kp = k * max(refine, 1)
    dev = torch.device("cuda", 0)
    shards = [IvfPqIndex(data.subset(np.nonzero(owner == g)[0])) for g in range(G)]
    ptrs = [ix.peer_thresholds_export(Q)[1] for ix in shards]
    for g, ix in enumerate(shards):
        ix.peer_thresholds_attach(G, g, ptrs=ptrs)
    pruned = 0
    for batch in range(3):
        nq = Q if batch != 1 else Q - 7  # a smaller batch in the middle
        g_d = torch.empty((G, nq, kp), dtype=torch.float32, device=dev)
        g_i = torch.empty((G, nq, kp), dtype=torch.int64, device=dev)
        g_c = torch.empty((G, nq), dtype=torch.int32, device=dev)
        g_e = torch.empty((G, nq, kp), dtype=torch.float32, device=dev)
        fd = torch.empty((nq, k), dtype=torch.float32, device=dev)
        fi = torch.empty((nq, k), dtype=torch.int64, device=dev)
        fc = torch.empty((nq,), dtype=torch.int32, device=dev)
        q = vecs[rng.integers(0, len(vecs), nq)] + 0.05 * rng.normal(size=(nq, d)).astype(np.float32)
        rd, ri, rc, _ = oracle.OracleIndex(data).search(q, k=k, nprobes=nprobes, refine_factor=refine)
        dq = torch.from_numpy(q).to(dev)
        order = list(range(G)) if batch % 2 == 0 else list(range(G - 1, -1, -1))
        for g in order:
            shards[g].search_device(dq.data_ptr(), nq, nprobes, k, refine, g_d[g].data_ptr(), g_i[g].data_ptr(),
                                    g_c[g].data_ptr(), d_out_exact=g_e[g].data_ptr())
            torch.cuda.synchronize()
        cnt = g_c.cpu().numpy()
        pruned += int((cnt[order[-1]] < kp).sum())
        merge_topk_device(0, g_d.data_ptr(), g_i.data_ptr(), g_e.data_ptr() if refine else None, g_c.data_ptr(), G, nq,
                          kp, kp, k, fd.data_ptr(), fi.data_ptr(), fc.data_ptr())
        torch.cuda.synchronize()
        _assert_same(fd.cpu().numpy(), fi.cpu().numpy().astype(np.uint64), fc.cpu().numpy(), rd[:, :k],
                     ri[:, :k], rc, f"shared bounds {metric} G={G} batch {batch}")
    assert pruned > 0, "the last shard of a batch never returned a shortened list: bounds were not shared"
    for ix in shards:
        ix.peer_thresholds_detach()
        ix.close()


@pytest.mark.parametrize("prepass", ["0", "1"])
@pytest.mark.parametrize("metric,d,M", [("l2", 128, 16), ("cosine", 768, 48), ("dot", 768, 48), ("cosine", 1536, 96),
                                        ("dot", 128, 16), ("l2", 64, 4)])
def test_table_prepass_equals_in_kernel_tables(oracle, metric, d, M, prepass, monkeypatch):
    """ADC tables built by lut_prepass_kernel (forced on) or by the scanning CTA (forced off) give the
    same bit-identical results as the oracle -- merged search, per-pair search and a shard with
    empty partitions (tables of pairs without local rows are skipped)."""
    from lance_b200 import IvfPqIndex, Query
    monkeypatch.setenv("LB200_LUT_PREPASS", prepass)
    sizes = [0, 300, 5, 0, 1000, 33, 2049, 0, 64, 7, 700, 4100, 9, 0, 0, 130]
    nlist = len(sizes)
    data = _ragged_index(metric, d, nlist, M, sizes, seed=d + M, keep_raw=True)
    o = oracle.OracleIndex(data)
    ix = IvfPqIndex(data)
    rng = np.random.default_rng(8)
    Q = 77
    q = rng.normal(size=(Q, d)).astype(np.float32)
    for nprobes, refine in ((1, 0), (5, 0), (nlist, 3)):
        rd, ri, rc, _ = o.search(q, k=7, nprobes=nprobes, refine_factor=refine)
        gd, gi, gc = ix.search(Query(key=q, k=7, nprobes=nprobes, refine_factor=refine or None))
        _assert_same(gd, gi, gc, rd[:, :gd.shape[1]], ri[:, :gd.shape[1]], rc, f"prepass={prepass} {metric} P={nprobes}")
    parts = np.tile(np.arange(nlist, dtype=np.uint32), (5, 1))
    key = q[:5] if metric != "cosine" else np.stack([oracle.normalize(x) for x in q[:5]])
    gd, gi, gc = ix.search_in_partition(parts, key, 6)
    for i in range(5):
        for p in range(nlist):
            rd1, ri1 = o.search_in_partition(p, key[i], 6)
            n = len(rd1)
            assert gc[i, p] == n
            assert gi[i, p, :n].tolist() == ri1.tolist() and _bits(gd[i, p, :n]).tolist() == _bits(rd1).tolist()
    ix.close()


@pytest.mark.parametrize("Q", [1, 3])
def test_single_query_on_large_partitions(oracle, Q):
    """Tiny batches cut large partitions into chunks to fill the GPU; the number of chunks must
    respect the merge capacity (P * chunks * k' <= 16384) -- a one-query search with k = 100 over
    50k-row partitions used to fail with LB200_EUNSUPPORTED."""
    from lance_b200 import IvfPqIndex, Query
    data = random_index(400_000, 128, 8, 16, "l2", seed=11)
    q = queries_near_centroids(data, Q, seed=12)
    rd, ri, rc, _ = oracle.OracleIndex(data).search(q, k=100, nprobes=8)
    ix = IvfPqIndex(data)
    gd, gi, gc = ix.search(Query(key=q, k=100, nprobes=8))
    _assert_same(gd, gi, gc, rd[:, :100], ri[:, :100], rc, f"single query Q={Q}")
    ix.close()


def test_many_probes_large_k_two_level_merge(oracle):
    """nprobes * k' > 16384 candidates per query (64 probes x k' = 300): the per-query merge runs in
    two levels; result identical to the oracle's single sort."""
    from lance_b200 import IvfPqIndex, Query
    rng = np.random.default_rng(21)
    vecs = rng.normal(size=(30000, 64)).astype(np.float32)
    data = build_ivf_pq(vecs, "l2", 64, 16, seed=3, keep_raw=True, ivf_iters=3, pq_iters=2, sample=4096)
    q = vecs[rng.integers(0, len(vecs), 9)] + 0.05 * rng.normal(size=(9, 64)).astype(np.float32)
    o = oracle.OracleIndex(data)
    ix = IvfPqIndex(data)
    for k, refine in ((100, 3), (300, 0)):
        rd, ri, rc, _ = o.search(q, k=k, nprobes=64, refine_factor=refine)
        gd, gi, gc = ix.search(Query(key=q, k=k, nprobes=64, refine_factor=refine or None))
        _assert_same(gd, gi, gc, rd[:, :gd.shape[1]], ri[:, :gd.shape[1]], rc, f"two-level merge k={k} refine={refine}")
    ix.close()


# ---- cfg 3 / cfg 5 SHAPES above toy size: M = 96 / 48, the real partition counts, >= 100k rows,
# so the QL = 2 / 16-warp / table pre-pass instantiations are checked on the shapes they are benchmarked on ------------
@pytest.mark.parametrize("name,N,d,nlist,M,metric,nprobes,refine,Q", [
    ("cfg3-shape", 120_000, 1536, 1024, 96, "cosine", 102, 5, 96),
    ("cfg5-shape", 400_000, 768, 2048, 48, "dot", 32, 0, 160),
    ("cfg4-shape-many-probes", 600_000, 128, 4096, 16, "l2", 32, 0, 300),
])
def test_baseline_config_shapes_bit_identical(oracle, name, N, d, nlist, M, metric, nprobes, refine, Q):
    from lance_b200 import IvfPqIndex, Query
    data = random_index(N, d, nlist, M, metric, seed=31, keep_raw=refine > 0)
    if metric == "cosine":                     # a cosine index holds L2 structures over normalised data
        data.centroids /= np.linalg.norm(data.centroids, axis=1, keepdims=True)
        data.codebook *= np.float32(0.02 / 16.0)
    q = queries_near_centroids(data, Q, seed=32, scale=0.05 if metric == "cosine" else None)
    ix = IvfPqIndex(data)
    gd, gi, gc = ix.search(Query(key=q, k=10, nprobes=nprobes, refine_factor=refine or None))
    st = ix.stats()
    assert st["scan_variant"] == 3 and st["scan_ql"] == (4 if M <= 48 else 2)
    o = oracle.OracleIndex(data)
    rd, ri, rc, rp = o.search(q, k=10, nprobes=nprobes, refine_factor=refine)
    _assert_same(gd, gi, gc, rd[:, :10], ri[:, :10], np.minimum(rc, 10), name)
    keys = np.stack([oracle.normalize(x) for x in q]) if metric == "cosine" else q
    assert ix.find_partitions(keys, nprobes).tolist() == rp.tolist(), f"{name}: probed partitions differ"
    # the same pairs through search_in_partition (per-pair lists, no shared bound): first 8 queries
    od, oi, oc = ix.search_in_partition(rp[:8], keys[:8], 10)
    for qi in range(8):
        for j in range(rp.shape[1]):
            pd, pr = o.search_in_partition(int(rp[qi, j]), keys[qi], 10)
            n = len(pd)
            assert oc[qi, j] == n and oi[qi, j, :n].tolist() == pr.tolist()
            assert _bits(od[qi, j, :n]).tolist() == _bits(pd).tolist()
    ix.close()


# ---- partition counts that are not a multiple of the prefix kernels' block (1024 threads x a run per thread), with
# empty partitions in the runs: offsets_kernel / queue_offsets_kernel (both queue orders) against the oracle ----------
@pytest.mark.parametrize("nlist,lpt", [(1500, "0"), (1500, "1"), (2049, "1"), (1023, "0")])
def test_ragged_partition_counts_through_worklist_kernels(oracle, nlist, lpt, monkeypatch):
    from lance_b200 import IvfPqIndex, Query
    monkeypatch.setenv("LB200_LPT", lpt)
    rng = np.random.default_rng(nlist)
    sizes = rng.integers(0, 90, size=nlist)
    sizes[rng.integers(0, nlist, size=nlist // 7)] = 0            # empty partitions inside the threads' runs
    sizes[[0, nlist - 1]] = [37, 0]
    data = _ragged_index("l2", 32, nlist, 16, sizes, seed=nlist + 1)
    q = queries_near_centroids(data, 48, seed=nlist + 2)
    ix = IvfPqIndex(data)
    o = oracle.OracleIndex(data)
    for nprobes, k in [(40, 10), (nlist, 5)]:
        gd, gi, gc = ix.search(Query(key=q, k=k, nprobes=nprobes))
        rd, ri, rc, _ = o.search(q, k=k, nprobes=nprobes, refine_factor=0)
        _assert_same(gd, gi, gc, rd[:, :k], ri[:, :k], np.minimum(rc, k), f"nlist={nlist} lpt={lpt} nprobes={nprobes}")
    ix.close()

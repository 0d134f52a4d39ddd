"""IVF_PQ over a Float16 column (lb200_index_create_f16 / lb200_search_f16, csrc/f16pq.cuh) against the CPU checker's f16
restatement (oracle/lance_oracle.c lo_search_batch_f16): probed partitions are implied by identical results; row ids
and distance bits must be identical."""
import numpy as np
import pytest

from lance_b200.synth import IvfPqData, build_ivf_pq

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _f16_index(metric, d, nlist, M, n, seed, scale=1.0):
    """A real-built index whose centroids / codebook / raw vectors are then rounded to f16 (what a Float16 column holds)."""
    rng = np.random.default_rng(seed)
    cents = rng.normal(size=(nlist, d)) * 2.0
    vecs = ((cents[rng.integers(0, nlist, n)] + rng.normal(size=(n, d))) * scale).astype(np.float16).astype(np.float32)
    data = build_ivf_pq(vecs, metric, nlist, M, seed=seed, keep_raw=True, ivf_iters=3, pq_iters=2, sample=4096)
    f = lambda a: np.asarray(a, dtype=np.float32).astype(np.float16).astype(np.float32)
    return IvfPqData(metric, d, nlist, M, f(data.centroids), f(data.codebook), data.part_offsets, data.codes, data.row_ids,
                     f(data.raw_vectors)), vecs


@pytest.mark.parametrize("metric,d,M", [("l2", 128, 16), ("l2", 256, 16), ("cosine", 96, 16), ("cosine", 512, 32), ("dot", 128, 16),
                                        ("dot", 768, 48), ("l2", 1536, 96)])
def test_f16_ivf_pq_search_parity(oracle, metric, d, M):
    from lance_b200 import IvfPqIndexF16, Query
    data, vecs = _f16_index(metric, d, 12, M, 6000, seed=d + M)
    rng = np.random.default_rng(3)
    Q = 24
    q = (vecs[rng.integers(0, len(vecs), Q)] + 0.05 * rng.normal(size=(Q, d))).astype(np.float16)
    ix, ox = IvfPqIndexF16(data), oracle.OracleIndexF16(data)
    try:
        probe = ox.search(q, 50, 12)[0]
        fin = probe[np.isfinite(probe)]
        lo, hi = float(np.quantile(fin, 0.2)), float(np.quantile(fin, 0.8))
        for k, nprobes, rf, lo_, hi_ in ((10, 4, 0, None, None), (1, 1, 0, None, None), (100, 12, 0, None, None),
                                         (10, 5, 3, None, None), (10, 6, 0, lo, hi), (5, 3, 2, None, hi)):
            rd, ri, rc, _ = ox.search(q, k, nprobes, rf, lo_, hi_)
            gd, gi, gc = ix.search(Query(q, k=k, nprobes=nprobes, refine_factor=rf or None, lower_bound=lo_, upper_bound=hi_))
            what = f"{metric} d={d} M={M} k={k} P={nprobes} rf={rf} [{lo_},{hi_})"
            assert gc.tolist() == rc.tolist(), what
            for j in range(Q):
                n = int(rc[j])
                assert gi[j, :n].tolist() == ri[j, :n].tolist(), f"{what} q{j}"
                assert _bits(gd[j, :n]).tolist() == _bits(rd[j, :n]).tolist(), f"{what} q{j}"
    finally:
        ix.close()


def test_f16_results_differ_from_the_f32_arithmetic_where_they_should(oracle):
    """The f16 path is not "f32 on widened inputs": the residual is rounded to f16 and the sub-vector sums run in
    l2_scalar order (dsub 8: sequential; f32 uses the l2_once tree).  On the same widened index the two checkers give
    different distance bits for most rows -- if the GPU path ever collapses to the f32 kernels, the parity test above
    fails, and this test shows it would be noticed."""
    data, vecs = _f16_index("l2", 128, 8, 16, 3000, seed=5)
    q = vecs[:16].astype(np.float16)
    d16 = oracle.OracleIndexF16(data).search(q, 10, 8)[0]
    d32 = oracle.OracleIndex(data).search(q.astype(np.float32), 10, 8)[0]
    assert (_bits(d16) != _bits(d32)).mean() > 0.3


def test_f16_handle_rejects_f32_entry_points_and_bad_shapes(oracle):
    from lance_b200 import IvfPqIndexF16, LanceB200Error
    import ctypes as C
    from lance_b200 import _lib
    data, vecs = _f16_index("l2", 64, 4, 16, 800, seed=9)
    ix = IvfPqIndexF16(data)
    q = vecs[:2].astype(np.float32)
    od = np.empty((2, 5), np.float32); oi = np.empty((2, 5), np.uint64); oc = np.empty(2, np.uint32)
    rc = _lib.lib().lb200_search(ix._h, q.ctypes.data_as(C.c_void_p), C.c_uint32(2), C.c_uint32(2), C.c_uint32(5), C.c_uint32(0), None, None,
                                 od.ctypes.data_as(C.c_void_p), oi.ctypes.data_as(C.c_void_p), oc.ctypes.data_as(C.c_void_p))
    assert rc != 0 and b"Float16" in _lib.lib().lb200_last_error()
    ix.close()
    bad = IvfPqData("l2", 64, 4, 8, data.centroids, np.zeros((8, 256, 8), np.float32), data.part_offsets,
                    np.zeros(800 * 8, np.uint8), data.row_ids, None)
    with pytest.raises(LanceB200Error, match="num_sub_vectors"):
        IvfPqIndexF16(bad)


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_f16_find_partitions_and_search_in_partition(oracle, metric):
    """The two calls the exec nodes make, on a Float16 handle: keys as given (normalised by the caller for cosine)."""
    from lance_b200 import IvfPqIndexF16
    data, vecs = _f16_index(metric, 128, 10, 16, 4000, seed=41)
    rng = np.random.default_rng(42)
    q = (vecs[rng.integers(0, len(vecs), 7)] + 0.05 * rng.normal(size=(7, 128))).astype(np.float16)
    keys = np.stack([oracle.normalize_f16(x) for x in q]) if metric == "cosine" else q
    ix, ox = IvfPqIndexF16(data), oracle.OracleIndexF16(data)
    try:
        ids, ds = ix.find_partitions(keys, 4, return_distances=True)
        for j in range(len(q)):
            ri, rd = ox.find_partitions(keys[j], 4)
            assert ids[j].tolist() == ri.tolist()
            assert _bits(ds[j]).tolist() == _bits(rd).tolist()
        parts = np.stack([np.roll(np.arange(10, dtype=np.uint32), j)[:3] for j in range(len(q))])
        for lo, hi in ((None, None), (None, float(np.median(ds)))):
            od, oi, oc = ix.search_in_partition(parts, keys, 15, lower_bound=lo, upper_bound=hi)
            for j in range(len(q)):
                for c in range(3):
                    rd, rr = ox.search_in_partition(int(parts[j, c]), keys[j], 15, lo, hi)
                    n = int(oc[j, c])
                    assert n == len(rr) and oi[j, c, :n].tolist() == rr.tolist(), (metric, j, c)
                    assert _bits(od[j, c, :n]).tolist() == _bits(rd).tolist()
    finally:
        ix.close()

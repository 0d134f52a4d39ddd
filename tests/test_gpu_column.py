"""Resident vector column + tensor-core candidate filter (lb200_column_*, csrc/flat_tc.cuh): results must equal the exact
flat kNN path (lb200_flat_topk, itself bit-identical to the oracle: test_flat_parity) -- indices, distance bits, counts --
whether the filter runs, falls back, or is not eligible."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _same(got, want):
    assert got[2].tolist() == want[2].tolist()
    assert got[1].tolist() == want[1].tolist()
    assert got[0].view(np.uint32).tolist() == want[0].view(np.uint32).tolist()


def _sift(rng, n, d):
    return np.clip(np.rint(rng.normal(64, 32, size=(n, d))), 0, 218).astype(np.float32)


@pytest.mark.parametrize("d,n,Q,k", [(128, 100_003, 9, 10), (128, 262_144, 300, 1), (64, 70_000, 128, 100), (192, 66_001, 5, 10),
                                     (256, 80_000, 130, 10), (128, 500_000, 256, 10)])
def test_column_filter_equals_exact_path(d, n, Q, k):
    from lance_b200 import VectorColumn, flat_knn
    rng = np.random.default_rng(d + Q)
    v = _sift(rng, n, d)
    v[1000:1100] = v[5]                                          # exact duplicates: ties in and around the top k
    q = np.concatenate([v[rng.integers(0, n, Q // 2 + 1)] + rng.normal(0, 2, size=(Q // 2 + 1, d)).astype(np.float32),
                        _sift(rng, Q, d)])[:Q]
    q[0] = v[5]
    col = VectorColumn(v)
    got = col.flat_knn(q, "l2", k)
    st = col.last_call()
    assert st["used_tc"] and not st["fell_back"], st
    assert st["candidates"] >= Q * k and st["candidates"] < Q * 4096
    _same(got, flat_knn(q, "l2", v, k))
    got2 = col.flat_knn(q[::-1].copy(), "l2", k)                 # the column's scratch is reused
    assert got2[1][::-1].tolist() == got[1].tolist()
    col.close()


def test_column_filter_on_embeddings_and_ordered_rows():
    """Unit-norm rows (distances in [0, 4], neighbours close together) stored CLUSTER BY CLUSTER: the thresholds come from a
    strided sample, so the row order must not matter."""
    from lance_b200 import VectorColumn, flat_knn
    rng = np.random.default_rng(7)
    d, n, Q = 128, 200_000, 64
    cent = rng.normal(size=(50, d)).astype(np.float32)
    lab = np.sort(rng.integers(0, 50, n))
    v = cent[lab] + 0.3 * rng.normal(size=(n, d)).astype(np.float32)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    q = v[rng.integers(0, n, Q)] + 0.01 * rng.normal(size=(Q, d)).astype(np.float32)
    col = VectorColumn(v)
    got = col.flat_knn(q, "l2", 10)
    assert col.last_call()["used_tc"]
    _same(got, flat_knn(q, "l2", v, 10))
    col.close()


def test_column_falls_back_when_the_bound_is_useless():
    """Rows = a huge common offset + small differences: the bf16 error bound dwarfs every distance gap, every row is a
    candidate, the buffers overflow -- and the exact path answers."""
    from lance_b200 import VectorColumn, flat_knn
    rng = np.random.default_rng(9)
    d, n, Q = 128, 100_000, 16
    v = (1000.0 + rng.normal(size=(n, d))).astype(np.float32)
    q = (1000.0 + rng.normal(size=(Q, d))).astype(np.float32)
    col = VectorColumn(v)
    got = col.flat_knn(q, "l2", 10)
    st = col.last_call()
    assert st["used_tc"] and st["fell_back"], st
    _same(got, flat_knn(q, "l2", v, 10))
    col.close()


def test_column_cases_the_filter_does_not_serve():
    from lance_b200 import VectorColumn, flat_knn
    rng = np.random.default_rng(11)
    d, n, Q = 128, 70_000, 12
    v = _sift(rng, n, d)
    q = _sift(rng, Q, d)
    col = VectorColumn(v)
    base = flat_knn(q, "l2", v, 50)[0]
    lo, hi = float(np.median(base)), float(base.max())
    _same(col.flat_knn(q, "l2", 10, lower_bound=lo, upper_bound=hi), flat_knn(q, "l2", v, 10, lower_bound=lo, upper_bound=hi))
    assert not col.last_call()["used_tc"]
    col.close()
    valid = np.packbits(rng.random(n) < 0.9, bitorder="little")
    colv = VectorColumn(v, validity=valid)
    _same(colv.flat_knn(q, "l2", 10), flat_knn(q, "l2", v, 10, validity=valid))
    colv.close()
    small = VectorColumn(v[:5000, :100].copy())                  # small, dim not a multiple of 64
    _same(small.flat_knn(q[:, :100].copy(), "l2", 10), flat_knn(q[:, :100].copy(), "l2", v[:5000, :100].copy(), 10))
    small.close()


def test_column_with_non_finite_values_uses_the_exact_path():
    """The filter's error bound assumes ordinary values: a NaN / inf row switches the column to the exact kernels,
    a NaN / inf query switches that call."""
    from lance_b200 import VectorColumn, flat_knn
    rng = np.random.default_rng(12)
    d, n, Q = 128, 70_000, 6
    v = _sift(rng, n, d)
    q = _sift(rng, Q, d)
    bad = v.copy()
    bad[17, 3] = np.nan
    bad[99, 0] = np.inf
    col = VectorColumn(bad)
    _same(col.flat_knn(q, "l2", 10), flat_knn(q, "l2", bad, 10))
    assert not col.last_call()["used_tc"]
    col.close()
    col = VectorColumn(v)
    qb = q.copy()
    qb[2, 5] = np.nan
    qb[4, 7] = -np.inf
    _same(col.flat_knn(qb, "l2", 10), flat_knn(qb, "l2", v, 10))
    assert col.last_call()["fell_back"]
    _same(col.flat_knn(q, "l2", 10), flat_knn(q, "l2", v, 10))
    st = col.last_call()
    assert st["used_tc"] and not st["fell_back"]
    col.close()


@pytest.mark.parametrize("metric", ["cosine", "dot"])
@pytest.mark.parametrize("d,n,Q,k", [(128, 100_003, 9, 10), (64, 150_000, 130, 100), (256, 70_000, 128, 1), (192, 262_144, 33, 10)])
def test_column_filter_cosine_and_dot(metric, d, n, Q, k):
    """The same bound-and-verify scheme for the other two metrics of compute_distance (flat.rs:28-86; the reference's flat
    benchmark runs all three): dot on the raw bf16 rows with E = e1 |q||y| + e0, cosine on rows normalised before the
    bf16 rounding with a constant E; exact re-score by the refine step's kernels.  Indices, distance bits and counts must
    equal the exact path's, including duplicates (ties) and a zero row (NaN cosine distance: never returned)."""
    from lance_b200 import VectorColumn, flat_knn
    rng = np.random.default_rng(d + Q + len(metric))
    v = (rng.normal(size=(n, d)) * rng.uniform(0.5, 2.0, size=(n, 1))).astype(np.float32)
    v[1000:1050] = v[5]
    v[77] = 0.0
    q = np.concatenate([v[rng.integers(0, n, Q // 2 + 1)] + 0.1 * rng.normal(size=(Q // 2 + 1, d)).astype(np.float32),
                        rng.normal(size=(Q, d)).astype(np.float32)])[:Q]
    q[0] = v[5]
    col = VectorColumn(v)
    got = col.flat_knn(q, metric, k)
    st = col.last_call()
    assert st["used_tc"] and not st["fell_back"], st
    assert st["candidates"] >= Q * k
    _same(got, flat_knn(q, metric, v, k))
    _same(col.flat_knn(q[:3], "l2", k), flat_knn(q[:3], "l2", v, k))     # the L2 form on the same column afterwards
    _same(col.flat_knn(q[::-1].copy(), metric, k), flat_knn(q[::-1].copy(), metric, v, k))
    col.close()


def test_column_large_query_batches_are_chunked():
    """More than 2048 queries per call: the column answers in chunks (bounded scratch); statistics add up."""
    from lance_b200 import VectorColumn, flat_knn
    rng = np.random.default_rng(77)
    d, n, Q, k = 64, 70_000, 2100, 5
    v = _sift(rng, n, d)
    q = _sift(rng, Q, d)
    col = VectorColumn(v)
    got = col.flat_knn(q, "l2", k)
    st = col.last_call()
    assert st["used_tc"] and not st["fell_back"] and st["candidates"] >= Q * k
    _same(got, flat_knn(q, "l2", v, k))
    col.close()

"""GPU parity of the 4-bit PQ path (SURVEY 8(a) row A5'': compute_pq_distance_4bit / quantize_distance_table,
rust/lance-index/src/vector/pq/distance.rs:138-229) against the oracle: bit-identical distances, identical row ids.
The scan between the f32 table and the dequantisation is integer arithmetic (u8 table, wrapping u8 pair sums on whole
16-row blocks, separate adds on the n % 16 tail), so exactness is the natural bar."""
import numpy as np
import pytest
import torch

from lance_b200.synth import IvfPqData

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from lance_b200 import _lib
    assert _lib.lib() is not None


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _index4(metric, d, nlist, M, sizes, seed, keep_raw=False, scale=1.0):
    rng = np.random.default_rng(seed)
    sizes = np.asarray(sizes, dtype=np.int64)
    N = int(sizes.sum())
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    return IvfPqData(metric, d, nlist, M, rng.normal(size=(nlist, d)).astype(np.float32),
                     (scale * rng.normal(size=(M, 16, d // M))).astype(np.float32), offs,
                     rng.integers(0, 256, size=N * M // 2, dtype=np.uint8),
                     rng.permutation(N * 3)[:N].astype(np.uint64),
                     rng.normal(size=(N, d)).astype(np.float32) if keep_raw else None, 4)


def _check_search(oracle, data, q, cases):
    """Every case through both device paths: the fused histogram-select kernel (default) and the distance-scratch +
    flat_select path (LB200_PQ4_SIMPLE=1, read per call)."""
    import os
    from lance_b200 import IvfPqIndex, Query
    ix, ox = IvfPqIndex(data), oracle.OracleIndex(data)
    try:
        for k, nprobes, rf, lo, hi in cases:
            rd, ri, rc, _ = ox.search(q, k, nprobes, rf, lo, hi)
            for simple in (False, True):
                os.environ.pop("LB200_PQ4_SIMPLE", None)
                if simple:
                    os.environ["LB200_PQ4_SIMPLE"] = "1"
                what = f"{'simple' if simple else 'fused'} k={k} P={nprobes} rf={rf} [{lo},{hi})"
                gd, gi, gc = ix.search(Query(q, k=k, nprobes=nprobes, refine_factor=rf or None, lower_bound=lo,
                                             upper_bound=hi))
                assert gc.tolist() == rc.tolist(), f"counts {what}"
                for j in range(len(q)):
                    n = int(rc[j])
                    assert gi[j, :n].tolist() == ri[j, :n].tolist(), f"row ids {what} q{j}"
                    assert _bits(gd[j, :n]).tolist() == _bits(rd[j, :n]).tolist(), f"distances {what} q{j}"
        assert ix.stats()["scan_variant"] == 4
    finally:
        os.environ.pop("LB200_PQ4_SIMPLE", None)
        ix.close()


@pytest.mark.parametrize("metric,d,M", [("l2", 128, 16), ("l2", 32, 2), ("l2", 30, 6), ("cosine", 96, 12),
                                        ("dot", 64, 8), ("dot", 768, 48), ("l2", 1536, 96)])
def test_search_4bit_parity_ragged(oracle, metric, d, M):
    sizes = [0, 1, 15, 16, 17, 100, 333, 0, 1000, 47, 256, 31]
    data = _index4(metric, d, len(sizes), M, sizes, seed=d + M, keep_raw=True)
    rng = np.random.default_rng(1)
    q = rng.normal(size=(13, d)).astype(np.float32)
    dall = oracle.OracleIndex(data).search(q, 50, len(sizes))[0]
    lo, hi = float(np.quantile(dall[np.isfinite(dall)], 0.1)), float(np.quantile(dall[np.isfinite(dall)], 0.7))
    _check_search(oracle, data, q, [(1, 1, 0, None, None), (10, 4, 0, None, None), (10, len(sizes), 0, None, None),
                                    (100, 6, 0, None, None), (10, 5, 0, lo, hi), (7, 5, 4, None, None),
                                    (10, len(sizes) + 3, 2, None, None)])


@pytest.mark.parametrize("metric", ["l2", "dot"])
def test_search_4bit_large_partitions_and_large_k(oracle, metric):
    """Partitions above 4096 rows take the early bound of the fused kernel's first pass; k' up to 512."""
    sizes = [5000, 1025, 1024, 2500, 3, 0, 7777]
    data = _index4(metric, 64, len(sizes), 16, sizes, seed=21, keep_raw=True)
    rng = np.random.default_rng(3)
    q = rng.normal(size=(9, 64)).astype(np.float32)
    dall = oracle.OracleIndex(data).search(q, 100, len(sizes))[0]
    lo, hi = float(np.quantile(dall, 0.3)), float(np.quantile(dall, 0.9))
    _check_search(oracle, data, q, [(10, 7, 0, None, None), (512, 3, 0, None, None), (64, 7, 8, None, None),
                                    (100, 7, 0, lo, hi), (1, 7, 0, hi, None), (5, 7, 0, None, lo)])


def test_search_4bit_many_ties_order_by_row_id(oracle):
    """A tiny codebook scale makes most quantised distances collide: the order inside a tie is by row id."""
    sizes = [500, 1700, 3000]
    data = _index4("l2", 16, 3, 2, sizes, seed=5)
    data.codes = (data.codes & 0x11).astype(np.uint8)          # 4 distinct codes per byte -> massive ties
    rng = np.random.default_rng(2)
    q = rng.normal(size=(6, 16)).astype(np.float32)
    _check_search(oracle, data, q, [(20, 3, 0, None, None), (128, 2, 0, None, None)])


def test_search_in_partition_4bit(oracle):
    from lance_b200 import IvfPqIndex
    sizes = [40, 0, 23, 160]
    data = _index4("l2", 32, 4, 8, sizes, seed=3)
    ix, ox = IvfPqIndex(data), oracle.OracleIndex(data)
    rng = np.random.default_rng(4)
    q = rng.normal(size=(5, 32)).astype(np.float32)
    parts = np.array([[0, 3], [1, 2], [3, 0], [2, 2], [3, 1]], dtype=np.uint32)
    od, oi, oc = ix.search_in_partition(parts, q, 25)
    for i in range(5):
        for j in range(2):
            rd, rr = ox.search_in_partition(int(parts[i, j]), q[i], 25)
            n = int(oc[i, j])
            assert n == len(rr) and oi[i, j, :n].tolist() == rr.tolist()
            assert _bits(od[i, j, :n]).tolist() == _bits(rd).tolist()
    ix.close()


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_search_in_partition_4bit_prefilter(oracle, metric):
    """flat/index.rs:106-122: with a prefilter the reference visits only the allowed rows and takes each distance from
    PQDistCalculator::distance(id) (pq/storage.rs:602-617) -- the unquantised f32 table, not the u8 one.  The handle
    keeps its rows sorted by row id, so the caller's bit order is permuted on upload."""
    from lance_b200 import IvfPqIndex
    sizes = [40, 0, 23, 160, 1000]
    data = _index4(metric, 32, 5, 8, sizes, seed=5)
    ix, ox = IvfPqIndex(data), oracle.OracleIndex(data)
    rng = np.random.default_rng(6)
    q = rng.normal(size=(6, 32)).astype(np.float32)
    parts = np.array([[0, 3], [4, 2], [3, 0], [2, 4], [3, 1], [4, 4]], dtype=np.uint32)
    masks = {p: rng.random(n) < frac for p, n, frac in ((0, 40, 0.5), (2, 23, 0.1), (3, 160, 0.9), (4, 1000, 0.02))}
    masks[2][:] = False                                                 # nothing allowed in partition 2

    def bits(p):
        m = masks.get(p)
        if m is None:
            return None
        b = np.zeros((len(m) + 63) // 64 + 1, dtype=np.uint64)
        on = np.nonzero(m)[0]
        np.bitwise_or.at(b, on >> 6, np.uint64(1) << (on & 63).astype(np.uint64))
        return b

    pf = ix.prefilter(masks)
    for lo, hi in ((None, None), (0.5, None), (None, 30.0)):
        for how in (masks, pf):
            od, oi, oc = ix.search_in_partition(parts, q, 12, lower_bound=lo, upper_bound=hi, prefilter=how)
            for i in range(len(q)):
                for j in range(2):
                    p = int(parts[i, j])
                    rd, rr = ox.search_in_partition(p, q[i], 12, lo, hi, bitmask=bits(p))
                    n = int(oc[i, j])
                    assert n == len(rr) and oi[i, j, :n].tolist() == rr.tolist(), (metric, lo, hi, i, j)
                    assert _bits(od[i, j, :n]).tolist() == _bits(rd).tolist()
    pf.close()
    ix.close()


@pytest.mark.parametrize("metric,refine,G", [("l2", 0, 2), ("cosine", 3, 3), ("dot", 0, 2)])
def test_sharded_4bit_search_merge_equals_single_index(oracle, metric, refine, G):
    """SURVEY 8(e) on 4-bit handles: every shard returns its merged k' list (+ exact distances when it refines),
    lb200_merge_topk_device makes the final cut; bit-identical to the unsharded search."""
    import torch
    from lance_b200 import IvfPqIndex
    from lance_b200.index import merge_topk_device
    from lance_b200.shard import assign_partitions
    sizes = [300, 0, 77, 512, 1, 950, 64, 200]
    data = _index4(metric, 32, len(sizes), 8, sizes, seed=7, keep_raw=True)
    rng = np.random.default_rng(8)
    Q, k, nprobes = 20, 10, 5
    q = rng.normal(size=(Q, 32)).astype(np.float32)
    rd, ri, rc, _ = oracle.OracleIndex(data).search(q, k=k, nprobes=nprobes, refine_factor=refine)
    owner = assign_partitions(np.asarray(sizes, dtype=np.int64), G)
    kp = k * max(refine, 1)
    dev = torch.device("cuda", 0)
    dq = torch.from_numpy(q).to(dev)
    g_d = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
    g_i = torch.empty((G, Q, kp), dtype=torch.int64, device=dev)
    g_c = torch.empty((G, Q), dtype=torch.int32, device=dev)
    g_e = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
    shards = [IvfPqIndex(data.subset(np.nonzero(owner == g)[0])) for g in range(G)]
    for g, ix in enumerate(shards):
        ix.search_device(dq.data_ptr(), Q, nprobes, k, refine, g_d[g].data_ptr(), g_i[g].data_ptr(), g_c[g].data_ptr(),
                         d_out_exact=g_e[g].data_ptr())
    torch.cuda.synchronize()
    fd = torch.empty((Q, k), dtype=torch.float32, device=dev)
    fi = torch.empty((Q, k), dtype=torch.int64, device=dev)
    fc = torch.empty((Q,), dtype=torch.int32, device=dev)
    merge_topk_device(0, g_d.data_ptr(), g_i.data_ptr(), g_e.data_ptr() if refine else None, g_c.data_ptr(), G, Q, kp, kp, k,
                      fd.data_ptr(), fi.data_ptr(), fc.data_ptr())
    torch.cuda.synchronize()
    fd, fi, fc = fd.cpu().numpy(), fi.cpu().numpy().astype(np.uint64), fc.cpu().numpy()
    for j in range(Q):
        n = int(rc[j])
        assert int(fc[j]) == n
        assert fi[j, :n].tolist() == ri[j, :n].tolist(), (metric, j)
        assert _bits(fd[j, :n]).tolist() == _bits(rd[j, :n]).tolist()
    for ix in shards:
        ix.close()


def test_4bit_real_built_index(oracle):
    """Codes produced by the oracle's transform_impl::<4> on residuals of clustered data; refine on raw vectors."""
    rng = np.random.default_rng(8)
    d, nlist, M, n = 64, 6, 16, 3000
    cents = rng.normal(scale=4.0, size=(nlist, d)).astype(np.float32)
    part = rng.integers(0, nlist, size=n)
    x = (cents[part] + rng.normal(size=(n, d))).astype(np.float32)
    order = np.argsort(part, kind="stable")
    x, part = x[order], part[order]
    cb = rng.normal(size=(M, 16, d // M)).astype(np.float32)
    packed = oracle.pq_quantize4("l2", cb, x - cents[part])     # [n][M/2]
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(np.bincount(part, minlength=nlist))
    codes = np.concatenate([packed[int(offs[p]):int(offs[p + 1])].T.reshape(-1) for p in range(nlist)])
    data = IvfPqData("l2", d, nlist, M, cents, cb, offs, codes, rng.permutation(n).astype(np.uint64), x, 4)
    q = (x[::250] + rng.normal(scale=0.1, size=x[::250].shape)).astype(np.float32)
    _check_search(oracle, data, q, [(10, 2, 0, None, None), (10, 3, 10, None, None), (1, nlist, 50, None, None)])


def test_4bit_create_errors():
    import ctypes as C
    from lance_b200 import IvfPqIndex, LanceB200Error
    data = _index4("l2", 30, 2, 6, [5, 5], seed=1)
    bad = IvfPqData(**{**data.__dict__})
    bad.M = 3
    bad.codebook = np.zeros((3, 16, 10), np.float32)
    with pytest.raises(LanceB200Error, match="divisible by 2"):
        IvfPqIndex(bad)


@pytest.mark.parametrize("metric", ["l2", "dot"])
def test_search_4bit_non_finite_tables(oracle, metric):
    """Queries whose ADC table is not finite (tests/test_oracle_pq4.py states what the reference's expressions give):
    an overflowing sub-vector makes qmax = +inf and every distance NaN (no rows returned); a NaN sub-vector is skipped
    by the min/max folds and quantised to 0, the rest of the table works as usual.  Both device paths, explicit
    partitions so that the coarse step (whose distances are NaN / inf too) does not decide what is compared."""
    import os
    from lance_b200 import IvfPqIndex
    sizes = [40, 100, 23, 333]
    data = _index4(metric, 32, 4, 8, sizes, seed=21)
    rng = np.random.default_rng(22)
    q = rng.normal(size=(4, 32)).astype(np.float32)
    q[1, 4:8] = 3e19            # (q - c)^2 or q . c overflows in sub-vector 1
    q[2, 8:12] = np.nan          # sub-vector 2 of the table is NaN
    q[3, :] = np.inf
    parts = np.tile(np.array([[3, 0, 1]], dtype=np.uint32), (4, 1))
    ix, ox = IvfPqIndex(data), oracle.OracleIndex(data)
    try:
        for simple in (False, True):
            os.environ.pop("LB200_PQ4_SIMPLE", None)
            if simple:
                os.environ["LB200_PQ4_SIMPLE"] = "1"
            od, oi, oc = ix.search_in_partition(parts, q, 15)
            for i in range(4):
                for j in range(3):
                    rd, rr = ox.search_in_partition(int(parts[i, j]), q[i], 15)
                    keep = ~np.isnan(rd)                      # FlatIndex::search keeps NaN rows only by accident of sort order
                    n = int(oc[i, j])
                    assert n == int(keep.sum()), (metric, simple, i, j, n, len(rr))
                    assert oi[i, j, :n].tolist() == rr[keep].tolist()
                    assert _bits(od[i, j, :n]).tolist() == _bits(rd[keep]).tolist()
        assert int(oc[0].min()) == 15
    finally:
        os.environ.pop("LB200_PQ4_SIMPLE", None)
        ix.close()

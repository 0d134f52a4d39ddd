"""ctypes binding of the CPU oracle (oracle/lance_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Nothing under lance_b200/ imports it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblance_oracle.so")

L2, COSINE, DOT = 0, 1, 2
METRICS = {"l2": L2, "cosine": COSINE, "dot": DOT}


def build(force=False):
    src = os.path.join(_HERE, "lance_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B"], stdout=subprocess.DEVNULL)
    return _SO


class _Index(C.Structure):
    _fields_ = [
        ("metric", C.c_int32), ("dim", C.c_uint32), ("nlist", C.c_uint32), ("M", C.c_uint32),
        ("centroids", C.c_void_p), ("codebook", C.c_void_p), ("part_offsets", C.c_void_p),
        ("codes", C.c_void_p), ("row_ids", C.c_void_p), ("raw_vectors", C.c_void_p),
        ("num_bits", C.c_uint32),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        f = _lib
        f.lo_l2.restype = C.c_float
        f.lo_dot.restype = C.c_float
        f.lo_dot_distance.restype = C.c_float
        f.lo_norm_l2.restype = C.c_float
        f.lo_cosine_fast.restype = C.c_float
        f.lo_pq_distance_one.restype = C.c_float
        f.lo_pq_distance_one_4bit.restype = C.c_float
        f.lo_find_partitions.restype = C.c_uint32
        f.lo_search_in_partition.restype = C.c_uint32
        f.lo_search_one.restype = C.c_uint32
        f.lo_flat_knn_one.restype = C.c_uint32
        f.lo_max_threads.restype = C.c_int
        for name in ("lo_l2_f16", "lo_dot_f16", "lo_norm_l2_f16", "lo_cosine_f16"):
            getattr(f, name).restype = C.c_float
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def l2(x, y):
    x, y = _f32(x), _f32(y)
    return float(lib().lo_l2(_p(x), _p(y), C.c_size_t(x.size)))


def dot(x, y):
    x, y = _f32(x), _f32(y)
    return float(lib().lo_dot(_p(x), _p(y), C.c_size_t(x.size)))


def norm_l2(x):
    x = _f32(x)
    return float(lib().lo_norm_l2(_p(x), C.c_size_t(x.size)))


def normalize(x):
    x = _f32(x)
    out = np.empty_like(x)
    lib().lo_normalize(_p(x), C.c_size_t(x.size), _p(out))
    return out


def distance_batch(metric, q, vectors):
    """DistanceType::arrow_batch_func: one query against [n][d] rows -> [n] f32."""
    q, vectors = _f32(q), _f32(vectors)
    n, d = vectors.shape
    out = np.empty(n, dtype=np.float32)
    lib().lo_distance_batch(C.c_int(METRICS[metric]), _p(q), _p(vectors), C.c_size_t(n),
                            C.c_size_t(d), _p(out))
    return out


def find_partitions(centroids, query, nprobes, metric):
    centroids, query = _f32(centroids), _f32(query)
    nlist, d = centroids.shape
    lim = min(nprobes, nlist)
    ids = np.empty(max(lim, 1), dtype=np.uint32)
    dists = np.empty(max(lim, 1), dtype=np.float32)
    m = METRICS[metric]
    n = lib().lo_find_partitions(_p(centroids), C.c_uint32(nlist), C.c_uint32(d), _p(query),
                                 C.c_uint32(nprobes), C.c_int(DOT if m == DOT else L2), _p(ids),
                                 _p(dists))
    return ids[:n], dists[:n]


def build_distance_table(metric, codebook, query):
    """codebook [M][256][dsub]; returns [M][256]."""
    codebook, query = _f32(codebook), _f32(query)
    M = codebook.shape[0]
    out = np.empty((M, 256), dtype=np.float32)
    m = METRICS[metric]
    lib().lo_build_distance_table(C.c_int(DOT if m == DOT else L2), _p(codebook), C.c_uint32(M),
                                  C.c_uint32(query.size), _p(query), _p(out))
    return out


def compute_pq_distance(table, codes_t):
    """table [M][256], codes_t transposed [M][n] u8 -> [n] f32."""
    table = _f32(table)
    codes_t = np.ascontiguousarray(codes_t, dtype=np.uint8)
    M, n = codes_t.shape
    out = np.empty(n, dtype=np.float32)
    lib().lo_compute_pq_distance(_p(table), C.c_uint32(M), _p(codes_t), C.c_size_t(n), _p(out))
    return out


def pq_distance_one(table, codes_t, i):
    table = _f32(table)
    codes_t = np.ascontiguousarray(codes_t, dtype=np.uint8)
    M, n = codes_t.shape
    return float(lib().lo_pq_distance_one(_p(table), C.c_uint32(M), _p(codes_t), C.c_size_t(n),
                                          C.c_size_t(i)))


def pq_quantize(metric, codebook, vecs):
    codebook, vecs = _f32(codebook), _f32(vecs)
    M = codebook.shape[0]
    n, d = vecs.shape
    out = np.empty((n, M), dtype=np.uint8)
    m = METRICS[metric]
    lib().lo_pq_quantize(C.c_int(DOT if m == DOT else L2), _p(codebook), C.c_uint32(M),
                         C.c_uint32(d), _p(vecs), C.c_size_t(n), _p(out))
    return out


# ---- 4-bit PQ (pq/distance.rs:138-229) -----------------------------------------------------
def build_distance_table4(metric, codebook, query):
    """codebook [M][16][dsub]; returns [M][16]."""
    codebook, query = _f32(codebook), _f32(query)
    M = codebook.shape[0]
    out = np.empty((M, 16), dtype=np.float32)
    m = METRICS[metric]
    lib().lo_build_distance_table4(C.c_int(DOT if m == DOT else L2), _p(codebook), C.c_uint32(M),
                                   C.c_uint32(query.size), _p(query), _p(out))
    return out


def quantize_distance_table(table):
    """quantize_distance_table: (qmin, qmax, u8 table [M][16])."""
    table = _f32(table)
    M = table.shape[0]
    qt = np.empty((M, 16), dtype=np.uint8)
    qmin, qmax = C.c_float(), C.c_float()
    lib().lo_quantize_distance_table(_p(table), C.c_uint32(M), C.byref(qmin), C.byref(qmax), _p(qt))
    return np.float32(qmin.value), np.float32(qmax.value), qt


def compute_pq_distance_4bit(table, codes_t):
    """table [M][16] f32, codes_t transposed packed [M/2][n] u8 -> [n] f32."""
    table = _f32(table)
    codes_t = np.ascontiguousarray(codes_t, dtype=np.uint8)
    M = table.shape[0]
    n = codes_t.shape[1]
    out = np.empty(n, dtype=np.float32)
    lib().lo_compute_pq_distance_4bit(_p(table), C.c_uint32(M), _p(codes_t), C.c_size_t(n), _p(out))
    return out


def pq_distance_one_4bit(table, codes_t, i):
    table = _f32(table)
    codes_t = np.ascontiguousarray(codes_t, dtype=np.uint8)
    return float(lib().lo_pq_distance_one_4bit(_p(table), C.c_uint32(table.shape[0]), _p(codes_t),
                                               C.c_size_t(codes_t.shape[1]), C.c_size_t(i)))


def pq_quantize4(metric, codebook, vecs):
    """transform_impl::<4>: codebook [M][16][dsub] -> packed row-major codes [n][M/2]."""
    codebook, vecs = _f32(codebook), _f32(vecs)
    M = codebook.shape[0]
    n, d = vecs.shape
    out = np.zeros((n, M // 2), dtype=np.uint8)
    m = METRICS[metric]
    lib().lo_pq_quantize4(C.c_int(DOT if m == DOT else L2), _p(codebook), C.c_uint32(M), C.c_uint32(d),
                          _p(vecs), C.c_size_t(n), _p(out))
    return out


class OracleIndex:
    """Mirror of IVFIndex<FlatIndex, ProductQuantizer> (lance/src/index/vector/ivf/v2.rs:76)
    over the exchange format of lance_b200.synth.IvfPqData."""

    def __init__(self, data):
        self.data = data
        self._keep = [
            _f32(data.centroids), _f32(data.codebook),
            np.ascontiguousarray(data.part_offsets, dtype=np.uint64),
            np.ascontiguousarray(data.codes, dtype=np.uint8),
            np.ascontiguousarray(data.row_ids, dtype=np.uint64),
            None if data.raw_vectors is None else _f32(data.raw_vectors),
        ]
        k = self._keep
        self.ix = _Index(METRICS[data.metric], data.dim, data.nlist, data.M, k[0].ctypes.data,
                         k[1].ctypes.data, k[2].ctypes.data, k[3].ctypes.data, k[4].ctypes.data,
                         None if k[5] is None else k[5].ctypes.data, int(getattr(data, "num_bits", 8)))

    def preprocess_query(self, part, key):
        key = _f32(key)
        out = np.empty_like(key)
        lib().lo_preprocess_query(C.byref(self.ix), C.c_uint32(part), _p(key), _p(out))
        return out

    def search_in_partition(self, part, key, kprime, lower=None, upper=None, bitmask=None):
        key = _f32(key)
        od = np.empty(max(kprime, 1), dtype=np.float32)
        oi = np.empty(max(kprime, 1), dtype=np.uint64)
        n = lib().lo_search_in_partition(
            C.byref(self.ix), C.c_uint32(part), _p(key), C.c_uint32(kprime),
            C.c_int(lower is not None), C.c_float(lower or 0.0), C.c_int(upper is not None),
            C.c_float(upper or 0.0), _p(bitmask), _p(od), _p(oi), None)
        return od[:n], oi[:n]

    def search(self, queries, k, nprobes, refine_factor=0, lower=None, upper=None, threads=0):
        """Batched VectorIndex::search (+ scanner refine when refine_factor>0).
        Returns (dist [Q][k'], rowid [Q][k'], count [Q], probes [Q][P])."""
        queries = _f32(queries)
        Q = queries.shape[0]
        kp = k * (refine_factor if refine_factor else 1)
        P = min(nprobes, self.data.nlist)
        od = np.full((Q, max(kp, 1)), np.inf, dtype=np.float32)
        oi = np.full((Q, max(kp, 1)), np.iinfo(np.uint64).max, dtype=np.uint64)
        oc = np.zeros(Q, dtype=np.uint32)
        pr = np.zeros((Q, max(nprobes, 1)), dtype=np.uint32)
        lib().lo_search_batch(
            C.byref(self.ix), _p(queries), C.c_uint32(Q), C.c_uint32(k), C.c_uint32(nprobes),
            C.c_uint32(refine_factor), C.c_int(lower is not None), C.c_float(lower or 0.0),
            C.c_int(upper is not None), C.c_float(upper or 0.0), C.c_int(threads), _p(od), _p(oi),
            _p(oc), _p(pr))
        return od, oi, oc, pr[:, :P]


def flat_knn(metric, queries, vectors, k, validity=None, row_ids=None, threads=0):
    queries, vectors = _f32(queries), _f32(vectors)
    Q, d = queries.shape
    n = vectors.shape[0]
    od = np.full((Q, k), np.inf, dtype=np.float32)
    oi = np.full((Q, k), np.iinfo(np.uint64).max, dtype=np.uint64)
    oc = np.zeros(Q, dtype=np.uint32)
    lib().lo_flat_knn_batch(C.c_int(METRICS[metric]), _p(queries), C.c_uint32(Q), _p(vectors),
                            _p(validity), _p(row_ids), C.c_size_t(n), C.c_size_t(d), C.c_uint32(k),
                            C.c_int(threads), _p(od), _p(oi), _p(oc))
    return od, oi, oc


def compute_membership(metric, centroids, data, threads=0):
    """KMeansAlgoFloat::compute_membership_and_loss (kmeans.rs:208-240): ids [n] u32 (0xffffffff = None), dists [n]."""
    centroids, data = _f32(centroids), _f32(data)
    n, d = data.shape
    ids = np.empty(n, dtype=np.uint32)
    dists = np.empty(n, dtype=np.float32)
    lib().lo_compute_membership(C.c_int(METRICS[metric]), _p(centroids), C.c_uint32(centroids.shape[0]), C.c_uint32(d), _p(data),
                                C.c_size_t(n), C.c_int(threads), _p(ids), _p(dists))
    return ids, dists


def kmeans_fit(metric, data, init_centroids, max_iters=50, tolerance=1e-4, threads=0):
    """KMeans::train_kmeans (kmeans.rs:440-525) from given initial centroids.  Returns (centroids, rounds, loss)."""
    data = _f32(data)
    cents = _f32(init_centroids).copy()
    loss = C.c_double(0)
    lib().lo_kmeans_fit.restype = C.c_uint32
    it = lib().lo_kmeans_fit(C.c_int(METRICS[metric]), _p(data), C.c_size_t(data.shape[0]), C.c_uint32(data.shape[1]),
                             C.c_uint32(cents.shape[0]), C.c_uint32(max_iters), C.c_double(tolerance), C.c_int(threads), _p(cents),
                             C.byref(loss))
    return cents, int(it), float(loss.value)


def max_threads():
    return int(lib().lo_max_threads())


# ---------------------------------------------------------------------------------------------
# The plan around the index: Scanner::knn / ann / flat_knn / knn_combined
# (rust/lance/src/dataset/scanner.rs:1670-1835,2066-2157), restated per query with the C oracle's
# arithmetic.  Test infrastructure like the rest of this package.
# ---------------------------------------------------------------------------------------------
def _topk_by_distance(dist, rowid, k):
    """SortExec(_distance ASC).with_fetch(k); equal distances ordered by row id (the oracle's own rule:
    the reference leaves tie order unpinned)."""
    order = np.lexsort((rowid, dist))[:k]
    return dist[order], rowid[order]


def flat_knn_rows(metric, key, vectors, row_ids, k, lower=None, upper=None):
    """Scanner::flat_knn (scanner.rs:2066-2136) for one key over rows that carry row ids."""
    vectors = _f32(vectors).reshape(-1, len(key))
    row_ids = np.asarray(row_ids, dtype=np.uint64)
    if vectors.shape[0] == 0:
        return np.zeros(0, np.float32), np.zeros(0, np.uint64)
    d = distance_batch(metric, key, vectors)
    keep = ~np.isnan(d)
    if lower is not None:
        keep &= d >= np.float32(lower)
    if upper is not None:
        keep &= d < np.float32(upper)
    return _topk_by_distance(d[keep], row_ids[keep], k)


def nearest_plan(deltas, take, key, k, nprobes, refine_factor=None, lower=None, upper=None,
                 unindexed_ids=None, unindexed_vectors=None, fast_search=False, allowed=None):
    """Scanner::knn for ONE key.  deltas: OracleIndex list (same metric); take(row_ids) -> vectors.
    allowed: optional row-id allow-list (the prefilter's RowIdMask): per probed partition the reference then runs the
    per-row path (flat/index.rs:106-122 -> DistCalculator::distance), and the flat scan is filtered."""
    key = _f32(key)
    metric = deltas[0].data.metric
    kp = k * (refine_factor or 1)
    ds, rs = [], []
    if allowed is not None:
        allowed = np.unique(np.asarray(allowed, dtype=np.uint64))
        if unindexed_ids is not None:
            keep = np.isin(np.asarray(unindexed_ids, dtype=np.uint64), allowed)
            unindexed_ids, unindexed_vectors = np.asarray(unindexed_ids)[keep], _f32(unindexed_vectors)[keep]
    for ix in deltas:          # ANNIvfSubIndexExec fan-out over the delta indices
        if allowed is not None:
            d = ix.data
            k2 = normalize(key) if metric == "cosine" else key
            probes, _ = find_partitions(d.centroids, k2, nprobes, metric)
            for p in probes.tolist():
                o, e = int(d.part_offsets[p]), int(d.part_offsets[p + 1])
                bits = np.zeros(((e - o + 63) // 64) + 1, dtype=np.uint64)
                on = np.nonzero(np.isin(d.row_ids[o:e], allowed))[0]
                np.bitwise_or.at(bits, on >> 6, np.uint64(1) << (on & 63).astype(np.uint64))
                pd, pr = ix.search_in_partition(p, k2, kp, lower, upper, bitmask=bits)
                ds.append(pd)
                rs.append(pr)
            continue
        od, oi, oc, _ = ix.search(key[None, :], kp, nprobes, 0, lower, upper)
        ds.append(od[0, :oc[0]])
        rs.append(oi[0, :oc[0]])
    ad, ar = _topk_by_distance(np.concatenate(ds), np.concatenate(rs), kp)   # ann(): fetch k * refine
    if refine_factor is not None:
        ad, ar = flat_knn_rows(metric, key, take(ar), ar, k, lower, upper)
    if fast_search or unindexed_ids is None or len(unindexed_ids) == 0:
        return ad, ar
    fd, fr = flat_knn_rows(metric, key, unindexed_vectors, unindexed_ids, k, lower, upper)
    lut = {int(r): i for i, r in enumerate(np.asarray(unindexed_ids, dtype=np.uint64).tolist())}
    fv = _f32(unindexed_vectors)[[lut[int(r)] for r in fr.tolist()]] if len(fr) else np.zeros((0, len(key)), np.float32)
    av = _f32(take(ar)).reshape(-1, len(key)) if len(ar) else np.zeros((0, len(key)), np.float32)
    return flat_knn_rows(metric, key, np.concatenate([fv, av]), np.concatenate([fr, ar]), k, lower, upper)


# ---- f16 vectors ---------------------------------------------------------------------------------
def distance_batch_f16(metric, q, vectors):
    """{l2,cosine,dot}_distance_arrow_batch on Float16 data, the reference's deterministic scalar fallback
    (lo_distance_batch_f16).  q [d] f16, vectors [n][d] f16 -> [n] f32."""
    q = np.ascontiguousarray(q, dtype=np.float16)
    vectors = np.ascontiguousarray(vectors, dtype=np.float16)
    n, d = vectors.shape
    out = np.empty(n, dtype=np.float32)
    lib().lo_distance_batch_f16(C.c_int(METRICS[metric]), _p(q), _p(vectors), C.c_size_t(n), C.c_size_t(d), _p(out))
    return out


_REF_SO = os.path.join(_HERE, "_ref", "libf16_ref.so")
_ref = None


def ref_f16():
    """oracle/_ref/libf16_ref.so: the REFERENCE's own f16 kernels (rust/lance-linalg/src/simd/f16.c compiled by
    `make -C oracle ref` with build.rs's flags), or None when it has not been built."""
    global _ref
    if _ref is None and os.path.exists(_REF_SO):
        _ref = C.CDLL(_REF_SO)
        for name in ("norm_l2_f16_avx2", "dot_f16_avx2", "l2_f16_avx2", "cosine_f16_avx2"):
            getattr(_ref, name).restype = C.c_float
    return _ref


def ref_distance_batch_f16(metric, q, vectors):
    """The same batch through the reference's C kernels, called the way l2.rs:150-155 / dot.rs:120-125 /
    cosine.rs:104-108 call them."""
    r = ref_f16()
    q = np.ascontiguousarray(q, dtype=np.float16)
    vectors = np.ascontiguousarray(vectors, dtype=np.float16)
    n, d = vectors.shape
    out = np.empty(n, dtype=np.float32)
    m = METRICS[metric]
    xn = r.norm_l2_f16_avx2(_p(q), C.c_uint32(d)) if m == COSINE else 0.0
    for i in range(n):
        y = vectors[i].ctypes.data_as(C.c_void_p)
        if m == L2:
            out[i] = r.l2_f16_avx2(_p(q), y, C.c_uint32(d))
        elif m == DOT:
            out[i] = 1.0 - r.dot_f16_avx2(_p(q), y, C.c_uint32(d))
        else:
            out[i] = r.cosine_f16_avx2(_p(q), C.c_float(xn), y, C.c_uint32(d))
    return out


# ---- IVF_PQ on Float16 vectors -------------------------------------------------------------------
class _IndexF16(C.Structure):
    _fields_ = [
        ("metric", C.c_int32), ("dim", C.c_uint32), ("nlist", C.c_uint32), ("M", C.c_uint32),
        ("centroids", C.c_void_p), ("codebook", C.c_void_p), ("part_offsets", C.c_void_p),
        ("codes", C.c_void_p), ("row_ids", C.c_void_p), ("raw_vectors", C.c_void_p),
    ]


def _f16(a):
    return np.ascontiguousarray(a, dtype=np.float16)


def normalize_f16(v):
    """kernels.rs:118-121 normalize::<f16> (lo_normalize_f16; parity unpinned, see the C header)."""
    v = _f16(v)
    out = np.empty_like(v)
    lib().lo_normalize_f16(_p(v), C.c_size_t(v.shape[-1]), _p(out))
    return out


class OracleIndexF16:
    """IVFIndex<FlatIndex, ProductQuantizer> over Float16 centroids / codebook / raw vectors (lo_search_batch_f16).
    data: IvfPqData whose centroids, codebook and raw_vectors hold values that are exactly representable in f16
    (they are converted with astype(float16))."""

    def __init__(self, data):
        self.data = data
        self._keep = [
            _f16(data.centroids), _f16(data.codebook), np.ascontiguousarray(data.part_offsets, dtype=np.uint64),
            np.ascontiguousarray(data.codes, dtype=np.uint8), np.ascontiguousarray(data.row_ids, dtype=np.uint64),
            None if data.raw_vectors is None else _f16(data.raw_vectors),
        ]
        k = self._keep
        self.ix = _IndexF16(METRICS[data.metric], data.dim, data.nlist, data.M, k[0].ctypes.data, k[1].ctypes.data,
                            k[2].ctypes.data, k[3].ctypes.data, k[4].ctypes.data, None if k[5] is None else k[5].ctypes.data)

    def find_partitions(self, key, nprobes):
        key = _f16(key)
        P = min(nprobes, self.data.nlist)
        ids = np.empty(max(P, 1), dtype=np.uint32)
        ds = np.empty(max(P, 1), dtype=np.float32)
        m = METRICS[self.data.metric]
        n = lib().lo_find_partitions_f16(self._keep[0].ctypes.data_as(C.c_void_p), C.c_uint32(self.data.nlist), C.c_uint32(self.data.dim),
                                         _p(key), C.c_uint32(nprobes), C.c_int(DOT if m == DOT else L2), _p(ids), _p(ds))
        return ids[:n], ds[:n]

    def search_in_partition(self, part, key, kprime, lower=None, upper=None):
        key = _f16(key)
        od = np.empty(max(kprime, 1), dtype=np.float32)
        oi = np.empty(max(kprime, 1), dtype=np.uint64)
        n = lib().lo_search_in_partition_f16(C.byref(self.ix), C.c_uint32(part), _p(key), C.c_uint32(kprime),
                                             C.c_int(lower is not None), C.c_float(lower or 0.0), C.c_int(upper is not None),
                                             C.c_float(upper or 0.0), None, _p(od), _p(oi), None)
        return od[:n], oi[:n]

    def search(self, queries, k, nprobes, refine_factor=0, lower=None, upper=None, threads=0):
        queries = _f16(queries)
        Q = queries.shape[0]
        kp = k * (refine_factor if refine_factor else 1)
        P = min(nprobes, self.data.nlist)
        od = np.full((Q, max(kp, 1)), np.inf, dtype=np.float32)
        oi = np.full((Q, max(kp, 1)), np.iinfo(np.uint64).max, dtype=np.uint64)
        oc = np.zeros(Q, dtype=np.uint32)
        pr = np.zeros((Q, max(nprobes, 1)), dtype=np.uint32)
        lib().lo_search_batch_f16(
            C.byref(self.ix), _p(queries), C.c_uint32(Q), C.c_uint32(k), C.c_uint32(nprobes), C.c_uint32(refine_factor),
            C.c_int(lower is not None), C.c_float(lower or 0.0), C.c_int(upper is not None), C.c_float(upper or 0.0),
            C.c_int(threads), _p(od), _p(oi), _p(oc), _p(pr))
        return od, oi, oc, pr[:, :P]


# ---- IVF_FLAT (IVFIndex<FlatIndex, FlatQuantizer>) -------------------------------------------------
def ivf_flat_search(data, query, k, nprobes, refine_factor=None, lower=None, upper=None):
    """VectorIndex::search on an IVF_FLAT index for ONE query (ivf/v2.rs:368-401), then the scanner's refine when
    refine_factor is given.  data: IvfPqData with raw_vectors [N][dim] in partition order (codes unused).
    Per probed partition FlatIndex::search (flat/index.rs:76-137) with FlatDistanceCal (flat/storage.rs:315-375):
    distance_type.func()(key, row) on the raw rows, key normalised for cosine, no residual."""
    query = _f32(query)
    metric = data.metric
    key = normalize(query) if metric == "cosine" else query
    kp = k * (refine_factor or 1)
    probes, _ = find_partitions(data.centroids, key, nprobes, metric)
    ds, rs = [], []
    for p in probes.tolist():
        o, e = int(data.part_offsets[p]), int(data.part_offsets[p + 1])
        d, r = flat_knn_rows(metric, key, data.raw_vectors[o:e], data.row_ids[o:e], kp, lower, upper)
        ds.append(d)
        rs.append(r)
    ad, ar = _topk_by_distance(np.concatenate(ds), np.concatenate(rs), kp) if ds else (np.zeros(0, np.float32), np.zeros(0, np.uint64))
    if refine_factor is not None:
        lut = {int(r): i for i, r in enumerate(np.asarray(data.row_ids, dtype=np.uint64).tolist())}
        vec = _f32(data.raw_vectors)[[lut[int(r)] for r in ar.tolist()]] if len(ar) else np.zeros((0, len(query)), np.float32)
        ad, ar = flat_knn_rows(metric, query, vec, ar, k, lower, upper)
    return ad, ar

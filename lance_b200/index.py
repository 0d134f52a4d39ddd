"""Host-side mirror of the reference's operator surface for the IVF_PQ query path.

Names, argument meaning and error behaviour follow `trait VectorIndex`
(rust/lance-index/src/vector.rs:117-217), `Query` (vector.rs:59-91), `DistanceType`
(rust/lance-linalg/src/distance.rs:36-43) and `compute_distance`
(rust/lance-index/src/vector/flat.rs:28-86); every method is a thin call into the C ABI
(include/lance_b200.h) and is batched over queries.
"""
import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _lib
from ._lib import Stats, check


class DistanceType:
    L2 = "l2"
    Cosine = "cosine"
    Dot = "dot"
    _CODE = {"l2": 0, "cosine": 1, "dot": 2}

    @staticmethod
    def code(m):
        try:
            return DistanceType._CODE[str(m).lower()]
        except KeyError:
            # DistanceType::try_from (distance.rs:87-103)
            raise ValueError(f"Metric type '{m}' is not supported") from None


@dataclass
class Query:
    """vector.rs:59-91.  `key` may be one vector [d] or a batch [Q][d]."""
    key: np.ndarray
    k: int = 10
    nprobes: int = 1
    refine_factor: Optional[int] = None
    lower_bound: Optional[float] = None
    upper_bound: Optional[float] = None
    column: str = "vector"
    metric_type: str = "l2"
    use_index: bool = True


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fptr(v):
    return None if v is None else C.byref(C.c_float(float(v)))


class IvfPqIndex:
    """Device-resident IVFIndex<FlatIndex, ProductQuantizer> (lance/src/index/vector/ivf/v2.rs:76)."""

    def __init__(self, data, device_id=0):
        L = _lib.lib()
        self._L = L
        self.metric = data.metric
        self.dim, self.nlist, self.M = int(data.dim), int(data.nlist), int(data.M)
        self.part_offsets = np.ascontiguousarray(data.part_offsets, dtype=np.uint64)
        self.has_raw = data.raw_vectors is not None
        self.device_id = device_id
        self.row_ids = np.ascontiguousarray(data.row_ids, dtype=np.uint64)   # host copy, index order (prefilter masks)
        h = C.c_void_p()
        cents, cb = _f32(data.centroids), _f32(data.codebook)
        codes = np.ascontiguousarray(data.codes, dtype=np.uint8)
        rids = np.ascontiguousarray(data.row_ids, dtype=np.uint64)
        raw = None if data.raw_vectors is None else _f32(data.raw_vectors)
        check(L.lb200_index_create(
            C.c_int32(device_id), C.c_int32(DistanceType.code(data.metric)), C.c_uint32(self.dim),
            C.c_uint32(self.nlist), _p(cents), C.c_uint32(self.M), C.c_uint32(int(getattr(data, "num_bits", 8))), _p(cb),
            _p(self.part_offsets), _p(codes), _p(rids), _p(raw), C.byref(h)))
        self._h = h

    @classmethod
    def _adopt(cls, handle, meta, device_id, has_raw):
        """Wrap a handle the library created itself (lb200_index_open)."""
        self = cls.__new__(cls)
        self._L = _lib.lib()
        self.metric, self.dim, self.nlist, self.M = meta.metric, int(meta.dim), int(meta.nlist), int(meta.M)
        self.part_offsets = np.ascontiguousarray(meta.part_offsets, dtype=np.uint64)
        self.has_raw, self.device_id, self._h = has_raw, device_id, handle
        self.row_ids = np.ascontiguousarray(meta.row_ids, dtype=np.uint64)
        return self

    def close(self):
        if getattr(self, "_h", None):
            self._L.lb200_index_destroy(self._h)
            self._h = None

    __del__ = close

    # -- VectorIndex surface ----------------------------------------------------------------
    def metric_type(self):
        return self.metric

    def find_partitions(self, key, nprobes, return_distances=False):
        """VectorIndex::find_partitions (vector.rs:146), batched.  Key used as given."""
        key = np.atleast_2d(_f32(key))
        self._check_dim(key)
        Q = key.shape[0]
        P = min(int(nprobes), self.nlist)
        ids = np.zeros((Q, P), dtype=np.uint32)
        d = np.zeros((Q, P), dtype=np.float32) if return_distances else None
        check(self._L.lb200_find_partitions(self._h, _p(key), C.c_uint32(Q), C.c_uint32(nprobes), _p(ids), _p(d)))
        return (ids, d) if return_distances else ids

    def _mask_table(self, prefilter, keep):
        """{partition id: bool array over the partition's rows} -> the [nlist] pointer table of the C ABI."""
        ptrs = (C.c_void_p * self.nlist)()
        for p, mask in prefilter.items():
            n = int(self.part_offsets[p + 1] - self.part_offsets[p])
            bits = np.zeros(((n + 63) // 64) * 8 + 8, dtype=np.uint8)
            packed = np.packbits(np.asarray(mask, dtype=bool)[:n], bitorder="little")
            bits[:packed.size] = packed
            keep.append(bits)
            ptrs[p] = bits.ctypes.data
        return ptrs

    def prefilter(self, masks):
        """Upload a prefilter once (lb200_prefilter_create): the Arc<dyn PreFilter> every search_in_partition
        call of a query shares (knn.rs:621).  masks: {partition id: bool array over that partition's rows}."""
        return Prefilter(self, masks)

    def search_in_partition(self, part_ids, key, k, lower_bound=None, upper_bound=None, prefilter=None):
        """VectorIndex::search_in_partition (vector.rs:165) for [Q][P] explicit partitions.
        prefilter: optional dict {partition id: bool/0-1 array over that partition's rows}, or a
        device-resident Prefilter made by self.prefilter(...)."""
        key = np.atleast_2d(_f32(key))
        self._check_dim(key)
        Q = key.shape[0]
        part_ids = np.ascontiguousarray(np.asarray(part_ids, dtype=np.uint32).reshape(Q, -1))
        P = part_ids.shape[1]
        od = np.full((Q, P, max(k, 1)), np.inf, dtype=np.float32)
        oi = np.full((Q, P, max(k, 1)), np.iinfo(np.uint64).max, dtype=np.uint64)
        oc = np.zeros((Q, P), dtype=np.uint32)
        bm_arr, keep = None, []
        if isinstance(prefilter, Prefilter):
            check(self._L.lb200_search_partitions_filtered(
                self._h, _p(key), C.c_uint32(Q), _p(part_ids), C.c_uint32(P), C.c_uint32(k), _fptr(lower_bound),
                _fptr(upper_bound), prefilter._h, _p(od), _p(oi), _p(oc)))
            return od, oi, oc
        if prefilter is not None:
            bm_arr = self._mask_table(prefilter, keep)
        check(self._L.lb200_search_partitions(
            self._h, _p(key), C.c_uint32(Q), _p(part_ids), C.c_uint32(P), C.c_uint32(k), _fptr(lower_bound),
            _fptr(upper_bound), bm_arr, _p(od), _p(oi), _p(oc)))
        return od, oi, oc

    def search(self, query: Query, out=None):
        """VectorIndex::search (vector.rs:133) + the scanner's refine when refine_factor is set.
        Returns (distances [Q][k_out], row ids [Q][k_out], counts [Q]).  `out` = caller-owned
        (f32 [Q][k_out], u64 [Q][k_out], u32 [Q]) arrays, e.g. views of pinned host memory."""
        key = np.atleast_2d(_f32(query.key))
        self._check_dim(key)
        if query.k <= 0:
            raise ValueError("k must be positive")  # scanner.rs:642 nearest() validation
        Q = key.shape[0]
        rf = int(query.refine_factor or 0)
        kout = query.k if (rf and self.has_raw) else query.k * max(rf, 1)
        if out is not None:
            od, oi, oc = out
            assert od.shape == (Q, kout) and od.dtype == np.float32 and od.flags.c_contiguous
            assert oi.shape == (Q, kout) and oi.dtype == np.uint64 and oi.flags.c_contiguous
            assert oc.shape == (Q,) and oc.dtype == np.uint32
        else:
            od = np.full((Q, kout), np.inf, dtype=np.float32)
            oi = np.full((Q, kout), np.iinfo(np.uint64).max, dtype=np.uint64)
            oc = np.zeros(Q, dtype=np.uint32)
        check(self._L.lb200_search(
            self._h, _p(key), C.c_uint32(Q), C.c_uint32(query.nprobes), C.c_uint32(query.k), C.c_uint32(rf),
            _fptr(query.lower_bound), _fptr(query.upper_bound), _p(od), _p(oi), _p(oc)))
        return od, oi, oc

    # -- device-pointer entry (benchmark / multi-GPU plumbing) -------------------------------
    def search_device(self, d_queries_ptr, Q, nprobes, k, refine_factor, d_out_dist, d_out_rowid, d_out_count,
                      d_out_exact=None, stream=None, lower=None, upper=None):
        check(self._L.lb200_search_device(
            self._h, C.c_void_p(d_queries_ptr), C.c_uint32(Q), C.c_uint32(nprobes), C.c_uint32(k),
            C.c_uint32(refine_factor), _fptr(lower), _fptr(upper), C.c_void_p(d_out_dist), C.c_void_p(d_out_rowid),
            C.c_void_p(d_out_count), C.c_void_p(d_out_exact) if d_out_exact else None,
            C.c_void_p(stream) if stream else None))

    def find_partitions_device(self, d_queries_ptr, Q, nprobes, d_out_part_ids, normalize=True, stream=None):
        check(self._L.lb200_find_partitions_device(
            self._h, C.c_void_p(d_queries_ptr), C.c_uint32(Q), C.c_uint32(nprobes), C.c_int32(1 if normalize else 0),
            C.c_void_p(d_out_part_ids), C.c_void_p(stream) if stream else None))

    def search_probed_device(self, d_queries_ptr, Q, d_part_ids, P, k, refine_factor, d_out_dist, d_out_rowid,
                             d_out_count, d_out_exact=None, stream=None, lower=None, upper=None):
        check(self._L.lb200_search_probed_device(
            self._h, C.c_void_p(d_queries_ptr), C.c_uint32(Q), C.c_void_p(d_part_ids), C.c_uint32(P), C.c_uint32(k),
            C.c_uint32(refine_factor), _fptr(lower), _fptr(upper), C.c_void_p(d_out_dist), C.c_void_p(d_out_rowid),
            C.c_void_p(d_out_count), C.c_void_p(d_out_exact) if d_out_exact else None,
            C.c_void_p(stream) if stream else None))

    def peer_thresholds_export(self, max_queries):
        """Allocate this shard's table of shared pruning bounds; returns (64-byte IPC handle, device address)."""
        h = (C.c_ubyte * 64)()
        ptr = C.c_uint64(0)
        check(self._L.lb200_peer_thresholds_export(self._h, C.c_uint32(max_queries), h, C.byref(ptr)))
        return bytes(h), int(ptr.value)

    def peer_thresholds_attach(self, world, rank, handles=None, ptrs=None):
        """handles: list of `world` 64-byte IPC handles (peers in other processes), or ptrs: list of
        `world` device addresses (peers in this process).  See include/lance_b200.h for the contract."""
        hb = pb = None
        if ptrs is not None:
            pb = (C.c_uint64 * world)(*[int(p) for p in ptrs])
        else:
            raw = b"".join(handles)
            assert len(raw) == 64 * world
            hb = (C.c_ubyte * len(raw)).from_buffer_copy(raw)
        check(self._L.lb200_peer_thresholds_attach(self._h, C.c_uint32(world), C.c_uint32(rank), hb, pb))

    def peer_thresholds_detach(self):
        check(self._L.lb200_peer_thresholds_detach(self._h))

    # -- multi-GPU: partition shards that exchange through peer memory (lb200_shard_*) ------------
    def shard_export(self, max_queries, max_nprobes, max_kprime):
        """Allocate this shard's exchange block; returns (128-byte blob for peers in other processes,
        (block address, bound-table address) for peers in this process)."""
        blob = (C.c_ubyte * 128)()
        ptrs = (C.c_uint64 * 2)()
        check(self._L.lb200_shard_export(self._h, C.c_uint32(max_queries), C.c_uint32(max_nprobes), C.c_uint32(max_kprime),
                                         blob, ptrs))
        return bytes(blob), (int(ptrs[0]), int(ptrs[1]))

    def shard_attach(self, world, rank, blobs=None, ptrs=None, root=0):
        """blobs: `world` 128-byte blobs (one process per GPU), or ptrs: `world` (block, table) address pairs."""
        bb = pb = None
        if ptrs is not None:
            flat = [int(x) for pair in ptrs for x in pair]
            pb = (C.c_uint64 * (2 * world))(*flat)
        else:
            raw = b"".join(blobs)
            assert len(raw) == 128 * world
            bb = (C.c_ubyte * len(raw)).from_buffer_copy(raw)
        check(self._L.lb200_shard_attach(self._h, C.c_uint32(world), C.c_uint32(rank), C.c_uint32(root), bb, pb))

    def shard_detach(self):
        check(self._L.lb200_shard_detach(self._h))

    def shard_check(self):
        check(self._L.lb200_shard_check(self._h))

    def shard_search_device(self, d_queries_ptr, Q, nprobes, k, refine_factor, d_out_dist, d_out_rowid, d_out_count,
                            stream=None, lower=None, upper=None):
        """One batch on this shard, asynchronous on `stream`; every rank calls it, the root gets the result."""
        check(self._L.lb200_shard_search_device(
            self._h, C.c_void_p(d_queries_ptr), C.c_uint32(Q), C.c_uint32(nprobes), C.c_uint32(k),
            C.c_uint32(refine_factor), _fptr(lower), _fptr(upper), C.c_void_p(d_out_dist), C.c_void_p(d_out_rowid),
            C.c_void_p(d_out_count), C.c_void_p(stream) if stream else None))

    def shard_search(self, query: Query, out=None):
        """Host-pointer form (blocking): every rank passes the full batch; returns (dist, ids, counts) -- meaningful
        on the root rank only."""
        key = np.atleast_2d(_f32(query.key))
        self._check_dim(key)
        Q, rf = key.shape[0], int(query.refine_factor or 0)
        if out is not None:
            od, oi, oc = out
        else:
            od = np.full((Q, query.k), np.inf, dtype=np.float32)
            oi = np.full((Q, query.k), np.iinfo(np.uint64).max, dtype=np.uint64)
            oc = np.zeros(Q, dtype=np.uint32)
        check(self._L.lb200_shard_search(
            self._h, _p(key), C.c_uint32(Q), C.c_uint32(query.nprobes), C.c_uint32(query.k), C.c_uint32(rf),
            _fptr(query.lower_bound), _fptr(query.upper_bound), _p(od), _p(oi), _p(oc)))
        return od, oi, oc

    def set_profiling(self, on=True):
        check(self._L.lb200_set_profiling(self._h, C.c_int32(1 if on else 0)))

    def set_scan_variant(self, v):
        check(self._L.lb200_set_scan_variant(self._h, C.c_int32(v)))

    def stats(self):
        s = Stats()
        check(self._L.lb200_get_stats(self._h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in Stats._fields_}

    def _check_dim(self, key):
        if key.shape[1] != self.dim:
            # scanner.rs:642-700: "query dim({}) doesn't match the column {} vector dim({})"
            raise ValueError(f"query dim({key.shape[1]}) doesn't match the index vector dim({self.dim})")


class IvfPqGroup:
    """A whole index sharded over several GPUs of ONE process (lb200_group_*): search() == IvfPqIndex.search()."""

    def __init__(self, data, device_ids, max_queries, max_nprobes, max_kprime, part_owner=None):
        L = _lib.lib()
        self._L, self.dim, self.has_raw = L, int(data.dim), data.raw_vectors is not None
        G = len(device_ids)
        devs = (C.c_int32 * G)(*[int(d) for d in device_ids])
        po = None if part_owner is None else np.ascontiguousarray(part_owner, dtype=np.uint32)
        cents, cb = _f32(data.centroids), _f32(data.codebook)
        codes = np.ascontiguousarray(data.codes, dtype=np.uint8)
        rids = np.ascontiguousarray(data.row_ids, dtype=np.uint64)
        offs = np.ascontiguousarray(data.part_offsets, dtype=np.uint64)
        raw = None if data.raw_vectors is None else _f32(data.raw_vectors)
        h = C.c_void_p()
        check(L.lb200_group_create(devs, C.c_uint32(G), C.c_int32(DistanceType.code(data.metric)), C.c_uint32(self.dim),
                                   C.c_uint32(int(data.nlist)), _p(cents), C.c_uint32(int(data.M)),
                                   C.c_uint32(int(getattr(data, "num_bits", 8))), _p(cb),
                                   _p(offs), _p(codes), _p(rids), _p(raw), _p(po), C.c_uint32(max_queries),
                                   C.c_uint32(max_nprobes), C.c_uint32(max_kprime), C.byref(h)))
        self._h = h

    def search(self, query: Query):
        key = np.atleast_2d(_f32(query.key))
        if key.shape[1] != self.dim:
            raise ValueError(f"query dim({key.shape[1]}) doesn't match the index vector dim({self.dim})")
        Q, rf = key.shape[0], int(query.refine_factor or 0)
        od = np.full((Q, query.k), np.inf, dtype=np.float32)
        oi = np.full((Q, query.k), np.iinfo(np.uint64).max, dtype=np.uint64)
        oc = np.zeros(Q, dtype=np.uint32)
        check(self._L.lb200_group_search(self._h, _p(key), C.c_uint32(Q), C.c_uint32(query.nprobes), C.c_uint32(query.k),
                                         C.c_uint32(rf), _fptr(query.lower_bound), _fptr(query.upper_bound), _p(od), _p(oi),
                                         _p(oc)))
        return od, oi, oc

    def close(self):
        if getattr(self, "_h", None):
            self._L.lb200_group_destroy(self._h)
            self._h = None

    __del__ = close


class IvfPqIndexF16:
    """IVF_PQ over a Float16 column (lb200_index_create_f16 / lb200_search_f16): centroids, codebook, raw vectors and keys
    are float16 arrays, as the reference holds them when the indexed column is Float16."""

    def __init__(self, data, device_id=0):
        L = _lib.lib()
        self._L = L
        self.metric = data.metric
        self.dim, self.nlist, self.M = int(data.dim), int(data.nlist), int(data.M)
        self.has_raw = data.raw_vectors is not None
        h = C.c_void_p()
        f16 = lambda a: np.ascontiguousarray(a, dtype=np.float16)
        cents, cb = f16(data.centroids), f16(data.codebook)
        offs = np.ascontiguousarray(data.part_offsets, dtype=np.uint64)
        codes = np.ascontiguousarray(data.codes, dtype=np.uint8)
        rids = np.ascontiguousarray(data.row_ids, dtype=np.uint64)
        raw = None if data.raw_vectors is None else f16(data.raw_vectors)
        check(L.lb200_index_create_f16(
            C.c_int32(device_id), C.c_int32(DistanceType.code(data.metric)), C.c_uint32(self.dim), C.c_uint32(self.nlist), _p(cents),
            C.c_uint32(self.M), _p(cb), _p(offs), _p(codes), _p(rids), _p(raw), C.byref(h)))
        self._h = h

    def search(self, query: "Query"):
        """VectorIndex::search for a batch of float16 keys; returns (dist [Q][k'], row ids [Q][k'], count [Q])."""
        q = np.ascontiguousarray(query.key, dtype=np.float16)
        if q.ndim == 1:
            q = q[None, :]
        Q = q.shape[0]
        rf = int(query.refine_factor or 0)
        stride = query.k if (rf and self.has_raw) else query.k * max(rf, 1)
        od = np.empty((Q, stride), dtype=np.float32)
        oi = np.empty((Q, stride), dtype=np.uint64)
        oc = np.empty(Q, dtype=np.uint32)
        check(self._L.lb200_search_f16(self._h, _p(q), C.c_uint32(Q), C.c_uint32(query.nprobes), C.c_uint32(query.k), C.c_uint32(rf),
                                       _fptr(query.lower_bound), _fptr(query.upper_bound), _p(od), _p(oi), _p(oc)))
        return od, oi, oc

    def find_partitions(self, keys, nprobes, return_distances=False):
        """VectorIndex::find_partitions for float16 keys (used as given)."""
        q = np.atleast_2d(np.ascontiguousarray(keys, dtype=np.float16))
        P = min(int(nprobes), self.nlist)
        ids = np.empty((q.shape[0], P), dtype=np.uint32)
        ds = np.empty((q.shape[0], P), dtype=np.float32) if return_distances else None
        check(self._L.lb200_find_partitions_f16(self._h, _p(q), C.c_uint32(q.shape[0]), C.c_uint32(nprobes), _p(ids), _p(ds)))
        return (ids, ds) if return_distances else ids

    def search_in_partition(self, part_ids, keys, k, lower_bound=None, upper_bound=None):
        """VectorIndex::search_in_partition for explicit (query, partition) pairs: (dist, row ids) [Q][P][k], count [Q][P]."""
        q = np.atleast_2d(np.ascontiguousarray(keys, dtype=np.float16))
        parts = np.ascontiguousarray(part_ids, dtype=np.uint32).reshape(q.shape[0], -1)
        Q, P = parts.shape
        od = np.full((Q, P, k), np.inf, dtype=np.float32)
        oi = np.full((Q, P, k), np.iinfo(np.uint64).max, dtype=np.uint64)
        oc = np.zeros((Q, P), dtype=np.uint32)
        check(self._L.lb200_search_partitions_f16(self._h, _p(q), C.c_uint32(Q), _p(parts), C.c_uint32(P), C.c_uint32(k),
                                                  _fptr(lower_bound), _fptr(upper_bound), _p(od), _p(oi), _p(oc)))
        return od, oi, oc

    def close(self):
        if getattr(self, "_h", None):
            self._L.lb200_index_destroy(self._h)
            self._h = None

    __del__ = close


class Prefilter:
    """Device-resident per-partition row bitmasks of one index handle (lb200_prefilter_*)."""

    def __init__(self, index, masks):
        self._L, self._index, keep = index._L, index, []
        h = C.c_void_p()
        check(self._L.lb200_prefilter_create(index._h, index._mask_table(masks, keep), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.lb200_prefilter_destroy(self._h)
            self._h = None

    __del__ = close


def merge_topk_device(device_id, d_dist, d_rowid, d_exact, d_count, G, Q, stride, kprime, k, d_out_dist,
                      d_out_rowid, d_out_count, lower=None, upper=None, stream=None):
    check(_lib.lib().lb200_merge_topk_device(
        C.c_int32(device_id), C.c_void_p(d_dist), C.c_void_p(d_rowid), C.c_void_p(d_exact) if d_exact else None,
        C.c_void_p(d_count), C.c_uint32(G), C.c_uint32(Q), C.c_uint32(stride), C.c_uint32(kprime), C.c_uint32(k),
        _fptr(lower), _fptr(upper), C.c_void_p(d_out_dist), C.c_void_p(d_out_rowid), C.c_void_p(d_out_count),
        C.c_void_p(stream) if stream else None))


def compute_distance(key, metric, vectors, validity=None, device_id=0):
    """compute_distance (flat.rs:28-86): `_distance` column for a batch; NaN where null.
    key [d] or [Q][d]; vectors [n][d]; validity: Arrow bitmap bytes (bit set = valid) or None."""
    key2 = np.atleast_2d(_f32(key))
    vectors = _f32(vectors)
    if vectors.ndim != 2 or key2.shape[1] != vectors.shape[1]:
        raise ValueError("dimension mismatch between key and vectors")
    Q, n = key2.shape[0], vectors.shape[0]
    out = np.zeros((Q, n), dtype=np.float32)
    v = None if validity is None else np.ascontiguousarray(validity, dtype=np.uint8)
    check(_lib.lib().lb200_flat_distance(C.c_int32(device_id), C.c_int32(DistanceType.code(metric)),
                                         C.c_uint32(vectors.shape[1]), _p(key2), C.c_uint32(Q), _p(vectors), _p(v),
                                         C.c_uint64(n), _p(out)))
    return out[0] if np.ndim(key) == 1 else out


def flat_knn(key, metric, vectors, k, validity=None, lower_bound=None, upper_bound=None, device_id=0):
    """Scanner::flat_knn (scanner.rs:2066-2136): KNNVectorDistanceExec -> bounds -> SortExec(fetch=k)
    -> NOT NULL.  Returns (distances [Q][k], row positions [Q][k], counts [Q])."""
    key2 = np.atleast_2d(_f32(key))
    vectors = _f32(vectors)
    if vectors.ndim != 2 or key2.shape[1] != vectors.shape[1]:
        raise ValueError("dimension mismatch between key and vectors")
    if k <= 0:
        raise ValueError("k must be positive")
    Q, n = key2.shape[0], vectors.shape[0]
    od = np.full((Q, k), np.inf, dtype=np.float32)
    oi = np.full((Q, k), np.iinfo(np.uint64).max, dtype=np.uint64)
    oc = np.zeros(Q, dtype=np.uint32)
    v = None if validity is None else np.ascontiguousarray(validity, dtype=np.uint8)
    check(_lib.lib().lb200_flat_topk(C.c_int32(device_id), C.c_int32(DistanceType.code(metric)),
                                     C.c_uint32(vectors.shape[1]), _p(key2), C.c_uint32(Q), _p(vectors), _p(v),
                                     C.c_uint64(n), C.c_uint32(k), _fptr(lower_bound), _fptr(upper_bound), _p(od),
                                     _p(oi), _p(oc)))
    return od, oi, oc


class VectorColumn:
    """A vector column resident on the device (lb200_column_*): flat_knn(...) without re-uploading the rows."""

    def __init__(self, vectors, validity=None, device_id=0):
        self._L = _lib.lib()
        v = _f32(vectors)
        if v.ndim != 2:
            raise ValueError("vectors must be [n][dim]")
        self.n, self.dim = v.shape
        m = None if validity is None else np.ascontiguousarray(validity, dtype=np.uint8)
        h = C.c_void_p()
        check(self._L.lb200_column_create(C.c_int32(device_id), C.c_uint32(self.dim), _p(v), _p(m), C.c_uint64(self.n), C.byref(h)))
        self._h = h

    def flat_knn(self, key, metric, k, lower_bound=None, upper_bound=None):
        key2 = np.atleast_2d(_f32(key))
        if key2.shape[1] != self.dim:
            raise ValueError("dimension mismatch between key and vectors")
        if k <= 0:
            raise ValueError("k must be positive")
        Q = key2.shape[0]
        od = np.full((Q, k), np.inf, dtype=np.float32)
        oi = np.full((Q, k), np.iinfo(np.uint64).max, dtype=np.uint64)
        oc = np.zeros(Q, dtype=np.uint32)
        check(self._L.lb200_column_flat_topk(self._h, C.c_int32(DistanceType.code(metric)), _p(key2), C.c_uint32(Q), C.c_uint32(k),
                                             _fptr(lower_bound), _fptr(upper_bound), _p(od), _p(oi), _p(oc)))
        return od, oi, oc

    def last_call(self):
        tc, cand, fb = C.c_uint32(0), C.c_uint64(0), C.c_uint32(0)
        check(self._L.lb200_column_last_call(self._h, C.byref(tc), C.byref(cand), C.byref(fb)))
        return {"used_tc": bool(tc.value), "candidates": int(cand.value), "fell_back": bool(fb.value)}

    def close(self):
        if getattr(self, "_h", None):
            self._L.lb200_column_destroy(self._h)
            self._h = None

    __del__ = close


def merge_topk(dist, rowid, count, k, exact=None, kprime=None, lower_bound=None, upper_bound=None, device_id=0):
    """Union of G result lists per query -> the k smallest by (distance,row id): the SortExec(fetch) above the
    fan-out over delta indices (scanner.rs:2139-2157).  dist/rowid [G][Q][stride], count [G][Q]."""
    dist = np.ascontiguousarray(dist, dtype=np.float32)
    rowid = np.ascontiguousarray(rowid, dtype=np.uint64)
    count = np.ascontiguousarray(count, dtype=np.uint32)
    if dist.ndim != 3 or rowid.shape != dist.shape or count.shape != dist.shape[:2]:
        raise ValueError("merge_topk: expected dist/rowid [G][Q][stride] and count [G][Q]")
    if k <= 0:
        raise ValueError("k must be positive")
    G, Q, stride = dist.shape
    ex = None if exact is None else np.ascontiguousarray(exact, dtype=np.float32)
    od = np.full((Q, k), np.inf, dtype=np.float32)
    oi = np.full((Q, k), np.iinfo(np.uint64).max, dtype=np.uint64)
    oc = np.zeros(Q, dtype=np.uint32)
    check(_lib.lib().lb200_merge_topk(
        C.c_int32(device_id), _p(dist), _p(rowid), _p(ex), _p(count), C.c_uint32(G), C.c_uint32(Q), C.c_uint32(stride),
        C.c_uint32(kprime or k), C.c_uint32(k), _fptr(lower_bound), _fptr(upper_bound), _p(od), _p(oi), _p(oc)))
    return od, oi, oc


def rerank(key, metric, cand_vectors, cand_rowids, cand_count, k, lower_bound=None, upper_bound=None, device_id=0):
    """flat_knn (scanner.rs:2066-2136) over per-query candidates fetched by the caller (TakeExec): exact distances
    with the original key, [lower,upper), top k by (distance,row id).  key [Q][d], cand_vectors [Q][C][d],
    cand_rowids [Q][C], cand_count [Q]."""
    key2 = np.atleast_2d(_f32(key))
    cv = _f32(cand_vectors)
    cr = np.ascontiguousarray(cand_rowids, dtype=np.uint64)
    cc = np.ascontiguousarray(cand_count, dtype=np.uint32)
    Q = key2.shape[0]
    if cv.ndim != 3 or cv.shape[0] != Q or cv.shape[2] != key2.shape[1] or cr.shape != cv.shape[:2] or cc.shape != (Q,):
        raise ValueError("rerank: expected cand_vectors [Q][C][d], cand_rowids [Q][C], cand_count [Q]")
    if k <= 0:
        raise ValueError("k must be positive")
    od = np.full((Q, k), np.inf, dtype=np.float32)
    oi = np.full((Q, k), np.iinfo(np.uint64).max, dtype=np.uint64)
    oc = np.zeros(Q, dtype=np.uint32)
    check(_lib.lib().lb200_rerank(
        C.c_int32(device_id), C.c_int32(DistanceType.code(metric)), C.c_uint32(key2.shape[1]), _p(key2), C.c_uint32(Q),
        _p(cv), _p(cr), _p(cc), C.c_uint32(cv.shape[1]), C.c_uint32(k), _fptr(lower_bound), _fptr(upper_bound),
        _p(od), _p(oi), _p(oc)))
    return od, oi, oc


def _f16(a):
    return np.ascontiguousarray(a, dtype=np.float16)


def compute_distance_f16(key, metric, vectors, validity=None, device_id=0):
    """compute_distance (flat.rs:28-86) on a Float16 column: key [d] or [Q][d] and vectors [n][d] as float16;
    returns f32 distances, NaN where null."""
    key2 = np.atleast_2d(_f16(key))
    vectors = _f16(vectors)
    if vectors.ndim != 2 or key2.shape[1] != vectors.shape[1]:
        raise ValueError("dimension mismatch between key and vectors")
    Q, n = key2.shape[0], vectors.shape[0]
    out = np.zeros((Q, n), dtype=np.float32)
    v = None if validity is None else np.ascontiguousarray(validity, dtype=np.uint8)
    check(_lib.lib().lb200_flat_distance_f16(C.c_int32(device_id), C.c_int32(DistanceType.code(metric)),
                                             C.c_uint32(vectors.shape[1]), _p(key2), C.c_uint32(Q), _p(vectors), _p(v),
                                             C.c_uint64(n), _p(out)))
    return out[0] if np.ndim(key) == 1 else out


def flat_knn_f16(key, metric, vectors, k, validity=None, lower_bound=None, upper_bound=None, device_id=0):
    """Scanner::flat_knn (scanner.rs:2066-2136) on a Float16 column."""
    key2 = np.atleast_2d(_f16(key))
    vectors = _f16(vectors)
    if vectors.ndim != 2 or key2.shape[1] != vectors.shape[1]:
        raise ValueError("dimension mismatch between key and vectors")
    if k <= 0:
        raise ValueError("k must be positive")
    Q, n = key2.shape[0], vectors.shape[0]
    od = np.full((Q, k), np.inf, dtype=np.float32)
    oi = np.full((Q, k), np.iinfo(np.uint64).max, dtype=np.uint64)
    oc = np.zeros(Q, dtype=np.uint32)
    v = None if validity is None else np.ascontiguousarray(validity, dtype=np.uint8)
    check(_lib.lib().lb200_flat_topk_f16(C.c_int32(device_id), C.c_int32(DistanceType.code(metric)),
                                         C.c_uint32(vectors.shape[1]), _p(key2), C.c_uint32(Q), _p(vectors), _p(v),
                                         C.c_uint64(n), C.c_uint32(k), _fptr(lower_bound), _fptr(upper_bound), _p(od),
                                         _p(oi), _p(oc)))
    return od, oi, oc


def normalize(key, device_id=0):
    """normalize_arrow (kernels.rs:119-142) of a batch of keys on the device (lb200_normalize)."""
    key2 = np.atleast_2d(_f32(key))
    out = np.empty_like(key2)
    check(_lib.lib().lb200_normalize(C.c_int32(device_id), C.c_uint32(key2.shape[1]), _p(key2), C.c_uint32(key2.shape[0]),
                                     _p(out)))
    return out

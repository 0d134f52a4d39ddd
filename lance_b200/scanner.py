"""Host-side mirror of `Scanner::nearest` for the vector-search plan (rust/lance/src/dataset/scanner.rs:1670-1835):
the callers either side of the IVF_PQ hot path — fan-out over the delta indices of a column,
the refine step with vectors fetched from the dataset (TakeExec), and `knn_combined`, which unions the ANN
result with a flat kNN over the fragments no index covers and re-scores everything exactly.

Every arithmetic step is a call into the C ABI (`lb200_search`, `lb200_merge_topk`, `lb200_rerank`,
`lb200_flat_topk`); this module only moves row ids and vectors between them, which is what the reference's
exec-node plumbing (UnionExec / TakeExec / project) does.  Batched over queries; results are per query what the
reference's plan produces for that query.
"""
from dataclasses import replace
from typing import Callable, Optional, Sequence

import numpy as np

from .index import IvfPqIndex, Query, flat_knn, merge_topk, normalize, rerank

_NONE = np.iinfo(np.uint64).max


class VectorDataset:
    """The part of `Dataset` the plan touches: one vector column addressed by row id, the delta indices built
    over it (all the `Index` entries sharing a name, `load_indices_by_name`), and the unindexed fragments
    (`Dataset::unindexed_fragments`)."""

    def __init__(self, take: Callable[[np.ndarray], np.ndarray], dim: int, indices: Sequence[IvfPqIndex] = (),
                 unindexed_row_ids: Optional[np.ndarray] = None, unindexed_vectors: Optional[np.ndarray] = None,
                 device_id: int = 0):
        self.take, self.dim, self.indices, self.device_id = take, int(dim), list(indices), device_id
        if (unindexed_row_ids is None) != (unindexed_vectors is None):
            raise ValueError("unindexed_row_ids and unindexed_vectors go together")
        self.un_ids = None if unindexed_row_ids is None else np.ascontiguousarray(unindexed_row_ids, dtype=np.uint64)
        self.un_vec = None if unindexed_vectors is None else np.ascontiguousarray(unindexed_vectors, dtype=np.float32)
        if self.un_ids is not None and (self.un_vec.ndim != 2 or self.un_vec.shape != (self.un_ids.size, self.dim)):
            raise ValueError("unindexed_vectors must be [len(unindexed_row_ids)][dim]")
        metrics = {ix.metric_type() for ix in self.indices}
        if len(metrics) > 1:
            raise ValueError(f"delta indices disagree on the metric: {sorted(metrics)}")

    # ------------------------------------------------------------------------------------------
    def _take(self, row_ids, count):
        """TakeExec (io/exec/take.rs:178) for [Q][C] row ids; slots past count[q] are not fetched."""
        Q, Cn = row_ids.shape
        out = np.zeros((Q, Cn, self.dim), dtype=np.float32)
        valid = np.arange(Cn)[None, :] < count[:, None]
        if valid.any():
            out[valid] = np.asarray(self.take(row_ids[valid]), dtype=np.float32)
        return out

    def _flat_over_unindexed(self, key, metric, k, lower, upper, un_ids=None, un_vec=None):
        un_ids = self.un_ids if un_ids is None else un_ids
        un_vec = self.un_vec if un_vec is None else un_vec
        od, op, oc = flat_knn(key, metric, un_vec, k, lower_bound=lower, upper_bound=upper, device_id=self.device_id)
        ids = np.full(op.shape, _NONE, dtype=np.uint64)
        valid = np.arange(k)[None, :] < oc[:, None]
        ids[valid] = un_ids[op[valid]]
        return od, ids, oc, op

    # ------------------------------------------------------------------------------------------
    def _ann_prefiltered(self, ix, key, kp, nprobes, lower, upper, allowed):
        """One delta index with a prefilter: what ANNIvfPartitionExec + ANNIvfSubIndexExec do (knn.rs:349-393,
        519-635): normalise the key (cosine), find_partitions, then search_in_partition per probed partition with the
        row-id mask (DatasetPreFilter, lance/src/index/prefilter.rs:227-290 -> per-partition bitmask of the C ABI)."""
        k2 = normalize(key, self.device_id) if ix.metric_type() == "cosine" else key
        probes = ix.find_partitions(k2, nprobes)                       # [Q][P]
        masks = {}
        for p in np.unique(probes).tolist():
            o, e = int(ix.part_offsets[p]), int(ix.part_offsets[p + 1])
            masks[p] = np.isin(ix.row_ids[o:e], allowed)
        od, oi, oc = ix.search_in_partition(probes, k2, kp, lower, upper, prefilter=masks)   # [Q][P][kp]
        return merge_topk(np.ascontiguousarray(od.transpose(1, 0, 2)), np.ascontiguousarray(oi.transpose(1, 0, 2)),
                          np.ascontiguousarray(oc.T), kp, device_id=self.device_id)

    def nearest(self, q: Query, fast_search: bool = False, allowed_row_ids=None):
        """Scanner::knn (scanner.rs:1670-1752).  Returns (distances [Q][k], row ids [Q][k], counts [Q]).
        allowed_row_ids: optional allow-list of row ids (the RowIdMask a prefilter evaluates to: deletions and
        scalar filter, prefilter.rs:26-51); rows outside it are skipped in the index scan and in the flat scan."""
        key = np.atleast_2d(np.ascontiguousarray(q.key, dtype=np.float32))
        if key.shape[1] != self.dim:
            raise ValueError(f"query dim({key.shape[1]}) doesn't match the column {q.column} vector dim({self.dim})")
        if q.k <= 0:
            raise ValueError("k must be positive")
        Q = key.shape[0]
        allowed = None if allowed_row_ids is None else np.unique(np.asarray(allowed_row_ids, dtype=np.uint64))
        un_ids, un_vec = self.un_ids, self.un_vec
        if allowed is not None and un_ids is not None:   # FilterExec on the flat scan (scanner.rs:1808-1813)
            keep = np.isin(un_ids, allowed)
            un_ids, un_vec = un_ids[keep], np.ascontiguousarray(un_vec[keep])
        if not (q.use_index and self.indices):
            # "No index found. use flat search" (scanner.rs:1734-1751) with the QUERY's metric
            if self.un_ids is None:
                raise ValueError("flat search needs the column's vectors (pass them as the unindexed rows)")
            od, ids, oc, _ = self._flat_over_unindexed(key, q.metric_type, q.k, q.lower_bound, q.upper_bound, un_ids, un_vec)
            return od, ids, oc
        if q.refine_factor == 0:
            raise ValueError("Refine factor can not be zero")                      # scanner.rs:1693-1698
        metric = self.indices[0].metric_type()   # distances must be comparable: the index's metric wins (1715-1721)
        rf = int(q.refine_factor or 1)
        kp = q.k * rf
        # ---- ann(): fan out over the deltas, SortExec(fetch = k * refine_factor) (scanner.rs:2139-2157) ----------
        sub = replace(q, key=key, k=kp, refine_factor=None)
        lists = []
        for ix in self.indices:
            if ix.has_raw:
                raise ValueError("VectorDataset refines through take(); build the delta indices without raw vectors")
            lists.append(ix.search(sub) if allowed is None else
                         self._ann_prefiltered(ix, key, kp, q.nprobes, q.lower_bound, q.upper_bound, allowed))
        if len(lists) == 1:
            ad, ai, ac = lists[0]
        else:
            ad, ai, ac = merge_topk(np.stack([l[0] for l in lists]), np.stack([l[1] for l in lists]),
                                    np.stack([l[2] for l in lists]), kp, device_id=self.device_id)
        # ---- refine: take + flat_knn on the candidates (scanner.rs:1706-1721) ------------------------------------
        vec = None
        if q.refine_factor is not None:
            vec = self._take(ai, ac)
            ad, ai, ac = rerank(key, metric, vec, ai, ac, q.k, q.lower_bound, q.upper_bound, self.device_id)
            vec = None   # re-fetched below only if the union needs it
        if fast_search or un_ids is None or un_ids.size == 0:
            return ad, ai, ac
        # ---- knn_combined (scanner.rs:1755-1835) -------------------------------------------------------------------
        fd, fi, fc, fpos = self._flat_over_unindexed(key, metric, q.k, q.lower_bound, q.upper_bound, un_ids, un_vec)
        kw = ad.shape[1]
        cand_ids = np.full((Q, q.k + kw), _NONE, dtype=np.uint64)
        cand_vec = np.zeros((Q, q.k + kw, self.dim), dtype=np.float32)
        cnt = (fc + ac).astype(np.uint32)
        ann_vec = self._take(ai, ac)
        for i in range(Q):   # UnionExec([topk_appended, knn_node]): appended rows first
            a, b = int(fc[i]), int(ac[i])
            cand_ids[i, :a] = fi[i, :a]
            cand_vec[i, :a] = un_vec[fpos[i, :a]]
            cand_ids[i, a:a + b] = ai[i, :b]
            cand_vec[i, a:a + b] = ann_vec[i, :b]
        return rerank(key, metric, cand_vec, cand_ids, cnt, q.k, q.lower_bound, q.upper_bound, self.device_id)

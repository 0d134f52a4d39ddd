"""Multi-GPU sharding of the IVF_PQ query path: one process per GPU.

IVF partitions are independent units (rust/lance/src/io/exec/knn.rs:519-635 searches them one by
one and merges), so they shard with no data-path collective except ONE exchange step:

  rank r:  coarse step on the full batch (replicated) -> scan of the probed partitions it owns
           -> local k' candidates per query            (lb200_search_device, shard mode)
  all:     all_gather of [Q][k'] (dist f32, row id i64, count i32[, exact f32])
  rank 0:  merge of the G lists == concat + sort_to_indices(limit k') of ivf/v2.rs:383-400,
           then the refine cut of scanner.rs:1706-1721     (lb200_merge_topk_device)

The result is bit-identical to the single-index search because every shard list is the exact
(distance, row id)-ordered top-k' of its rows and top-k'(union) is contained in the union of the
shards' top-k'.  This module is host-side plumbing only (torch.distributed); all compute is in
liblance_b200.so.
"""
import numpy as np


def assign_partitions(sizes, world):
    """Greedy bin-packing (largest first) of partitions onto ranks by row count, so the code
    bytes per GPU are balanced.  Deterministic: every rank computes the same owner map."""
    sizes = np.asarray(sizes, dtype=np.int64)
    order = np.argsort(-sizes, kind="stable")
    load = np.zeros(world, dtype=np.int64)
    owner = np.zeros(len(sizes), dtype=np.int64)
    for p in order:
        r = int(np.argmin(load))
        owner[p] = r
        load[r] += sizes[p]
    return owner


def shard_of(data, rank, world):
    """The IvfPqData shard of `rank`: all nlist partitions, the ones it does not own empty;
    centroids and codebook replicated."""
    sizes = np.diff(data.part_offsets.astype(np.int64))
    owner = assign_partitions(sizes, world)
    return data.subset(np.nonzero(owner == rank)[0]), owner


def gather_candidates(dist, local_dist, local_rowid, local_count, local_exact=None):
    """The exchange step: all_gather of the per-rank candidate lists.  Tensors may live on CPU
    (gloo, tests) or on the GPU (nccl, bench).  Returns [G][Q][k'] / [G][Q] tensors."""
    import torch
    world = dist.get_world_size()

    def ag(t):
        out = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        dist.all_gather([out[g] for g in range(world)], t.contiguous())  # views: lands in `out`
        return out

    g_dist, g_rowid, g_count = ag(local_dist), ag(local_rowid), ag(local_count)
    g_exact = ag(local_exact) if local_exact is not None else None
    return g_dist, g_rowid, g_count, g_exact


class ShardedIvfPq:
    """One rank's view of a partition-sharded index: `search` returns the global top-k on rank 0."""

    def __init__(self, index, dist, device):
        import torch
        self.ix, self.dist, self.dev = index, dist, device
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self._torch = torch

    def search_device(self, d_queries, Q, nprobes, k, refine_factor, stream):
        """d_queries: CUDA tensor [Q][dim].  Returns (dist [Q][k], rowid [Q][k], count [Q]) CUDA
        tensors on rank 0, None elsewhere.  Everything is ordered on `stream`."""
        from .index import merge_topk_device
        torch = self._torch
        kp = k * max(refine_factor, 1)
        od = torch.empty((Q, kp), dtype=torch.float32, device=self.dev)
        oi = torch.empty((Q, kp), dtype=torch.int64, device=self.dev)
        oc = torch.empty((Q,), dtype=torch.int32, device=self.dev)
        ex = torch.empty((Q, kp), dtype=torch.float32, device=self.dev)
        self.ix.search_device(d_queries.data_ptr(), Q, nprobes, k, refine_factor, od.data_ptr(), oi.data_ptr(),
                              oc.data_ptr(), d_out_exact=ex.data_ptr(), stream=stream.cuda_stream)
        g_od, g_oi, g_oc, g_ex = gather_candidates(self.dist, od, oi, oc, ex if refine_factor else None)
        if self.rank != 0:
            return None
        fd = torch.empty((Q, k), dtype=torch.float32, device=self.dev)
        fi = torch.empty((Q, k), dtype=torch.int64, device=self.dev)
        fc = torch.empty((Q,), dtype=torch.int32, device=self.dev)
        merge_topk_device(self.dev.index, g_od.data_ptr(), g_oi.data_ptr(), g_ex.data_ptr() if refine_factor else None,
                          g_oc.data_ptr(), self.world, Q, kp, kp, k, fd.data_ptr(), fi.data_ptr(), fc.data_ptr(),
                          stream=stream.cuda_stream)
        return fd, fi, fc

"""Multi-GPU sharding of the IVF_PQ query path: one process per GPU.

IVF partitions are independent units (rust/lance/src/io/exec/knn.rs:519-635 searches them one by
one and merges), so they shard with no data-path collective except ONE exchange step:

  rank r:  coarse step (sharded by queries, probe lists all-gathered) -> scan of the probed
           partitions it owns -> local k' candidates per query   (lb200_search_probed_device)
           While scanning, the ranks publish each query's running k'-th-best bound into their
           peers' tables over NVLink peer memory (lb200_peer_thresholds_*), so a row that cannot
           reach the merged top-k' is skipped on every GPU, not only on the one that knows.
  all:     all_gather of [Q][k'] (dist f32, row id i64, count i32[, exact f32])
  rank 0:  merge of the G lists == concat + sort_to_indices(limit k') of ivf/v2.rs:383-400,
           then the refine cut of scanner.rs:1706-1721     (lb200_merge_topk_device)

The result is bit-identical to the single-index search because every shard list holds, in
(distance, row id) order, every one of its rows that can be in the merged top-k': its exact local
top-k', minus only rows above a bound some shard has PROVED (k' rows at or below it exist), and
top-k'(union) is contained in the union of the shards' top-k'.  This module is host-side plumbing only (torch.distributed); all compute is in
liblance_b200.so.
"""
import numpy as np


def assign_partitions(sizes, world, weights=None):
    """Greedy bin-packing (largest first) of partitions onto ranks.  Deterministic: every rank computes the same
    owner map.  By default the balanced quantity is the row count (code bytes per GPU); with `weights` (e.g.
    probe_load_weights) it is rows x expected probes, i.e. the SCAN work per GPU -- in high dimension a few
    "hub" centroids are among the nprobes nearest of many more queries than the others (cfg 4: probes per
    partition have std = mean), so equal bytes are not equal work (r02: 6.9 % extra on the busiest of 8 shards)."""
    sizes = np.asarray(sizes, dtype=np.int64)
    w = sizes.astype(np.float64) if weights is None else sizes.astype(np.float64) * np.asarray(weights, dtype=np.float64)
    order = np.argsort(-w, kind="stable")
    load = np.zeros(world, dtype=np.float64)
    owner = np.zeros(len(sizes), dtype=np.int64)
    for p in order:
        r = int(np.argmin(load))
        owner[p] = r
        load[r] += w[p]
    return owner


def probe_load_weights(centroids, metric, nprobes, chunk=1024):
    """Expected probe frequency of every partition, estimated from the index alone: how often a partition is among
    the `nprobes` nearest of the OTHER centroids (queries follow the data, the data follows the centroids).  Uses the
    coarse metric of find_partitions (dot -> largest inner product, else L2; a cosine index holds normalised
    centroids).  +1 so that no partition weighs nothing."""
    c = np.ascontiguousarray(centroids, dtype=np.float32)
    nlist = c.shape[0]
    P = min(int(nprobes), nlist)
    counts = np.zeros(nlist, dtype=np.int64)
    c2 = (c * c).sum(1)
    for s in range(0, nlist, chunk):
        x = c[s:s + chunk]
        score = -(x @ c.T) if metric == "dot" else c2[None, :] - 2.0 * (x @ c.T)
        idx = np.argpartition(score, P - 1, axis=1)[:, :P] if P < nlist else np.tile(np.arange(nlist), (len(x), 1))
        counts += np.bincount(idx.ravel(), minlength=nlist)
    return counts.astype(np.float64) + 1.0


def shard_of(data, rank, world, nprobes=None):
    """The IvfPqData shard of `rank`: all nlist partitions, the ones it does not own empty;
    centroids and codebook replicated.  nprobes: balance scan work (rows x expected probes) instead of rows."""
    sizes = np.diff(data.part_offsets.astype(np.int64))
    weights = None if nprobes is None else probe_load_weights(data.centroids, data.metric, nprobes)
    owner = assign_partitions(sizes, world, weights)
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
    """One rank's view of a partition-sharded index.  `prepare` fixes the batch shape (buffers are
    allocated once), `search_device` runs one batch and leaves the global top-k on rank 0.

    Per batch: (1) coarse step sharded by QUERIES -- rank r selects the probes of its Q/G slice,
    one all_gather makes the [Q][P] probe lists identical everywhere; (2) every rank scans the
    probed partitions it owns; (3) all_gather of the local k' lists; (4) merge on rank 0."""

    def __init__(self, index, dist, device, shard_coarse=True):
        import torch
        self.ix, self.dist, self.dev = index, dist, device
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.shard_coarse = shard_coarse and self.world > 1
        self._torch = torch

    def prepare(self, Q, nprobes, k, refine_factor):
        torch, dev, G = self._torch, self.dev, self.world
        self.Q, self.P, self.k, self.refine = Q, min(nprobes, self.ix.nlist), k, refine_factor
        kp = self.kp = k * max(refine_factor, 1)
        self.qs = (Q + G - 1) // G                      # queries per rank in the coarse step
        self.l_probes = torch.zeros((self.qs, self.P), dtype=torch.int32, device=dev)
        self.g_probes = torch.zeros((G * self.qs, self.P), dtype=torch.int32, device=dev)
        self.od = torch.empty((Q, kp), dtype=torch.float32, device=dev)
        self.oi = torch.empty((Q, kp), dtype=torch.int64, device=dev)
        self.oc = torch.empty((Q,), dtype=torch.int32, device=dev)
        self.ex = torch.empty((Q, kp), dtype=torch.float32, device=dev)
        self.g_od = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
        self.g_oi = torch.empty((G, Q, kp), dtype=torch.int64, device=dev)
        self.g_oc = torch.empty((G, Q), dtype=torch.int32, device=dev)
        self.g_ex = torch.empty((G, Q, kp), dtype=torch.float32, device=dev)
        self.fd = torch.empty((Q, k), dtype=torch.float32, device=dev)
        self.fi = torch.empty((Q, k), dtype=torch.int64, device=dev)
        self.fc = torch.empty((Q,), dtype=torch.int32, device=dev)
        self.l_q = torch.zeros((self.qs, self.ix.dim), dtype=torch.float32, device=dev)
        self.g_q = torch.zeros((G * self.qs, self.ix.dim), dtype=torch.float32, device=dev)
        self._share_bounds(Q)

    def _share_bounds(self, Q):
        """Give every shard a table of per-query pruning bounds that its peers update from inside
        their scan kernels (remote atomic-min over NVLink peer memory; include/lance_b200.h,
        lb200_peer_thresholds_*).  One all_gather of the 64-byte CUDA IPC handles at set-up."""
        import os
        torch, G = self._torch, self.world
        self.shared_bounds = False
        if G == 1 or os.environ.get("LB200_NO_PEER_THR"):
            return
        handle, _ = self.ix.peer_thresholds_export(Q)
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=self.dev)
        allh = torch.empty((G, 64), dtype=torch.uint8, device=self.dev)
        self.dist.all_gather_into_tensor(allh, mine)
        handles = [bytes(allh[g].cpu().tolist()) for g in range(G)]
        self.ix.peer_thresholds_attach(G, self.rank, handles=handles)
        self.dist.barrier()  # every table exists and is cleared before anyone's first batch publishes
        self.shared_bounds = True

    def search_host(self, h_queries, stream):
        """End-to-end entry: h_queries = pinned host f32 tensor [Q][dim], the same on every rank.  Each rank
        copies only ITS Q/G slice over PCIe and the slices are all-gathered over NVLink (every rank needs
        every query for the tables of its pairs), then search_device."""
        lo = min(self.rank * self.qs, self.Q)
        n = min(self.qs, self.Q - lo)
        if n > 0:
            self.l_q[:n].copy_(h_queries[lo:lo + n], non_blocking=True)
        self.dist.all_gather_into_tensor(self.g_q, self.l_q)
        return self.search_device(self.g_q[:self.Q], stream)

    def search_device(self, d_queries, stream):
        """d_queries: CUDA f32 tensor [Q][dim].  Everything is ordered on `stream` (which must be
        torch's current stream so that NCCL orders after our kernels).  Returns (fd, fi, fc) --
        valid on rank 0 only."""
        from .index import merge_topk_device
        dist, ix, sp = self.dist, self.ix, stream.cuda_stream
        Q, P, k, rf = self.Q, self.P, self.k, self.refine
        if self.shard_coarse:
            lo = min(self.rank * self.qs, Q)
            n = min(self.qs, Q - lo)
            if n > 0:
                ix.find_partitions_device(d_queries[lo:].data_ptr(), n, P, self.l_probes.data_ptr(), normalize=True,
                                          stream=sp)
            dist.all_gather_into_tensor(self.g_probes, self.l_probes)
            ix.search_probed_device(d_queries.data_ptr(), Q, self.g_probes.data_ptr(), P, k, rf, self.od.data_ptr(),
                                    self.oi.data_ptr(), self.oc.data_ptr(), d_out_exact=self.ex.data_ptr(), stream=sp)
        else:
            ix.search_device(d_queries.data_ptr(), Q, P, k, rf, self.od.data_ptr(), self.oi.data_ptr(),
                             self.oc.data_ptr(), d_out_exact=self.ex.data_ptr(), stream=sp)
        dist.all_gather_into_tensor(self.g_od, self.od)
        dist.all_gather_into_tensor(self.g_oi, self.oi)
        dist.all_gather_into_tensor(self.g_oc, self.oc)
        if rf:
            dist.all_gather_into_tensor(self.g_ex, self.ex)
        if self.rank == 0:
            merge_topk_device(self.dev.index, self.g_od.data_ptr(), self.g_oi.data_ptr(),
                              self.g_ex.data_ptr() if rf else None, self.g_oc.data_ptr(), self.world, Q, self.kp,
                              self.kp, k, self.fd.data_ptr(), self.fi.data_ptr(), self.fc.data_ptr(), stream=sp)
        return self.fd, self.fi, self.fc


class PeerShardedIvfPq:
    """One rank's view of a partition-sharded index whose whole exchange lives in liblance_b200.so (lb200_shard_*):
    probe lists and candidate lists move between the GPUs through NVLink peer memory, ranks meet through epoch flags
    -- no collective per batch.  torch.distributed is used ONCE, at set-up, to hand the 128-byte blobs around (any
    channel would do)."""

    def __init__(self, index, dist, device):
        import torch
        self.ix, self.dist, self.dev, self._torch = index, dist, device, torch
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

    def prepare(self, Q, nprobes, k, refine_factor):
        torch, G = self._torch, self.world
        self.Q, self.P, self.k, self.refine = Q, min(nprobes, self.ix.nlist), k, refine_factor
        blob, _ = self.ix.shard_export(Q, self.P, k * max(refine_factor, 1))
        mine = torch.tensor(list(blob), dtype=torch.uint8, device=self.dev)
        allb = torch.empty((G, 128), dtype=torch.uint8, device=self.dev)
        self.dist.all_gather_into_tensor(allb, mine)
        self.ix.shard_attach(G, self.rank, blobs=[bytes(allb[g].cpu().tolist()) for g in range(G)])
        self.dist.barrier()   # every block exists and is mapped before anyone's first batch stores into it
        self.fd = torch.empty((Q, k), dtype=torch.float32, device=self.dev)
        self.fi = torch.empty((Q, k), dtype=torch.int64, device=self.dev)
        self.fc = torch.empty((Q,), dtype=torch.int32, device=self.dev)

    def search_device(self, d_queries, stream):
        """d_queries: CUDA f32 [Q][dim], the same on every rank.  Asynchronous on `stream`; (fd, fi, fc) on rank 0."""
        self.ix.shard_search_device(d_queries.data_ptr(), self.Q, self.P, self.k, self.refine, self.fd.data_ptr(),
                                    self.fi.data_ptr(), self.fc.data_ptr(), stream=stream.cuda_stream)
        return self.fd, self.fi, self.fc

    def search_host(self, h_queries_np, out):
        """Blocking host-pointer entry (lb200_shard_search): only this rank's slice of the batch crosses PCIe."""
        from .index import Query
        return self.ix.shard_search(Query(key=h_queries_np, k=self.k, nprobes=self.P, refine_factor=self.refine or None), out=out)

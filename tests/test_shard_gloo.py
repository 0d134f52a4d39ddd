"""N>1 path on CPU (gloo, world size 2): partition sharding + the one exchange step + merge.

Each rank searches ITS shard with the CPU oracle (the checker standing in for the GPU kernels,
which need a device), the ranks all_gather their local k' lists with the same
`lance_b200.shard.gather_candidates` the bench uses over NCCL, and rank 0 merges with a numpy
restatement of lb200_merge_topk_device's rule.  The merged result must be bit-identical to the
oracle's search of the UNSHARDED index -- the property the multi-GPU design rests on
(DESIGN.md section 5).
"""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _key(d):
    b = np.ascontiguousarray(d, dtype=np.float32).view(np.uint32).astype(np.uint64)
    return np.where(b & 0x80000000, (~b) & 0xFFFFFFFF, b | 0x80000000)


def _merge(g_d, g_i, g_c, g_ex, kprime, k):
    """[G][Q][k'] lists -> per query: k' smallest by (dist,row id); if exact given, re-rank those by
    (exact,row id) and keep k (scanner.rs:1706-1721), else keep min(k,k')."""
    G, Q, _ = g_d.shape
    out_d = np.full((Q, k), np.inf, np.float32)
    out_i = np.full((Q, k), -1, np.int64)
    out_c = np.zeros(Q, np.int32)
    for q in range(Q):
        ds = np.concatenate([g_d[g, q, :g_c[g, q]] for g in range(G)])
        ids = np.concatenate([g_i[g, q, :g_c[g, q]] for g in range(G)])
        order = np.lexsort((ids.astype(np.uint64), _key(ds)))[:kprime]
        if g_ex is not None:
            ex = np.concatenate([g_ex[g, q, :g_c[g, q]] for g in range(G)])[order]
            ids2 = ids[order]
            o2 = np.lexsort((ids2.astype(np.uint64), _key(ex)))[:k]
            n = len(o2)
            out_d[q, :n], out_i[q, :n] = ex[o2], ids2[o2]
        else:
            order = order[:k]
            n = len(order)
            out_d[q, :n], out_i[q, :n] = ds[order], ids[order]
        out_c[q] = n
    return out_d, out_i, out_c


def _worker(rank, world, port, metric, refine, ret):
    import torch
    import torch.distributed as dist
    import oracle
    from lance_b200.shard import gather_candidates, shard_of
    from lance_b200.synth import build_ivf_pq
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(3)
        d, nlist, M, k, nprobes = 32, 12, 8, 5, 6
        vecs = rng.normal(size=(3000, d)).astype(np.float32)
        data = build_ivf_pq(vecs, metric, nlist, M, seed=1, keep_raw=True, ivf_iters=3, pq_iters=2, sample=2048)
        q = vecs[rng.integers(0, len(vecs), 17)] + 0.05 * rng.normal(size=(17, d)).astype(np.float32)
        shard, owner = shard_of(data, rank, world)
        assert sorted(np.unique(owner).tolist()) == list(range(world))
        kp = k * max(refine, 1)
        ld, li, lc, probes = oracle.OracleIndex(shard).search(q, k=kp, nprobes=nprobes)
        ld = np.ascontiguousarray(ld[:, :kp], np.float32)
        li = np.ascontiguousarray(li[:, :kp]).astype(np.int64)
        ex = None
        if refine:
            # exact index-metric distance of the local candidates on the shard's raw vectors
            pos = {int(r): i for i, r in enumerate(shard.row_ids)}
            ex = np.full_like(ld, np.nan)
            for qi in range(len(q)):
                for j in range(int(lc[qi])):
                    v = shard.raw_vectors[pos[int(li[qi, j])]][None, :]
                    ex[qi, j] = oracle.distance_batch(metric, q[qi], v)[0]
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a))
        g_d, g_i, g_c, g_ex = gather_candidates(dist, t(ld), t(li), t(lc.astype(np.int32)), t(ex) if refine else None)
        # every rank sees identical probe lists (replicated coarse step)
        pl = [torch.empty_like(t(probes.astype(np.int64))) for _ in range(world)]
        dist.all_gather(pl, t(probes.astype(np.int64)))
        assert all(torch.equal(pl[0], x) for x in pl)
        if rank == 0:
            md, mi, mc = _merge(g_d.numpy(), g_i.numpy(), g_c.numpy(), g_ex.numpy() if refine else None, kp, k)
            rd, ri, rc, _ = oracle.OracleIndex(data).search(q, k=k, nprobes=nprobes, refine_factor=refine)
            ok = (mc.tolist() == rc.tolist())
            for qi in range(len(q)):
                n = int(rc[qi])
                ok = ok and mi[qi, :n].tolist() == ri[qi, :n].astype(np.int64).tolist()
                ok = ok and md[qi, :n].view(np.uint32).tolist() == np.ascontiguousarray(rd[qi, :n], np.float32).view(np.uint32).tolist()
            ret.put(bool(ok))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("metric,refine", [("l2", 0), ("cosine", 3), ("dot", 0)])
def test_two_rank_shard_merge_equals_single_index(oracle, metric, refine):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, metric, refine, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert ret.get(timeout=5) is True


def test_assign_partitions_balanced():
    from lance_b200.shard import assign_partitions
    rng = np.random.default_rng(0)
    sizes = (rng.lognormal(0, 0.5, 4096) * 24000).astype(np.int64)
    for world in (1, 2, 4, 8):
        owner = assign_partitions(sizes, world)
        load = np.bincount(owner, weights=sizes, minlength=world)
        assert load.max() / load.mean() < 1.01

"""Synthetic IVF_PQ index builder (host side, numpy/torch-CPU).

Produces the exchange format that both the CUDA library (lb200_index_create) and the CPU
oracle consume: the in-memory image of what Lance's v3 IVF_PQ index holds after
`IvfQuantizationStorage::load_partition` (rust/lance-index/src/vector/storage.rs:271-298):
per-partition *transposed* codes [M][n_p] (pq/storage.rs:357-377,455-470), row ids,
IVF centroids (ivf/storage.rs) and the PQ codebook [M][256][dsub] (pq/utils.rs:54-71).

Index *training quality* is irrelevant to parity (both sides search the same index); a few
Lloyd iterations keep recall meaningful.  Distributions follow SURVEY.md 8(d).
"""
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch


@dataclass
class IvfPqData:
    metric: str               # "l2" | "cosine" | "dot"
    dim: int
    nlist: int
    M: int
    centroids: np.ndarray     # [nlist][dim] f32
    codebook: np.ndarray      # [M][256][dim/M] f32
    part_offsets: np.ndarray  # [nlist+1] u64 (rows)
    codes: np.ndarray         # u8, concat over p of [M][n_p] (transposed)
    row_ids: np.ndarray       # [N] u64, index order
    raw_vectors: Optional[np.ndarray] = None  # [N][dim] f32, index order (for refine)
    num_bits: int = 8         # 4: codebook [M][16][dim/M], codes concat over p of [M/2][n_p] (two codes per byte)

    @property
    def N(self):
        return int(self.part_offsets[-1])

    @property
    def code_bytes(self):
        """bytes of PQ code per row (pq/storage.rs:172-176)."""
        return self.M // 2 if self.num_bits == 4 else self.M

    def part_codes(self, p):
        o, e = int(self.part_offsets[p]), int(self.part_offsets[p + 1])
        cb = self.code_bytes
        return self.codes[o * cb:e * cb].reshape(cb, e - o)

    def subset(self, parts):
        """Index restricted to the given partitions (others become empty): the per-GPU
        shard of SURVEY 8(e).  Centroids/codebook are replicated."""
        keep = np.zeros(self.nlist, dtype=bool)
        keep[np.asarray(parts, dtype=np.int64)] = True
        sizes = np.diff(self.part_offsets.astype(np.int64)) * keep
        offs = np.zeros(self.nlist + 1, dtype=np.uint64)
        offs[1:] = np.cumsum(sizes)
        codes, rids, raws = [], [], []
        for p in np.nonzero(keep)[0]:
            o, e = int(self.part_offsets[p]), int(self.part_offsets[p + 1])
            codes.append(self.codes[o * self.code_bytes:e * self.code_bytes])
            rids.append(self.row_ids[o:e])
            if self.raw_vectors is not None:
                raws.append(self.raw_vectors[o:e])
        cat = lambda xs, dt, shp: (np.concatenate(xs) if xs else np.zeros(shp, dtype=dt))
        return IvfPqData(self.metric, self.dim, self.nlist, self.M, self.centroids, self.codebook,
                         offs, cat(codes, np.uint8, (0,)), cat(rids, np.uint64, (0,)),
                         None if self.raw_vectors is None else cat(raws, np.float32, (0, self.dim)), self.num_bits)


def _assign(x, cents, chunk=65536):
    """argmin_c ||x - c||^2 (expanded form; training-time only, not on the parity path)."""
    c2 = (cents * cents).sum(1)
    out = torch.empty(x.shape[0], dtype=torch.int64)
    for s in range(0, x.shape[0], chunk):
        xs = x[s:s + chunk]
        d = c2[None, :] - 2.0 * (xs @ cents.T)
        out[s:s + chunk] = d.argmin(1)
    return out


def kmeans(x, k, iters, seed):
    g = torch.Generator().manual_seed(seed)
    n = x.shape[0]
    if n >= k:
        cents = x[torch.randperm(n, generator=g)[:k]].clone()
    else:
        cents = x[torch.randint(0, n, (k,), generator=g)].clone()
        cents += 1e-3 * torch.randn(cents.shape, generator=g)
    for _ in range(iters):
        a = _assign(x, cents)
        sums = torch.zeros_like(cents).index_add_(0, a, x)
        cnt = torch.bincount(a, minlength=k).to(x.dtype)
        nz = cnt > 0
        cents[nz] = sums[nz] / cnt[nz, None]
    return cents


def build_ivf_pq(vectors, metric, nlist, M, seed=0, ivf_iters=8, pq_iters=6, sample=65536,
                 keep_raw=False, row_ids=None):
    """Train + encode an IVF_PQ index over `vectors` [N][d] f32 (numpy)."""
    torch.manual_seed(seed)
    raw = np.ascontiguousarray(vectors, dtype=np.float32)
    N, d = raw.shape
    assert d % M == 0
    dsub = d // M
    x = torch.from_numpy(raw)
    if metric == "cosine":  # a cosine index stores L2 structures over normalised data
        x = x / x.norm(dim=1, keepdim=True).clamp_min(1e-30)
    g = torch.Generator().manual_seed(seed + 1)
    samp = x[torch.randperm(N, generator=g)[:min(sample, N)]]
    cents = kmeans(samp, nlist, ivf_iters, seed + 2)
    if metric == "dot":
        part = (x @ cents.T).argmax(1) if N <= 262144 else torch.cat(
            [(x[s:s + 65536] @ cents.T).argmax(1) for s in range(0, N, 65536)])
        resid = x
    else:
        part = _assign(x, cents)
        resid = x - cents[part]
    rs = resid[torch.randperm(N, generator=g)[:min(sample, N)]]
    codebook = torch.empty(M, 256, dsub)
    codes = torch.empty(N, M, dtype=torch.uint8)
    for m in range(M):
        cb = kmeans(rs[:, m * dsub:(m + 1) * dsub].contiguous(), 256, pq_iters, seed + 10 + m)
        codebook[m] = cb
        codes[:, m] = _assign(resid[:, m * dsub:(m + 1) * dsub].contiguous(), cb).to(torch.uint8)
    order = torch.argsort(part, stable=True)
    sizes = torch.bincount(part, minlength=nlist).numpy().astype(np.uint64)
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    codes_sorted = codes[order].numpy()
    flat = np.empty(N * M, dtype=np.uint8)
    for p in range(nlist):
        o, e = int(offs[p]), int(offs[p + 1])
        flat[o * M:e * M] = codes_sorted[o:e].T.reshape(-1)
    rid = order.numpy().astype(np.uint64)
    if row_ids is not None:
        rid = np.asarray(row_ids, dtype=np.uint64)[order.numpy()]
    return IvfPqData(metric, d, nlist, M, cents.numpy().copy(), codebook.numpy().copy(), offs, flat,
                     rid, raw[order.numpy()].copy() if keep_raw else None)


def sift_like(N, d, seed, nblobs=256, centres=None):
    """SIFT-shaped data (SURVEY 8d cfg 1): Gaussian blobs, clipped to [0,218], integer-valued."""
    rng = np.random.default_rng(seed)
    if centres is None:
        centres = np.random.default_rng(1234).uniform(0, 128, size=(nblobs, d)).astype(np.float32)
    which = rng.integers(0, centres.shape[0], size=N)
    x = centres[which] + rng.normal(0, 16, size=(N, d)).astype(np.float32)
    return np.rint(np.clip(x, 0, 218)).astype(np.float32)


def embedding_like(N, d, seed, rank=64):
    """OpenAI-embedding-shaped data (cfg 3): unit-norm, low-rank + isotropic noise."""
    rng = np.random.default_rng(seed)
    basis = np.random.default_rng(2025).normal(size=(rank, d)).astype(np.float32) / np.sqrt(d)
    x = rng.normal(size=(N, rank)).astype(np.float32) @ basis
    x += 0.05 * rng.normal(size=(N, d)).astype(np.float32) / np.sqrt(d)
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)


def gaussian(N, d, seed):
    return np.random.default_rng(seed).normal(size=(N, d)).astype(np.float32)


def random_index(N, d, nlist, M, metric, seed, sigma=0.5, keep_raw=False):
    """bigann-scale index without training (cfg 4/5): centroids ~ SIFT-like, partition sizes
    log-normal(sigma) normalised to N, codes uniform u8 (scan cost is value independent),
    row ids a seeded permutation, codebook Gaussian scaled to the residual magnitude."""
    rng = np.random.default_rng(seed)
    if metric == "dot":
        cents = rng.normal(size=(nlist, d)).astype(np.float32)
        codebook = rng.normal(size=(M, 256, d // M)).astype(np.float32)
    else:
        cents = rng.uniform(0, 128, size=(nlist, d)).astype(np.float32)
        codebook = rng.normal(0, 16, size=(M, 256, d // M)).astype(np.float32)
    w = rng.lognormal(0.0, sigma, size=nlist)
    sizes = np.floor(w / w.sum() * N).astype(np.int64)
    sizes[: N - sizes.sum()] += 1
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    codes = rng.integers(0, 256, size=N * M, dtype=np.uint8)
    rid = rng.permutation(N).astype(np.uint64)
    raw = rng.normal(size=(N, d)).astype(np.float32) if keep_raw else None
    return IvfPqData(metric, d, nlist, M, cents, codebook, offs, codes, rid, raw)


def queries_near_centroids(data, Q, seed, scale=None):
    """Queries drawn around random IVF centroids so probe lists are realistic."""
    rng = np.random.default_rng(seed)
    c = data.centroids[rng.integers(0, data.nlist, size=Q)]
    s = scale if scale is not None else (16.0 if data.metric != "dot" else 0.5)
    return (c + rng.normal(0, s, size=c.shape)).astype(np.float32)

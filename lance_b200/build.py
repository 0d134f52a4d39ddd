"""GPU index build: host-side mirror of the reference's IVF_PQ build for the steps that now run behind the C ABI
(include/lance_b200.h "GPU index BUILD"): k-means training of the IVF centroids and the PQ codebooks, partition
assignment, residuals, PQ encoding.  Mirrors `lance.vector.train_ivf_centroids_on_accelerator` /
`compute_partitions` / `train_pq_codebook_on_accelerator` (python/python/lance/vector.py:134-700, the reference's torch
path) and the Rust builder (rust/lance-index/src/vector/ivf/builder.rs, pq/builder.rs): sample_rate 256 rows per
centroid, k-means max 50 rounds / tolerance 1e-4, cosine = normalise + L2, PQ trained on residuals for L2 / cosine.

torch is used for device buffers, sampling and the final re-ordering only; every distance, argmin, mean and code comes
from liblance_b200.so.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check
from .index import DistanceType, _f32, _p
from .synth import IvfPqData


def kmeans_train(vectors, k, metric="l2", max_iters=50, tolerance=1e-4, seed=0, init_centroids=None, device_id=0):
    """KMeans::new_with_params (kmeans.rs:528-561) on host arrays.  Returns (centroids [k][d], rounds, loss)."""
    x = _f32(vectors)
    if x.ndim != 2:
        raise ValueError("vectors must be [n][dim]")
    init = None if init_centroids is None else _f32(init_centroids)
    out = np.empty((k, x.shape[1]), dtype=np.float32)
    it, loss = C.c_uint32(0), C.c_double(0)
    check(_lib.lib().lb200_kmeans_train(
        C.c_int32(device_id), C.c_int32(DistanceType.code(metric)), C.c_uint32(x.shape[1]), _p(x), C.c_uint64(x.shape[0]),
        C.c_uint32(k), C.c_uint32(max_iters), C.c_float(tolerance), C.c_uint64(seed), _p(init), _p(out), C.byref(it), C.byref(loss)))
    return out, int(it.value), float(loss.value)


def compute_partitions(centroids, vectors, metric="l2", device_id=0, return_distances=False):
    """compute_partitions (kmeans.rs:721-770): ids [n] u32, 0xffffffff where the reference yields None."""
    c, x = _f32(centroids), _f32(vectors)
    ids = np.empty(x.shape[0], dtype=np.uint32)
    d = np.empty(x.shape[0], dtype=np.float32) if return_distances else None
    check(_lib.lib().lb200_compute_partitions(
        C.c_int32(device_id), C.c_int32(DistanceType.code(metric)), C.c_uint32(x.shape[1]), _p(c), C.c_uint32(c.shape[0]), _p(x),
        C.c_uint64(x.shape[0]), _p(ids), _p(d)))
    return (ids, d) if return_distances else ids


def pq_encode(codebook, vectors, metric="l2", device_id=0):
    """ProductQuantizer::transform (pq.rs:120-190): codes [n][M] u8.  codebook [M][256][dsub]."""
    cb, x = _f32(codebook), _f32(vectors)
    M = cb.shape[0]
    out = np.empty((x.shape[0], M), dtype=np.uint8)
    check(_lib.lib().lb200_pq_encode(C.c_int32(device_id), C.c_int32(DistanceType.code(metric)), C.c_uint32(x.shape[1]),
                                     C.c_uint32(M), _p(cb), _p(x), C.c_uint64(x.shape[0]), _p(out)))
    return out


def build_ivf_pq_gpu(vectors, metric, nlist, M, seed=0, sample_rate=256, max_iters=50, tolerance=1e-4, keep_raw=False,
                     row_ids=None, device_id=0, init_centroids=None, init_codebook=None):
    """Train + encode an IVF_PQ index on the GPU.  vectors: numpy [N][d] f32 or a CUDA torch tensor.
    Returns (IvfPqData, timings dict)."""
    import time

    import torch
    L = _lib.lib()
    dev = torch.device("cuda", device_id)
    t0 = time.perf_counter()
    x = vectors if isinstance(vectors, torch.Tensor) else torch.from_numpy(_f32(vectors))
    x = x.to(dev, dtype=torch.float32).contiguous()
    N, d = x.shape
    if d % M:
        raise ValueError(f"dimension {d} is not divisible by num_sub_vectors {M}")
    mcode = DistanceType.code(metric)
    xi = x
    if metric == "cosine":      # a cosine index stores L2 structures over normalised rows (ivf/builder.rs)
        xi = x / x.norm(dim=1, keepdim=True).clamp_min(1e-30)
    train_metric = 2 if metric == "dot" else 0
    g = torch.Generator(device="cpu").manual_seed(seed)
    t = {"upload_s": time.perf_counter() - t0}

    def ptr(tn):
        return C.c_void_p(tn.data_ptr())

    # ---- IVF centroids: k-means on sample_rate * nlist rows ----
    t0 = time.perf_counter()
    ns = min(N, sample_rate * nlist)
    samp = xi[torch.randperm(N, generator=g)[:ns].to(dev)].contiguous() if ns < N else xi
    cents = np.empty((nlist, d), dtype=np.float32)
    it, loss = C.c_uint32(0), C.c_double(0)
    init = None if init_centroids is None else _f32(init_centroids)
    check(L.lb200_kmeans_train_device(C.c_int32(device_id), C.c_int32(train_metric), C.c_uint32(d), ptr(samp), C.c_uint64(ns),
                                      C.c_uint32(nlist), C.c_uint32(max_iters), C.c_float(tolerance), C.c_uint64(seed), _p(init),
                                      _p(cents), C.byref(it), C.byref(loss)))
    t["ivf_train_s"], t["ivf_rounds"] = time.perf_counter() - t0, int(it.value)
    # ---- partition of every row ----
    t0 = time.perf_counter()
    d_cents = torch.from_numpy(cents).to(dev)
    part = torch.empty(N, dtype=torch.int32, device=dev)
    check(L.lb200_compute_partitions_device(C.c_int32(device_id), C.c_int32(train_metric), C.c_uint32(d), ptr(d_cents),
                                            C.c_uint32(nlist), ptr(xi), C.c_uint64(N), ptr(part), None, None))
    t["assign_s"] = time.perf_counter() - t0
    # ---- residuals (L2 / cosine), PQ codebook on a 256 * 256 row sample, codes ----
    t0 = time.perf_counter()
    resid = xi
    if metric != "dot":
        resid = torch.empty_like(xi)
        check(L.lb200_residuals_device(C.c_int32(device_id), C.c_uint32(d), ptr(xi), ptr(part), ptr(d_cents), C.c_uint64(N), ptr(resid), None))
    npq = min(N, sample_rate * 256)
    rs = resid[torch.randperm(N, generator=g)[:npq].to(dev)].contiguous() if npq < N else resid
    codebook = np.empty((M, 256, d // M), dtype=np.float32)
    icb = None if init_codebook is None else _f32(init_codebook)
    torch.cuda.synchronize(dev)
    check(L.lb200_pq_train_device(C.c_int32(device_id), C.c_int32(mcode), C.c_uint32(d), C.c_uint32(M), ptr(rs), C.c_uint64(npq),
                                  C.c_uint32(max_iters), C.c_float(tolerance), C.c_uint64(seed + 1000), _p(icb), _p(codebook)))
    t["pq_train_s"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    d_cb = torch.from_numpy(codebook).to(dev)
    codes = torch.empty((N, M), dtype=torch.uint8, device=dev)
    check(L.lb200_pq_encode_device(C.c_int32(device_id), C.c_int32(mcode), C.c_uint32(d), C.c_uint32(M), ptr(d_cb), ptr(resid),
                                   C.c_uint64(N), ptr(codes), None))
    torch.cuda.synchronize(dev)
    t["encode_s"] = time.perf_counter() - t0
    # ---- arrange by partition: codes transposed per partition [M][n_p] (pq/storage.rs:357-377), row ids ----
    t0 = time.perf_counter()
    del resid
    part64 = part.to(torch.int64)
    valid = part >= 0                                          # int32 view: None (0xffffffff) reads as -1
    order = torch.argsort(torch.where(valid, part64, torch.full_like(part64, nlist)), stable=True)
    nvalid = int(valid.sum().item())
    order = order[:nvalid]
    sizes = torch.bincount(part64[order], minlength=nlist)
    offs = torch.zeros(nlist + 1, dtype=torch.int64, device=dev)
    offs[1:] = torch.cumsum(sizes, 0)
    p_sorted = part64[order]
    local = torch.arange(nvalid, device=dev) - offs[p_sorted]
    flat = torch.empty(nvalid * M, dtype=torch.uint8, device=dev)
    base = offs[p_sorted] * M + local
    n_p = sizes[p_sorted]
    cs = codes[order]
    for m in range(M):                                          # dest = o_p * M + m * n_p + i
        flat[base + m * n_p] = cs[:, m]
    rid = order.cpu().numpy().astype(np.uint64)
    if row_ids is not None:
        rid = np.asarray(row_ids, dtype=np.uint64)[order.cpu().numpy()]
    raw = x[order].cpu().numpy() if keep_raw else None
    data = IvfPqData(metric, d, nlist, M, cents, codebook, offs.cpu().numpy().astype(np.uint64), flat.cpu().numpy(), rid, raw)
    t["arrange_s"] = time.perf_counter() - t0
    return data, t

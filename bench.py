#!/usr/bin/env python
"""bench.py -- IVF_PQ batched query throughput on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload cfg4|cfg1|cfg5|cfg3]

A "step" is one pass of the whole hot path (coarse probe selection -> ADC tables -> PQ code
scan -> top-k -> merge [-> refine]) over one batch of synthetic queries.  Default workload is
BASELINE.json's target config: bigann-shaped 100M x 128, IVF4096, PQ16, L2, k=10, nprobes=32,
10k queries (fits one B200).  With N>1 the IVF partitions are sharded over the ranks
(SURVEY 8e): each rank scans its partitions, one NCCL all_gather of the local top-k lists and a
device-side merge produce the result; total work is fixed, so scaling is "strong".

Prints ONE JSON line (rank 0).  `--impl reference` times the CPU oracle port of the reference
path (the Rust reference cannot be built in this image) on all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (N, dim, nlist, M, metric, nprobes, k, refine, Q, description)
    "cfg4": (100_000_000, 128, 4096, 16, "l2", 32, 10, 0, 10_000,
             "bigann-shaped 100Mx128 f32, IVF4096,PQ16, L2, k=10, nprobes=32, 10k queries (BASELINE configs[3])"),
    "cfg1": (1_000_000, 128, 256, 16, "l2", 20, 10, 0, 10_000,
             "SIFT-shaped 1Mx128 f32, IVF256,PQ16, L2, k=10, nprobes=20, 10k queries (BASELINE configs[0])"),
    "cfg5": (10_000_000, 768, 2048, 48, "dot", 32, 10, 0, 4096,
             "10Mx768 dot IVF2048,PQ48, k=10, nprobes=32, 4096 queries (BASELINE configs[4])"),
    "cfg3": (1_000_000, 1536, 1024, 96, "cosine", 102, 10, 5, 1000,
             "dbpedia-openai-shaped 1Mx1536 cosine IVF1024,PQ96, k=10, refine_factor=5, nprobes=102 (BASELINE configs[2])"),
    "tiny": (200_000, 128, 64, 16, "l2", 8, 10, 0, 512, "tiny CI workload"),
    # SURVEY 8(a) row A5'': the same shape as cfg1 with 4-bit PQ codes (not a BASELINE config)
    "cfg1_pq4": (1_000_000, 128, 256, 16, "l2", 20, 10, 0, 10_000,
                 "SIFT-shaped 1Mx128 f32, IVF256,PQ16 4-bit, L2, k=10, nprobes=20, 10k queries (cfg1 shape, num_bits=4)"),
}
WORKLOADS["cfg1_flat"] = (1_000_000, 128, 256, 16, "l2", 20, 10, 0, 2_000,
                          "SIFT-shaped 1Mx128 f32, IVF256 IVF_FLAT (raw rows, no quantizer), L2, k=10, nprobes=20, 2000 queries (cfg1 shape)")
PQ_BITS = {"cfg1_pq4": 4, "cfg1_flat": 32}
_BITS = 8  # set by main() from the workload


def partition_sizes(N, nlist, seed, sigma=0.5):
    rng = np.random.default_rng(seed)
    w = rng.lognormal(0.0, sigma, size=nlist)
    sizes = np.floor(w / w.sum() * N).astype(np.int64)
    sizes[: N - sizes.sum()] += 1
    return sizes


from lance_b200.shard import assign_partitions, probe_load_weights  # noqa: E402  (greedy bin-packing)


def make_shard(wl, seed, rank, world, keep_raw=False):
    """Synthetic index (SURVEY 8d cfg 4 recipe): log-normal partition sizes, uniform u8 codes
    (scan cost is value independent), seeded per partition so every rank builds only its shard."""
    from lance_b200.synth import IvfPqData
    N, d, nlist, M, metric = wl[0], wl[1], wl[2], wl[3], wl[4]
    ncent, cbytes = (16, M // 2) if _BITS == 4 else (256, M)
    sizes = partition_sizes(N, nlist, seed)
    rng = np.random.default_rng(seed + 1)
    if metric == "dot":
        cents = rng.normal(size=(nlist, d)).astype(np.float32)
        codebook = rng.normal(size=(M, ncent, d // M)).astype(np.float32)
    elif metric == "cosine":
        cents = rng.normal(size=(nlist, d)).astype(np.float32)
        cents /= np.linalg.norm(cents, axis=1, keepdims=True)
        codebook = (rng.normal(size=(M, ncent, d // M)) * 0.02).astype(np.float32)
    else:
        cents = rng.uniform(0, 128, size=(nlist, d)).astype(np.float32)
        codebook = rng.normal(0, 16, size=(M, ncent, d // M)).astype(np.float32)
    # partitions -> ranks: balanced on rows x expected probes (estimated from the centroids alone, shard.py)
    weights = None if (world == 1 or os.environ.get("LB200_BALANCE") == "rows") else probe_load_weights(cents, metric, wl[5])
    owner = assign_partitions(sizes, world, weights)
    goffs = np.zeros(nlist + 1, dtype=np.int64)
    goffs[1:] = np.cumsum(sizes)
    mine = owner == rank
    lsizes = sizes * mine
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(lsizes)
    n_local = int(offs[-1])
    codes = np.empty(n_local * cbytes, dtype=np.uint8)
    rids = np.empty(n_local, dtype=np.uint64)
    raw = np.empty((n_local, d), dtype=np.float32) if keep_raw else None
    for p in np.nonzero(mine)[0]:
        prng = np.random.default_rng([seed, int(p)])
        o, n = int(offs[p]), int(sizes[p])
        codes[o * cbytes:(o + n) * cbytes] = prng.integers(0, 256, size=n * cbytes, dtype=np.uint8)
        gpos = np.arange(goffs[p], goffs[p] + n, dtype=np.uint64)
        rids[o:o + n] = (gpos * np.uint64(2654435761)) % np.uint64(1 << 32)  # bijection on u32
        if keep_raw:
            raw[o:o + n] = cents[p] + prng.normal(0, 0.05 if metric == "cosine" else 8.0, size=(n, d)).astype(np.float32)
    data = IvfPqData(metric, d, nlist, M, cents, codebook, offs, codes, rids, raw, _BITS)
    return data, sizes, owner


def make_queries(wl, cents, seed):
    Q, metric = wl[8], wl[4]
    rng = np.random.default_rng(seed + 2)
    c = cents[rng.integers(0, cents.shape[0], size=Q)]
    s = 0.05 if metric == "cosine" else (0.5 if metric == "dot" else 16.0)
    return (c + rng.normal(0, s, size=c.shape)).astype(np.float32)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_run(wl, data, queries, nq, threads, keep=False):
    """The reference's CPU algorithm (oracle port) on a bounded sample of the workload.  keep=True also returns
    the oracle's outputs (dist [nq][k'], row ids, counts, probed partitions) for the parity block."""
    import oracle
    o = oracle.OracleIndex(data)
    k, nprobes, refine = wl[6], wl[5], wl[7]
    t0 = time.perf_counter()
    res = o.search(queries[:nq], k=k, nprobes=nprobes, refine_factor=refine, threads=threads)
    dt = time.perf_counter() - t0
    return (nq / dt, dt, res) if keep else (nq / dt, dt)


def parity_block(ref, gpu, k, gpu_probes=None, tol=1e-5):
    """Compare the GPU results of the benchmarked batch with the oracle's on the same queries of the same index
    (north star: identical probed partitions; row ids identical except distance ties within 1e-5 relative;
    distances within 1e-5 relative).  ref = (dist, rowid, count, probes) from the oracle, gpu = (dist, rowid, count)."""
    rd, ri, rc, rp = ref
    gd, gi, gc = gpu
    nq = rd.shape[0]
    rd, ri = rd[:, :k], ri[:, :k]
    gd, gi, gc = gd[:nq, :k], gi[:nq, :k].astype(np.uint64), gc[:nq].astype(np.int64)
    rc = np.minimum(rc.astype(np.int64), k)
    valid = np.arange(k)[None, :] < rc[:, None]
    out = {"queries_checked": int(nq), "tolerance_rel": tol,
           "count_mismatches": int(np.count_nonzero(gc != rc))}
    if gpu_probes is not None:
        gp = np.asarray(gpu_probes)[:nq]
        out["probe_set_mismatches"] = int(sum(set(a.tolist()) != set(b.tolist()) for a, b in zip(gp, rp)))
        out["probe_order_mismatches"] = int(np.count_nonzero((gp != rp).any(axis=1)))
    idm = (gi != ri) & valid
    scale = np.maximum(np.abs(rd), np.float32(1e-30))
    with np.errstate(invalid="ignore", over="ignore"):
        rel = np.where(valid, np.abs(gd.astype(np.float64) - rd.astype(np.float64)) / scale, 0.0)
    rel = np.nan_to_num(rel, nan=0.0, posinf=np.inf)
    # an id mismatch is an explained tie when the two rows' distances agree within the tolerance
    explained = idm & (rel <= tol)
    out["id_mismatches"] = int(np.count_nonzero(idm))
    out["explained_ties"] = int(np.count_nonzero(explained))
    out["unexplained_mismatches"] = int(np.count_nonzero(idm & ~explained)) + out["count_mismatches"]
    out["max_rel_err"] = float(rel.max()) if rel.size else 0.0
    out["distance_bits_identical"] = bool(np.array_equal(gd.view(np.uint32)[valid], rd.view(np.uint32)[valid]))
    out["checker"] = "oracle port (oracle/lance_oracle.c) on the same index and the first queries of the same batch"
    return out


def run_reference(args, wl):
    """--impl reference: the CPU path on the host cores, same config/metric/unit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    seed = 77
    data, sizes, _ = make_shard(wl, seed, 0, 1, keep_raw=wl[7] > 0)
    queries = make_queries(wl, data.centroids, seed)
    cores = oracle.max_threads()
    # bounded sample per step: sized from a probe so the whole run stays within a few minutes
    qps0, _ = cpu_reference_run(wl, data, queries, min(64, len(queries)), cores)
    budget_s = 120.0 / max(1, args.steps + args.warmup)
    nq = int(max(16, min(len(queries), qps0 * min(budget_s, 10.0))))
    for _ in range(args.warmup):
        cpu_reference_run(wl, data, queries, nq, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_run(wl, data, queries, nq, cores)
    dt = time.perf_counter() - t0
    qps = nq * args.steps / dt
    line = {
        "impl": "reference", "metric": "IVF_PQ queries/sec (batched)", "value": qps, "unit": "queries/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl[9], "queries_per_step": nq},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": "port",
                         "sample": f"{nq} queries per step of the same workload, queries spread over {cores} threads"},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def _result_crc(od, oi, oc):
    """CRC of the last batch's results (counts, row ids, distance bits of the valid entries): the same
    value at every N says the sharded result is bit-identical to the single-GPU one."""
    import zlib
    valid = np.arange(od.shape[1])[None, :] < oc.astype(np.int64)[:, None]
    c = zlib.crc32(np.ascontiguousarray(oc).tobytes())
    c = zlib.crc32(np.ascontiguousarray(oi[valid]).tobytes(), c)
    return zlib.crc32(np.ascontiguousarray(od[valid]).tobytes(), c)


def run_flat(args):
    """--workload cfg2: flat brute-force L2 kNN 1M x 128, k=10 (KNNVectorDistanceExec path, no index; BASELINE
    configs[1]).  A step = one batch of 256 queries against the device-resident column (lb200_column_*): tensor-core
    candidate filter (tcgen05 + TMA, csrc/flat_tc.cuh) + exact fp32 re-score, results bit-identical to the exact path."""
    import ctypes as C
    import torch
    from lance_b200 import _lib
    from lance_b200._lib import check
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    import torch.distributed as dist
    from lance_b200.index import merge_topk_device
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N, d, k, Q = 1_000_000, 128, 10, args.queries or 256
    fmetric = args.flat_metric          # "also run cosine and dot for coverage" (benchmarks/flat/benchmark.py:68)
    fcode = {"l2": 0, "cosine": 1, "dot": 2}[fmetric]
    rng = np.random.default_rng(1234)
    vecs = np.clip(np.rint(rng.normal(64, 32, size=(N, d))), 0, 218).astype(np.float32)  # SIFT-like integers
    queries = np.clip(np.rint(rng.normal(64, 32, size=(Q, d))), 0, 218).astype(np.float32)
    # multi-GPU ("flat kNN shards by rows the same way"): rank r keeps rows [lo, hi) resident, every rank
    # scores the whole query batch against its rows, one NCCL all_gather of the [Q][k] lists, merge on rank 0
    lo, hi = rank * N // world, (rank + 1) * N // world
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    L = _lib.lib()
    col = C.c_void_p()
    mine = np.ascontiguousarray(vecs[lo:hi])
    check(L.lb200_column_create(C.c_int32(local), C.c_uint32(d), mine.ctypes.data_as(C.c_void_p), None, C.c_uint64(hi - lo), C.byref(col)))
    d_q = torch.from_numpy(queries).to(dev)
    od = torch.empty((Q, k), dtype=torch.float32, device=dev)
    oi = torch.empty((Q, k), dtype=torch.int64, device=dev)
    oc = torch.empty((Q,), dtype=torch.int32, device=dev)
    if world > 1:
        g_d = torch.empty((world, Q, k), dtype=torch.float32, device=dev)
        g_i = torch.empty((world, Q, k), dtype=torch.int64, device=dev)
        g_c = torch.empty((world, Q), dtype=torch.int32, device=dev)
        l_d, l_i, l_c = torch.empty_like(od), torch.empty_like(oi), torch.empty_like(oc)
    h_q = torch.from_numpy(queries).pin_memory()
    h_od = torch.empty((Q, k), dtype=torch.float32).pin_memory()
    h_oi = torch.empty((Q, k), dtype=torch.int64).pin_memory()
    h_oc = torch.empty((Q,), dtype=torch.int32).pin_memory()

    def step():
        if world == 1:
            check(L.lb200_column_flat_topk_device(col, C.c_int32(fcode), C.c_void_p(d_q.data_ptr()), C.c_uint32(Q), C.c_uint32(k), None, None,
                                                  C.c_void_p(od.data_ptr()), C.c_void_p(oi.data_ptr()), C.c_void_p(oc.data_ptr()),
                                                  C.c_void_p(stream.cuda_stream)))
            return
        check(L.lb200_column_flat_topk_device(col, C.c_int32(fcode), C.c_void_p(d_q.data_ptr()), C.c_uint32(Q), C.c_uint32(k), None, None,
                                              C.c_void_p(l_d.data_ptr()), C.c_void_p(l_i.data_ptr()), C.c_void_p(l_c.data_ptr()),
                                              C.c_void_p(stream.cuda_stream)))
        l_i.add_(lo)                                       # local row index -> row index in the whole column
        dist.all_gather_into_tensor(g_d.view(-1), l_d.view(-1))
        dist.all_gather_into_tensor(g_i.view(-1), l_i.view(-1))
        dist.all_gather_into_tensor(g_c.view(-1), l_c.view(-1))
        if rank == 0:
            merge_topk_device(local, g_d.data_ptr(), g_i.data_ptr(), None, g_c.data_ptr(), world, Q, k, k, k,
                              od.data_ptr(), oi.data_ptr(), oc.data_ptr(), stream=stream.cuda_stream)

    def e2e_step():  # host queries in, host results out (H2D + D2H inside the timed region)
        if world == 1:
            check(L.lb200_column_flat_topk(col, C.c_int32(fcode), C.c_void_p(h_q.data_ptr()), C.c_uint32(Q), C.c_uint32(k), None, None,
                                           C.c_void_p(h_od.data_ptr()), C.c_void_p(h_oi.data_ptr()), C.c_void_p(h_oc.data_ptr())))
            return
        d_q.copy_(h_q, non_blocking=True)
        step()
        if rank == 0:
            h_od.copy_(od, non_blocking=True); h_oi.copy_(oi, non_blocking=True); h_oc.copy_(oc, non_blocking=True)
        stream.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        ms = float(tmax.item())
    clocks = sampler.stop()
    need_repeat = not clocks.get("samples")
    if world > 1:                   # the same decision on every rank: the repeat holds collectives
        flag = torch.tensor([1 if need_repeat else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        need_repeat = bool(flag.item())
    if need_repeat:   # a 10 ms timed region is shorter than nvidia-smi's sampling period: sample a repeat of it
        sampler = ClockSampler(local)
        sampler.start()
        for _ in range(30 if world > 1 else 1):            # (the same count on every rank: step() holds collectives)
            t0 = time.perf_counter()
            while world == 1 and time.perf_counter() - t0 < 1.5:
                for _ in range(20):
                    step()
                torch.cuda.synchronize()
            if world > 1:
                for _ in range(100):
                    step()
                torch.cuda.synchronize()
        clocks = sampler.stop()
        clocks["note"] = "sampled over a 1.5 s repeat of the same steps right after the timed region (which is too short to sample)"
    tc, cand, fb = C.c_uint32(0), C.c_uint64(0), C.c_uint32(0)
    check(L.lb200_column_last_call(col, C.byref(tc), C.byref(cand), C.byref(fb)))
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    ms_e2e = (time.perf_counter() - t0) * 1e3
    # the exact path alone on the same column, for the line's own A/B
    os.environ["LB200_FLAT_NO_TC"] = "1"
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    ms_exact = (time.perf_counter() - t0) * 1e3 / 3
    del os.environ["LB200_FLAT_NO_TC"]
    step()
    torch.cuda.synchronize()
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peaks = json.load(open(peaks_path)) if os.path.exists(peaks_path) else {}
    peak_hbm = float(peaks.get("hbm_gbs", 6650.0))
    qps = Q * args.steps / (ms * 1e-3)
    step_ms = ms / args.steps
    # the filter kernel streams the bf16 rows once per 128-query block: its algorithmic bytes
    qblocks = (Q + 127) // 128
    bytes_step = N * d * 2 * qblocks
    line = {"metric": "flat kNN queries/sec (batched)", "value": qps, "unit": "queries/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (results); bf16 tensor-core candidate filter + exact f32 re-score",
            "data": "synthetic",
            "config": {"workload": f"flat brute-force {fmetric.upper()} kNN 1Mx128 f32, k=10, {Q} queries per step (BASELINE configs[1]"
                                   + ("" if fmetric == "l2" else "; metric varied as the reference's flat benchmark does") + ")",
                       "l2_policy": "inputs larger than L2 (256 MB of bf16 rows + 512 MB of f32 rows)",
                       "path": "lb200_column_flat_topk_device (resident column)" + ("" if world == 1 else
                               f"; rows sharded over {world} GPUs, NCCL all_gather of the [Q][k] lists, lb200_merge_topk_device on rank 0"),
                       "filter": {"used_tc": bool(tc.value), "fell_back": bool(fb.value), "candidates_per_query": cand.value / Q},
                       "exact_path_ms_per_step": ms_exact},
            "roofline": {"bound": "hbm", "achieved": bytes_step / (step_ms * 1e-3) / 1e9, "peak": peak_hbm, "unit": "GB/s",
                         "frac": bytes_step / (step_ms * 1e-3) / 1e9 / peak_hbm, "traffic": None,
                         "kernel": "whole step (coarse_dist_l2_slab_kernel over the 16k-row sample, ftc_kth_kernel, flat_tc_filter_kernel, ftc_rescore_l2_kernel, ftc_prune_kernel, merge_lists_kernel)",
                         "note": "bytes = N*d*2 (bf16 rows) per 128-query block of the filter kernel; flops = Q*N*d*2 on tcgen05"},
            "e2e": {"value": Q * e2e_steps / (ms_e2e * 1e-3), "unit": "queries/s", "h2d_bytes_per_step": int(queries.nbytes),
                    "d2h_bytes_per_step": Q * k * 12 + Q * 4, "ms_per_step": ms_e2e / e2e_steps,
                    "path": "lb200_column_flat_topk (host-pointer C ABI on the resident column; queries H2D, results D2H inside)"
                            if world == 1 else "pinned H2D of the queries on every rank + the sharded step + D2H of the merged result on rank 0"},
            "gpu_launches": 8 * args.steps,  # sample distances, k-th select, query bf16, filter, clamp, exact re-score, prune, merge
            "clocks": clocks}
    if rank != 0:
        L.lb200_column_destroy(col)
        dist.destroy_process_group()
        return
    if not args.no_cpu_baseline:
        import oracle
        cores = oracle.max_threads()
        nq = min(32, Q)
        t0 = time.perf_counter()
        fres = oracle.flat_knn(fmetric, queries[:nq], vecs, k, threads=cores)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": nq / dt, "unit": "queries/s", "cores": cores, "kind": "port",
                                "sample": f"first {nq} queries, {dt:.1f} s, oracle port on {cores} threads"}
        torch.cuda.synchronize()
        line["parity"] = parity_block((fres[0], fres[1], fres[2], None),
                                      (od.cpu().numpy(), oi.cpu().numpy().view(np.uint64), oc.cpu().numpy()), k)
    print(json.dumps(line), flush=True)
    L.lb200_column_destroy(col)
    if world > 1:
        dist.destroy_process_group()


def clustered_lowrank(N, d, seed, dev, nblobs=4096, rank=16, chunk=2_000_000):
    """Synthetic vectors WITH neighbourhood structure, generated on the device: blob centres U[0,128)^d, inside a blob a
    `rank`-dimensional Gaussian cloud (one random basis shared by all blobs, sigma 12) plus a little isotropic noise
    (sigma 1).  Isotropic blobs (SURVEY 8(d)'s first recipe) make all neighbours of a 128-d point nearly equidistant,
    so recall on them says nothing (r01: 0.17); here a point's neighbours differ clearly in distance, as in real data."""
    import torch
    g = torch.Generator(device=dev).manual_seed(1234)              # the structure (centres, basis) is fixed
    centres = torch.rand((nblobs, d), generator=g, device=dev) * 128.0
    basis = torch.randn((rank, d), generator=g, device=dev) / (rank ** 0.5)
    g2 = torch.Generator(device=dev).manual_seed(seed)             # the draws depend on the seed
    out = torch.empty((N, d), dtype=torch.float32, device=dev)
    for s in range(0, N, chunk):
        n = min(chunk, N - s)
        which = torch.randint(0, nblobs, (n,), generator=g2, device=dev)
        z = torch.randn((n, rank), generator=g2, device=dev) * 12.0
        out[s:s + n] = centres[which] + z @ basis + torch.randn((n, d), generator=g2, device=dev)
    return out


def recall_block(wl, local, max_rows=10_000_000, nq=1000):
    """recall@10 of the path on a REAL-BUILT index (the metric is "queries/sec @ recall@10"; pattern of
    python/python/lance/torch/bench_utils.py:50-133): vectors with neighbourhood structure -> GPU index build
    (lance_b200/build.py) with the workload's nlist / M / metric -> search with the workload's nprobes / k / refine ->
    exact ground truth from the flat kernel (lb200_flat_topk_device) on the same vectors."""
    import ctypes as C
    import torch
    from lance_b200 import IvfPqIndex, Query, _lib
    from lance_b200.build import build_ivf_pq_gpu
    N, d, nlist, M, metric, nprobes, k, refine, Q, desc = wl
    dev = torch.device("cuda", local)
    n = min(N, max_rows)
    t0 = time.perf_counter()
    x = clustered_lowrank(n, d, 99, dev)
    q = clustered_lowrank(nq, d, 4321, dev)
    if metric == "dot":   # inner-product search is used on embedding-like rows: zero mean, unit length
        for t_ in (x, q):
            t_.sub_(64.0)
            t_.div_(t_.norm(dim=1, keepdim=True).clamp_min(1e-30))
    torch.cuda.synchronize(dev)
    t_gen = time.perf_counter() - t0
    data, tb = build_ivf_pq_gpu(x, metric, nlist, M, seed=7, max_iters=10, keep_raw=refine > 0, device_id=local)
    ix = IvfPqIndex(data, device_id=local)
    hq = q.cpu().numpy()
    qy = Query(key=hq, k=k, nprobes=nprobes, refine_factor=refine or None)
    gd, gi, gc = ix.search(qy)
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        ix.search(qy)
    qps = nq * reps / (time.perf_counter() - t0)
    # exact ground truth with the flat kernel, queries in tiles that keep the [tile][n] distance block small
    L = _lib.lib()
    od = torch.empty((nq, k), dtype=torch.float32, device=dev)
    oi = torch.empty((nq, k), dtype=torch.int64, device=dev)
    oc = torch.empty((nq,), dtype=torch.int32, device=dev)
    mcode = {"l2": 0, "cosine": 1, "dot": 2}[metric]
    t0 = time.perf_counter()
    _lib.check(L.lb200_flat_topk_device(C.c_int32(local), C.c_int32(mcode), C.c_uint32(d), C.c_void_p(q.data_ptr()), C.c_uint32(nq),
                                        C.c_void_p(x.data_ptr()), None, C.c_uint64(n), C.c_uint32(k), C.c_void_p(od.data_ptr()),
                                        C.c_void_p(oi.data_ptr()), C.c_void_p(oc.data_ptr()), None))
    torch.cuda.synchronize(dev)
    t_truth = time.perf_counter() - t0
    truth = oi.cpu().numpy()
    hits = [len(set(gi[i, :int(gc[i])].tolist()) & set(truth[i].tolist())) for i in range(nq)]
    # a parity spot check on this index too: the oracle on the first queries
    import oracle
    ns = 64
    rd, ri, rc, _ = oracle.OracleIndex(data).search(hq[:ns], k=k, nprobes=nprobes, refine_factor=refine, threads=oracle.max_threads())
    par = parity_block((rd, ri, rc, None), (gd, gi, gc), k)
    sizes = np.diff(data.part_offsets.astype(np.int64))
    ix.close()
    return {"recall_at_10": float(np.mean(hits)) / k, "queries": nq, "rows": n, "k": k, "nprobes": nprobes, "refine_factor": refine,
            "index": f"real-built on the GPU (lance_b200/build.py): IVF{nlist},PQ{M}, {metric}, k-means <= 10 rounds, "
                     f"partition sizes {int(sizes.min())}..{int(sizes.max())}",
            "data": "clustered low-rank synthetic (4096 blobs, 16-dim clouds + noise" + ("; centred and unit-normalised for dot" if metric == "dot" else "") + "), queries = held-out draws",
            "ground_truth": "exact kNN over the same rows with lb200_flat_topk_device",
            "qps_on_this_index": qps, "parity_vs_oracle": {"queries_checked": ns, "unexplained_mismatches": par["unexplained_mismatches"],
                                                            "distance_bits_identical": par["distance_bits_identical"]},
            "seconds": {"generate": round(t_gen, 2), "build": {kk: (round(v, 2) if isinstance(v, float) else v) for kk, v in tb.items()},
                        "ground_truth": round(t_truth, 2)}}


class _DropinBackend(__import__("ctypes").Structure):
    import ctypes as _C
    _fields_ = [("kind", _C.c_int32), ("handle", _C.c_void_p), ("fn_find", _C.c_void_p), ("fn_search", _C.c_void_p),
                ("metric", _C.c_int32), ("dim", _C.c_uint32), ("nlist", _C.c_uint32), ("centroids", _C.c_void_p)]


def _dropin_lib():
    import ctypes as C
    path = os.path.join(ROOT, "examples", "libdropin_host.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "examples")])
    lib = C.CDLL(path)
    lib.dropin_run.restype = C.c_int
    return lib


def _dropin_call(lib, be, queries, nprobes, k, threads, clients):
    import ctypes as C
    Q = queries.shape[0]
    od = np.full((Q, k), np.inf, dtype=np.float32)
    oi = np.full((Q, k), np.iinfo(np.uint64).max, dtype=np.uint64)
    oc = np.zeros(Q, dtype=np.uint32)
    lat = np.zeros(Q, dtype=np.float64)
    wall = C.c_double(0)
    q = np.ascontiguousarray(queries, dtype=np.float32)
    rc = lib.dropin_run(C.byref(be), q.ctypes.data_as(C.c_void_p), C.c_uint32(Q), C.c_uint32(nprobes), C.c_uint32(k),
                        C.c_int(threads), C.c_int(clients), od.ctypes.data_as(C.c_void_p), oi.ctypes.data_as(C.c_void_p),
                        oc.ctypes.data_as(C.c_void_p), lat.ctypes.data_as(C.c_void_p), C.byref(wall))
    if rc:
        from lance_b200 import _lib
        raise RuntimeError(f"dropin_run failed: {rc}: {_lib.lib().lb200_last_error().decode()}")
    return od, oi, oc, lat, wall.value


def run_dropin(args, wl):
    """--mode dropin: the call pattern of the reference's exec nodes (knn.rs:349-393,606-634) through the C ABI:
    per query ONE lb200_find_partitions(Q=1), then nprobes x lb200_search_partitions(Q=1,P=1) with T host threads in
    flight, then a host-side merge -- driven by examples/dropin_host.cpp -- and the CPU oracle port driven the SAME
    way beside it (the "reference-shaped" baseline of SURVEY 8(d): partitions of one query fanned out over
    T = nproc - 2 threads, lance-core/src/utils/tokio.rs:12-31)."""
    import ctypes as C
    import torch
    import oracle
    from lance_b200 import IvfPqIndex, Query, _lib
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    N, d, nlist, M, metric, nprobes, k, refine, Q, desc = wl
    seed = 77
    data, sizes, _ = make_shard(wl, seed, 0, 1, keep_raw=False)
    queries = make_queries(wl, data.centroids, seed)
    ix = IvfPqIndex(data, device_id=0)
    L = _lib.lib()
    lib = _dropin_lib()
    nproc = os.cpu_count() or 1
    T = max(1, nproc - 2)                                   # get_num_compute_intensive_cpus()
    mcode = {"l2": 0, "cosine": 1, "dot": 2}[metric]
    gbe = _DropinBackend(0, ix._h, C.cast(L.lb200_find_partitions, C.c_void_p), C.cast(L.lb200_search_partitions, C.c_void_p),
                         mcode, d, nlist, None)
    nq = min(Q, args.queries or 2000)
    kq = k * max(refine, 1)
    _dropin_call(lib, gbe, queries[:64], nprobes, kq, T, 1)                     # warm-up (contexts, buffers)
    sampler = ClockSampler(0)
    sampler.start()
    gd, gi, gc, glat, gwall = _dropin_call(lib, gbe, queries[:nq], nprobes, kq, T, 1)
    clocks = sampler.stop()
    st = ix.stats()
    per_call_launches = st["kernel_launches"]
    # several queries in flight (concurrent users of one index handle)
    clients = min(8, T)
    _, _, _, _, gwall_c = _dropin_call(lib, gbe, queries[:nq], nprobes, kq, T, clients)
    # the batched entry point with Q = 1 beside it (host buffers, blocking)
    one = np.empty(min(nq, 500))
    for i in range(-20, len(one)):
        t0 = time.perf_counter()
        ix.search(Query(key=queries[max(i, 0)], k=k, nprobes=nprobes, refine_factor=None))
        if i >= 0:
            one[i] = (time.perf_counter() - t0) * 1e3
    # the CPU oracle port driven the same way, bounded sample
    o = oracle.OracleIndex(data)
    olib = oracle.lib()
    obe = _DropinBackend(1, C.addressof(o.ix), C.cast(olib.lo_find_partitions, C.c_void_p),
                         C.cast(olib.lo_search_in_partition, C.c_void_p), mcode, d, nlist,
                         data.centroids.ctypes.data_as(C.c_void_p))
    _, _, _, l0, w0 = _dropin_call(lib, obe, queries[:8], nprobes, kq, T, 1)
    ncpu = int(max(16, min(nq, 12.0 / max(w0 / 8, 1e-6))))
    cd, ci, cc, clat, cwall = _dropin_call(lib, obe, queries[:ncpu], nprobes, kq, T, 1)
    par = parity_block((cd, ci, cc, None), (gd, gi, gc), kq)
    par["checker"] = "oracle port driven through the same drop-in host (examples/dropin_host.cpp), same queries"
    pct = lambda a, p: float(np.percentile(a, p))
    line = {
        "metric": "IVF_PQ queries/sec (drop-in call pattern: find_partitions + nprobes x search_in_partition)",
        "mode": "dropin", "value": nq / gwall, "unit": "queries/s", "n_gpus": 1, "steps": 1, "warmup": 1,
        "ms_per_step": gwall * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": desc, "queries": nq, "k": k, "nprobes": nprobes, "threads_in_flight": T,
                   "host": "examples/dropin_host.cpp (std::thread pool; mirrors knn.rs:349-393,606-634)",
                   "calls_per_query": 1 + nprobes},
        "latency_ms": {"p50": pct(glat, 50), "p99": pct(glat, 99), "mean": float(glat.mean()),
                       "lb200_search_Q1_p50": pct(one, 50), "lb200_search_Q1_p99": pct(one, 99),
                       "ratio_p50_vs_lb200_search_Q1": pct(glat, 50) / pct(one, 50)},
        "concurrent_clients": {"clients": clients, "value": nq / gwall_c, "unit": "queries/s"},
        "e2e": {"value": nq / gwall, "unit": "queries/s", "h2d_bytes_per_step": int(nq * (1 + nprobes) * d * 4),
                "d2h_bytes_per_step": int(nq * (nprobes * 4 + nprobes * (kq * 12 + 4))),
                "path": "lb200_find_partitions(Q=1) + nprobes x lb200_search_partitions(Q=1,P=1), host buffers"},
        "gpu_launches": int(nq * (2 + nprobes * per_call_launches)),
        "cpu_baseline": {"value": ncpu / cwall, "unit": "queries/s", "cores": T, "kind": "port",
                         "sample": f"first {ncpu} queries, {cwall:.1f} s; reference-shaped: one query at a time, its "
                                   f"{nprobes} partitions fanned out over {T} threads (nproc {nproc} - 2)",
                         "latency_ms": {"p50": pct(clat, 50), "p99": pct(clat, 99)}},
        "parity": par, "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    ix.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("LB200_WORKLOAD", "cfg4"), choices=sorted(WORKLOADS) + ["cfg2"])
    ap.add_argument("--mode", default="batch", choices=["batch", "dropin"],
                    help="dropin: the exec nodes' call pattern (per query find_partitions + nprobes x search_in_partition)")
    ap.add_argument("--scan-variant", type=int, default=0)
    ap.add_argument("--queries", type=int, default=0,
                    help="override the workload's queries per step (batch-size sweep of BASELINE configs[4])")
    ap.add_argument("--nprobes", type=int, default=0, help="override the workload's nprobes (SURVEY 8: cfg 3 is run at 1, 20 and 102)")
    ap.add_argument("--flat-metric", default="l2", choices=["l2", "cosine", "dot"], help="metric of --workload cfg2")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-recall", action="store_true", help="skip the real-built-index recall@10 block")
    ap.add_argument("--shard-of", type=int, default=0,
                    help="diagnostic: single process holding rank 0's shard of an N-way partition split")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.workload == "cfg2":
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": "cfg2 reference arm: use --workload cfg4 (headline)"}))
            return
        run_flat(args)
        return
    wl = WORKLOADS[args.workload]
    global _BITS
    _BITS = PQ_BITS.get(args.workload, 8)
    if args.queries:
        wl = wl[:8] + (args.queries, wl[9].replace("%d queries" % wl[8], "%d queries" % args.queries)
                       .replace("10k queries", "%d queries" % args.queries))
    if args.nprobes:
        wl = wl[:5] + (args.nprobes,) + wl[6:9] + (wl[9] + f" [nprobes overridden: {args.nprobes}]",)
    if args.mode == "dropin":
        if args.impl == "reference" or _BITS != 8:
            print(json.dumps({"impl": args.impl, "unavailable": "--mode dropin runs both arms itself (8-bit IVF_PQ only)"}))
            return
        run_dropin(args, wl)
        return
    if args.impl == "reference":
        if _BITS == 32:
            print(json.dumps({"impl": "reference", "unavailable": "cfg1_flat has no oracle arm: use --workload cfg4 (headline)"}))
            return
        run_reference(args, wl)
        return

    import torch
    import torch.distributed as dist
    from lance_b200 import IvfPqIndex
    from lance_b200.index import merge_topk_device

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    seed = 77
    N, d, nlist, M, metric, nprobes, k, refine, Q, desc = wl
    kprime = k * max(refine, 1)
    t_build = time.perf_counter()
    data, sizes, owner = make_shard(wl, seed, rank, args.shard_of or world, keep_raw=refine > 0 or _BITS == 32)
    queries = make_queries(wl, data.centroids, seed)
    ix = IvfPqIndex(data, device_id=local)
    if args.scan_variant:
        ix.set_scan_variant(args.scan_variant)
    t_build = time.perf_counter() - t_build

    stream = torch.cuda.Stream(device=dev)  # everything (our kernels, copies, NCCL) is ordered on it
    torch.cuda.set_stream(stream)
    sp = stream.cuda_stream
    d_q = torch.from_numpy(queries).to(dev)
    shard_mode = world > 1
    ks = kprime if shard_mode else (k if refine else kprime)
    d_od = torch.empty((Q, ks), dtype=torch.float32, device=dev)
    d_oi = torch.empty((Q, ks), dtype=torch.int64, device=dev)
    d_oc = torch.empty((Q,), dtype=torch.int32, device=dev)
    d_ex = torch.empty((Q, ks), dtype=torch.float32, device=dev) if shard_mode else None
    peer_exchange = shard_mode and os.environ.get("LB200_EXCHANGE", "peer") == "peer"
    if shard_mode and peer_exchange:
        # the whole exchange inside the library (lb200_shard_*): NVLink peer stores + epoch flags, no collective
        from lance_b200.shard import PeerShardedIvfPq
        sh = PeerShardedIvfPq(ix, dist, dev)
        sh.prepare(Q, nprobes, k, refine)
        f_od, f_oi, f_oc = sh.fd, sh.fi, sh.fc
    elif shard_mode:
        from lance_b200.shard import ShardedIvfPq
        sh = ShardedIvfPq(ix, dist, dev, shard_coarse=not os.environ.get("LB200_REPLICATED_COARSE"))
        sh.prepare(Q, nprobes, k, refine)
        f_od, f_oi, f_oc = sh.fd, sh.fi, sh.fc
    h_q = torch.from_numpy(queries).pin_memory()
    h_od = torch.empty((Q, k if (refine or shard_mode) else kprime), dtype=torch.float32).pin_memory()
    h_oi = torch.empty(h_od.shape, dtype=torch.int64).pin_memory()
    h_oc = torch.empty((Q,), dtype=torch.int32).pin_memory()

    def device_step():
        if shard_mode:
            sh.search_device(d_q, stream)
        else:
            ix.search_device(d_q.data_ptr(), Q, nprobes, k, refine, d_od.data_ptr(), d_oi.data_ptr(), d_oc.data_ptr(),
                             stream=sp)

    from lance_b200 import Query
    hq_np = h_q.numpy()
    e2e_out = (h_od.numpy(), h_oi.numpy().view(np.uint64), h_oc.numpy().view(np.uint32))

    def e2e_host_abi_step():
        # the call a Lance host makes: lb200_search with HOST buffers (pinned), blocking; the H2D of
        # the queries and the D2H of the results happen inside the library call
        ix.search(Query(key=hq_np, k=k, nprobes=nprobes, refine_factor=refine or None), out=e2e_out)

    def e2e_step():
        if not shard_mode:
            return e2e_host_abi_step()
        if peer_exchange:   # lb200_shard_search: host buffers in, host results out (rank 0), blocking
            sh.search_host(hq_np, e2e_out)
            return
        sh.search_host(h_q, stream)
        if rank == 0:
            src = (f_od, f_oi, f_oc) if shard_mode else (d_od, d_oi, d_oc)
            h_od.copy_(src[0], non_blocking=True)
            h_oi.copy_(src[1], non_blocking=True)
            h_oc.copy_(src[2], non_blocking=True)
        stream.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(args.warmup):
        device_step()
    barrier()
    ix.set_profiling(True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    scan_ms, scan_bytes, launches = [], [], 0

    def profiled_step():
        nonlocal launches
        device_step()
        st = ix.stats()
        scan_ms.append(st["scan_ms"])
        scan_bytes.append(st["scan_code_bytes"])
        if shard_mode and peer_exchange:   # + coarse (2) + publish + signal/wait x2 (+ root: wait + merge)
            launches += st["kernel_launches"] + 6 + (2 if rank == 0 else 0)
        else:
            launches += st["kernel_launches"] + ((3 if sh.shard_coarse else 0) + (1 if rank == 0 else 0) if shard_mode else 0)

    ms_total = timed(profiled_step, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    st = ix.stats()
    ix.set_profiling(False)
    for _ in range(2):
        e2e_step()
    e2e_steps = max(3, min(args.steps, 10))
    if shard_mode:
        ms_e2e = timed(e2e_step, e2e_steps)
    else:
        # blocking host call: host wall clock between two device syncs is the end-to-end time
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        ms_e2e = (time.perf_counter() - t0) * 1e3

    # scan bytes / time across ranks (sum of bytes, max of time): whole-job scan bandwidth
    agg = torch.tensor([float(np.mean(scan_ms)), float(np.mean(scan_bytes))], device=dev, dtype=torch.float64)
    if world > 1:
        tmax = agg[:1].clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        bsum = agg[1:].clone()
        dist.all_reduce(bsum, op=dist.ReduceOp.SUM)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    else:
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    my_scan_ms = float(np.mean(scan_ms))
    my_bytes = float(np.mean(scan_bytes))
    achieved = my_bytes / (my_scan_ms * 1e-3) / 1e9 if my_scan_ms > 0 else 0.0
    # which instantiation ran (the library's rule, api_search.inl): pre-built tables or mostly-cold items
    # (shards, few probes) -> the BOOT=true instantiation of scan3_kernel; otherwise the call-free BOOT=false one
    avg_np = float(np.mean(sizes[sizes > 0])) if len(sizes) else 0.0
    pre = metric == "dot" or 256.0 * (d // M) > avg_np
    if os.environ.get("LB200_LUT_PREPASS"):
        pre = os.environ["LB200_LUT_PREPASS"] != "0"
    pre = pre and (d // M) in (8, 16)
    kname = "scan_kernel"
    if st["scan_variant"] == 4:
        kname = ("scan4_kernel + flat_select_kernel" if os.environ.get("LB200_PQ4_SIMPLE") else "scan4f_kernel") + " (4-bit PQ)"
    if st["scan_variant"] == 32:
        kname = "scanflat_kernel + flat_select_kernel (IVF_FLAT; bytes = n_p*dim*4 per pair)"
    if st["scan_variant"] == 3:
        kname = "scan3_kernel<BOOT=%s>" % ("true" if (shard_mode or nprobes < 16) else "false")
        if pre:
            kname += " + lut_prepass_kernel"
    traffic, traffic_src, lsu_pct = None, None, None
    tpath = os.path.join(ROOT, "profiles", "ncu_scan_traffic.json")
    if os.path.exists(tpath):
        for rec in json.load(open(tpath)):
            if rec["workload"] == args.workload and rec["scan_variant"] == st["scan_variant"] and rec["n_gpus"] == world:
                traffic, traffic_src = rec["dram_bytes_per_launch"] / 1e9, rec["source"]
                lsu_pct = rec.get("lsu_pipe_pct")
    qps = Q * args.steps / (ms_total * 1e-3)
    e2e_qps = Q * e2e_steps / (ms_e2e * 1e-3)
    line = {
        "metric": "IVF_PQ queries/sec (batched)", "value": qps, "unit": "queries/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "queries_per_step": Q, "k": k, "nprobes": nprobes, "refine_factor": refine,
                   "parallelism": f"ivf-partitions sharded over {world} GPU(s)",
                   "exchange": ("peer memory inside liblance_b200.so (lb200_shard_*), 0 collectives per batch" if peer_exchange
                                else "NCCL all_gather from the host") if shard_mode else None,
                   "l2_policy": "inputs larger than L2 (PQ codes %.2f GB per GPU)" % (data.codes.nbytes / 1e9),
                   "scan_variant": st["scan_variant"], "scan_queries_per_cta": st["scan_ql"],
                   "index_build_s": round(t_build, 1)},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_unit": "GB per launch (ncu dram read+write)",
                     # the physical figures beside the algorithmic one.  dram_frac = ncu DRAM bytes of one
                     # launch / the live kernel time / peak; lsu_frac = the SM load/store data pipe (shared-memory table
                     # gather + code stream through L1), the unit that actually saturates first (ncu, same capture)
                     "dram_frac": (traffic * 1e9 / (my_scan_ms * 1e-3) / 1e9 / peak) if (traffic and my_scan_ms > 0) else None,
                     "lsu_frac": (lsu_pct / 100.0) if lsu_pct is not None else None,
                     "traffic_source": traffic_src, "kernel": kname + " (rank 0)",
                     "peak_source": peak_src, "scan_ms_per_launch": my_scan_ms,
                     "algorithmic_bytes_per_launch": my_bytes,
                     "note": "algorithmic bytes = sum over (query,partition) pairs of n_p*M code bytes; QL queries "
                             "share each code load, so the figure can exceed DRAM traffic (SURVEY 8d)"},
        "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes_per_step": int(h_q.numel() * 4),
                "d2h_bytes_per_step": int(h_od.numel() * 4 + h_oi.numel() * 8 + h_oc.numel() * 4),
                "ms_per_step": ms_e2e / e2e_steps,
                "path": ("lb200_shard_search on every rank (host buffers; own Q/N slice over PCIe, peers' slices and probe lists over NVLink peer memory, lists into rank 0's gather buffer, merge, D2H on rank 0; no collective)" if peer_exchange else
                         "per-rank pinned H2D of Q/N queries + NCCL all_gather of queries + sharded search + NCCL all_gather + merge, D2H on rank 0") if shard_mode else
                        "lb200_search (host-pointer C ABI, pinned host buffers, blocking; host wall clock)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "result_crc32": _result_crc(h_od.numpy(), h_oi.numpy(), h_oc.numpy()),
    }
    if world > 1:
        line["roofline"]["all_ranks_achieved"] = float(bsum.item()) / (float(tmax.item()) * 1e-3) / 1e9
    if world == 1 and not args.no_cpu_baseline and _BITS != 32:  # the C oracle index has no IVF_FLAT mode (tests use oracle.ivf_flat_search)
        import oracle
        cores = oracle.max_threads()
        qps0, _ = cpu_reference_run(wl, data, queries, min(32, Q), cores)
        nq = int(max(32, min(Q, qps0 * 15.0)))
        cqps, cdt, ref = cpu_reference_run(wl, data, queries, nq, cores, keep=True)
        line["cpu_baseline"] = {"value": cqps, "unit": "queries/s", "cores": cores, "kind": "port",
                                "sample": f"first {nq} queries of the same batch, {cdt:.1f} s, queries spread over "
                                          f"{cores} threads (oracle port of the Rust reference)",
                                "note": "kinder to the CPU than the reference: heap top-k per partition where "
                                        "flat/index.rs:124 sorts all n_p candidates; queries spread over all threads"}
        # parity at the benchmarked size: the oracle's outputs for those queries against the GPU results of the
        # same batch (the last lb200_search call above), plus the probed partitions through lb200_find_partitions
        from lance_b200 import normalize as _normalize
        keys = _normalize(queries[:nq], device_id=local) if metric == "cosine" else queries[:nq]
        gp = ix.find_partitions(keys, nprobes)
        line["parity"] = parity_block(ref, (h_od.numpy(), h_oi.numpy().view(np.uint64), h_oc.numpy()), k, gpu_probes=gp)
    if world == 1 and not args.no_cpu_baseline and _BITS == 32:
        # IVF_FLAT: the oracle's per-query restatement (oracle.ivf_flat_search: find_partitions + flat kNN per probed
        # partition, Python loop over the partitions, C distance kernels), a small sample on one thread
        import oracle
        nq = min(64, Q)
        rd = np.full((nq, k), np.inf, dtype=np.float32)
        ri = np.full((nq, k), np.iinfo(np.uint64).max, dtype=np.uint64)
        rc = np.zeros(nq, dtype=np.uint32)
        rp = np.zeros((nq, min(nprobes, nlist)), dtype=np.uint32)
        t0 = time.perf_counter()
        for j in range(nq):
            d_, r_ = oracle.ivf_flat_search(data, queries[j], k, nprobes, refine or None)
            rd[j, :len(d_)], ri[j, :len(r_)], rc[j] = d_, r_, len(r_)
        cdt = time.perf_counter() - t0
        for j in range(nq):
            key = oracle.normalize(queries[j]) if metric == "cosine" else queries[j]
            rp[j] = oracle.find_partitions(data.centroids, key, nprobes, metric)[0]
        line["cpu_baseline"] = {"value": nq / cdt, "unit": "queries/s", "cores": 1, "kind": "port",
                                "sample": f"first {nq} queries of the same batch, {cdt:.1f} s, one thread (oracle.ivf_flat_search: "
                                          "Python loop over the probed partitions around the C distance kernels)"}
        gp = ix.find_partitions(queries[:nq], nprobes)
        line["parity"] = parity_block((rd, ri, rc, rp), (h_od.numpy(), h_oi.numpy().view(np.uint64), h_oc.numpy()), k, gpu_probes=gp)
    if world == 1 and _BITS == 8 and not args.no_recall and not args.no_cpu_baseline and args.workload != "tiny":
        ix.close()   # free the benchmark index first
        rb = recall_block(wl, local)
        line["recall"] = rb
        line["recall_at_10"] = rb["recall_at_10"]
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

// api_group.inl -- the multi-GPU form of the path behind the C ABI (SURVEY 8(e), 8(b) "lb200_group_search";
// included by api.cu).
//
// IVF partitions shard over G GPUs; every shard is an lb200_index built from its partitions only.  Per batch a
// shard (1) selects the probes of ITS slice of the queries, (2) PUBLISHES the slice's probe lists (and, on the
// host-pointer entry, the slice's query vectors) into every peer's exchange block with plain stores over
// NVLink/NVSwitch peer memory, (3) scans the probed partitions it owns -- sharing pruning bounds with the peers
// from inside the scan kernel (lb200_peer_thresholds_*), (4) writes its k' merged candidates straight into the
// ROOT shard's gather buffer over peer memory, and the root merges the G lists.  There is no collective and
// no host round trip per batch: ranks meet through epoch flags in each other's memory (one flag per rank and
// phase, written after a system-scope fence, awaited by a one-thread kernel on the consumer's stream).
//
// Two hosts are served by the same device protocol:
//   * one process per GPU (torchrun, MPI, ...): lb200_shard_export -> the host exchanges the opaque blobs any
//     way it likes -> lb200_shard_attach (CUDA IPC) -> lb200_shard_search[_device] on every rank;
//   * one process driving all GPUs (what a Rust host embedding the library does): lb200_group_create /
//     lb200_group_search / lb200_group_destroy, which shard the index by a partition -> device map and attach the
//     shards through same-process peer pointers.
// Results are bit-identical to the single-index search (every shard list holds every one of
// its rows that can be in the merged top-k').

namespace {

constexpr uint32_t SHARD_MAX = LB200_MAX_PEERS + 1;
constexpr uint32_t SHARD_BLOB = 128;  // bytes per rank: [64 IPC handle of the exchange block][64 of the bound table]

struct PeerBases {
  char* base[SHARD_MAX];
};

// src[0, bytes) -> every peer's block at off (16-byte granules; bytes % 16 == 0).  Grid-stride.
__global__ void shard_publish_kernel(PeerBases peers, uint32_t world, uint32_t self, size_t off, const uint4* __restrict__ src,
                                     size_t n16, int include_self) {
  for (uint32_t r = 0; r < world; r++) {
    if (r == self && !include_self) continue;
    uint4* dst = reinterpret_cast<uint4*>(peers.base[r] + off);
    if (dst == src) continue;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
  }
  __threadfence_system();
}

// flags[phase][self] = epoch in every listed peer's block (root only when only_root).
__global__ void shard_signal_kernel(PeerBases peers, uint32_t world, uint32_t self, size_t off_flags, uint32_t phase,
                                    uint32_t epoch, int root /* < 0: all ranks */) {
  const uint32_t r = threadIdx.x;
  if (r >= world || (root >= 0 && r != (uint32_t)root)) return;
  __threadfence_system();
  volatile uint32_t* f = reinterpret_cast<volatile uint32_t*>(peers.base[r] + off_flags) + phase * SHARD_MAX + self;
  *f = epoch;
  __threadfence_system();
}

// One thread per rank spins until flags[phase][r] >= epoch (epochs only grow; wrap-safe compare).  A peer that
// never arrives (crashed process) must not hang the GPU: after `timeout_ns` the kernel gives up and records it.
__global__ void shard_wait_kernel(const uint32_t* flags_base, uint32_t world, uint32_t phase, uint32_t epoch,
                                  unsigned long long timeout_ns, uint32_t* err) {
  const uint32_t r = threadIdx.x;
  if (r >= world) return;
  const volatile uint32_t* f = flags_base + phase * SHARD_MAX + r;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    if ((int32_t)(*f - epoch) >= 0) break;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) {
      atomicOr(err, 1u << (r & 31));
      break;
    }
    __nanosleep(200);
  }
  __threadfence_system();
}

static size_t pad256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace

struct ShardComm {
  uint32_t world = 0, rank = 0, root = 0;
  uint32_t qmax = 0, pmax = 0, kmax = 0, dim = 0;
  char* local = nullptr;
  size_t bytes = 0;
  PeerBases peers = {};
  void* ipc[SHARD_MAX] = {};
  bool attached = false;
  uint32_t epoch = 0;
  // offsets inside every rank's block (identical on all ranks)
  size_t o_flags = 0, o_err = 0, o_q = 0, o_probes[2] = {}, o_gd = 0, o_gr = 0, o_ge = 0, o_gc = 0;
  size_t o_td = 0, o_te = 0, o_tr = 0, o_tc = 0;  // root merge scratch
  void layout(uint32_t world_max) {
    size_t o = 0;
    auto put = [&](size_t b) { size_t r = o; o += pad256(b); return r; };
    o_flags = put((size_t)4 * SHARD_MAX * 4);
    o_err = put(256);
    o_q = put((size_t)qmax * dim * 4);
    o_probes[0] = put((size_t)qmax * pmax * 4);
    o_probes[1] = put((size_t)qmax * pmax * 4);
    const size_t cells = (size_t)world_max * qmax * kmax;
    o_gd = put(cells * 4);
    o_gr = put(cells * 8);
    o_ge = put(cells * 4);
    o_gc = put((size_t)world_max * qmax * 4);
    o_td = put((size_t)qmax * kmax * 4);
    o_te = put((size_t)qmax * kmax * 4);
    o_tr = put((size_t)qmax * kmax * 8);
    o_tc = put((size_t)qmax * 4);
    bytes = o;
  }
};

static void shard_comm_free(lb200_index* ix) {
  ShardComm* sc = ix->comm;
  if (!sc) return;
  cudaSetDevice(ix->device);
  for (auto& p : sc->ipc)
    if (p) cudaIpcCloseMemHandle(p);
  cudaFree(sc->local);
  delete sc;
  ix->comm = nullptr;
}

extern "C" int32_t lb200_shard_export(lb200_index* ix, uint32_t max_queries, uint32_t max_nprobes, uint32_t max_kprime,
                                      void* blob_out_128, uint64_t* dev_ptrs_out_2) {
  if (!ix || !max_queries || !max_nprobes || !max_kprime) return set_err(LB200_EINVAL, "null index or zero capacity");
  if (max_kprime > 512) return set_err(LB200_EUNSUPPORTED, "k*refine_factor=%u > 512", max_kprime);
  CK(cudaSetDevice(ix->device));
  lb200_peer_thresholds_detach(ix);
  shard_comm_free(ix);
  ShardComm* sc = new ShardComm();
  sc->qmax = max_queries;
  sc->pmax = std::min(max_nprobes, ix->nlist);
  sc->kmax = max_kprime;
  sc->dim = ix->dim;
  sc->layout(SHARD_MAX);
  cudaError_t e = cudaMalloc((void**)&sc->local, sc->bytes);  // plain cudaMalloc: exportable over CUDA IPC
  if (e != cudaSuccess) {
    delete sc;
    return set_err(LB200_ENOMEM, "cudaMalloc(%zu) for the shard exchange block failed: %s", sc->bytes, cudaGetErrorString(e));
  }
  CK(cudaMemset(sc->local, 0, sc->o_q));  // flags + error word
  ix->comm = sc;
  unsigned char blob[SHARD_BLOB] = {};
  uint64_t thr_ptr = 0;
  int rc = lb200_peer_thresholds_export(ix, max_queries, blob_out_128 ? blob + 64 : nullptr, &thr_ptr);
  if (rc) return rc;
  if (blob_out_128) {
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, sc->local));
    memcpy(blob, &h, 64);
    memcpy(blob_out_128, blob, SHARD_BLOB);
  }
  if (dev_ptrs_out_2) {
    dev_ptrs_out_2[0] = (uint64_t)(uintptr_t)sc->local;
    dev_ptrs_out_2[1] = thr_ptr;
  }
  return 0;
}

static int shard_warmup(lb200_index* ix);
extern "C" int32_t lb200_shard_attach(lb200_index* ix, uint32_t world, uint32_t rank, uint32_t root,
                                      const void* blobs_or_null, const uint64_t* same_process_ptrs_or_null) {
  if (!ix || !ix->comm) return set_err(LB200_EINVAL, "call lb200_shard_export first");
  if (world < 1 || world > SHARD_MAX || rank >= world || root >= world)
    return set_err(LB200_EINVAL, "world=%u rank=%u root=%u out of range (at most %u shards)", world, rank, root, SHARD_MAX);
  if (!blobs_or_null && !same_process_ptrs_or_null) return set_err(LB200_EINVAL, "no peer handles given");
  CK(cudaSetDevice(ix->device));
  ShardComm* sc = ix->comm;
  for (auto& p : sc->ipc)
    if (p) { cudaIpcCloseMemHandle(p); p = nullptr; }
  sc->world = world;
  sc->rank = rank;
  sc->root = root;
  unsigned char thr_handles[SHARD_MAX * 64] = {};
  uint64_t thr_ptrs[SHARD_MAX] = {};
  for (uint32_t r = 0; r < world; r++) {
    if (r == rank) {
      sc->peers.base[r] = sc->local;
      continue;
    }
    if (same_process_ptrs_or_null) {
      sc->peers.base[r] = reinterpret_cast<char*>((uintptr_t)same_process_ptrs_or_null[2 * r]);
      thr_ptrs[r] = same_process_ptrs_or_null[2 * r + 1];
      if (!sc->peers.base[r]) return set_err(LB200_EINVAL, "peer %u: null exchange block", r);
      // a peer on another device of this process: enable direct access (idempotent)
      cudaPointerAttributes at;
      if (cudaPointerGetAttributes(&at, sc->peers.base[r]) == cudaSuccess && at.device != ix->device) {
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, ix->device, at.device));
        if (!can) return set_err(LB200_EUNSUPPORTED, "device %d cannot access peer device %d", ix->device, at.device);
        cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
          return set_err(LB200_ECUDA, "cudaDeviceEnablePeerAccess(%d) failed: %s", at.device, cudaGetErrorString(e));
        cudaGetLastError();
      }
    } else {
      cudaIpcMemHandle_t h;
      memcpy(&h, (const char*)blobs_or_null + (size_t)r * SHARD_BLOB, 64);
      void* p = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) return set_err(LB200_ECUDA, "cudaIpcOpenMemHandle(peer %u) failed: %s", r, cudaGetErrorString(e));
      sc->ipc[r] = p;
      sc->peers.base[r] = (char*)p;
      memcpy(thr_handles + (size_t)r * 64, (const char*)blobs_or_null + (size_t)r * SHARD_BLOB + 64, 64);
    }
  }
  // Warm-up before the first real batch: one full-capacity batch through this shard ALONE (world = 1: no flags,
  // no peer stores), with and without refine.  It loads every kernel module the batches will launch and sizes the
  // call context's scratch, so that a real batch never has to load a module or allocate while its stream holds a
  // kernel that waits for a peer (lazy module loading may synchronise with running kernels: a dead-lock when the
  // peer is driven by this very host thread, a time-out otherwise).
  int rc = shard_warmup(ix);
  if (rc) return rc;
  sc->world = world;
  sc->rank = rank;
  sc->root = root;
  rc = lb200_peer_thresholds_attach(ix, world, rank, same_process_ptrs_or_null ? nullptr : thr_handles,
                                    same_process_ptrs_or_null ? thr_ptrs : nullptr);
  if (rc) return rc;
  sc->attached = true;
  sc->epoch = 0;
  return 0;
}

extern "C" int32_t lb200_shard_detach(lb200_index* ix) {
  if (!ix) return set_err(LB200_EINVAL, "null index");
  lb200_peer_thresholds_detach(ix);
  shard_comm_free(ix);
  return 0;
}

namespace {

// Push [off, off + nb) of the local block into the same place of every peer's block.
static int shard_publish(const ShardComm* sc, size_t off, size_t nb, cudaStream_t st) {
  if (sc->world <= 1 || !nb) return 0;
  if ((off % 16) || (nb % 16)) {  // odd row sizes: byte-exact copies through the copy engines
    for (uint32_t r = 0; r < sc->world; r++)
      if (r != sc->rank) CK(cudaMemcpyAsync(sc->peers.base[r] + off, sc->local + off, nb, cudaMemcpyDefault, st));
    return 0;
  }
  const size_t n16 = nb / 16;
  shard_publish_kernel<<<(uint32_t)std::max<size_t>(1, std::min<size_t>(64, (n16 + 255) / 256)), 256, 0, st>>>(
      sc->peers, sc->world, sc->rank, off, reinterpret_cast<const uint4*>(sc->local + off), n16, 0);
  CK(cudaGetLastError());
  return 0;
}

// Enqueue one batch on this shard.  queries: device [Q][dim] (full batch) when !host, else host [Q][dim]
// (every rank holds the batch; only the own slice crosses PCIe, the rest arrives from the peers over NVLink).
// Root: results to d_out_* (device) -- the caller copies them out.  Asynchronous on st.
static int shard_enqueue(lb200_index* ix, CallCtx& cx, cudaStream_t st, const float* queries, bool host, uint32_t Q,
                         uint32_t nprobes, uint32_t k, uint32_t refine_factor, const float* lower_or_null,
                         const float* upper_or_null, float* d_out_dist, uint64_t* d_out_rowid, uint32_t* d_out_count) {
  ShardComm* sc = ix->comm;
  if (!sc || !sc->attached) return set_err(LB200_EINVAL, "shard is not attached (lb200_shard_export / lb200_shard_attach)");
  const uint32_t P = std::min(nprobes, ix->nlist), kprime = k * std::max(refine_factor, 1u);
  if (k == 0 || nprobes == 0) return set_err(LB200_EINVAL, "k and nprobes must be > 0");
  if (Q > sc->qmax || P > sc->pmax || kprime > sc->kmax)
    return set_err(LB200_EINVAL, "batch (Q=%u, nprobes=%u, k'=%u) exceeds the exported capacity (%u, %u, %u)", Q, P, kprime,
                   sc->qmax, sc->pmax, sc->kmax);
  if (!Q) return 0;
  const uint32_t G = sc->world, me = sc->rank, dim = ix->dim;
  const uint32_t epoch = ++sc->epoch, par = epoch & 1;
  const uint32_t qs = (Q + G - 1) / G, lo = std::min(me * qs, Q), n = std::min(qs, Q - lo);
  char* L = sc->local;
  float* lq = reinterpret_cast<float*>(L + sc->o_q);
  uint32_t* lprobes = reinterpret_cast<uint32_t*>(L + sc->o_probes[par]);
  const float* d_queries = queries;
  if (host) {  // own slice over PCIe into the exchange block, then to every peer over NVLink
    if (n) {
      const size_t nb = (size_t)n * dim * 4;
      if (nb <= LB200_PIN_BYTES) {
        int rc = cx.pin_reserve(nb);
        if (rc) return rc;
        memcpy(cx.h_pin, queries + (size_t)lo * dim, nb);
        CK(cudaMemcpyAsync(lq + (size_t)lo * dim, cx.h_pin, nb, cudaMemcpyHostToDevice, st));
      } else {
        CK(cudaMemcpyAsync(lq + (size_t)lo * dim, queries + (size_t)lo * dim, nb, cudaMemcpyHostToDevice, st));
      }
      int rcp = shard_publish(sc, sc->o_q + (size_t)lo * dim * 4, nb, st);
      if (rcp) return rcp;
    }
    d_queries = lq;
  }
  // ---- (1) probes of the own slice, straight into the local table ----
  if (n) {
    size_t need = Arena::pad((size_t)n * dim * 4) + Arena::pad((size_t)n * ix->nlist * 4) + 1024;
    int rc = cx.arena.reserve(need);
    if (rc) return rc;
    cx.arena.reset();
    float* d_keys = cx.arena.take<float>((size_t)n * dim);
    float* d_D = cx.arena.take<float>((size_t)n * ix->nlist);
    const float* keys = d_queries + (size_t)lo * dim;
    uint32_t launches = 0;
    if (ix->metric == LB200_COSINE) {
      normalize_kernel<<<(n * 32 + 255) / 256, 256, 0, st>>>(keys, d_keys, n, dim);
      CK(cudaGetLastError());
      keys = d_keys;
    }
    rc = coarse_device(ix, keys, n, P, d_D, lprobes + (size_t)lo * P, nullptr, st, &launches);
    if (rc) return rc;
    // ---- (2) publish the slice's probe lists ----
    rc = shard_publish(sc, sc->o_probes[par] + (size_t)lo * P * 4, (size_t)n * P * 4, st);
    if (rc) return rc;
  }
  uint32_t* d_err = reinterpret_cast<uint32_t*>(L + sc->o_err);
  if (G > 1) {
    shard_signal_kernel<<<1, 32, 0, st>>>(sc->peers, G, me, sc->o_flags, 0, epoch, -1);
    shard_wait_kernel<<<1, 32, 0, st>>>(reinterpret_cast<const uint32_t*>(L + sc->o_flags), G, 0, epoch, 4000000000ull, d_err);
    CK(cudaGetLastError());
  }
  // ---- (3) scan of the owned partitions; (4) lists straight into the root's gather buffer ----
  char* R = sc->peers.base[sc->root];
  const size_t cell = (size_t)me * Q * kprime;
  SearchArgs a = {};
  a.d_queries = d_queries; a.Q = Q; a.d_part_ids = lprobes; a.P = P; a.k = k; a.refine_factor = refine_factor;
  a.normalize = true; a.shard_mode = true;
  a.has_lower = lower_or_null != nullptr; a.has_upper = upper_or_null != nullptr;
  a.lower = lower_or_null ? *lower_or_null : 0.f; a.upper = upper_or_null ? *upper_or_null : 0.f;
  a.d_out_dist = reinterpret_cast<float*>(R + sc->o_gd) + cell;
  a.d_out_rowid = reinterpret_cast<uint64_t*>(R + sc->o_gr) + cell;
  a.d_out_exact = reinterpret_cast<float*>(R + sc->o_ge) + cell;
  a.d_out_count = reinterpret_cast<uint32_t*>(R + sc->o_gc) + (size_t)me * Q;
  int rc = search_pipeline(ix, cx, a, st);
  if (rc) return rc;
  if (G > 1) {
    shard_signal_kernel<<<1, 32, 0, st>>>(sc->peers, G, me, sc->o_flags, 1, epoch, (int)sc->root);
    CK(cudaGetLastError());
  }
  if (me != sc->root) return 0;
  if (G > 1) {
    shard_wait_kernel<<<1, 32, 0, st>>>(reinterpret_cast<const uint32_t*>(L + sc->o_flags), G, 1, epoch, 4000000000ull, d_err);
    CK(cudaGetLastError());
  }
  // ---- root: merge of the G lists == concat + sort_to_indices(limit k') (ivf/v2.rs:383-400), then the refine cut ----
  const float* g_d = reinterpret_cast<const float*>(L + sc->o_gd);
  const uint64_t* g_r = reinterpret_cast<const uint64_t*>(L + sc->o_gr);
  const float* g_e = reinterpret_cast<const float*>(L + sc->o_ge);
  const uint32_t* g_c = reinterpret_cast<const uint32_t*>(L + sc->o_gc);
  if (!refine_factor) {
    return launch_merge(g_d, g_r, nullptr, nullptr, g_c, Q, G, kprime, kprime, (size_t)Q * kprime, 1, Q, std::min(k, kprime), k,
                        d_out_dist, d_out_rowid, nullptr, nullptr, d_out_count, st);
  }
  float* t_d = reinterpret_cast<float*>(L + sc->o_td);
  float* t_e = reinterpret_cast<float*>(L + sc->o_te);
  uint64_t* t_r = reinterpret_cast<uint64_t*>(L + sc->o_tr);
  uint32_t* t_c = reinterpret_cast<uint32_t*>(L + sc->o_tc);
  rc = launch_merge(g_d, g_r, nullptr, g_e, g_c, Q, G, kprime, kprime, (size_t)Q * kprime, 1, Q, kprime, kprime, t_d, t_r, nullptr,
                    t_e, t_c, st);
  if (rc) return rc;
  rerank_kernel<<<(Q * 32 + 127) / 128, 128, 0, st>>>(t_e, t_r, t_c, Q, kprime, k, a.has_lower, a.lower, a.has_upper, a.upper,
                                                    d_out_dist, d_out_rowid, d_out_count);
  CK(cudaGetLastError());
  return 0;
}

}  // namespace

static int shard_warmup(lb200_index* ix) {
  ShardComm* sc = ix->comm;
  const PeerBases saved = sc->peers;
  const uint32_t w0 = sc->world, r0 = sc->rank, root0 = sc->root;
  sc->world = 1; sc->rank = 0; sc->root = 0;
  sc->peers.base[0] = sc->local;
  sc->attached = true;
  int rc = 0;
  {
    ScanPlan plan;
    if (!plan_scan(ix, sc->qmax, sc->pmax, sc->kmax, &plan) && plan.sc.variant == 3) preload_scan3(plan.sc.ql, plan.sc.dsub_t);
    CtxLease ls;
    rc = ls.acquire(ix, nullptr);
    cudaStream_t st = ls.st;
    float* zq = reinterpret_cast<float*>(sc->local + sc->o_q);  // warm-up queries in the block itself: the centroids
    for (uint32_t q0 = 0; !rc && q0 < sc->qmax; q0 += ix->nlist) {
      const size_t rows = std::min<size_t>(ix->nlist, sc->qmax - q0);
      if (cudaMemcpyAsync(zq + (size_t)q0 * sc->dim, ix->d_centroids, rows * sc->dim * 4, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
        rc = set_err(LB200_ECUDA, "warm-up copy failed");
    }
    // outputs of the warm-up land in gather slot 1 (world = 1 uses slot 0 only)
    const size_t cell = (size_t)sc->qmax * sc->kmax;
    float* od = reinterpret_cast<float*>(sc->local + sc->o_gd) + cell;
    uint64_t* orow = reinterpret_cast<uint64_t*>(sc->local + sc->o_gr) + cell;
    uint32_t* oc = reinterpret_cast<uint32_t*>(sc->local + sc->o_gc) + sc->qmax;
    if (!rc) rc = shard_enqueue(ix, *ls.cx, st, zq, false, sc->qmax, sc->pmax, sc->kmax, 0, nullptr, nullptr, od, orow, oc);
    if (!rc && ix->d_raw && sc->kmax >= 2)
      rc = shard_enqueue(ix, *ls.cx, st, zq, false, sc->qmax, sc->pmax, sc->kmax / 2, 2, nullptr, nullptr, od, orow, oc);
    // the exchange kernels themselves
    if (!rc) {
      shard_publish_kernel<<<1, 32, 0, st>>>(sc->peers, 1, 0, sc->o_q, reinterpret_cast<const uint4*>(zq), 0, 0);
      shard_signal_kernel<<<1, 32, 0, st>>>(sc->peers, 1, 0, sc->o_flags, 3, 0, -1);
      shard_wait_kernel<<<1, 32, 0, st>>>(reinterpret_cast<const uint32_t*>(sc->local + sc->o_flags), 1, 3, 0, 1000ull,
                                          reinterpret_cast<uint32_t*>(sc->local + sc->o_err));
      if (cudaGetLastError() != cudaSuccess) rc = set_err(LB200_ECUDA, "warm-up of the exchange kernels failed");
    }
    if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = set_err(LB200_ECUDA, "warm-up batch failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  sc->peers = saved;
  sc->world = w0; sc->rank = r0; sc->root = root0;
  sc->attached = false;
  sc->epoch = 0;
  return rc;
}

namespace {

static int shard_check_err(lb200_index* ix) {
  uint32_t e = 0;
  CK(cudaMemcpy(&e, ix->comm->local + ix->comm->o_err, 4, cudaMemcpyDeviceToHost));
  if (e) {
    cudaMemset(ix->comm->local + ix->comm->o_err, 0, 4);
    return set_err(LB200_ECUDA, "shard exchange timed out waiting for peer ranks (mask 0x%x)", e);
  }
  return 0;
}

}  // namespace

extern "C" int32_t lb200_shard_search_device(lb200_index* ix, const float* d_queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                                             uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                                             float* d_out_dist, uint64_t* d_out_rowid, uint32_t* d_out_count, void* stream) {
  int rc = check_queries(ix, d_queries, Q);
  if (rc) return rc;
  if (ix->comm && ix->comm->attached && ix->comm->rank == ix->comm->root && Q && (!d_out_dist || !d_out_rowid || !d_out_count))
    return set_err(LB200_EINVAL, "null output on the root shard");
  CtxLease ls;
  if ((rc = ls.acquire(ix, (cudaStream_t)stream))) return rc;
  rc = shard_enqueue(ix, *ls.cx, ls.st, d_queries, false, Q, nprobes, k, refine_factor, lower_or_null, upper_or_null, d_out_dist,
                     d_out_rowid, d_out_count);
  if (rc) return rc;
  if (!stream) {
    CK(cudaStreamSynchronize(ls.st));
    return shard_check_err(ix);
  }
  return 0;
}

extern "C" int32_t lb200_shard_search(lb200_index* ix, const float* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                                      uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                                      float* out_dist, uint64_t* out_rowid, uint32_t* out_count) {
  int rc = check_queries(ix, queries, Q);
  if (rc) return rc;
  if (!ix->comm || !ix->comm->attached) return set_err(LB200_EINVAL, "shard is not attached");
  const bool root = ix->comm->rank == ix->comm->root;
  if (root && Q && (!out_dist || !out_rowid || !out_count)) return set_err(LB200_EINVAL, "null output on the root shard");
  if (!Q) return 0;
  CtxLease ls;
  if ((rc = ls.acquire(ix, nullptr))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;
  const size_t no = (size_t)Q * k;
  const size_t o_or = Arena::pad(no * 4), o_oc = o_or + Arena::pad(no * 8), total = o_oc + Arena::pad((size_t)Q * 4);
  if ((rc = cx.stage.reserve(total))) return rc;
  char* b = cx.stage.base;
  rc = shard_enqueue(ix, cx, st, queries, true, Q, nprobes, k, refine_factor, lower_or_null, upper_or_null,
                     reinterpret_cast<float*>(b), reinterpret_cast<uint64_t*>(b + o_or), reinterpret_cast<uint32_t*>(b + o_oc));
  if (rc) return rc;
  if (root) {
    CK(cudaMemcpyAsync(out_dist, b, no * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(out_rowid, b + o_or, no * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(out_count, b + o_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  }
  CK(cudaStreamSynchronize(st));
  return shard_check_err(ix);
}

// ---- one process, G devices ---------------------------------------------------------------------------
struct lb200_group {
  std::vector<lb200_index*> shards;
  std::vector<uint32_t> owner;  // [nlist] partition -> shard
  uint32_t qmax = 0, pmax = 0, kmax = 0, dim = 0;
};

extern "C" void lb200_group_destroy(lb200_group* g) {
  if (!g) return;
  for (lb200_index* ix : g->shards) {
    if (!ix) continue;
    lb200_shard_detach(ix);
    lb200_index_destroy(ix);
  }
  delete g;
}

extern "C" int32_t lb200_group_create(const int32_t* device_ids, uint32_t G, int32_t metric, uint32_t dim, uint32_t nlist,
                                      const float* centroids, uint32_t M, uint32_t num_bits, const float* codebook,
                                      const uint64_t* part_offsets, const uint8_t* codes, const uint64_t* row_ids,
                                      const float* raw_vectors_or_null, const uint32_t* part_owner_or_null,
                                      uint32_t max_queries, uint32_t max_nprobes, uint32_t max_kprime, lb200_group** out) {
  if (!out) return set_err(LB200_EINVAL, "out is null");
  *out = nullptr;
  if (!device_ids || !G || G > SHARD_MAX) return set_err(LB200_EINVAL, "G=%u out of range [1,%u]", G, SHARD_MAX);
  if (num_bits != 8 && num_bits != 4 && num_bits != 32) return set_err(LB200_EUNSUPPORTED, "num_bits=%u", num_bits);
  // bytes of one row's codes: 8-bit PQ M, 4-bit PQ M / 2, IVF_FLAT (num_bits 32) none -- its rows are the raw vectors
  const size_t cbytes = num_bits == 8 ? M : num_bits == 4 ? M / 2 : 0;
  if (!nlist || !part_offsets || !M || !dim) return set_err(LB200_EINVAL, "bad index arguments");
  lb200_group* g = nullptr;
  try {
    g = new lb200_group();
    g->owner.assign(nlist, 0);
    if (part_owner_or_null) {
      for (uint32_t p = 0; p < nlist; p++) {
        if (part_owner_or_null[p] >= G) { delete g; return set_err(LB200_EINVAL, "partition %u assigned to shard %u >= G", p, part_owner_or_null[p]); }
        g->owner[p] = part_owner_or_null[p];
      }
    } else {
      // greedy bin-packing, largest first, on rows x expected probes: in high dimension a few "hub"
      // centroids are among the nprobes nearest of many more queries than the others, so equal rows are not
      // equal scan work.  The probe frequency is estimated from the index alone -- how often a partition is among
      // the max_nprobes nearest of the other centroids -- with the library's own find_partitions on an empty
      // handle (same coarse metric as the searches).
      std::vector<double> w(nlist, 1.0);
      {
        std::vector<uint64_t> zoffs(nlist + 1, 0);
        lb200_index* tmp = nullptr;
        const uint32_t Pw = std::min(std::max(max_nprobes, 1u), nlist);
        int rc = lb200_index_create(device_ids[0], metric, dim, nlist, centroids, M, num_bits, codebook, zoffs.data(), nullptr,
                                    nullptr, nullptr, &tmp);
        if (!rc) {
          std::vector<uint32_t> ids((size_t)nlist * Pw);
          rc = lb200_find_partitions(tmp, centroids, nlist, Pw, ids.data(), nullptr);
          if (!rc)
            for (uint32_t id : ids) w[id] += 1.0;
          lb200_index_destroy(tmp);
        }
        if (rc) { delete g; return rc; }
      }
      std::vector<uint32_t> order(nlist);
      for (uint32_t p = 0; p < nlist; p++) {
        order[p] = p;
        w[p] *= (double)(part_offsets[p + 1] - part_offsets[p]);
      }
      std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return w[a] > w[b]; });
      std::vector<double> load(G, 0.0);
      for (uint32_t p : order) {
        uint32_t r = (uint32_t)(std::min_element(load.begin(), load.end()) - load.begin());
        g->owner[p] = r;
        load[r] += w[p];
      }
    }
    std::vector<uint64_t> ptrs(2 * (size_t)G, 0);
    for (uint32_t r = 0; r < G; r++) {
      // the shard's view: all nlist partitions, the ones it does not own empty
      std::vector<uint64_t> offs(nlist + 1, 0);
      for (uint32_t p = 0; p < nlist; p++) offs[p + 1] = offs[p] + (g->owner[p] == r ? part_offsets[p + 1] - part_offsets[p] : 0);
      const uint64_t n = offs[nlist];
      std::vector<uint8_t> sc(n * cbytes);
      std::vector<uint64_t> sr(n);
      std::vector<float> sv(raw_vectors_or_null ? n * dim : 0);
      for (uint32_t p = 0; p < nlist; p++) {
        if (g->owner[p] != r) continue;
        const uint64_t o = part_offsets[p], np = part_offsets[p + 1] - o;
        if (cbytes) memcpy(sc.data() + offs[p] * cbytes, codes + o * cbytes, np * cbytes);
        memcpy(sr.data() + offs[p], row_ids + o, np * 8);
        if (raw_vectors_or_null) memcpy(sv.data() + offs[p] * dim, raw_vectors_or_null + o * dim, np * dim * 4);
      }
      lb200_index* ix = nullptr;
      int rc = lb200_index_create(device_ids[r], metric, dim, nlist, centroids, M, num_bits, codebook, offs.data(), sc.data(),
                                  sr.data(), raw_vectors_or_null ? sv.data() : nullptr, &ix);
      if (rc) { lb200_group_destroy(g); return rc; }
      g->shards.push_back(ix);
      rc = lb200_shard_export(ix, max_queries, max_nprobes, max_kprime, nullptr, &ptrs[2 * (size_t)r]);
      if (rc) { lb200_group_destroy(g); return rc; }
    }
    for (uint32_t r = 0; r < G; r++) {
      int rc = lb200_shard_attach(g->shards[r], G, r, 0, nullptr, ptrs.data());
      if (rc) { lb200_group_destroy(g); return rc; }
    }
  } catch (const std::bad_alloc&) {
    lb200_group_destroy(g);
    return set_err(LB200_ENOMEM, "out of host memory while sharding the index");
  }
  g->qmax = max_queries; g->pmax = max_nprobes; g->kmax = max_kprime; g->dim = dim;
  *out = g;
  return 0;
}

extern "C" int32_t lb200_group_search(lb200_group* g, const float* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                                      uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                                      float* out_dist, uint64_t* out_rowid, uint32_t* out_count) {
  if (!g || g->shards.empty()) return set_err(LB200_EINVAL, "null group");
  if (Q && (!queries || !out_dist || !out_rowid || !out_count)) return set_err(LB200_EINVAL, "null argument");
  if (!Q) return 0;
  const uint32_t G = (uint32_t)g->shards.size();
  for (lb200_index* ix : g->shards)  // profiling synchronises a shard's stream mid-batch: its peers are not launched yet
    if (ix->profiling) return set_err(LB200_EINVAL, "profiling is not available through lb200_group_search");
  // enqueue every shard's part (asynchronous), the root last so that its D2H follows; then wait for all
  std::vector<CtxLease> ls(G);
  const size_t no = (size_t)Q * k;
  const size_t o_or = Arena::pad(no * 4), o_oc = o_or + Arena::pad(no * 8), total = o_oc + Arena::pad((size_t)Q * 4);
  char* b = nullptr;
  for (uint32_t i = 0; i < G; i++) {
    const uint32_t r = (i + 1) % G;  // 1, 2, ..., G-1, 0 (the root)
    lb200_index* ix = g->shards[r];
    int rc = ls[r].acquire(ix, nullptr);
    if (rc) return rc;
    if (r == 0) {
      if ((rc = ls[r].cx->stage.reserve(total))) return rc;
      b = ls[r].cx->stage.base;
    }
    rc = shard_enqueue(ix, *ls[r].cx, ls[r].st, queries, true, Q, nprobes, k, refine_factor, lower_or_null, upper_or_null,
                       reinterpret_cast<float*>(b), reinterpret_cast<uint64_t*>(b ? b + o_or : nullptr),
                       reinterpret_cast<uint32_t*>(b ? b + o_oc : nullptr));
    if (rc) return rc;
  }
  cudaStream_t st0 = ls[0].st;
  CK(cudaSetDevice(g->shards[0]->device));
  CK(cudaMemcpyAsync(out_dist, b, no * 4, cudaMemcpyDeviceToHost, st0));
  CK(cudaMemcpyAsync(out_rowid, b + o_or, no * 8, cudaMemcpyDeviceToHost, st0));
  CK(cudaMemcpyAsync(out_count, b + o_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost, st0));
  for (uint32_t r = 0; r < G; r++) {
    CK(cudaSetDevice(g->shards[r]->device));
    CK(cudaStreamSynchronize(ls[r].st));
  }
  for (uint32_t r = 0; r < G; r++) {
    int rc = shard_check_err(g->shards[r]);
    if (rc) return rc;
  }
  return 0;
}

// Error word of the exchange (a peer that never arrived within the time-out); clears it.
extern "C" int32_t lb200_shard_check(lb200_index* ix) {
  if (!ix || !ix->comm) return set_err(LB200_EINVAL, "shard is not exported");
  CK(cudaSetDevice(ix->device));
  return shard_check_err(ix);
}

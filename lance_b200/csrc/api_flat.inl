// api_flat.inl -- flat kNN entry points (included by api.cu).

namespace {

static int flat_check(int32_t metric, uint32_t dim, const void* q, uint32_t Q, const void* v, uint64_t n) {
  if (metric != LB200_L2 && metric != LB200_COSINE && metric != LB200_DOT) return set_err(LB200_EINVAL, "unknown metric %d", metric);
  if (!dim) return set_err(LB200_EINVAL, "dim must be > 0");
  if ((Q && !q) || (n && !v)) return set_err(LB200_EINVAL, "null argument");
  if (n >= (1ull << 32)) return set_err(LB200_EUNSUPPORTED, "n >= 2^32 rows");
  return 0;
}

static int flat_dist_launch(int metric, const float* d_q, uint32_t Q, const float* d_v, const uint8_t* d_valid,
                            uint64_t n, uint32_t dim, float* d_out, size_t out_stride, cudaStream_t st) {
  if (!Q || !n) return 0;
  // L2, d % 16 == 0: the tiled kernel of the coarse step (same l2_scalar<16> order) reads the rows once
  // per 32 queries instead of once per 4
  const size_t tsm = ((size_t)dim * CQ_PITCH + (size_t)CQ_QB * dim) * 4;
  int dev = 0, optin = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const size_t ssm2 = (2 * (size_t)CD_SLAB + (size_t)CD_QB * dim) * 4;
  if (metric == LB200_L2 && dim % 64 == 0 && ssm2 <= (size_t)optin && Q >= 4 && !getenv("LB200_COARSE_SIMPLE") &&
      !getenv("LB200_COARSE_TILED")) {
    CK(cudaFuncSetAttribute(coarse_dist_l2_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm2));
    dim3 g2((uint32_t)((n + 127) / 128), (Q + CD_QB - 1) / CD_QB);
    coarse_dist_l2_slab_kernel<<<g2, 512, ssm2, st>>>(d_q, d_v, Q, (uint32_t)n, dim, d_out, out_stride, d_valid);
    CK(cudaGetLastError());
    return 0;
  }
  // dot, d % 64 == 0, no nulls: the slab kernel of the dot coarse step (dot_distance in the same f32x16 order)
  if (metric == LB200_DOT && dim % 64 == 0 && !d_valid && ssm2 <= (size_t)optin && Q >= 4 && n < (1ull << 32) &&
      !getenv("LB200_COARSE_SIMPLE")) {
    CK(cudaFuncSetAttribute(coarse_dist_dot_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm2));
    dim3 g2((uint32_t)((n + 127) / 128), (Q + CD_QB - 1) / CD_QB);
    coarse_dist_dot_slab_kernel<<<g2, 512, ssm2, st>>>(d_q, d_v, Q, (uint32_t)n, dim, d_out, out_stride);
    CK(cudaGetLastError());
    return 0;
  }
  if (metric == LB200_L2 && dim % 16 == 0 && dim >= 32 && tsm <= (size_t)optin && Q >= 8 && !getenv("LB200_COARSE_SIMPLE")) {
    CK(cudaFuncSetAttribute(coarse_dist_l2_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm));
    dim3 g2((uint32_t)((n + 127) / 128), (Q + CQ_QB - 1) / CQ_QB);
    coarse_dist_l2_tiled_kernel<<<g2, 128, tsm, st>>>(d_q, d_v, Q, (uint32_t)n, dim, d_out, out_stride, d_valid);
    CK(cudaGetLastError());
    return 0;
  }
  dim3 grid((uint32_t)((n + FLAT_THREADS - 1) / FLAT_THREADS), (Q + FLAT_TQ - 1) / FLAT_TQ);
  if (grid.y > 65535) return set_err(LB200_EINVAL, "query tile too large");
  size_t smem = (size_t)FLAT_TQ * dim * 4 + FLAT_TQ * 4;
  if (smem > 48 * 1024) CK(cudaFuncSetAttribute(flat_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  flat_dist_kernel<<<grid, FLAT_THREADS, smem, st>>>(metric, d_q, Q, d_v, d_valid, n, dim, d_out, out_stride);
  CK(cudaGetLastError());
  return 0;
}

struct FlatScratch {
  std::mutex mu;
  void* p = nullptr;
  size_t cap = 0;
};
static FlatScratch& flat_scratch(int dev) {
  static FlatScratch tab[64];
  return tab[(dev < 0 ? 0 : dev) & 63];
}

// queries are processed in tiles so that the [tile][n] distance block stays L2-resident
static int flat_topk_device_impl(int metric, uint32_t dim, const float* d_q, uint32_t Q, const float* d_v,
                                 const uint8_t* d_valid, uint64_t n, uint32_t k, bool has_lower, float lower,
                                 bool has_upper, float upper, float* d_od, uint64_t* d_oi, uint32_t* d_oc,
                                 cudaStream_t st) {
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (k > 1024) return set_err(LB200_EUNSUPPORTED, "k=%u > 1024", k);
  if (!Q) return 0;
  if (n == 0) {  // an empty batch: empty result (the reference returns an empty RecordBatch), not a scratch request
    fill_u32_kernel<<<(uint32_t)(((size_t)Q * k + 255) / 256), 256, 0, st>>>(reinterpret_cast<uint32_t*>(d_od), 0x7f800000u, (size_t)Q * k);
    CK(cudaMemsetAsync(d_oi, 0xff, (size_t)Q * k * 8, st));
    CK(cudaMemsetAsync(d_oc, 0, (size_t)Q * 4, st));
    CK(cudaGetLastError());
    return 0;
  }
  uint32_t tile = (uint32_t)std::max<uint64_t>(FLAT_TQ, std::min<uint64_t>(Q, (128ull << 20) / std::max<uint64_t>(1, n * 4)));
  tile = (tile + FLAT_TQ - 1) / FLAT_TQ * FLAT_TQ;
  size_t stride = (size_t)std::max<uint64_t>(n, 1);
  uint32_t kp2 = std::max(2u, next_pow2(k));
  size_t smem = (size_t)kp2 * 8;
  // segments per row: ~32k rows each, bounded by what the merge kernel sorts per query
  uint32_t nseg = (uint32_t)std::min<uint64_t>({(n + 32767) / 32768, 128ull, 16384ull / k});
  nseg = std::max(nseg, 1u);
  const uint64_t seg_len = (n + nseg - 1) / nseg;
  nseg = (uint32_t)((n + seg_len - 1) / std::max<uint64_t>(seg_len, 1));
  nseg = std::max(nseg, 1u);
  // Scratch (distance block + per-segment candidates) comes from a grow-only per-device cache: a
  // cudaMalloc/cudaFree pair per call costs up to tens of milliseconds (r01: 15 -> 195 ms per 256-query
  // batch on one box).  Calls on one device serialise on the cache's mutex.
  const size_t cells = nseg > 1 ? (size_t)tile * nseg * k : 0;
  const size_t b_D = Arena::pad((size_t)tile * stride * 4);
  const size_t need = b_D + Arena::pad(cells * 4) + Arena::pad(cells * 8) + Arena::pad((size_t)tile * nseg * 4) + 1024;
  int dev = 0;
  CK(cudaGetDevice(&dev));
  FlatScratch& fs = flat_scratch(dev);
  std::lock_guard<std::mutex> lk(fs.mu);
  if (fs.cap < need) {
    if (fs.p) CK(cudaFree(fs.p));
    fs.p = nullptr;
    fs.cap = 0;
    cudaError_t e = cudaMalloc(&fs.p, need);
    if (e != cudaSuccess) return set_err(LB200_ENOMEM, "flat kNN scratch of %zu bytes: %s", need, cudaGetErrorString(e));
    fs.cap = need;
  }
  float* d_D = (float*)fs.p;
  float* t_d = nullptr;
  uint64_t* t_i = nullptr;
  uint32_t* t_c = nullptr;
  if (nseg > 1) {
    t_d = (float*)((char*)fs.p + b_D);
    t_i = (uint64_t*)((char*)t_d + Arena::pad(cells * 4));
    t_c = (uint32_t*)((char*)t_i + Arena::pad(cells * 8));
  }
  for (uint32_t q0 = 0; q0 < Q; q0 += tile) {
    uint32_t qn = std::min(tile, Q - q0);
    int rc = flat_dist_launch(metric, d_q + (size_t)q0 * dim, qn, d_v, d_valid, n, dim, d_D, stride, st);
    if (rc) return rc;
    if (nseg == 1) {
      flat_select_kernel<<<qn, 256, smem, st>>>(d_D, stride, n, 1, std::max<uint64_t>(n, 1), k, kp2, has_lower, lower,
                                                has_upper, upper, d_od + (size_t)q0 * k, d_oi + (size_t)q0 * k,
                                                d_oc + q0);
      CK(cudaGetLastError());
    } else {
      flat_select_kernel<<<qn * nseg, 256, smem, st>>>(d_D, stride, n, nseg, seg_len, k, kp2, has_lower, lower,
                                                       has_upper, upper, t_d, t_i, t_c);
      CK(cudaGetLastError());
      rc = launch_merge(t_d, t_i, nullptr, nullptr, t_c, qn, nseg, k, (size_t)nseg * k, k, nseg, 1, k, k,
                        d_od + (size_t)q0 * k, d_oi + (size_t)q0 * k, nullptr, nullptr, d_oc + q0, st);
      if (rc) return rc;
    }
  }
  CK(cudaStreamSynchronize(st));  // the scratch is shared: finish before the next call may take it
  return 0;
}

}  // namespace

extern "C" int32_t lb200_flat_distance(int32_t device_id, int32_t metric, uint32_t dim, const float* queries,
                                       uint32_t Q, const float* vectors, const uint8_t* validity_or_null,
                                       uint64_t n, float* out) {
  int rc = flat_check(metric, dim, queries, Q, vectors, n);
  if (rc) return rc;
  rc = select_device(device_id);
  if (rc) return rc;
  if (!Q || !n) return 0;
  if (!out) return set_err(LB200_EINVAL, "null output");
  HostStage hs;
  size_t bq = Arena::pad((size_t)Q * dim * 4), bv = Arena::pad((size_t)n * dim * 4), bm = Arena::pad((size_t)(n + 7) / 8),
         bo = Arena::pad((size_t)Q * n * 4);
  CK(cudaMalloc(&hs.p, bq + bv + bm + bo));
  char* b = (char*)hs.p;
  float* d_q = (float*)b; b += bq;
  float* d_v = (float*)b; b += bv;
  uint8_t* d_m = (uint8_t*)b; b += bm;
  float* d_o = (float*)b;
  cudaStream_t st = nullptr;
  CK(cudaMemcpyAsync(d_q, queries, (size_t)Q * dim * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_v, vectors, (size_t)n * dim * 4, cudaMemcpyHostToDevice, st));
  if (validity_or_null) CK(cudaMemcpyAsync(d_m, validity_or_null, (size_t)(n + 7) / 8, cudaMemcpyHostToDevice, st));
  for (uint32_t q0 = 0; q0 < Q; q0 += 65535 * FLAT_TQ) {
    uint32_t qn = std::min<uint32_t>(65535 * FLAT_TQ, Q - q0);
    rc = flat_dist_launch(metric, d_q + (size_t)q0 * dim, qn, d_v, validity_or_null ? d_m : nullptr, n, dim,
                          d_o + (size_t)q0 * n, (size_t)n, st);
    if (rc) return rc;
  }
  CK(cudaMemcpyAsync(out, d_o, (size_t)Q * n * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int32_t lb200_flat_topk_device(int32_t device_id, int32_t metric, uint32_t dim, const float* d_queries,
                                          uint32_t Q, const float* d_vectors, const uint8_t* d_validity_or_null,
                                          uint64_t n, uint32_t k, float* d_out_dist, uint64_t* d_out_idx,
                                          uint32_t* d_out_count, void* stream) {
  int rc = flat_check(metric, dim, d_queries, Q, d_vectors, n);
  if (rc) return rc;
  rc = select_device(device_id);
  if (rc) return rc;
  return flat_topk_device_impl(metric, dim, d_queries, Q, d_vectors, d_validity_or_null, n, k, false, 0.f, false, 0.f,
                               d_out_dist, d_out_idx, d_out_count, (cudaStream_t)stream);
}

extern "C" int32_t lb200_flat_topk(int32_t device_id, int32_t metric, uint32_t dim, const float* queries, uint32_t Q,
                                   const float* vectors, const uint8_t* validity_or_null, uint64_t n, uint32_t k,
                                   const float* lower_or_null, const float* upper_or_null, float* out_dist,
                                   uint64_t* out_idx, uint32_t* out_count) {
  int rc = flat_check(metric, dim, queries, Q, vectors, n);
  if (rc) return rc;
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  rc = select_device(device_id);
  if (rc) return rc;
  if (!Q) return 0;
  if (!out_dist || !out_idx || !out_count) return set_err(LB200_EINVAL, "null output");
  HostStage hs;
  size_t bq = Arena::pad((size_t)Q * dim * 4), bv = Arena::pad((size_t)std::max<uint64_t>(n, 1) * dim * 4),
         bm = Arena::pad((size_t)(n + 7) / 8 + 1), bod = Arena::pad((size_t)Q * k * 4), boi = Arena::pad((size_t)Q * k * 8),
         boc = Arena::pad((size_t)Q * 4);
  CK(cudaMalloc(&hs.p, bq + bv + bm + bod + boi + boc));
  char* b = (char*)hs.p;
  float* d_q = (float*)b; b += bq;
  float* d_v = (float*)b; b += bv;
  uint8_t* d_m = (uint8_t*)b; b += bm;
  float* d_od = (float*)b; b += bod;
  uint64_t* d_oi = (uint64_t*)b; b += boi;
  uint32_t* d_oc = (uint32_t*)b;
  cudaStream_t st = nullptr;
  CK(cudaMemcpyAsync(d_q, queries, (size_t)Q * dim * 4, cudaMemcpyHostToDevice, st));
  if (n) CK(cudaMemcpyAsync(d_v, vectors, (size_t)n * dim * 4, cudaMemcpyHostToDevice, st));
  if (validity_or_null && n) CK(cudaMemcpyAsync(d_m, validity_or_null, (size_t)(n + 7) / 8, cudaMemcpyHostToDevice, st));
  rc = flat_topk_device_impl(metric, dim, d_q, Q, d_v, validity_or_null ? d_m : nullptr, n, k, lower_or_null != nullptr,
                             lower_or_null ? *lower_or_null : 0.f, upper_or_null != nullptr,
                             upper_or_null ? *upper_or_null : 0.f, d_od, d_oi, d_oc, st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out_dist, d_od, (size_t)Q * k * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_idx, d_oi, (size_t)Q * k * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_count, d_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

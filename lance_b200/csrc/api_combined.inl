// api_combined.inl -- host-pointer entry points for the steps the scanner runs AROUND one index
// (included by api.cu): the union of several delta indices' results and the flat re-rank of a
// candidate set whose vectors the caller fetched (refine on the drop-in path, knn_combined).

namespace {

__global__ void iota_u64_kernel(uint64_t* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = i;
}

}  // namespace

extern "C" int32_t lb200_merge_topk(int32_t device_id, const float* dist, const uint64_t* rowid,
                                    const float* exact_or_null, const uint32_t* count, uint32_t G, uint32_t Q,
                                    uint32_t stride, uint32_t kprime, uint32_t k, const float* lower_or_null,
                                    const float* upper_or_null, float* out_dist, uint64_t* out_rowid,
                                    uint32_t* out_count) {
  int rc = select_device(device_id);
  if (rc) return rc;
  if (!G || !Q) return 0;
  if (!dist || !rowid || !count || !out_dist || !out_rowid || !out_count) return set_err(LB200_EINVAL, "null argument");
  if (!stride || !k) return set_err(LB200_EINVAL, "stride and k must be > 0");
  const size_t n = (size_t)G * Q * stride;
  const size_t b4 = Arena::pad(n * 4), b8 = Arena::pad(n * 8), bc = Arena::pad((size_t)G * Q * 4);
  const size_t bod = Arena::pad((size_t)Q * k * 4), boi = Arena::pad((size_t)Q * k * 8), boc = Arena::pad((size_t)Q * 4);
  HostStage hs;
  CK(cudaMalloc(&hs.p, b4 * 2 + b8 + bc + bod + boi + boc));
  char* b = (char*)hs.p;
  float* d_d = (float*)b; b += b4;
  float* d_e = (float*)b; b += b4;
  uint64_t* d_r = (uint64_t*)b; b += b8;
  uint32_t* d_c = (uint32_t*)b; b += bc;
  float* d_od = (float*)b; b += bod;
  uint64_t* d_oi = (uint64_t*)b; b += boi;
  uint32_t* d_oc = (uint32_t*)b;
  cudaStream_t st = nullptr;
  CK(cudaMemcpyAsync(d_d, dist, n * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_r, rowid, n * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_c, count, (size_t)G * Q * 4, cudaMemcpyHostToDevice, st));
  if (exact_or_null) CK(cudaMemcpyAsync(d_e, exact_or_null, n * 4, cudaMemcpyHostToDevice, st));
  rc = lb200_merge_topk_device(device_id, d_d, d_r, exact_or_null ? d_e : nullptr, d_c, G, Q, stride, kprime, k,
                               lower_or_null, upper_or_null, d_od, d_oi, d_oc, (void*)st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out_dist, d_od, (size_t)Q * k * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_rowid, d_oi, (size_t)Q * k * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_count, d_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int32_t lb200_rerank(int32_t device_id, int32_t metric, uint32_t dim, const float* queries, uint32_t Q,
                                const float* cand_vectors, const uint64_t* cand_rowids, const uint32_t* cand_count,
                                uint32_t C, uint32_t k, const float* lower_or_null, const float* upper_or_null,
                                float* out_dist, uint64_t* out_rowid, uint32_t* out_count) {
  if (metric != LB200_L2 && metric != LB200_COSINE && metric != LB200_DOT) return set_err(LB200_EINVAL, "unknown metric %d", metric);
  if (!dim) return set_err(LB200_EINVAL, "dim must be > 0");
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (C > 4096) return set_err(LB200_EUNSUPPORTED, "candidate stride %u > 4096", C);
  int rc = select_device(device_id);
  if (rc) return rc;
  if (!Q) return 0;
  if (!queries || !cand_count || !out_dist || !out_rowid || !out_count || (C && (!cand_vectors || !cand_rowids)))
    return set_err(LB200_EINVAL, "null argument");
  for (uint32_t q = 0; q < Q; q++)
    if (cand_count[q] > C) return set_err(LB200_EINVAL, "cand_count[%u]=%u exceeds the candidate stride %u", q, cand_count[q], C);
  const size_t n = (size_t)Q * std::max(C, 1u);
  const size_t bq = Arena::pad((size_t)Q * dim * 4), bv = Arena::pad(n * dim * 4), b8 = Arena::pad(n * 8), b4 = Arena::pad(n * 4);
  const size_t bc = Arena::pad((size_t)Q * 4), bod = Arena::pad((size_t)Q * k * 4), boi = Arena::pad((size_t)Q * k * 8);
  HostStage hs;
  CK(cudaMalloc(&hs.p, bq + bv + b8 * 2 + b4 + bc * 2 + bod + boi));
  char* b = (char*)hs.p;
  float* d_q = (float*)b; b += bq;
  float* d_v = (float*)b; b += bv;
  uint64_t* d_r = (uint64_t*)b; b += b8;
  uint64_t* d_pos = (uint64_t*)b; b += b8;
  float* d_e = (float*)b; b += b4;
  uint32_t* d_c = (uint32_t*)b; b += bc;
  uint32_t* d_oc = (uint32_t*)b; b += bc;
  float* d_od = (float*)b; b += bod;
  uint64_t* d_oi = (uint64_t*)b;
  cudaStream_t st = nullptr;
  CK(cudaMemcpyAsync(d_q, queries, (size_t)Q * dim * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_c, cand_count, (size_t)Q * 4, cudaMemcpyHostToDevice, st));
  if (C) {
    CK(cudaMemcpyAsync(d_v, cand_vectors, (size_t)Q * C * dim * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_r, cand_rowids, (size_t)Q * C * 8, cudaMemcpyHostToDevice, st));
    iota_u64_kernel<<<(uint32_t)(((size_t)Q * C + 255) / 256), 256, 0, st>>>(d_pos, (size_t)Q * C);
    CK(cudaGetLastError());
    // exact distances with the ORIGINAL query (flat.rs:28-86), then bounds + NOT NULL + order by (distance,row id)
    refine_distance_kernel<<<(uint32_t)(((size_t)Q * C + 127) / 128), 128, 0, st>>>(metric, d_q, d_v, dim, d_pos, d_c, Q, C,
                                                                                  d_e);
    CK(cudaGetLastError());
  }
  rerank_kernel<<<(Q * 32 + 127) / 128, 128, 0, st>>>(d_e, d_r, d_c, Q, C, k, lower_or_null != nullptr,
                                                    lower_or_null ? *lower_or_null : 0.f, upper_or_null != nullptr,
                                                    upper_or_null ? *upper_or_null : 0.f, d_od, d_oi, d_oc);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out_dist, d_od, (size_t)Q * k * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_rowid, d_oi, (size_t)Q * k * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_count, d_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

// normalize_arrow (rust/lance-linalg/src/kernels.rs:119-142) for a batch of keys: what the ANN exec nodes apply to the
// key of a cosine index before find_partitions / search_in_partition (rust/lance/src/io/exec/knn.rs:366-369,616-619).
extern "C" int32_t lb200_normalize(int32_t device_id, uint32_t dim, const float* queries, uint32_t Q, float* out) {
  if (!dim) return set_err(LB200_EINVAL, "dim must be > 0");
  int rc = select_device(device_id);
  if (rc) return rc;
  if (!Q) return 0;
  if (!queries || !out) return set_err(LB200_EINVAL, "null argument");
  HostStage hs;
  const size_t b = Arena::pad((size_t)Q * dim * 4);
  CK(cudaMalloc(&hs.p, 2 * b));
  float* d_in = (float*)hs.p;
  float* d_out = (float*)((char*)hs.p + b);
  cudaStream_t st = nullptr;
  CK(cudaMemcpyAsync(d_in, queries, (size_t)Q * dim * 4, cudaMemcpyHostToDevice, st));
  normalize_kernel<<<(Q * 32 + 255) / 256, 256, 0, st>>>(d_in, d_out, Q, dim);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out, d_out, (size_t)Q * dim * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

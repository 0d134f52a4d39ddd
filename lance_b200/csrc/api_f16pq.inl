// api_f16pq.inl -- IVF_PQ handles over Float16 vectors (included by api.cu).
// Replaces the reference's generic path with T = f16: kmeans_find_partitions::<f16> (kmeans.rs:581-586),
// PQDistCalculator::new over a Float16 codebook (pq/storage.rs:500-512) -> build_distance_table_{l2,dot}::<f16>
// (pq/distance.rs:17-85), residual as a Float16 array (v2.rs:285-301), refine on the Float16 column.
// Kernels and what differs from f32: csrc/f16pq.cuh.  Everything after the ADC table is the f32 pipeline.

namespace {

// IEEE binary16 -> f32 on the host (exact)
static float half_bits_to_float(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, exp = (h >> 10) & 31u, man = h & 1023u;
  uint32_t b;
  if (exp == 0) {
    if (man == 0) b = sign;
    else {  // subnormal: normalise
      int e = -1;
      uint32_t m = man;
      do { m <<= 1; e++; } while (!(m & 1024u));
      b = sign | ((uint32_t)(127 - 15 - e) << 23) | ((m & 1023u) << 13);
    }
  } else if (exp == 31) b = sign | 0x7f800000u | (man << 13);
  else b = sign | ((exp + 112u) << 23) | (man << 13);
  float f;
  memcpy(&f, &b, 4);
  return f;
}

}  // namespace

extern "C" int32_t lb200_index_create_f16(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist, const uint16_t* centroids,
                                          uint32_t num_sub_vectors, const uint16_t* codebook, const uint64_t* part_offsets,
                                          const uint8_t* codes, const uint64_t* row_ids, const uint16_t* raw_vectors_or_null,
                                          lb200_index** out) {
  if (!out) return set_err(LB200_EINVAL, "out is null");
  *out = nullptr;
  if (!centroids || !codebook || !part_offsets) return set_err(LB200_EINVAL, "null index buffers");
  if (!dim || !nlist || !num_sub_vectors || dim % num_sub_vectors)
    return set_err(LB200_EINVAL, "bad shape dim=%u nlist=%u M=%u", dim, nlist, num_sub_vectors);
  if (num_sub_vectors % 16)
    return set_err(LB200_EUNSUPPORTED, "Float16 IVF_PQ handles need num_sub_vectors %% 16 == 0 (got %u)", num_sub_vectors);
  std::vector<float> c32, cb32;
  try {
    c32.resize((size_t)nlist * dim);
    cb32.resize((size_t)256 * dim);
  } catch (...) {
    return set_err(LB200_ENOMEM, "out of host memory widening the index");
  }
  for (size_t i = 0; i < c32.size(); i++) c32[i] = half_bits_to_float(centroids[i]);
  for (size_t i = 0; i < cb32.size(); i++) cb32[i] = half_bits_to_float(codebook[i]);
  lb200_index* ix = nullptr;
  int rc = lb200_index_create(device_id, metric, dim, nlist, c32.data(), num_sub_vectors, 8, cb32.data(), part_offsets, codes, row_ids,
                              nullptr, &ix);
  if (rc) return rc;
  ix->f16 = true;
  if (!ix->d_codes3) {
    lb200_index_destroy(ix);
    return set_err(LB200_EUNSUPPORTED, "Float16 IVF_PQ handles need the skew-stream code layout");
  }
  const uint64_t N = part_offsets[nlist];
  if (raw_vectors_or_null && N) {
    cudaError_t e = cudaMalloc(&ix->d_raw16, (size_t)N * dim * 2);
    if (e == cudaSuccess) e = cudaMemcpy(ix->d_raw16, raw_vectors_or_null, (size_t)N * dim * 2, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      lb200_index_destroy(ix);
      return set_err(e == cudaErrorMemoryAllocation ? LB200_ENOMEM : LB200_ECUDA, "raw f16 vectors: %s", cudaGetErrorString(e));
    }
  }
  *out = ix;
  return 0;
}

// VectorIndex::search (ivf/v2.rs:368-401) + the scanner's refine for a batch of Float16 keys (IEEE binary16 bit patterns).
// Outputs as lb200_search: [Q][k] when refined, else [Q][k * max(refine_factor, 1)].
extern "C" int32_t lb200_search_f16(lb200_index* ix, const uint16_t* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                                    uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                                    float* out_dist, uint64_t* out_rowid, uint32_t* out_count) {
  int rc = check_queries(ix, queries, Q);
  if (rc) return rc;
  if (!ix->f16) return set_err(LB200_EINVAL, "lb200_search_f16 needs a handle made by lb200_index_create_f16");
  if (Q && (!out_dist || !out_rowid || !out_count)) return set_err(LB200_EINVAL, "null output");
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (nprobes == 0) return set_err(LB200_EINVAL, "nprobes must be > 0");
  if (!Q) return 0;
  const bool refined = refine_factor > 0 && ix->d_raw16;
  if (refine_factor > 0 && !ix->d_raw16) return set_err(LB200_EINVAL, "refine_factor>0 needs raw vectors in the index handle");
  const uint32_t ostride = refined ? k : k * std::max(refine_factor, 1u);
  CtxLease ls;
  if ((rc = ls.acquire(ix, nullptr))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;
  const uint32_t dim = ix->dim;
  // sub-batches bound the table scratch (one table per pair for L2 / cosine)
  const uint32_t P = std::min(nprobes, ix->nlist);
  const size_t per_q = (size_t)(ix->metric == LB200_DOT ? 1 : P) * ix->M * 1024;
  const uint32_t qmax = (uint32_t)std::max<size_t>(1, std::min<size_t>(Q, ((size_t)8 << 30) / per_q));
  const size_t b_q16 = Arena::pad((size_t)qmax * dim * 2), b_f32 = Arena::pad((size_t)qmax * dim * 4), no = (size_t)qmax * ostride;
  const size_t o_orig = b_q16, o_key = o_orig + b_f32, o_od = o_key + b_f32, o_or = o_od + Arena::pad(no * 4),
               o_oc = o_or + Arena::pad(no * 8), total = o_oc + Arena::pad((size_t)qmax * 4);
  if ((rc = cx.stage.reserve(total))) return rc;
  char* b = cx.stage.base;
  for (uint32_t q0 = 0; q0 < Q; q0 += qmax) {
    const uint32_t qn = std::min(qmax, Q - q0);
    CK(cudaMemcpyAsync(b, queries + (size_t)q0 * dim, (size_t)qn * dim * 2, cudaMemcpyHostToDevice, st));
    f16_keys_kernel<<<(qn + 127) / 128, 128, 0, st>>>(reinterpret_cast<const __half*>(b), qn, dim, ix->metric == LB200_COSINE,
                                                     reinterpret_cast<float*>(b + o_orig), reinterpret_cast<float*>(b + o_key));
    CK(cudaGetLastError());
    SearchArgs a = {};
    a.d_queries = reinterpret_cast<const float*>(b + o_orig);   // the ORIGINAL keys, widened: what the refine step reads
    a.d_keys_override = reinterpret_cast<const float*>(b + o_key);
    a.Q = qn; a.nprobes = nprobes; a.k = k; a.refine_factor = refine_factor;
    a.normalize = false;
    a.has_lower = lower_or_null != nullptr; a.has_upper = upper_or_null != nullptr;
    a.lower = lower_or_null ? *lower_or_null : 0.f; a.upper = upper_or_null ? *upper_or_null : 0.f;
    a.d_out_dist = reinterpret_cast<float*>(b + o_od);
    a.d_out_rowid = reinterpret_cast<uint64_t*>(b + o_or);
    a.d_out_count = reinterpret_cast<uint32_t*>(b + o_oc);
    if ((rc = search_pipeline(ix, cx, a, st))) return rc;
    CK(cudaMemcpyAsync(out_dist + (size_t)q0 * ostride, b + o_od, (size_t)qn * ostride * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(out_rowid + (size_t)q0 * ostride, b + o_or, (size_t)qn * ostride * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(out_count + q0, b + o_oc, (size_t)qn * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  }
  return 0;
}

// VectorIndex::find_partitions for Float16 keys (kmeans_find_partitions::<f16>): keys used as given -- the exec node
// has normalised them for a cosine index (knn.rs:366-369), in f16 arithmetic.
extern "C" int32_t lb200_find_partitions_f16(lb200_index* ix, const uint16_t* queries, uint32_t Q, uint32_t nprobes,
                                             uint32_t* out_part_ids, float* out_part_dists_or_null) {
  int rc = check_queries(ix, queries, Q);
  if (rc) return rc;
  if (!ix->f16) return set_err(LB200_EINVAL, "lb200_find_partitions_f16 needs a handle made by lb200_index_create_f16");
  if (!out_part_ids) return set_err(LB200_EINVAL, "null output");
  const uint32_t P = std::min(nprobes, ix->nlist);
  if (!Q || !P) return 0;
  CtxLease ls;
  if ((rc = ls.acquire(ix, nullptr))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;
  const uint32_t dim = ix->dim;
  const size_t np = (size_t)Q * P * 4;
  const size_t need = Arena::pad((size_t)Q * dim * 2) + 2 * Arena::pad((size_t)Q * dim * 4) + Arena::pad((size_t)Q * ix->nlist * 4) +
                      2 * Arena::pad(np) + 4096;
  if ((rc = cx.arena.reserve(need))) return rc;
  cx.arena.reset();
  __half* d_q16 = cx.arena.take<__half>((size_t)Q * dim);
  float* d_orig = cx.arena.take<float>((size_t)Q * dim);
  float* d_key = cx.arena.take<float>((size_t)Q * dim);
  float* d_D = cx.arena.take<float>((size_t)Q * ix->nlist);
  uint32_t* d_p = cx.arena.take<uint32_t>((size_t)Q * P);
  float* d_pd = cx.arena.take<float>((size_t)Q * P);
  uint32_t launches = 0;
  CK(cudaMemcpyAsync(d_q16, queries, (size_t)Q * dim * 2, cudaMemcpyHostToDevice, st));
  f16_keys_kernel<<<(Q + 127) / 128, 128, 0, st>>>(d_q16, Q, dim, 0, d_orig, d_key);
  CK(cudaGetLastError());
  if ((rc = coarse_device(ix, d_key, Q, P, d_D, d_p, d_pd, st, &launches))) return rc;
  CK(cudaMemcpyAsync(out_part_ids, d_p, np, cudaMemcpyDeviceToHost, st));
  if (out_part_dists_or_null) CK(cudaMemcpyAsync(out_part_dists_or_null, d_pd, np, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

// VectorIndex::search_in_partition for explicit (query, partition) pairs on a Float16 handle: keys used as given,
// one list of up to k' rows per pair (out [Q][P][k'], counts [Q][P]); [lower, upper) as lb200_search_partitions.
extern "C" int32_t lb200_search_partitions_f16(lb200_index* ix, const uint16_t* queries, uint32_t Q, const uint32_t* part_ids,
                                               uint32_t P, uint32_t k_prime, const float* lower_or_null, const float* upper_or_null,
                                               float* out_dist, uint64_t* out_rowid, uint32_t* out_count) {
  int rc = check_queries(ix, queries, Q);
  if (rc) return rc;
  if (!ix->f16) return set_err(LB200_EINVAL, "lb200_search_partitions_f16 needs a handle made by lb200_index_create_f16");
  if (Q && P && (!part_ids || !out_dist || !out_rowid || !out_count)) return set_err(LB200_EINVAL, "null argument");
  if (k_prime == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (!Q || !P) return 0;
  const size_t npairs = (size_t)Q * P;
  if (npairs >= (1ull << 31)) return set_err(LB200_EUNSUPPORTED, "Q*P too large");
  for (size_t i = 0; i < npairs; i++)
    if (part_ids[i] >= ix->nlist) return set_err(LB200_EINVAL, "partition centroid %u does not exist", part_ids[i]);
  CtxLease ls;
  if ((rc = ls.acquire(ix, nullptr))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;
  const uint32_t dim = ix->dim;
  const size_t b_q16 = Arena::pad((size_t)Q * dim * 2), b_f32 = Arena::pad((size_t)Q * dim * 4), b_p = Arena::pad(npairs * 4),
               no = npairs * k_prime;
  const size_t o_orig = b_q16, o_key = o_orig + b_f32, o_p = o_key + b_f32, o_od = o_p + b_p, o_or = o_od + Arena::pad(no * 4),
               o_oc = o_or + Arena::pad(no * 8), total = o_oc + Arena::pad(npairs * 4);
  if ((rc = cx.stage.reserve(total))) return rc;
  char* b = cx.stage.base;
  CK(cudaMemcpyAsync(b, queries, (size_t)Q * dim * 2, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(b + o_p, part_ids, npairs * 4, cudaMemcpyHostToDevice, st));
  f16_keys_kernel<<<(Q + 127) / 128, 128, 0, st>>>(reinterpret_cast<const __half*>(b), Q, dim, 0, reinterpret_cast<float*>(b + o_orig),
                                                  reinterpret_cast<float*>(b + o_key));
  CK(cudaGetLastError());
  SearchArgs a = {};
  a.d_queries = reinterpret_cast<const float*>(b + o_orig);
  a.d_keys_override = reinterpret_cast<const float*>(b + o_key);
  a.Q = Q; a.d_part_ids = reinterpret_cast<const uint32_t*>(b + o_p); a.P = P; a.k = k_prime; a.refine_factor = 0;
  a.normalize = false; a.per_pair_output = true;
  a.has_lower = lower_or_null != nullptr; a.has_upper = upper_or_null != nullptr;
  a.lower = lower_or_null ? *lower_or_null : 0.f; a.upper = upper_or_null ? *upper_or_null : 0.f;
  a.d_out_dist = reinterpret_cast<float*>(b + o_od);
  a.d_out_rowid = reinterpret_cast<uint64_t*>(b + o_or);
  a.d_out_count = reinterpret_cast<uint32_t*>(b + o_oc);
  if ((rc = search_pipeline(ix, cx, a, st))) return rc;
  CK(cudaMemcpyAsync(out_dist, b + o_od, no * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_rowid, b + o_or, no * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_count, b + o_oc, npairs * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

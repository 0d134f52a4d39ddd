// api_build.inl -- GPU index build behind the C ABI (SURVEY 8(f)-4 "GPU index build"; included by api.cu).
// k-means training (IVF centroids, PQ codebooks), partition assignment, residuals and PQ encoding -- the steps
// rust/lance-linalg/src/kmeans.rs:208-300,440-525, lance-index/src/vector/pq.rs:120-190 and the reference's torch path
// (python/python/lance/torch/kmeans.py, lance/vector.py:134-700) run on the CPU / through torch.  Assignment uses the
// query path's own coarse kernels, so a vector is encoded into exactly the partition the search will look for it in.
// Deterministic: given the initial centroids every step is bit-identical to the CPU restatement the tests hold
// (lo_kmeans_fit, lo_pq_quantize; tests/test_gpu_build.py).  The random initialisation is
// unpinned in the reference (SmallRng::from_entropy); here it is a seeded choice of k distinct rows.

#include <random>

#include "build.cuh"

namespace {

struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  int alloc(size_t bytes) {
    cudaError_t e = cudaMalloc(&p, std::max<size_t>(bytes, 16));
    if (e != cudaSuccess) return set_err(LB200_ENOMEM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    return 0;
  }
  template <typename T> T* as() { return reinterpret_cast<T*>(p); }
};

// argmin_value_float (kernels.rs:79-89) keeps None when no distance is < +inf (NaN / +inf rows)
__global__ void mark_none_kernel(uint32_t* __restrict__ ids, float* __restrict__ dists, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (!(dists[i] < __int_as_float(0x7f800000))) {
    ids[i] = 0xffffffffu;
    dists[i] = 0.0f;
  }
}

static uint32_t assign_tile_rows(uint64_t n, uint32_t k) {
  // <= 128k rows per tile: the coarse kernels put the query tiles on grid.y (<= 65535 blocks of >= 4 queries)
  return (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>({n, 131072ull, std::max<uint64_t>(1024, (256ull << 20) / ((uint64_t)k * 4))}));
}

// ids[i] = argmin_c dist(vec_i, centroid_c) (first minimum; 0xffffffff = None), dists[i] = that distance.
// d_D: scratch of assign_tile_rows(n, k) * k floats.
static int assign_device(int device, int metric, uint32_t dim, const float* d_cent, uint32_t k, const float* d_vec, uint64_t n,
                         uint32_t* d_ids, float* d_dists, float* d_D, cudaStream_t st) {
  lb200_index tix;  // the coarse step reads only these fields
  tix.device = device;
  tix.metric = metric == LB200_DOT ? LB200_DOT : LB200_L2;  // cosine: callers normalise, then L2 (ivf/v2.rs:403-411)
  tix.dim = dim;
  tix.nlist = k;
  tix.d_centroids = const_cast<float*>(d_cent);
  int optin = 0;
  CK(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
  tix.smem_optin = (size_t)optin;
  const uint32_t tile = assign_tile_rows(n, k);
  uint32_t launches = 0;
  for (uint64_t r0 = 0; r0 < n; r0 += tile) {
    const uint32_t q = (uint32_t)std::min<uint64_t>(tile, n - r0);
    int rc = coarse_device(&tix, d_vec + r0 * dim, q, 1, d_D, d_ids + r0, d_dists + r0, st, &launches);
    if (rc) return rc;
  }
  mark_none_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, st>>>(d_ids, d_dists, n);
  CK(cudaGetLastError());
  return 0;
}

// KMeans::train_kmeans (kmeans.rs:440-525) on device rows; d_cent: in = initial centroids, out = trained.
static int kmeans_fit_device(int device, int metric, uint32_t dim, const float* d_vec, uint64_t n, uint32_t k, uint32_t max_iters,
                             double tolerance, float* d_cent, uint32_t* out_iters, double* out_loss, cudaStream_t st) {
  if (n >= (1ull << 32)) return set_err(LB200_EUNSUPPORTED, "k-means on %llu rows >= 2^32", (unsigned long long)n);
  DevBuf b_ids, b_dists, b_D, b_order, b_off;
  int rc;
  if ((rc = b_ids.alloc(n * 4)) || (rc = b_dists.alloc(n * 4)) || (rc = b_D.alloc((size_t)assign_tile_rows(n, k) * k * 4)) ||
      (rc = b_order.alloc(n * 4)) || (rc = b_off.alloc(((size_t)k + 1) * 4)))
    return rc;
  std::vector<uint32_t> ids(n), order(n), off(k + 1), fill(k);
  std::vector<float> dists(n), cent;
  std::vector<uint64_t> cnts(k);
  const float one_eps = 1.0f + 1.1920929e-7f;
  double loss = 1.7976931348623157e308, last_loss = 0.0;
  uint32_t it = 1;
  for (; it <= std::max(max_iters, 1u); it++) {
    if ((rc = assign_device(device, metric, dim, d_cent, k, d_vec, n, b_ids.as<uint32_t>(), b_dists.as<float>(), b_D.as<float>(), st)))
      return rc;
    CK(cudaMemcpyAsync(ids.data(), b_ids.p, n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(dists.data(), b_dists.p, n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    last_loss = 0.0;  // sequential f64 sum of the f32 minima, in data order
    for (uint64_t i = 0; i < n; i++) last_loss += (double)dists[i];
    std::fill(cnts.begin(), cnts.end(), 0);
    for (uint64_t i = 0; i < n; i++)
      if (ids[i] != 0xffffffffu) cnts[ids[i]]++;
    uint32_t acc = 0;
    bool empty = false;
    for (uint32_t c = 0; c < k; c++) {
      off[c] = acc;
      fill[c] = acc;
      acc += (uint32_t)cnts[c];
      empty = empty || cnts[c] == 0;
    }
    off[k] = acc;
    for (uint64_t i = 0; i < n; i++)  // stable: members of a cluster in ascending row order
      if (ids[i] != 0xffffffffu) order[fill[ids[i]]++] = (uint32_t)i;
    CK(cudaMemcpyAsync(b_order.p, order.data(), n * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(b_off.p, off.data(), ((size_t)k + 1) * 4, cudaMemcpyHostToDevice, st));
    kmeans_update_kernel<<<k, (uint32_t)std::min<uint32_t>(256, std::max<uint32_t>(32, (dim + 31) / 32 * 32)), 0, st>>>(
        d_vec, dim, b_order.as<uint32_t>(), b_off.as<uint32_t>(), d_cent);
    CK(cudaGetLastError());
    if (empty) {  // split_clusters (kmeans.rs:135-148): k * dim floats through the host, only when a cluster is empty
      cent.resize((size_t)k * dim);
      CK(cudaMemcpyAsync(cent.data(), d_cent, cent.size() * 4, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      for (uint32_t i = 0; i < k; i++) {
        if (cnts[i] != 0) continue;
        uint32_t largest = 0;
        uint64_t mx = 0;
        for (uint32_t j = 0; j < k; j++)
          if (cnts[j] > mx) { mx = cnts[j]; largest = j; }
        cnts[i] = cnts[largest] / 2;
        cnts[largest] /= 2;
        for (uint32_t j = 0; j < dim; j++) {
          cent[(size_t)i * dim + j] = cent[(size_t)largest * dim + j] * one_eps;
          cent[(size_t)largest * dim + j] /= one_eps;
        }
      }
      CK(cudaMemcpyAsync(d_cent, cent.data(), cent.size() * 4, cudaMemcpyHostToDevice, st));
    }
    CK(cudaStreamSynchronize(st));
    if (std::fabs(loss - last_loss) / last_loss < tolerance) break;
    loss = last_loss;
  }
  if (out_iters) *out_iters = std::min(it, std::max(max_iters, 1u));
  if (out_loss) *out_loss = last_loss;
  return 0;
}

// k distinct rows of [n][dim] (host or device memory) -> d_cent, by a seeded partial Fisher-Yates
static int kmeans_random_init(const float* vec, bool vec_on_device, uint64_t n, uint32_t dim, uint32_t k, uint64_t seed, float* d_cent,
                              cudaStream_t st) {
  std::vector<uint32_t> idx(n);
  for (uint64_t i = 0; i < n; i++) idx[i] = (uint32_t)i;
  std::mt19937_64 rng(seed);
  for (uint32_t i = 0; i < k; i++) {
    const uint64_t j = i + rng() % (n - i);
    std::swap(idx[i], idx[j]);
    CK(cudaMemcpyAsync(d_cent + (size_t)i * dim, vec + (size_t)idx[i] * dim, (size_t)dim * 4,
                       vec_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
  }
  CK(cudaStreamSynchronize(st));
  return 0;
}

static int build_check(int32_t metric, uint32_t dim, uint64_t n, uint32_t k) {
  if (metric != LB200_L2 && metric != LB200_COSINE && metric != LB200_DOT) return set_err(LB200_EINVAL, "unknown metric %d", metric);
  if (!dim || !k) return set_err(LB200_EINVAL, "dim and k must be > 0");
  if (k > 8192) return set_err(LB200_EUNSUPPORTED, "k=%u > 8192 centroids", k);
  if (n >= (1ull << 32)) return set_err(LB200_EUNSUPPORTED, "n >= 2^32 rows");
  return 0;
}

}  // namespace

extern "C" int32_t lb200_kmeans_train_device(int32_t device_id, int32_t metric, uint32_t dim, const float* d_vectors, uint64_t n,
                                             uint32_t k, uint32_t max_iters, float tolerance, uint64_t seed,
                                             const float* init_centroids_or_null /*host*/, float* out_centroids /*host*/,
                                             uint32_t* out_iters_or_null, double* out_loss_or_null) {
  int rc = build_check(metric, dim, n, k);
  if (rc) return rc;
  if (!d_vectors || !out_centroids) return set_err(LB200_EINVAL, "null argument");
  // KMeans::new_with_params (kmeans.rs:533-541)
  if (n < k) return set_err(LB200_EINVAL, "KMeans: training does not have sufficient data points: n(%llu) is smaller than k(%u)", (unsigned long long)n, k);
  if ((rc = select_device(device_id))) return rc;
  DevBuf b_cent;
  if ((rc = b_cent.alloc((size_t)k * dim * 4))) return rc;
  cudaStream_t st = nullptr;
  if (init_centroids_or_null) CK(cudaMemcpyAsync(b_cent.p, init_centroids_or_null, (size_t)k * dim * 4, cudaMemcpyHostToDevice, st));
  else if ((rc = kmeans_random_init(d_vectors, true, n, dim, k, seed, b_cent.as<float>(), st))) return rc;
  try {
    rc = kmeans_fit_device(device_id, metric, dim, d_vectors, n, k, max_iters, (double)tolerance, b_cent.as<float>(), out_iters_or_null,
                           out_loss_or_null, st);
  } catch (const std::bad_alloc&) {
    return set_err(LB200_ENOMEM, "out of host memory in k-means");
  }
  if (rc) return rc;
  CK(cudaMemcpy(out_centroids, b_cent.p, (size_t)k * dim * 4, cudaMemcpyDeviceToHost));
  return 0;
}

extern "C" int32_t lb200_kmeans_train(int32_t device_id, int32_t metric, uint32_t dim, const float* vectors, uint64_t n, uint32_t k,
                                      uint32_t max_iters, float tolerance, uint64_t seed, const float* init_centroids_or_null,
                                      float* out_centroids, uint32_t* out_iters_or_null, double* out_loss_or_null) {
  int rc = build_check(metric, dim, n, k);
  if (rc) return rc;
  if (!vectors || !out_centroids) return set_err(LB200_EINVAL, "null argument");
  if ((rc = select_device(device_id))) return rc;
  DevBuf b_vec;
  if ((rc = b_vec.alloc((size_t)n * dim * 4))) return rc;
  CK(cudaMemcpy(b_vec.p, vectors, (size_t)n * dim * 4, cudaMemcpyHostToDevice));
  return lb200_kmeans_train_device(device_id, metric, dim, b_vec.as<float>(), n, k, max_iters, tolerance, seed, init_centroids_or_null,
                                   out_centroids, out_iters_or_null, out_loss_or_null);
}

extern "C" int32_t lb200_compute_partitions_device(int32_t device_id, int32_t metric, uint32_t dim, const float* d_centroids, uint32_t k,
                                                   const float* d_vectors, uint64_t n, uint32_t* d_out_ids,
                                                   float* d_out_dists_or_null, void* stream) {
  int rc = build_check(metric, dim, n, k);
  if (rc) return rc;
  if (!n) return 0;
  if (!d_centroids || !d_vectors || !d_out_ids) return set_err(LB200_EINVAL, "null argument");
  if ((rc = select_device(device_id))) return rc;
  DevBuf b_D, b_dists;
  if ((rc = b_D.alloc((size_t)assign_tile_rows(n, k) * k * 4))) return rc;
  float* dists = d_out_dists_or_null;
  if (!dists) {
    if ((rc = b_dists.alloc(n * 4))) return rc;
    dists = b_dists.as<float>();
  }
  cudaStream_t st = (cudaStream_t)stream;
  if ((rc = assign_device(device_id, metric, dim, d_centroids, k, d_vectors, n, d_out_ids, dists, b_D.as<float>(), st))) return rc;
  CK(cudaStreamSynchronize(st));  // the scratch is freed on return
  return 0;
}

extern "C" int32_t lb200_compute_partitions(int32_t device_id, int32_t metric, uint32_t dim, const float* centroids, uint32_t k,
                                            const float* vectors, uint64_t n, uint32_t* out_ids, float* out_dists_or_null) {
  int rc = build_check(metric, dim, n, k);
  if (rc) return rc;
  if (!n) return 0;
  if (!centroids || !vectors || !out_ids) return set_err(LB200_EINVAL, "null argument");
  if ((rc = select_device(device_id))) return rc;
  DevBuf b_c, b_v, b_i, b_d;
  if ((rc = b_c.alloc((size_t)k * dim * 4)) || (rc = b_v.alloc((size_t)n * dim * 4)) || (rc = b_i.alloc(n * 4)) || (rc = b_d.alloc(n * 4)))
    return rc;
  CK(cudaMemcpy(b_c.p, centroids, (size_t)k * dim * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(b_v.p, vectors, (size_t)n * dim * 4, cudaMemcpyHostToDevice));
  if ((rc = lb200_compute_partitions_device(device_id, metric, dim, b_c.as<float>(), k, b_v.as<float>(), n, b_i.as<uint32_t>(),
                                            b_d.as<float>(), nullptr)))
    return rc;
  CK(cudaMemcpy(out_ids, b_i.p, n * 4, cudaMemcpyDeviceToHost));
  if (out_dists_or_null) CK(cudaMemcpy(out_dists_or_null, b_d.p, n * 4, cudaMemcpyDeviceToHost));
  return 0;
}

extern "C" int32_t lb200_residuals_device(int32_t device_id, uint32_t dim, const float* d_vectors, const uint32_t* d_part_ids,
                                          const float* d_centroids, uint64_t n, float* d_out, void* stream) {
  if (!dim) return set_err(LB200_EINVAL, "dim must be > 0");
  if (!n) return 0;
  if (!d_vectors || !d_part_ids || !d_centroids || !d_out) return set_err(LB200_EINVAL, "null argument");
  int rc = select_device(device_id);
  if (rc) return rc;
  const uint64_t tot = n * dim;
  residual_kernel<<<(uint32_t)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_vectors, d_part_ids, d_centroids, n, dim, d_out);
  CK(cudaGetLastError());
  return 0;
}

extern "C" int32_t lb200_pq_encode_device(int32_t device_id, int32_t metric, uint32_t dim, uint32_t M, const float* d_codebook,
                                          const float* d_vectors, uint64_t n, uint8_t* d_out_codes, void* stream) {
  if (metric != LB200_L2 && metric != LB200_COSINE && metric != LB200_DOT) return set_err(LB200_EINVAL, "unknown metric %d", metric);
  if (!dim || !M || dim % M) return set_err(LB200_EINVAL, "dimension %u is not divisible by num_sub_vectors %u", dim, M);
  if (!n) return 0;
  if (!d_codebook || !d_vectors || !d_out_codes) return set_err(LB200_EINVAL, "null argument");
  if (M > 65535) return set_err(LB200_EUNSUPPORTED, "num_sub_vectors %u > 65535", M);
  int rc = select_device(device_id);
  if (rc) return rc;
  const uint32_t dsub = dim / M;
  const size_t smem = (size_t)256 * dsub * 4;
  const int residual = metric != LB200_DOT;  // L2 / cosine codebooks quantise with l2, dot with 1 - x.y (pq.rs:120-190)
  dim3 grid((uint32_t)((n + 255) / 256), M);
  cudaStream_t st = (cudaStream_t)stream;
#define LB200_ENC(D)                                                                                            \
  do {                                                                                                          \
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute(pq_encode_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    pq_encode_kernel<D><<<grid, 256, smem, st>>>(d_vectors, d_codebook, n, dim, M, dsub, residual, d_out_codes); \
  } while (0)
  if (dsub == 8) LB200_ENC(8);
  else if (dsub == 16) LB200_ENC(16);
  else LB200_ENC(0);
#undef LB200_ENC
  CK(cudaGetLastError());
  return 0;
}

extern "C" int32_t lb200_pq_encode(int32_t device_id, int32_t metric, uint32_t dim, uint32_t M, const float* codebook,
                                   const float* vectors, uint64_t n, uint8_t* out_codes) {
  if (!dim || !M || dim % M) return set_err(LB200_EINVAL, "dimension %u is not divisible by num_sub_vectors %u", dim, M);
  if (!n) return 0;
  if (!codebook || !vectors || !out_codes) return set_err(LB200_EINVAL, "null argument");
  int rc = select_device(device_id);
  if (rc) return rc;
  DevBuf b_cb, b_v, b_o;
  if ((rc = b_cb.alloc((size_t)256 * dim * 4)) || (rc = b_v.alloc((size_t)n * dim * 4)) || (rc = b_o.alloc((size_t)n * M))) return rc;
  CK(cudaMemcpy(b_cb.p, codebook, (size_t)256 * dim * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(b_v.p, vectors, (size_t)n * dim * 4, cudaMemcpyHostToDevice));
  if ((rc = lb200_pq_encode_device(device_id, metric, dim, M, b_cb.as<float>(), b_v.as<float>(), n, b_o.as<uint8_t>(), nullptr))) return rc;
  CK(cudaMemcpy(out_codes, b_o.p, (size_t)n * M, cudaMemcpyDeviceToHost));
  return 0;
}

// PQ codebook training (lance-index/src/vector/pq/builder.rs: one k-means with k = 256 per sub-space on the
// sub-vector slices of the training rows).  init_codebook_or_null: [M][256][dim/M] host (tests), else seeded rows.
extern "C" int32_t lb200_pq_train_device(int32_t device_id, int32_t metric, uint32_t dim, uint32_t M, const float* d_vectors, uint64_t n,
                                         uint32_t max_iters, float tolerance, uint64_t seed, const float* init_codebook_or_null,
                                         float* out_codebook /*host [M][256][dim/M]*/) {
  if (!dim || !M || dim % M) return set_err(LB200_EINVAL, "dimension %u is not divisible by num_sub_vectors %u", dim, M);
  if (!d_vectors || !out_codebook) return set_err(LB200_EINVAL, "null argument");
  if (n < 256) return set_err(LB200_EINVAL, "KMeans: training does not have sufficient data points: n(%llu) is smaller than k(256)", (unsigned long long)n);
  int rc = select_device(device_id);
  if (rc) return rc;
  const uint32_t dsub = dim / M;
  DevBuf b_slice;
  if ((rc = b_slice.alloc((size_t)n * dsub * 4))) return rc;
  for (uint32_t m = 0; m < M; m++) {
    const uint64_t tot = n * dsub;
    slice_columns_kernel<<<(uint32_t)((tot + 255) / 256), 256>>>(d_vectors, n, dim, m * dsub, dsub, b_slice.as<float>());
    CK(cudaGetLastError());
    rc = lb200_kmeans_train_device(device_id, metric, dsub, b_slice.as<float>(), n, 256, max_iters, tolerance, seed + m,
                                   init_codebook_or_null ? init_codebook_or_null + (size_t)m * 256 * dsub : nullptr,
                                   out_codebook + (size_t)m * 256 * dsub, nullptr, nullptr);
    if (rc) return rc;
  }
  return 0;
}

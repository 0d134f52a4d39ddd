// api_noalloc.inl -- lb200_flat_topk_device_noalloc: flat kNN top-k that allocates NOTHING (included by api.cu).
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is
// lb200_flat_topk_device (compute_distance + SortExec(fetch = k) + NOT NULL, flat.rs:28-86, scanner.rs:2066-2136)
// with the device scratch -- the [Q][n] distance block and the per-segment candidate lists -- taken from a file-scope
// `static __device__` array instead of the grow-only cudaMalloc'd cache of api_flat.inl.  Same kernels, same order of
// operations, so the results are bit-identical to lb200_flat_topk_device.
//
//   * LB200_NOALLOC_SCRATCH_BYTES (32 MiB) is the limit: a call whose scratch need exceeds it fails with LB200_EINVAL
//     and a message that states both numbers.  Every GPU that loads the library holds the array until exit.
//   * In HOST code the array's name is a host-side stand-in, not a device address: the address is resolved once per
//     device, with that device current, by cudaGetSymbolAddress (below) and it is that pointer the kernels get.
//   * The entry point calls no cudaMalloc / cudaFree / allocating library routine, is asynchronous on `stream` and
//     does not synchronise.  (The first call on a device resolves the symbol address: make one call outside a graph
//     capture.)
//   * Every call shares the ONE array, which keeps the last call's contents; what must start at zero (the segment
//     counts) is cleared by the call itself.  Calls that overlap on the GPU -- two streams, two host threads -- RACE
//     on it: the caller must order them.

constexpr size_t LB200_NOALLOC_SCRATCH_BYTES = (size_t)32 << 20;
static __device__ __align__(256) unsigned char g_noalloc_scratch[LB200_NOALLOC_SCRATCH_BYTES];

namespace {

struct NoAllocDev {
  std::mutex mu;
  unsigned char* ptr[64] = {};
};
static NoAllocDev& noalloc_dev() {
  static NoAllocDev d;
  return d;
}

// Device address of g_noalloc_scratch on `dev` (which must be the current device); resolved once per device.
static int noalloc_scratch_ptr(int dev, unsigned char** out) {
  if (dev < 0 || dev >= 64) return set_err(LB200_EINVAL, "device_id %d out of range for the no-allocation scratch", dev);
  NoAllocDev& t = noalloc_dev();
  std::lock_guard<std::mutex> lk(t.mu);
  if (!t.ptr[dev]) {
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return set_err(LB200_ECUDA, "cudaGetSymbolAddress(no-allocation scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
    t.ptr[dev] = p;
  }
  *out = t.ptr[dev];
  return 0;
}

}  // namespace

extern "C" uint64_t lb200_noalloc_scratch_bytes(void) { return (uint64_t)LB200_NOALLOC_SCRATCH_BYTES; }

extern "C" int32_t lb200_flat_topk_device_noalloc(int32_t device_id, int32_t metric, uint32_t dim, const float* d_queries,
                                                  uint32_t Q, const float* d_vectors, const uint8_t* d_validity_or_null,
                                                  uint64_t n, uint32_t k, float* d_out_dist, uint64_t* d_out_idx,
                                                  uint32_t* d_out_count, void* stream) {
  int rc = flat_check(metric, dim, d_queries, Q, d_vectors, n);
  if (rc) return rc;
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (k > 1024) return set_err(LB200_EUNSUPPORTED, "k=%u > 1024", k);
  rc = select_device(device_id);
  if (rc) return rc;
  if (!Q) return 0;
  if (!d_out_dist || !d_out_idx || !d_out_count) return set_err(LB200_EINVAL, "null output");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {  // an empty batch: empty result
    fill_u32_kernel<<<(uint32_t)(((size_t)Q * k + 255) / 256), 256, 0, st>>>(reinterpret_cast<uint32_t*>(d_out_dist), 0x7f800000u,
                                                                             (size_t)Q * k);
    CK(cudaMemsetAsync(d_out_idx, 0xff, (size_t)Q * k * 8, st));
    CK(cudaMemsetAsync(d_out_count, 0, (size_t)Q * 4, st));
    CK(cudaGetLastError());
    return 0;
  }
  // the same segmentation as flat_topk_device_impl, all Q queries in one tile
  const size_t stride = (size_t)n;
  const uint32_t kp2 = std::max(2u, next_pow2(k));
  const size_t smem = (size_t)kp2 * 8;
  uint32_t nseg = (uint32_t)std::min<uint64_t>({(n + 32767) / 32768, 128ull, 16384ull / k});
  nseg = std::max(nseg, 1u);
  const uint64_t seg_len = (n + nseg - 1) / nseg;
  nseg = std::max((uint32_t)((n + seg_len - 1) / seg_len), 1u);
  const size_t cells = nseg > 1 ? (size_t)Q * nseg * k : 0;
  const size_t b_D = Arena::pad((size_t)Q * stride * 4), b_td = Arena::pad(cells * 4), b_ti = Arena::pad(cells * 8),
               b_tc = Arena::pad((size_t)Q * nseg * 4);
  const size_t need = b_D + b_td + b_ti + b_tc;
  if ((uint64_t)Q * n > (LB200_NOALLOC_SCRATCH_BYTES / 4) || need > LB200_NOALLOC_SCRATCH_BYTES)
    return set_err(LB200_EINVAL,
                   "lb200_flat_topk_device_noalloc: Q=%u x n=%llu (k=%u) needs %zu bytes of scratch, the fixed scratch holds %zu "
                   "(about Q*n*4 bytes: split the batch or use lb200_flat_topk_device)",
                   Q, (unsigned long long)n, k, need, LB200_NOALLOC_SCRATCH_BYTES);
  unsigned char* base = nullptr;
  if ((rc = noalloc_scratch_ptr(device_id, &base))) return rc;
  float* d_D = reinterpret_cast<float*>(base);
  float* t_d = reinterpret_cast<float*>(base + b_D);
  uint64_t* t_i = reinterpret_cast<uint64_t*>(base + b_D + b_td);
  uint32_t* t_c = reinterpret_cast<uint32_t*>(base + b_D + b_td + b_ti);
  // the array keeps the last call's contents: the segment counts are what a merge would trust, so they start at zero
  CK(cudaMemsetAsync(t_c, 0, (size_t)Q * nseg * 4, st));
  rc = flat_dist_launch(metric, d_queries, Q, d_vectors, d_validity_or_null, n, dim, d_D, stride, st);
  if (rc) return rc;
  if (nseg == 1) {
    flat_select_kernel<<<Q, 256, smem, st>>>(d_D, stride, n, 1, n, k, kp2, 0, 0.f, 0, 0.f, d_out_dist, d_out_idx, d_out_count);
    CK(cudaGetLastError());
    return 0;
  }
  flat_select_kernel<<<Q * nseg, 256, smem, st>>>(d_D, stride, n, nseg, seg_len, k, kp2, 0, 0.f, 0, 0.f, t_d, t_i, t_c);
  CK(cudaGetLastError());
  return launch_merge(t_d, t_i, nullptr, nullptr, t_c, Q, nseg, k, (size_t)nseg * k, k, nseg, 1, k, k, d_out_dist, d_out_idx,
                      nullptr, nullptr, d_out_count, st);
}

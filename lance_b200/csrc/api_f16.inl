// api_f16.inl -- flat kNN on Float16 vectors (included by api.cu; SURVEY 8(f)-4, the B200 analogue of the reference's
// rust/lance-linalg/src/simd/f16.c).  Rows and keys are IEEE binary16 in HBM (half the bytes of the f32 path); every
// element is widened to f32 in registers and the sums follow the reference's deterministic scalar fallbacks
// (l2_scalar::<f16,f32,16>, dot_scalar::<f16,f32,32>, norm_l2_impl::<f16,f32,32>, cosine_scalar), so results are
// bit-identical to that path and within rounding of the -ffast-math C kernels.

#include <cuda_fp16.h>

namespace {

// dot_scalar::<f16,f32,32> (dot.rs:33-61) with x already widened (shared memory) and y in f16
__device__ __forceinline__ float dot32_f16(const float* __restrict__ x, const __half* __restrict__ y, int d) {
  const int nchunk = d >> 5, rem0 = nchunk << 5;
  float s = 0.0f;
  for (int i = rem0; i < d; i++) s = __fadd_rn(s, __fmul_rn(x[i], __half2float(y[i])));
  float sums[32];
#pragma unroll
  for (int i = 0; i < 32; i++) sums[i] = 0.0f;
  for (int c = 0; c < nchunk; c++) {
#pragma unroll
    for (int i = 0; i < 32; i++) sums[i] = __fadd_rn(sums[i], __fmul_rn(x[c * 32 + i], __half2float(y[c * 32 + i])));
  }
  float t = 0.0f;
#pragma unroll
  for (int i = 0; i < 32; i++) t = __fadd_rn(t, sums[i]);
  return __fadd_rn(s, t);
}
// dot(y, y) of cosine_scalar (cosine.rs:233-238)
__device__ __forceinline__ float sq32_f16(const __half* __restrict__ y, int d) {
  const int nchunk = d >> 5, rem0 = nchunk << 5;
  float s = 0.0f;
  for (int i = rem0; i < d; i++) { const float v = __half2float(y[i]); s = __fadd_rn(s, __fmul_rn(v, v)); }
  float sums[32];
#pragma unroll
  for (int i = 0; i < 32; i++) sums[i] = 0.0f;
  for (int c = 0; c < nchunk; c++) {
#pragma unroll
    for (int i = 0; i < 32; i++) { const float v = __half2float(y[c * 32 + i]); sums[i] = __fadd_rn(sums[i], __fmul_rn(v, v)); }
  }
  float t = 0.0f;
#pragma unroll
  for (int i = 0; i < 32; i++) t = __fadd_rn(t, sums[i]);
  return __fadd_rn(s, t);
}
// l2_scalar::<f16,f32,16> (l2.rs:65-99)
__device__ __forceinline__ float l2_16_f16(const float* __restrict__ x, const __half* __restrict__ y, int d) {
  const int nchunk = d >> 4, rem0 = nchunk << 4;
  float s = 0.0f;
  for (int i = rem0; i < d; i++) { const float df = __fsub_rn(x[i], __half2float(y[i])); s = __fadd_rn(s, __fmul_rn(df, df)); }
  float sums[16];
#pragma unroll
  for (int i = 0; i < 16; i++) sums[i] = 0.0f;
  for (int c = 0; c < nchunk; c++) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      const float df = __fsub_rn(x[c * 16 + i], __half2float(y[c * 16 + i]));
      sums[i] = __fadd_rn(sums[i], __fmul_rn(df, df));
    }
  }
  float t = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; i++) t = __fadd_rn(t, sums[i]);
  return __fadd_rn(s, t);
}

// out[q][row]; one thread per row, FLAT_TQ keys per CTA widened once into shared memory; NaN for null rows
__global__ void __launch_bounds__(FLAT_THREADS)
flat_dist_f16_kernel(int metric, const __half* __restrict__ queries, uint32_t Q, const __half* __restrict__ vectors,
                     const uint8_t* __restrict__ validity, uint64_t n, uint32_t d, float* __restrict__ out,
                     size_t out_stride) {
  extern __shared__ float sq16[];  // [TQ][d] widened keys + [TQ] norms
  float* snorm = sq16 + FLAT_TQ * d;
  const uint32_t q0 = blockIdx.y * FLAT_TQ;
  for (uint32_t i = threadIdx.x; i < FLAT_TQ * d; i += FLAT_THREADS) {
    const uint32_t qq = q0 + i / d;
    sq16[i] = qq < Q ? __half2float(queries[(size_t)qq * d + (i % d)]) : 0.0f;
  }
  __syncthreads();
  if (threadIdx.x < FLAT_TQ) {  // norm_l2_impl::<f16,f32,32> of the key (cosine_batch default impl)
    float nrm = 0.0f;
    if (metric == 1) {
      const float* x = sq16 + threadIdx.x * d;
      const int nchunk = d >> 5, rem0 = nchunk << 5;
      float s = 0.0f;
      for (uint32_t i = rem0; i < d; i++) s = __fadd_rn(s, __fmul_rn(x[i], x[i]));
      float t = 0.0f;
      for (int l = 0; l < 32; l++) {
        float a = 0.0f;
        for (int c = 0; c < nchunk; c++) a = __fadd_rn(a, __fmul_rn(x[c * 32 + l], x[c * 32 + l]));
        t = __fadd_rn(t, a);
      }
      nrm = __fsqrt_rn(__fadd_rn(s, t));
    }
    snorm[threadIdx.x] = nrm;
  }
  __syncthreads();
  const uint64_t row = (uint64_t)blockIdx.x * FLAT_THREADS + threadIdx.x;
  if (row >= n) return;
  const bool valid = !validity || ((validity[row >> 3] >> (row & 7)) & 1);
  const __half* y = vectors + row * d;
  const float ysq = (valid && metric == 1) ? sq32_f16(y, (int)d) : 0.0f;
#pragma unroll
  for (int t = 0; t < FLAT_TQ; t++) {
    if (q0 + t >= Q) break;
    float v = __int_as_float(0x7fc00000);
    if (valid) {
      const float* x = sq16 + t * d;
      if (metric == 0) v = l2_16_f16(x, y, (int)d);
      else if (metric == 2) v = __fsub_rn(1.0f, dot32_f16(x, y, (int)d));
      else v = __fsub_rn(1.0f, __fdiv_rn(dot32_f16(x, y, (int)d), __fmul_rn(snorm[t], __fsqrt_rn(ysq))));
    }
    out[(size_t)(q0 + t) * out_stride + row] = v;
  }
}

static int flat_f16_launch(int metric, const __half* d_q, uint32_t Q, const __half* d_v, const uint8_t* d_valid, uint64_t n,
                           uint32_t dim, float* d_out, size_t out_stride, cudaStream_t st) {
  if (!Q || !n) return 0;
  const size_t smem = (size_t)FLAT_TQ * dim * 4 + FLAT_TQ * 4;
  if (smem > 48 * 1024) CK(cudaFuncSetAttribute(flat_dist_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (uint32_t q0 = 0; q0 < Q; q0 += 65535 * FLAT_TQ) {
    const uint32_t qn = std::min<uint32_t>(65535 * FLAT_TQ, Q - q0);
    dim3 grid((uint32_t)((n + FLAT_THREADS - 1) / FLAT_THREADS), (qn + FLAT_TQ - 1) / FLAT_TQ);
    flat_dist_f16_kernel<<<grid, FLAT_THREADS, smem, st>>>(metric, d_q + (size_t)q0 * dim, qn, d_v, d_valid, n, dim,
                                                           d_out + (size_t)q0 * out_stride, out_stride);
    CK(cudaGetLastError());
  }
  return 0;
}

}  // namespace

extern "C" int32_t lb200_flat_distance_f16(int32_t device_id, int32_t metric, uint32_t dim, const uint16_t* queries,
                                           uint32_t Q, const uint16_t* vectors, const uint8_t* validity_or_null,
                                           uint64_t n, float* out) {
  int rc = flat_check(metric, dim, queries, Q, vectors, n);
  if (rc) return rc;
  rc = select_device(device_id);
  if (rc) return rc;
  if (!Q || !n) return 0;
  if (!out) return set_err(LB200_EINVAL, "null output");
  HostStage hs;
  const size_t bq = Arena::pad((size_t)Q * dim * 2), bv = Arena::pad((size_t)n * dim * 2), bm = Arena::pad((size_t)(n + 7) / 8),
               bo = Arena::pad((size_t)Q * n * 4);
  CK(cudaMalloc(&hs.p, bq + bv + bm + bo));
  char* b = (char*)hs.p;
  __half* d_q = (__half*)b; b += bq;
  __half* d_v = (__half*)b; b += bv;
  uint8_t* d_m = (uint8_t*)b; b += bm;
  float* d_o = (float*)b;
  cudaStream_t st = nullptr;
  CK(cudaMemcpyAsync(d_q, queries, (size_t)Q * dim * 2, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(d_v, vectors, (size_t)n * dim * 2, cudaMemcpyHostToDevice, st));
  if (validity_or_null) CK(cudaMemcpyAsync(d_m, validity_or_null, (size_t)(n + 7) / 8, cudaMemcpyHostToDevice, st));
  rc = flat_f16_launch(metric, d_q, Q, d_v, validity_or_null ? d_m : nullptr, n, dim, d_o, (size_t)n, st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out, d_o, (size_t)Q * n * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int32_t lb200_flat_topk_f16(int32_t device_id, int32_t metric, uint32_t dim, const uint16_t* queries, uint32_t Q,
                                       const uint16_t* vectors, const uint8_t* validity_or_null, uint64_t n, uint32_t k,
                                       const float* lower_or_null, const float* upper_or_null, float* out_dist,
                                       uint64_t* out_idx, uint32_t* out_count) {
  int rc = flat_check(metric, dim, queries, Q, vectors, n);
  if (rc) return rc;
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (k > 1024) return set_err(LB200_EUNSUPPORTED, "k=%u > 1024", k);
  rc = select_device(device_id);
  if (rc) return rc;
  if (!Q) return 0;
  if (!out_dist || !out_idx || !out_count) return set_err(LB200_EINVAL, "null output");
  // keys in tiles so that the [tile][n] distance block stays around 256 MB
  const uint64_t stride = std::max<uint64_t>(n, 1);
  const uint32_t tile = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(Q, (256ull << 20) / (stride * 4)));
  HostStage hs;
  const size_t bq = Arena::pad((size_t)Q * dim * 2), bv = Arena::pad((size_t)stride * dim * 2), bm = Arena::pad((size_t)(n + 7) / 8 + 1),
               bD = Arena::pad((size_t)tile * stride * 4), bod = Arena::pad((size_t)Q * k * 4), boi = Arena::pad((size_t)Q * k * 8),
               boc = Arena::pad((size_t)Q * 4);
  CK(cudaMalloc(&hs.p, bq + bv + bm + bD + bod + boi + boc));
  char* b = (char*)hs.p;
  __half* d_q = (__half*)b; b += bq;
  __half* d_v = (__half*)b; b += bv;
  uint8_t* d_m = (uint8_t*)b; b += bm;
  float* d_D = (float*)b; b += bD;
  float* d_od = (float*)b; b += bod;
  uint64_t* d_oi = (uint64_t*)b; b += boi;
  uint32_t* d_oc = (uint32_t*)b;
  cudaStream_t st = nullptr;
  CK(cudaMemcpyAsync(d_q, queries, (size_t)Q * dim * 2, cudaMemcpyHostToDevice, st));
  if (n) CK(cudaMemcpyAsync(d_v, vectors, (size_t)n * dim * 2, cudaMemcpyHostToDevice, st));
  if (validity_or_null && n) CK(cudaMemcpyAsync(d_m, validity_or_null, (size_t)(n + 7) / 8, cudaMemcpyHostToDevice, st));
  const uint32_t kp2 = std::max(2u, next_pow2(k));
  for (uint32_t q0 = 0; q0 < Q; q0 += tile) {
    const uint32_t qn = std::min(tile, Q - q0);
    rc = flat_f16_launch(metric, d_q + (size_t)q0 * dim, qn, d_v, validity_or_null ? d_m : nullptr, n, dim, d_D, stride, st);
    if (rc) return rc;
    flat_select_kernel<<<qn, 256, (size_t)kp2 * 8, st>>>(d_D, stride, n, 1, stride, k, kp2, lower_or_null != nullptr,
                                                         lower_or_null ? *lower_or_null : 0.f, upper_or_null != nullptr,
                                                         upper_or_null ? *upper_or_null : 0.f, d_od + (size_t)q0 * k,
                                                         d_oi + (size_t)q0 * k, d_oc + q0);
    CK(cudaGetLastError());
  }
  CK(cudaMemcpyAsync(out_dist, d_od, (size_t)Q * k * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_idx, d_oi, (size_t)Q * k * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_count, d_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

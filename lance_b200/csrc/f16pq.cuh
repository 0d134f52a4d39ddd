// f16pq.cuh -- IVF_PQ on Float16 vectors: the kernels that differ from the f32 path.
//
// The reference's code for this path is its generic code with T = f16 (kmeans.rs:581-586, pq/storage.rs:500-512,
// pq/distance.rs:17-85): every f16 element is widened to f32 (`as_()`) and the sums are f32, so the handle keeps
// centroids and codebook as EXACT f32 widenings and the f32 kernels do the arithmetic.  What changes:
//   * summation ORDER: f16::l2 is l2_scalar::<f16,f32,16> for every length (f32 special-cases 8 and 16 with l2_once),
//     f16::dot is dot_scalar::<f16,f32,32> (f32 uses its f32x16/f32x8 SIMD order)            -> table + coarse(dot)
//   * the residual is an f16 array: key - centroid rounded to f16 (v2.rs:294, half's Sub)       -> table
//   * normalize::<f16> for cosine keys runs in half's software arithmetic (kernels.rs:118-121)  -> keys
//   * refine reads the raw f16 rows (half the bytes) with the f16 flat-kNN arithmetic           -> refine
// The code scan, the selects and the merges are the f32 path's own kernels: from the table on nothing is typed.
#pragma once
#include <cuda_fp16.h>

#include "dist.cuh"

namespace lb200 {

// dot_scalar::<f16,f32,32> (dot.rs:33-61) on widened values: remainder first, 32 lane sums, lanes added in order
__device__ __forceinline__ float dot_scalar32(const float* __restrict__ x, const float* __restrict__ y, int d) {
  const int nchunk = d >> 5, rem0 = nchunk << 5;
  float s = 0.0f;
  for (int i = rem0; i < d; i++) s = __fadd_rn(s, __fmul_rn(x[i], y[i]));
  float t = 0.0f;
  for (int l = 0; l < 32; l++) {
    float a = 0.0f;
    for (int c = 0; c < nchunk; c++) a = __fadd_rn(a, __fmul_rn(x[c * 32 + l], y[c * 32 + l]));
    t = __fadd_rn(t, a);
  }
  return __fadd_rn(s, t);
}

// keys of a batch: optional normalize::<f16> (x.powi(2) per element rounded to f16, f32 accumulator rounded once,
// sqrt and the divisions each rounded to f16), then the exact f32 widening the other kernels read.
// One thread per query (a sequential chain by definition).
__global__ void f16_keys_kernel(const __half* __restrict__ q, uint32_t Q, uint32_t dim, int normalize, float* __restrict__ orig32,
                                float* __restrict__ key32) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Q) return;
  const __half* x = q + (size_t)i * dim;
  float nrm = 1.0f;
  if (normalize) {
    float ss = 0.0f;
    for (uint32_t j = 0; j < dim; j++) {
      const float v = __half2float(x[j]);
      ss = __fadd_rn(ss, __half2float(__float2half_rn(__fmul_rn(v, v))));
    }
    const float s16 = __half2float(__float2half_rn(ss));
    nrm = __half2float(__float2half_rn(__fsqrt_rn(s16)));
  }
  for (uint32_t j = 0; j < dim; j++) {
    const float v = __half2float(x[j]);
    orig32[(size_t)i * dim + j] = v;
    key32[(size_t)i * dim + j] = normalize ? __half2float(__float2half_rn(__fdiv_rn(v, nrm))) : v;
  }
}

// find_partitions distances in the f16 order, one thread per (query, centroid): metric 2 -> 1 - dot_scalar32, else
// l2_scalar16.  Only used where the f32 kernels' order differs (dot; L2 with dim 8 / 16).
__global__ void coarse_dist_f16order_kernel(int dot, const float* __restrict__ keys, const float* __restrict__ cents, uint32_t Q,
                                            uint32_t nlist, uint32_t dim, float* __restrict__ D) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (uint64_t)Q * nlist) return;
  const uint32_t q = (uint32_t)(i / nlist), c = (uint32_t)(i - (uint64_t)q * nlist);
  const float* x = keys + (size_t)q * dim;
  const float* y = cents + (size_t)c * dim;
  D[i] = dot ? __fsub_rn(1.0f, dot_scalar32(x, y, (int)dim)) : l2_scalar16(x, y, (int)dim);
}

// ADC tables of a batch, layout per unit [M/16][256][16] (the scan kernel's order for one query, scan3.cuh lut_fetch).
// unit = (query, probe) pair when `probes` is given (residual metrics), else a query (dot).
// One block per unit; the residual sits in shared memory, rounded to f16 and widened again.
__global__ void __launch_bounds__(256)
lut_prepass_f16_kernel(const float* __restrict__ keys, const float* __restrict__ centroids, const float* __restrict__ codebook,
                       const uint32_t* __restrict__ probes, const uint64_t* __restrict__ part_off, uint32_t P, uint32_t dim,
                       uint32_t M, uint32_t dsub, float* __restrict__ out, uint32_t* __restrict__ out_max) {
  extern __shared__ float f16_resid[];  // [dim]
  const uint32_t u = blockIdx.x, t = threadIdx.x;
  const bool residual = probes != nullptr;
  uint32_t part = 0;
  if (residual) {
    part = probes[u];
    if (part_off[part + 1] == part_off[part]) return;  // an empty partition has no work item and needs no table
  }
  const float* key = keys + (size_t)(residual ? u / P : u) * dim;
  for (uint32_t i = t; i < dim; i += 256)
    f16_resid[i] = residual ? __half2float(__float2half_rn(__fsub_rn(key[i], centroids[(size_t)part * dim + i]))) : key[i];
  __syncthreads();
  uint32_t mb = 0;
  for (uint32_t e = t; e < M * 256; e += 256) {
    const uint32_t m = e >> 8, code = e & 255u;
    const float* x = f16_resid + (size_t)m * dsub;
    const float* y = codebook + (size_t)e * dsub;
    const float v = residual ? l2_scalar16(x, y, (int)dsub) : __fsub_rn(1.0f, dot_scalar32(x, y, (int)dsub));
    out[(size_t)u * M * 256 + (size_t)(m >> 4) * 4096 + code * 16 + (m & 15)] = v;
    mb = max(mb, __float_as_uint(v) & 0x7fffffffu);
  }
  if (mb >= 0x7a800000u) atomicMax(out_max + u, mb);  // as lut_prepass_kernel: only magnitudes that forbid the FMA row switch
}

// refine on raw f16 rows with the ORIGINAL key: {l2,cosine,dot}_distance_arrow_batch on Float16 data (the flat-kNN
// arithmetic of api_f16.inl: l2_scalar::<f16,f32,16>, dot_scalar::<f16,f32,32>, norm_l2_impl::<f16,f32,32>).
__global__ void refine_distance_f16_kernel(int metric, const float* __restrict__ orig32, const __half* __restrict__ raw, uint32_t dim,
                                           const uint64_t* __restrict__ pos, const uint32_t* __restrict__ cnt, uint32_t Q,
                                           uint32_t kprime, float* __restrict__ exact) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= Q * kprime) return;
  const uint32_t q = e / kprime, r = e - q * kprime;
  float out = __int_as_float(0x7fc00000);
  if (r < cnt[q]) {
    const float* x = orig32 + (size_t)q * dim;
    const __half* y = raw + (size_t)pos[e] * dim;
    const int d = (int)dim, nchunk = d >> 5, rem0 = nchunk << 5;
    if (metric == 0) {
      const int nc16 = d >> 4, r16 = nc16 << 4;
      float s = 0.0f;
      for (int i = r16; i < d; i++) { const float df = __fsub_rn(x[i], __half2float(y[i])); s = __fadd_rn(s, __fmul_rn(df, df)); }
      float t = 0.0f;
      for (int l = 0; l < 16; l++) {
        float a = 0.0f;
        for (int c = 0; c < nc16; c++) { const float df = __fsub_rn(x[c * 16 + l], __half2float(y[c * 16 + l])); a = __fadd_rn(a, __fmul_rn(df, df)); }
        t = __fadd_rn(t, a);
      }
      out = __fadd_rn(s, t);
    } else {
      // three dot_scalar32 sums at once: x.y, y.y, x.x
      float sxy = 0.0f, syy = 0.0f, sxx = 0.0f;
      for (int i = rem0; i < d; i++) {
        const float yv = __half2float(y[i]);
        sxy = __fadd_rn(sxy, __fmul_rn(x[i], yv)); syy = __fadd_rn(syy, __fmul_rn(yv, yv)); sxx = __fadd_rn(sxx, __fmul_rn(x[i], x[i]));
      }
      float txy = 0.0f, tyy = 0.0f, txx = 0.0f;
      for (int l = 0; l < 32; l++) {
        float a = 0.0f, b = 0.0f, c2 = 0.0f;
        for (int c = 0; c < nchunk; c++) {
          const float xv = x[c * 32 + l], yv = __half2float(y[c * 32 + l]);
          a = __fadd_rn(a, __fmul_rn(xv, yv)); b = __fadd_rn(b, __fmul_rn(yv, yv)); c2 = __fadd_rn(c2, __fmul_rn(xv, xv));
        }
        txy = __fadd_rn(txy, a); tyy = __fadd_rn(tyy, b); txx = __fadd_rn(txx, c2);
      }
      const float xy = __fadd_rn(sxy, txy);
      if (metric == 2) out = __fsub_rn(1.0f, xy);
      else out = __fsub_rn(1.0f, __fdiv_rn(xy, __fmul_rn(__fsqrt_rn(__fadd_rn(sxx, txx)), __fsqrt_rn(__fadd_rn(syy, tyy)))));
    }
  }
  exact[e] = out;
}

}  // namespace lb200

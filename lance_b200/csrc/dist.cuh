// dist.cuh -- f32 distance arithmetic in the reference's own operation order.
//
// Every function reproduces, rounding for rounding, the x86_64/haswell build of
// rust/lance-linalg (see file:line on each).  __fadd_rn/__fmul_rn/__fsub_rn are used
// everywhere so that nvcc never contracts a*b+c into an FMA (the Rust code is compiled
// without contraction); __fmaf_rn only where the reference calls multiply_add.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lb200 {

// f32::total_cmp as an unsigned key (rust/lance-index/src/vector/graph.rs:54-74):
// monotone map float -> u32, NaN(+) after +inf.
__host__ __device__ __forceinline__ uint32_t ordered_key(float f) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(f);
#else
  uint32_t b;
  memcpy(&b, &f, 4);
#endif
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float key_to_float(uint32_t k) {
  uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}

// simd/f32.rs:203-218  f32x8::reduce_sum (AVX2 permute/hadd order)
__device__ __forceinline__ float reduce_sum8(const float* s) {
  float a0 = __fadd_rn(s[0], s[4]), a1 = __fadd_rn(s[1], s[5]);
  float a2 = __fadd_rn(s[2], s[6]), a3 = __fadd_rn(s[3], s[7]);
  return __fadd_rn(__fadd_rn(a0, a2), __fadd_rn(a1, a3));
}
// simd/f32.rs:625-644  f32x16::reduce_sum (two __m256 added first)
__device__ __forceinline__ float reduce_sum16(const float* s) {
  float y[8];
#pragma unroll
  for (int i = 0; i < 8; i++) y[i] = __fadd_rn(s[i], s[i + 8]);
  return reduce_sum8(y);
}

// distance/l2.rs:65-99  l2_scalar::<f32,f32,16>
__device__ __forceinline__ float l2_scalar16(const float* __restrict__ x,
                                             const float* __restrict__ y, int d) {
  int nchunk = d >> 4, rem0 = nchunk << 4;
  float s = 0.0f;
  for (int i = rem0; i < d; i++) {
    float df = __fsub_rn(x[i], y[i]);
    s = __fadd_rn(s, __fmul_rn(df, df));
  }
  float sums[16];
#pragma unroll
  for (int i = 0; i < 16; i++) sums[i] = 0.0f;
  for (int c = 0; c < nchunk; c++) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      float df = __fsub_rn(x[c * 16 + i], y[c * 16 + i]);
      sums[i] = __fadd_rn(sums[i], __fmul_rn(df, df));
    }
  }
  float t = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; i++) t = __fadd_rn(t, sums[i]);
  return __fadd_rn(s, t);
}

// distance/l2.rs:207-216  l2_once::<f32x8,8>, <f32x16,16>
__device__ __forceinline__ float l2_once8(const float* x, const float* y) {
  float s[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float df = __fsub_rn(x[i], y[i]);
    s[i] = __fmul_rn(df, df);
  }
  return reduce_sum8(s);
}
__device__ __forceinline__ float l2_once16(const float* x, const float* y) {
  float s[16];
#pragma unroll
  for (int i = 0; i < 16; i++) {
    float df = __fsub_rn(x[i], y[i]);
    s[i] = __fmul_rn(df, df);
  }
  return reduce_sum16(s);
}
// distance/l2.rs:177-195 f32::l2_batch dispatch on the dimension
__device__ __forceinline__ float l2_batch_elem(const float* x, const float* y, int d) {
  if (d == 8) return l2_once8(x, y);
  if (d == 16) return l2_once16(x, y);
  return l2_scalar16(x, y, d);
}

// distance/dot.rs:141-195  Dot for f32
__device__ __forceinline__ float dot_f32(const float* __restrict__ x, const float* __restrict__ y,
                                         int d) {
  int n64 = d >> 6, off64 = n64 << 6;
  int n8 = (d - off64) >> 3, off8 = off64 + (n8 << 3);
  float sum = 0.0f;
  for (int i = off8; i < d; i++) sum = __fadd_rn(sum, __fmul_rn(x[i], y[i]));
  float sum8[8];
#pragma unroll
  for (int i = 0; i < 8; i++) sum8[i] = 0.0f;
  for (int c = 0; c < n8; c++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      sum8[i] = __fadd_rn(sum8[i], __fmul_rn(x[off64 + c * 8 + i], y[off64 + c * 8 + i]));
  }
  float sum16[16];
#pragma unroll
  for (int i = 0; i < 16; i++) sum16[i] = 0.0f;
  for (int c = 0; c < n64; c++) {
    const float* xc = x + c * 64;
    const float* yc = y + c * 64;
#pragma unroll
    for (int i = 0; i < 16; i++) {
      float p1 = __fmul_rn(xc[i], yc[i]), p2 = __fmul_rn(xc[16 + i], yc[16 + i]);
      float p3 = __fmul_rn(xc[32 + i], yc[32 + i]), p4 = __fmul_rn(xc[48 + i], yc[48 + i]);
      sum16[i] = __fadd_rn(sum16[i], __fadd_rn(__fadd_rn(p1, p2), __fadd_rn(p3, p4)));
    }
  }
  return __fadd_rn(__fadd_rn(reduce_sum16(sum16), reduce_sum8(sum8)), sum);
}
// distance/dot.rs:71-73
__device__ __forceinline__ float dot_distance(const float* x, const float* y, int d) {
  return __fsub_rn(1.0f, dot_f32(x, y, d));
}

// distance/norm_l2.rs:102-126  norm_l2_impl::<f32,f32,16>
__device__ __forceinline__ float norm_l2(const float* __restrict__ v, int d) {
  int nchunk = d >> 4, rem0 = nchunk << 4;
  float sum = 0.0f;
  for (int i = rem0; i < d; i++) sum = __fadd_rn(sum, __fmul_rn(v[i], v[i]));
  float sums[16];
#pragma unroll
  for (int i = 0; i < 16; i++) sums[i] = 0.0f;
  for (int c = 0; c < nchunk; c++) {
#pragma unroll
    for (int i = 0; i < 16; i++)
      sums[i] = __fadd_rn(sums[i], __fmul_rn(v[c * 16 + i], v[c * 16 + i]));
  }
  float t = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; i++) t = __fadd_rn(t, sums[i]);
  return __fsqrt_rn(__fadd_rn(sum, t));
}

// distance/cosine.rs:143-174  cosine_fast for f32 (multiply_add == FMA)
__device__ __forceinline__ float cosine_fast(const float* __restrict__ x, float x_norm,
                                             const float* __restrict__ y, int d) {
  int unrolled = d / 16 * 16, aligned = d / 8 * 8;
  float yn16[16], xy16[16], yn8[8], xy8[8];
#pragma unroll
  for (int i = 0; i < 16; i++) yn16[i] = xy16[i] = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; i++) yn8[i] = xy8[i] = 0.0f;
  for (int c = 0; c < unrolled; c += 16) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      float xv = x[c + i], yv = y[c + i];
      xy16[i] = __fmaf_rn(xv, yv, xy16[i]);
      yn16[i] = __fmaf_rn(yv, yv, yn16[i]);
    }
  }
  for (int c = unrolled; c < aligned; c += 8) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      float xv = x[c + i], yv = y[c + i];
      xy8[i] = __fmaf_rn(xv, yv, xy8[i]);
      yn8[i] = __fmaf_rn(yv, yv, yn8[i]);
    }
  }
  float tn = norm_l2(y + aligned, d - aligned);
  float y_norm = __fadd_rn(__fadd_rn(reduce_sum16(yn16), reduce_sum8(yn8)), __fmul_rn(tn, tn));
  float xy = __fadd_rn(__fadd_rn(reduce_sum16(xy16), reduce_sum8(xy8)),
                       dot_f32(x + aligned, y + aligned, d - aligned));
  return __fsub_rn(1.0f, __fdiv_rn(__fdiv_rn(xy, x_norm), __fsqrt_rn(y_norm)));
}
// distance/cosine.rs:128-141  cosine_once (dimension 8 / 16 rows in cosine_batch :201-224)
__device__ __forceinline__ float cosine_once(const float* x, float x_norm, const float* y, int n) {
  float y2[16], xy[16];
#pragma unroll
  for (int i = 0; i < 16; i++) {
    if (i < n) {
      y2[i] = __fmul_rn(y[i], y[i]);
      xy[i] = __fmul_rn(x[i], y[i]);
    }
  }
  float sxy = n == 8 ? reduce_sum8(xy) : reduce_sum16(xy);
  float sy2 = n == 8 ? reduce_sum8(y2) : reduce_sum16(y2);
  return __fsub_rn(1.0f, __fdiv_rn(__fdiv_rn(sxy, x_norm), __fsqrt_rn(sy2)));
}

// DistanceType::arrow_batch_func (distance.rs:56-63): one (query,row) element of
// l2_distance_arrow_batch / cosine_distance_arrow_batch / dot_distance_arrow_batch.
// x_norm = norm_l2(x), only read for cosine.
__device__ __forceinline__ float exact_distance(int metric, const float* x, float x_norm,
                                                const float* y, int d) {
  if (metric == 0) return l2_batch_elem(x, y, d);
  if (metric == 2) return dot_distance(x, y, d);
  if (d == 8 || d == 16) return cosine_once(x, x_norm, y, d);
  return cosine_fast(x, x_norm, y, d);
}

// Packed FP32 (Blackwell FADD2/FMUL2, PTX .f32x2): two IEEE round-to-nearest operations per
// instruction, bit-identical per half to the scalar ops.  ptxas 12.9 CONTRACTS a mul.rn.f32x2 that
// feeds an add.rn.f32x2 into one FFMA2 (single rounding) although both carry .rn, so products are
// always unpacked and summed first with scalar __fadd_rn; only sub, mul and add-of-sums are packed.
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t pk2(float a, float b) {
  f32x2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk2(f32x2_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ f32x2_t add2(f32x2_t a, f32x2_t b) {
  f32x2_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2_t sub2(f32x2_t a, f32x2_t b) {
  f32x2_t r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2_t mul2(f32x2_t a, f32x2_t b) {
  f32x2_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2_t fma2(f32x2_t a, f32x2_t b, f32x2_t c) {
  f32x2_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
// two packed -0.0f that the compiler cannot see through (fma2(d, d, -0) == mul.rn exactly, and cannot be contracted into
// a following add2; a literal -0 would be folded back into a mul): read through the constant bank at run time
static __constant__ unsigned long long g_negzero2 = 0x8000000080000000ull;

}  // namespace lb200

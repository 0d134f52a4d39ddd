// coarse.cuh -- K0 normalise + K1 coarse (query x centroid) distances and top-nprobes select.
//
// Replaces rust/lance-linalg/src/kmeans.rs:624-644 kmeans_find_partitions (via
// ivf/storage.rs:92-104) and rust/lance-linalg/src/kernels.rs:119-122 normalize.
#pragma once
#include "dist.cuh"

namespace lb200 {

// kernels.rs:119-122: l2_norm = sqrt(sum_i x_i^2) with a plain sequential f32 sum, then x / norm.
// One warp per query; lane 0 does the (order-sensitive) sum, all lanes divide.
__global__ void normalize_kernel(const float* __restrict__ q, float* __restrict__ out, uint32_t Q,
                                 uint32_t d) {
  uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= Q) return;
  const float* x = q + (size_t)w * d;
  float s = 0.0f;
  if (lane == 0)
    for (uint32_t i = 0; i < d; i++) s = __fadd_rn(s, __fmul_rn(x[i], x[i]));
  float n = __fsqrt_rn(__shfl_sync(0xffffffffu, s, 0));
  for (uint32_t i = lane; i < d; i += 32) out[(size_t)w * d + i] = __fdiv_rn(x[i], n);
}

// Coarse distances.  Block = 128 threads = 128 centroids x TQ queries; the query tile lives in
// shared memory (warp-broadcast reads), each thread streams its own centroid row.
// metric_l2: true -> l2_distance_batch (l2.rs:237 -> f32::l2_batch), false -> dot_distance_batch.
constexpr int COARSE_TQ = 4;
constexpr int COARSE_THREADS = 128;

template <bool L2>
__global__ void __launch_bounds__(COARSE_THREADS)
coarse_dist_kernel(const float* __restrict__ queries, const float* __restrict__ centroids,
                   uint32_t Q, uint32_t nlist, uint32_t d, float* __restrict__ out /*[Q][nlist]*/) {
  extern __shared__ float sq[];  // [TQ][d]
  uint32_t q0 = blockIdx.y * COARSE_TQ;
  uint32_t c = blockIdx.x * COARSE_THREADS + threadIdx.x;
  for (uint32_t i = threadIdx.x; i < COARSE_TQ * d; i += COARSE_THREADS) {
    uint32_t qq = q0 + i / d;
    sq[i] = qq < Q ? queries[(size_t)qq * d + (i % d)] : 0.0f;
  }
  __syncthreads();
  if (c >= nlist) return;
  const float* y = centroids + (size_t)c * d;
  float res[COARSE_TQ];
  if (L2) {
    if (d == 8 || d == 16) {
#pragma unroll
      for (int t = 0; t < COARSE_TQ; t++) res[t] = l2_batch_elem(sq + t * d, y, d);
    } else {
      // l2_scalar<16> for TQ queries at once (centroid values loaded once per chunk)
      uint32_t nchunk = d >> 4, rem0 = nchunk << 4;
      float s[COARSE_TQ], sums[COARSE_TQ][16];
#pragma unroll
      for (int t = 0; t < COARSE_TQ; t++) {
        s[t] = 0.0f;
#pragma unroll
        for (int i = 0; i < 16; i++) sums[t][i] = 0.0f;
      }
      for (uint32_t i = rem0; i < d; i++) {
        float yv = y[i];
#pragma unroll
        for (int t = 0; t < COARSE_TQ; t++) {
          float df = __fsub_rn(sq[t * d + i], yv);
          s[t] = __fadd_rn(s[t], __fmul_rn(df, df));
        }
      }
      for (uint32_t ch = 0; ch < nchunk; ch++) {
        float yv[16];
        if ((d & 3) == 0) {
          const float4* y4 = reinterpret_cast<const float4*>(y + ch * 16);
#pragma unroll
          for (int i = 0; i < 4; i++) {
            float4 v = __ldg(y4 + i);
            yv[4 * i] = v.x; yv[4 * i + 1] = v.y; yv[4 * i + 2] = v.z; yv[4 * i + 3] = v.w;
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; i++) yv[i] = y[ch * 16 + i];
        }
#pragma unroll
        for (int t = 0; t < COARSE_TQ; t++) {
          const float* xq = sq + t * d + ch * 16;
#pragma unroll
          for (int i = 0; i < 16; i++) {
            float df = __fsub_rn(xq[i], yv[i]);
            sums[t][i] = __fadd_rn(sums[t][i], __fmul_rn(df, df));
          }
        }
      }
#pragma unroll
      for (int t = 0; t < COARSE_TQ; t++) {
        float tt = 0.0f;
#pragma unroll
        for (int i = 0; i < 16; i++) tt = __fadd_rn(tt, sums[t][i]);
        res[t] = __fadd_rn(s[t], tt);
      }
    }
  } else {
#pragma unroll
    for (int t = 0; t < COARSE_TQ; t++) res[t] = dot_distance(sq + t * d, y, d);
  }
#pragma unroll
  for (int t = 0; t < COARSE_TQ; t++)
    if (q0 + t < Q) out[(size_t)(q0 + t) * nlist + c] = res[t];
}

// Block-wide bitonic sort of n (power of two) u64 keys in shared memory, ascending.
__device__ __forceinline__ void block_bitonic_sort_u64(unsigned long long* s, uint32_t n) {
  for (uint32_t k = 2; k <= n; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        uint32_t ixj = i ^ j;
        if (ixj > i) {
          unsigned long long a = s[i], b = s[ixj];
          bool up = (i & k) == 0;
          if ((a > b) == up) {
            s[i] = b;
            s[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
}

// sort_to_indices(dists, None, Some(nprobes)) (kmeans.rs:643; arrow-ord 53.3.0): ascending by
// f32 total order; equal distances by centroid id (the reference's order is unstable there).
// One block per query; npow2 = next pow2 >= nlist, <= 8192.
__global__ void __launch_bounds__(256)
coarse_select_kernel(const float* __restrict__ dist /*[Q][nlist]*/, uint32_t nlist, uint32_t npow2,
                     uint32_t P, uint32_t* __restrict__ out_ids /*[Q][P]*/,
                     float* __restrict__ out_dists /*[Q][P] or null*/) {
  extern __shared__ unsigned long long skeys[];
  uint32_t q = blockIdx.x;
  const float* drow = dist + (size_t)q * nlist;
  for (uint32_t i = threadIdx.x; i < npow2; i += blockDim.x)
    skeys[i] = i < nlist ? (((unsigned long long)ordered_key(drow[i]) << 32) | i) : ~0ull;
  __syncthreads();
  block_bitonic_sort_u64(skeys, npow2);
  for (uint32_t i = threadIdx.x; i < P; i += blockDim.x) {
    unsigned long long kv = skeys[i];
    out_ids[(size_t)q * P + i] = (uint32_t)kv;
    if (out_dists) out_dists[(size_t)q * P + i] = key_to_float((uint32_t)(kv >> 32));
  }
}

}  // namespace lb200

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

// Fast path of the same selection for P <= 256 (every BASELINE config): instead of sorting all
// nlist keys, bound the P-th smallest from above by the P-th smallest of 256 strided group
// minima, collect the (few) keys under that bound and sort only those.  Keys are (ordered
// distance << 32 | centroid id): unique, so the result is the same total order as the full sort.
constexpr uint32_t CSEL_CAP = 1024;
__global__ void __launch_bounds__(256)
coarse_select2_kernel(const float* __restrict__ dist /*[Q][nlist]*/, uint32_t nlist, uint32_t P,
                      uint32_t* __restrict__ out_ids, float* __restrict__ out_dists) {
  __shared__ unsigned long long cand[CSEL_CAP];
  __shared__ unsigned long long mins[256];
  __shared__ unsigned long long sT;
  __shared__ uint32_t scnt;
  const uint32_t q = blockIdx.x, t = threadIdx.x;
  const float* drow = dist + (size_t)q * nlist;
  unsigned long long m = ~0ull;
  for (uint32_t i = t; i < nlist; i += 256) {
    unsigned long long kv = ((unsigned long long)ordered_key(drow[i]) << 32) | i;
    m = kv < m ? kv : m;
  }
  mins[t] = m;
  if (t == 0) scnt = 0;
  __syncthreads();
  {
    uint32_t r = 0;
    for (uint32_t j = 0; j < 256; j++) r += mins[j] < m ? 1u : 0u;
    if (r == P - 1 && m != ~0ull) sT = m;  // unique keys -> exactly one thread has rank P-1
  }
  __syncthreads();
  unsigned long long T = sT;
  for (uint32_t i = t; i < nlist; i += 256) {
    unsigned long long kv = ((unsigned long long)ordered_key(drow[i]) << 32) | i;
    if (kv <= T) {
      uint32_t pos = atomicAdd(&scnt, 1u);
      if (pos < CSEL_CAP) cand[pos] = kv;
    }
  }
  __syncthreads();
  uint32_t cnt = scnt;
  if (cnt > CSEL_CAP) {
    // adversarial layout: find the exact P-th smallest key by bitwise binary search, re-collect
    unsigned long long X = 0;
    for (int bit = 63; bit >= 0; --bit) {
      unsigned long long c2 = X | (1ull << bit);
      uint32_t c = 0;
      for (uint32_t i = t; i < nlist; i += 256) {
        unsigned long long kv = ((unsigned long long)ordered_key(drow[i]) << 32) | i;
        c += kv < c2 ? 1u : 0u;
      }
      __syncthreads();
      if (t == 0) scnt = 0;
      __syncthreads();
      atomicAdd(&scnt, c);
      __syncthreads();
      if (scnt < P) X = c2;  // largest X with #(kv < X) < P  ==  the P-th smallest key
    }
    __syncthreads();
    if (t == 0) scnt = 0;
    __syncthreads();
    for (uint32_t i = t; i < nlist; i += 256) {
      unsigned long long kv = ((unsigned long long)ordered_key(drow[i]) << 32) | i;
      if (kv <= X) cand[atomicAdd(&scnt, 1u)] = kv;  // exactly P <= 256 entries
    }
    __syncthreads();
    cnt = scnt;
  }
  uint32_t n2 = 2;
  while (n2 < cnt) n2 <<= 1;
  for (uint32_t i = cnt + t; i < n2; i += 256) cand[i] = ~0ull;
  __syncthreads();
  block_bitonic_sort_u64(cand, n2);
  for (uint32_t i = t; i < P; i += 256) {
    unsigned long long kv = cand[i];
    out_ids[(size_t)q * P + i] = (uint32_t)kv;
    if (out_dists) out_dists[(size_t)q * P + i] = key_to_float((uint32_t)(kv >> 32));
  }
}

// Fast L2 coarse distances for d % 16 == 0, 32 <= d (l2_scalar<16> path, l2.rs:65-99): the block's
// 128 centroid rows are staged ONCE in shared memory (pitch 129 floats: conflict-free both for the
// coalesced fill and for the lane=centroid reads) and reused for CQ_QB queries, 4 at a time, so the
// centroid matrix crosses L2 Q/CQ_QB times instead of Q/4 times.  Same operation order as
// coarse_dist_kernel: 16 lane accumulators per (query, centroid), chunks ascending, lanes summed
// 0..15, then 0 + t.
constexpr int CQ_QB = 32;
constexpr int CQ_PITCH = 129;
__global__ void __launch_bounds__(128)
coarse_dist_l2_tiled_kernel(const float* __restrict__ queries, const float* __restrict__ centroids,
                            uint32_t Q, uint32_t nlist, uint32_t d, float* __restrict__ out, size_t out_stride,
                            const uint8_t* __restrict__ validity /*Arrow bitmap or null (flat path)*/) {
  extern __shared__ float sm[];
  float* tile = sm;                         // [d][129]
  float* sq = sm + (size_t)d * CQ_PITCH;    // [CQ_QB][d]
  const uint32_t c0 = blockIdx.x * 128, q0 = blockIdx.y * CQ_QB, t = threadIdx.x;
  const uint32_t lane = t & 31, w = t >> 5;
  // fill: warp w takes centroid rows w, w+4, ...; lanes read consecutive elements (coalesced)
  for (uint32_t r = w; r < 128; r += 4) {
    const uint32_t c = c0 + r;
    for (uint32_t e = lane; e < d; e += 32)
      tile[(size_t)e * CQ_PITCH + r] = c < nlist ? __ldg(centroids + (size_t)c * d + e) : 0.0f;
  }
  for (uint32_t i = t; i < CQ_QB * d; i += 128) {
    uint32_t qq = q0 + i / d;
    sq[i] = qq < Q ? __ldg(queries + (size_t)qq * d + (i % d)) : 0.0f;
  }
  __syncthreads();
  const uint32_t c = c0 + t;
  const uint32_t nchunk = d >> 4;
#pragma unroll 1
  for (uint32_t g = 0; g < CQ_QB / 4; g++) {
    if (q0 + g * 4 >= Q) break;
    float sums[4][16];
#pragma unroll
    for (int k = 0; k < 4; k++)
#pragma unroll
      for (int i = 0; i < 16; i++) sums[k][i] = 0.0f;
    const float* xq = sq + (size_t)g * 4 * d;
#pragma unroll 1
    for (uint32_t ch = 0; ch < nchunk; ch++) {
      float yv[16];
#pragma unroll
      for (int i = 0; i < 16; i++) yv[i] = tile[(size_t)(ch * 16 + i) * CQ_PITCH + t];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const float4* x4 = reinterpret_cast<const float4*>(xq + (size_t)k * d + ch * 16);
#pragma unroll
        for (int i4 = 0; i4 < 4; i4++) {
          const float4 xv = x4[i4];  // warp-wide broadcast
          const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
          for (int j = 0; j < 4; j++) {
            float df = __fsub_rn(xs[j], yv[i4 * 4 + j]);
            sums[k][i4 * 4 + j] = __fadd_rn(sums[k][i4 * 4 + j], __fmul_rn(df, df));
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      float tt = 0.0f;
#pragma unroll
      for (int i = 0; i < 16; i++) tt = __fadd_rn(tt, sums[k][i]);
      const uint32_t qq = q0 + g * 4 + k;
      if (qq < Q && c < nlist) {
        const bool valid = !validity || ((validity[c >> 3] >> (c & 7)) & 1);
        out[(size_t)qq * out_stride + c] = valid ? __fadd_rn(0.0f, tt) : __int_as_float(0x7fc00000);
      }
    }
  }
}

// Fast dot-metric coarse distances for d % 64 == 0 (Dot for f32, dot.rs:141-195: 16 lane accumulators,
// per 64-chunk sum16[i] += (p1 + p2) + (p3 + p4), p_j = x[16j+i]*y[16j+i]; no 8-lane or scalar tail).
// Block = 512 threads = 128 centroids x 4 query groups of 4 queries.  The centroid tile is streamed
// through shared memory in 64-dimension slabs (pitch 129, double-buffered: the next slab is fetched
// into registers while the current one is consumed), so the centroid matrix crosses L2 once per 16
// queries and each thread keeps only its 4 x 16 accumulators across slabs -- any d fits.
constexpr int CD_QB = 16;
constexpr int CD_SLAB = 64 * CQ_PITCH;  // floats per slab buffer
__global__ void __launch_bounds__(512, 1)
coarse_dist_dot_slab_kernel(const float* __restrict__ queries, const float* __restrict__ centroids, uint32_t Q,
                            uint32_t nlist, uint32_t d, float* __restrict__ out, size_t out_stride) {
  extern __shared__ float sm[];
  float* tile = sm;                 // [2][64][129]
  float* sq = sm + 2 * CD_SLAB;     // [CD_QB][d]
  const uint32_t c0 = blockIdx.x * 128, q0 = blockIdx.y * CD_QB, t = threadIdx.x;
  const uint32_t lane = t & 31, w = t >> 5, ct = t & 127, qg = t >> 7;
  const uint32_t nslab = d >> 6;
  for (uint32_t i = t; i < CD_QB * d; i += 512) {
    uint32_t qq = q0 + i / d;
    sq[i] = qq < Q ? __ldg(queries + (size_t)qq * d + (i % d)) : 0.0f;
  }
  // slab fetch: warp w takes rows w, w+16, ... (8 rows); lanes read elements lane, lane+32 (coalesced)
  float pre[16];
  auto fetch = [&](uint32_t slab) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint32_t c = c0 + w + 16 * i;
      const float* src = centroids + (size_t)c * d + slab * 64 + lane;
      pre[2 * i] = c < nlist ? __ldg(src) : 0.0f;
      pre[2 * i + 1] = c < nlist ? __ldg(src + 32) : 0.0f;
    }
  };
  auto stash = [&](float* buf) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      buf[(size_t)lane * CQ_PITCH + w + 16 * i] = pre[2 * i];
      buf[(size_t)(lane + 32) * CQ_PITCH + w + 16 * i] = pre[2 * i + 1];
    }
  };
  fetch(0);
  stash(tile);
  __syncthreads();
  float acc[4][16];
#pragma unroll
  for (int k = 0; k < 4; k++)
#pragma unroll
    for (int i = 0; i < 16; i++) acc[k][i] = 0.0f;
  const float* xq = sq + (size_t)qg * 4 * d;
#pragma unroll 1
  for (uint32_t s = 0; s < nslab; s++) {
    const float* buf = tile + (s & 1) * CD_SLAB;
    if (s + 1 < nslab) fetch(s + 1);
#pragma unroll
    for (int i4 = 0; i4 < 4; i4++) {
      float y[4][4];
#pragma unroll
      for (int j = 0; j < 4; j++)
#pragma unroll
        for (int r = 0; r < 4; r++) y[j][r] = buf[(size_t)(16 * j + 4 * i4 + r) * CQ_PITCH + ct];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        float pj[4][4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const float4 xv = *reinterpret_cast<const float4*>(xq + (size_t)k * d + s * 64 + 16 * j + 4 * i4);
          pj[j][0] = __fmul_rn(xv.x, y[j][0]);
          pj[j][1] = __fmul_rn(xv.y, y[j][1]);
          pj[j][2] = __fmul_rn(xv.z, y[j][2]);
          pj[j][3] = __fmul_rn(xv.w, y[j][3]);
        }
#pragma unroll
        for (int r = 0; r < 4; r++)
          acc[k][4 * i4 + r] = __fadd_rn(acc[k][4 * i4 + r],
                                         __fadd_rn(__fadd_rn(pj[0][r], pj[1][r]), __fadd_rn(pj[2][r], pj[3][r])));
      }
    }
    if (s + 1 < nslab) stash(tile + ((s + 1) & 1) * CD_SLAB);
    __syncthreads();
  }
  const uint32_t c = c0 + ct;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint32_t qq = q0 + qg * 4 + k;
    float z8[8];
#pragma unroll
    for (int i = 0; i < 8; i++) z8[i] = 0.0f;
    const float dot = __fadd_rn(__fadd_rn(reduce_sum16(acc[k]), reduce_sum8(z8)), 0.0f);
    if (qq < Q && c < nlist) out[(size_t)qq * out_stride + c] = __fsub_rn(1.0f, dot);
  }
}

// The same slab pipeline for L2 (l2_scalar<16>, l2.rs:65-99; d % 64 == 0): 16 lane accumulators per
// (query, row), 16-element chunks in ascending order, lanes summed 0..15, then 0 + t -- the operation
// order of coarse_dist_l2_tiled_kernel, with 16 warps per CTA sharing each staged slab instead of 4
// (the whole-row tile of that kernel leaves room for only 8 warps per SM).  Serves the coarse step and
// the flat kNN distances (validity = Arrow bitmap or null; null rows yield NaN).
__global__ void __launch_bounds__(512, 1)
coarse_dist_l2_slab_kernel(const float* __restrict__ queries, const float* __restrict__ rows, uint32_t Q,
                           uint32_t n, uint32_t d, float* __restrict__ out, size_t out_stride,
                           const uint8_t* __restrict__ validity) {
  extern __shared__ float sm[];
  float* tile = sm;                 // [2][64][129]
  float* sq = sm + 2 * CD_SLAB;     // [CD_QB][d]
  const uint32_t c0 = blockIdx.x * 128, q0 = blockIdx.y * CD_QB, t = threadIdx.x;
  const uint32_t lane = t & 31, w = t >> 5, ct = t & 127, qg = t >> 7;
  const uint32_t nslab = d >> 6;
  for (uint32_t i = t; i < CD_QB * d; i += 512) {
    uint32_t qq = q0 + i / d;
    sq[i] = qq < Q ? __ldg(queries + (size_t)qq * d + (i % d)) : 0.0f;
  }
  float pre[16];
  auto fetch = [&](uint32_t slab) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint32_t c = c0 + w + 16 * i;
      const float* src = rows + (size_t)c * d + slab * 64 + lane;
      pre[2 * i] = c < n ? __ldg(src) : 0.0f;
      pre[2 * i + 1] = c < n ? __ldg(src + 32) : 0.0f;
    }
  };
  auto stash = [&](float* buf) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      buf[(size_t)lane * CQ_PITCH + w + 16 * i] = pre[2 * i];
      buf[(size_t)(lane + 32) * CQ_PITCH + w + 16 * i] = pre[2 * i + 1];
    }
  };
  fetch(0);
  stash(tile);
  __syncthreads();
  // lane accumulators in packed pairs (lanes 2j, 2j+1): sub, square and add run as FADD2 / FFMA2 / FADD2 -- each half is
  // the scalar IEEE operation (the square is fma(d, d, -0) == mul.rn, see dist.cuh), half the issue slots of the
  // scalar form (this kernel is fp32-issue bound: r01 ncu)
  f32x2_t acc2[4][8];
#pragma unroll
  for (int k = 0; k < 4; k++)
#pragma unroll
    for (int i = 0; i < 8; i++) acc2[k][i] = 0ull;
  const f32x2_t nz = g_negzero2;
  const float* xq = sq + (size_t)qg * 4 * d;
#pragma unroll 1
  for (uint32_t s = 0; s < nslab; s++) {
    const float* buf = tile + (s & 1) * CD_SLAB;
    if (s + 1 < nslab) fetch(s + 1);
#pragma unroll
    for (int ch = 0; ch < 4; ch++) {
#pragma unroll
      for (int i4 = 0; i4 < 4; i4++) {
        float y[4];
#pragma unroll
        for (int r = 0; r < 4; r++) y[r] = buf[(size_t)(16 * ch + 4 * i4 + r) * CQ_PITCH + ct];
        const f32x2_t y01 = pk2(y[0], y[1]), y23 = pk2(y[2], y[3]);
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const float4 xv = *reinterpret_cast<const float4*>(xq + (size_t)k * d + s * 64 + 16 * ch + 4 * i4);
          const f32x2_t d01 = sub2(pk2(xv.x, xv.y), y01), d23 = sub2(pk2(xv.z, xv.w), y23);
          acc2[k][2 * i4] = add2(acc2[k][2 * i4], fma2(d01, d01, nz));
          acc2[k][2 * i4 + 1] = add2(acc2[k][2 * i4 + 1], fma2(d23, d23, nz));
        }
      }
    }
    if (s + 1 < nslab) stash(tile + ((s + 1) & 1) * CD_SLAB);
    __syncthreads();
  }
  float acc[4][16];
#pragma unroll
  for (int k = 0; k < 4; k++)
#pragma unroll
    for (int i = 0; i < 8; i++) upk2(acc2[k][i], acc[k][2 * i], acc[k][2 * i + 1]);
  const uint32_t c = c0 + ct;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint32_t qq = q0 + qg * 4 + k;
    float tt = 0.0f;
#pragma unroll
    for (int i = 0; i < 16; i++) tt = __fadd_rn(tt, acc[k][i]);
    if (qq < Q && c < n) {
      const bool valid = !validity || ((validity[c >> 3] >> (c & 7)) & 1);
      out[(size_t)qq * out_stride + c] = valid ? __fadd_rn(0.0f, tt) : __int_as_float(0x7fc00000);
    }
  }
}

}  // namespace lb200

// build.cuh -- kernels of the GPU index BUILD (last item): k-means centroid update, PQ encoding,
// column slicing.  The assignment step (argmin over centroid distances) reuses the query path's coarse kernels
// (coarse.cuh), so build and search agree bit for bit on which partition a vector belongs to.
// Replaces python/python/lance/torch/kmeans.py + lance/vector.py:134-700 (the reference's torch build path) and
// rust/lance-linalg/src/kmeans.rs (the CPU one) for the steps below.
#pragma once
#include "scan.cuh"

namespace lb200 {

// KMeansAlgoFloat::to_kmeans (kmeans.rs:242-300): centroid c = (f32 sum of its members IN DATA ORDER) / (count as
// f32); an empty cluster becomes NaN (the host re-seeds it, split_clusters).  `order` lists the member rows of every
// cluster in ascending row order (stable counting sort on the host), off[c]..off[c+1].  One block per cluster, one
// thread per dimension: the threads of a block read one member row together (coalesced), every (c, j) sum is a
// single sequential chain exactly like the reference's.
__global__ void __launch_bounds__(256)
kmeans_update_kernel(const float* __restrict__ data, uint32_t dim, const uint32_t* __restrict__ order,
                     const uint32_t* __restrict__ off, float* __restrict__ centroids) {
  const uint32_t c = blockIdx.x;
  const uint32_t lo = off[c], hi = off[c + 1];
  for (uint32_t j = threadIdx.x; j < dim; j += blockDim.x) {
    float s = 0.0f;
    uint32_t r = lo;
    // four independent loads in flight, added in order
    for (; r + 4 <= hi; r += 4) {
      const float a0 = data[(size_t)order[r] * dim + j], a1 = data[(size_t)order[r + 1] * dim + j],
                  a2 = data[(size_t)order[r + 2] * dim + j], a3 = data[(size_t)order[r + 3] * dim + j];
      s = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(s, a0), a1), a2), a3);
    }
    for (; r < hi; r++) s = __fadd_rn(s, data[(size_t)order[r] * dim + j]);
    centroids[(size_t)c * dim + j] = hi > lo ? __fdiv_rn(s, (float)(hi - lo)) : __int_as_float(0x7fc00000);
  }
}

// [n][dim] -> [n][w] : columns [c0, c0 + w) of every row (the sub-vector slice one PQ codebook is trained on)
__global__ void slice_columns_kernel(const float* __restrict__ src, uint64_t n, uint32_t dim, uint32_t c0, uint32_t w,
                                     float* __restrict__ dst) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * w) return;
  const uint64_t r = i / w;
  const uint32_t j = (uint32_t)(i - r * w);
  dst[i] = src[r * dim + c0 + j];
}

// x - centroid[part] per row (residual of ivf/v2.rs:285-301 at build time, arrow_arith::numeric::sub: plain f32
// subtraction); part == 0xffffffff (no partition) leaves the row as it is.
__global__ void residual_kernel(const float* __restrict__ x, const uint32_t* __restrict__ part,
                                const float* __restrict__ centroids, uint64_t n, uint32_t dim, float* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * dim) return;
  const uint64_t r = i / dim;
  const uint32_t j = (uint32_t)(i - r * dim);
  const uint32_t p = part[r];
  out[i] = p == 0xffffffffu ? x[i] : __fsub_rn(x[i], centroids[(size_t)p * dim + j]);
}

// ProductQuantizer::transform (lance-index/src/vector/pq.rs:120-190; argmin per sub-space, tested by pq.rs:595-632):
// code[row][m] = argmin_j dist(sub-vector m of the row, codebook[m][j]) with l2_distance_batch / dot_distance_batch
// (the arithmetic of the ADC table: lut_elem) and argmin_value_float (strict <, first minimum, NaN skipped).
// grid = (ceil(n / 256), M); the block's codebook slice [256][dsub] sits in shared memory and is read as a broadcast.
template <int DSUB>
__global__ void __launch_bounds__(256)
pq_encode_kernel(const float* __restrict__ vecs, const float* __restrict__ codebook, uint64_t n, uint32_t dim, uint32_t M,
                 uint32_t dsub, int residual, uint8_t* __restrict__ out /*[n][M]*/) {
  extern __shared__ float cb[];  // [256][dsub]
  const uint32_t m = blockIdx.y;
  for (uint32_t i = threadIdx.x; i < 256 * dsub; i += 256) cb[i] = codebook[(size_t)m * 256 * dsub + i];
  __syncthreads();
  const uint64_t row = (uint64_t)blockIdx.x * 256 + threadIdx.x;
  if (row >= n) return;
  const float* xp = vecs + row * dim + (size_t)m * dsub;
  float x[DSUB > 0 ? DSUB : 1];
  if (DSUB > 0) {
#pragma unroll
    for (int i = 0; i < DSUB; i++) x[i] = xp[i];
  }
  float best = __int_as_float(0x7f800000);
  uint32_t bi = 0;
  for (uint32_t j = 0; j < 256; j++) {
    const float v = DSUB > 0 ? lut_elem<DSUB>(residual != 0, x, cb + (size_t)j * DSUB, dsub)
                             : lut_elem<0>(residual != 0, xp, cb + (size_t)j * dsub, dsub);
    if (v < best) {
      best = v;
      bi = j;
    }
  }
  out[row * M + m] = (uint8_t)bi;
}

}  // namespace lb200

// scan4.cuh -- 4-bit PQ: ADC table, u8 quantisation of the table and the code scan
// (rust/lance-index/src/vector/pq/distance.rs:138-229 compute_pq_distance_4bit / quantize_distance_table).
//
// One CTA per (query, probed partition) pair.  Per pair the reference
//   1. builds T[m][j], j < 16 (build_distance_table_*_impl::<4>, distance.rs:24,60),
//   2. quantises it to u8: qmin = min T, qmax = max over ADJACENT sub-space pairs of (max T[m] + max T[m+1]),
//      entry = ceil((T - qmin) * 255 / (qmax - qmin)) as u8 (saturating),
//   3. per code byte (two sub-spaces: low nibble 2j, high nibble 2j+1) looks both entries up; rows in whole
//      16-row blocks add them as WRAPPING u8 (u8x16 `+` is _mm_add_epi8 on x86_64, simd/u8.rs:279-287), the
//      n % 16 tail rows add them separately; sums are widened to f32 and accumulated,
//   4. dequantises: d * (qmax - qmin) / 255 + qmin.
// Everything between the table and step 4 is integer arithmetic, so the result is bit-identical by construction;
// the f32 steps use explicit round-to-nearest intrinsics in the reference's order (no FMA contraction).
//
// Layout: codes4 = concat over partitions of [M/2][n_p] bytes (the reference's transposed 4-bit storage), rows of a
// partition sorted by row id at create time so that (distance, local index) order == (distance, row id) order.
// Step 3 depends on a row's POSITION in the reference's order (block or tail), so tail4[row] keeps that bit.
// This is the first correct path for 4-bit PQ (SURVEY 8(a) row A5''): distances of a pair are written to an HBM
// scratch row and selected by flat_select_kernel, as the reference's distance_all + sort does; it is not tuned.
#pragma once
#include <cstdint>

#include "dist.cuh"

namespace lb200 {

__device__ __forceinline__ uint8_t sat_u8(float v) {  // Rust `f32 as u8`
  if (!(v == v)) return 0;
  if (v <= 0.0f) return 0;
  if (v >= 255.0f) return 255;
  return (uint8_t)v;
}

__global__ void __launch_bounds__(256)
scan4_kernel(int metric, const float* __restrict__ keys, const uint32_t* __restrict__ probes, uint32_t P, uint32_t dim,
             uint32_t M, const float* __restrict__ centroids, const float* __restrict__ codebook /*[M][16][dsub]*/,
             const uint64_t* __restrict__ part_off, const uint8_t* __restrict__ codes4,
             const uint8_t* __restrict__ tail4, float* __restrict__ out, size_t out_stride,
             unsigned long long* __restrict__ stats) {
  extern __shared__ float sm4[];
  float* r = sm4;                 // [dim] residual (or the key for dot)
  float* T = sm4 + dim;           // [M][16]
  float* smax = T + (size_t)M * 16;  // [M]
  uint8_t* qt = (uint8_t*)(smax + M);  // [M][16]
  __shared__ float s_red[8];
  __shared__ float s_qmin, s_qmax;
  const uint32_t pair = blockIdx.x, q = pair / P, p = probes[pair], tid = threadIdx.x;
  const uint32_t dsub = dim / M;
  const float* key = keys + (size_t)q * dim;
  const float* c = centroids + (size_t)p * dim;
  for (uint32_t i = tid; i < dim; i += 256) r[i] = metric == 2 ? key[i] : __fsub_rn(key[i], c[i]);  // v2.rs:285-301
  __syncthreads();
  float lmin = __int_as_float(0x7f800000);
  for (uint32_t e = tid; e < M * 16; e += 256) {
    const uint32_t m = e >> 4;
    const float* cb = codebook + (size_t)e * dsub;
    float v = metric == 2 ? dot_distance(r + (size_t)m * dsub, cb, (int)dsub) : l2_batch_elem(r + (size_t)m * dsub, cb, (int)dsub);
    T[e] = v;
    lmin = fminf(lmin, v);
  }
  for (int o = 16; o; o >>= 1) lmin = fminf(lmin, __shfl_xor_sync(0xffffffffu, lmin, o));
  if ((tid & 31) == 0) s_red[tid >> 5] = lmin;
  __syncthreads();
  for (uint32_t m = tid; m < M; m += 256) {
    float mx = __int_as_float(0xff800000);
    for (int j = 0; j < 16; j++) mx = fmaxf(mx, T[m * 16 + j]);
    smax[m] = mx;
  }
  __syncthreads();
  if (tid == 0) {
    float mn = s_red[0];
    for (int w = 1; w < 8; w++) mn = fminf(mn, s_red[w]);
    float mx = __int_as_float(0xff800000);
    for (uint32_t m = 0; m + 1 < M; m++) mx = fmaxf(mx, __fadd_rn(smax[m], smax[m + 1]));
    s_qmin = mn;
    s_qmax = mx;
  }
  __syncthreads();
  const float qmin = s_qmin, range = __fsub_rn(s_qmax, s_qmin);
  for (uint32_t e = tid; e < M * 16; e += 256)
    qt[e] = sat_u8(ceilf(__fdiv_rn(__fmul_rn(__fsub_rn(T[e], qmin), 255.0f), range)));
  __syncthreads();
  const uint64_t off = part_off[p], n = part_off[p + 1] - off;
  const uint8_t* cp = codes4 + off * (M / 2);
  float* orow = out + (size_t)pair * out_stride;
  for (uint64_t i = tid; i < out_stride; i += 256) {
    float v = __int_as_float(0x7fc00000);
    if (i < n) {
      uint32_t acc = 0;
      const bool blk = !tail4[off + i];  // was this row in a whole 16-row block of the reference's order?
      for (uint32_t j = 0; j < M / 2; j++) {
        const uint8_t b = cp[(size_t)j * n + i];
        const uint32_t lo = qt[(2 * j) * 16 + (b & 15)], hi = qt[(2 * j + 1) * 16 + (b >> 4)];
        acc += blk ? ((lo + hi) & 255u) : (lo + hi);
      }
      v = __fadd_rn(__fdiv_rn(__fmul_rn((float)acc, range), 255.0f), qmin);
    }
    orow[i] = v;
  }
  if (stats && tid == 0) {
    atomicAdd(&stats[0], (unsigned long long)n * (M / 2));
    atomicAdd(&stats[1], 1ull);
  }
}

// local index of a selected row -> row id and position in the index (for the merge and the refine)
__global__ void gather4_kernel(const uint32_t* __restrict__ probes, const uint64_t* __restrict__ part_off,
                               const uint64_t* __restrict__ row_ids, const uint64_t* __restrict__ sel_idx,
                               const uint32_t* __restrict__ sel_cnt, size_t npairs, uint32_t kprime,
                               uint64_t* __restrict__ out_rowid, uint64_t* __restrict__ out_pos) {
  size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= npairs * kprime) return;
  size_t pair = e / kprime;
  uint32_t j = (uint32_t)(e - pair * kprime);
  uint64_t rid = ~0ull, pos = 0;
  if (j < sel_cnt[pair]) {
    pos = part_off[probes[pair]] + sel_idx[e];
    rid = row_ids[pos];
  }
  out_rowid[e] = rid;
  if (out_pos) out_pos[e] = pos;
}

}  // namespace lb200

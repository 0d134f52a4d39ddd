// flat.cuh -- K7 flat brute-force kNN (KNNVectorDistanceExec path, BASELINE config 2).
//
// Replaces compute_distance (rust/lance-index/src/vector/flat.rs:28-86) ->
// {l2,cosine,dot}_distance_arrow_batch (rust/lance-linalg/src/distance/l2.rs:288,
// cosine.rs:315, dot.rs:270) and the SortExec(fetch=k) + NOT NULL filter that follows it
// (rust/lance/src/dataset/scanner.rs:2066-2136).
#pragma once
#include "coarse.cuh"

namespace lb200 {

constexpr int FLAT_TQ = 4;
constexpr int FLAT_THREADS = 128;

// out[q][row] = metric distance (exact reference order); NaN for null rows.
__global__ void __launch_bounds__(FLAT_THREADS)
flat_dist_kernel(int metric, const float* __restrict__ queries, uint32_t Q, const float* __restrict__ vectors,
                 const uint8_t* __restrict__ validity, uint64_t n, uint32_t d, float* __restrict__ out,
                 size_t out_stride) {
  extern __shared__ float sq[];  // [TQ][d] + [TQ] norms
  float* snorm = sq + FLAT_TQ * d;
  const uint32_t q0 = blockIdx.y * FLAT_TQ;
  for (uint32_t i = threadIdx.x; i < FLAT_TQ * d; i += FLAT_THREADS) {
    uint32_t qq = q0 + i / d;
    sq[i] = qq < Q ? queries[(size_t)qq * d + (i % d)] : 0.0f;
  }
  __syncthreads();
  if (threadIdx.x < FLAT_TQ) snorm[threadIdx.x] = metric == 1 ? norm_l2(sq + threadIdx.x * d, (int)d) : 0.0f;
  __syncthreads();
  uint64_t row = (uint64_t)blockIdx.x * FLAT_THREADS + threadIdx.x;
  if (row >= n) return;
  const bool valid = !validity || ((validity[row >> 3] >> (row & 7)) & 1);
  const float* y = vectors + row * d;
#pragma unroll
  for (int t = 0; t < FLAT_TQ; t++) {
    if (q0 + t >= Q) break;
    float v = __int_as_float(0x7fc00000);
    if (valid) v = exact_distance(metric, sq + t * d, snorm[t], y, (int)d);
    out[(size_t)(q0 + t) * out_stride + row] = v;
  }
}

// Per-query top-k of a distance row by (total-order key, row index): 3-level radix select on
// the 32-bit ordered key, then an unordered gather of keys < T plus an index-ordered gather of
// the ties, then a small bitonic sort.  One block (256 threads) per query.  k <= 1024.
__global__ void __launch_bounds__(256)
flat_select_kernel(const float* __restrict__ dist, size_t dist_stride, uint64_t n_total, uint32_t nseg,
                   uint64_t seg_len, uint32_t k, uint32_t kpow2,
                   int has_lower, float lower, int has_upper, float upper, float* __restrict__ out_dist,
                   uint64_t* __restrict__ out_idx, uint32_t* __restrict__ out_cnt) {
  extern __shared__ unsigned long long sl[];  // [kpow2] result list
  __shared__ uint32_t hist[8 * 256];
  __shared__ uint32_t s_prefix, s_mask, s_kk, s_less, s_fill, s_tiebase, s_wsum[8];
  // block = (query, segment of the row): a row is cut into nseg segments so that a small query tile
  // still fills the GPU; the per-segment lists are merged afterwards (merge_lists_kernel)
  const uint32_t q = blockIdx.x;                       // output slot [query][segment]
  const uint64_t off = (uint64_t)(blockIdx.x % nseg) * seg_len;
  const uint64_t n = min(seg_len, n_total - off);
  const float* drow = dist + (size_t)(blockIdx.x / nseg) * dist_stride + off;
  auto admissible = [&](float v) {
    bool ok = v == v;
    if (has_lower) ok = ok && v >= lower;
    if (has_upper) ok = ok && v < upper;
    return ok;
  };
  if (threadIdx.x == 0) {
    s_prefix = 0;
    s_mask = 0;
    s_kk = k;  // rank still to be located inside the current prefix bucket
    s_less = 0;
  }
  __syncthreads();
  // Radix select of the k-th smallest admissible key, 8 bits per pass.  The top bits of a row's distances are nearly
  // all alike, so equal digits are first counted in registers (runs) and then added to a per-warp histogram: the
  // first version (12-bit digits, one shared histogram, thread 0 walking 4096 bins) spent its time in same-address
  // atomics and in that serial walk (cfg1_flat: 1.6 ms of select per 2.6 ms of scan).
  bool have_kth = true;
  const uint32_t lane_ = threadIdx.x & 31, w_ = threadIdx.x >> 5;
  uint32_t(*whist)[256] = reinterpret_cast<uint32_t(*)[256]>(hist);   // [8 warps][256]
#pragma unroll 1
  for (int shift = 24; shift >= 0 && have_kth; shift -= 8) {
    for (uint32_t i = threadIdx.x; i < 8 * 256; i += 256) hist[i] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix, mask = s_mask;
    uint32_t run_d = 0xffffffffu, run_c = 0;
    for (uint64_t i = threadIdx.x; i < n; i += 256) {
      const float v = drow[i];
      if (!admissible(v)) continue;
      const uint32_t key = ordered_key(v);
      if ((key & mask) != prefix) continue;
      const uint32_t dg = (key >> shift) & 255u;
      if (dg == run_d) run_c++;
      else {
        if (run_c) atomicAdd(&whist[w_][run_d], run_c);
        run_d = dg;
        run_c = 1;
      }
    }
    if (run_c) atomicAdd(&whist[w_][run_d], run_c);
    __syncthreads();
    uint32_t tot = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) tot += whist[j][threadIdx.x];
    __syncthreads();
    hist[threadIdx.x] = tot;
    __syncthreads();
    if (w_ == 0) {  // lane owns bins 8 lane .. 8 lane + 7
      uint32_t c8[8], sum = 0;
#pragma unroll
      for (int j = 0; j < 8; j++) { c8[j] = hist[lane_ * 8 + j]; sum += c8[j]; }
      uint32_t incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane_ >= (uint32_t)o) incl += v;
      }
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      const uint32_t kk = s_kk;
      uint32_t before = incl - sum;
      __syncwarp();
      if (total < kk) {
        if (lane_ == 0) s_kk = 0xffffffffu;  // fewer than k admissible rows: take everything
      } else if (before < kk && kk <= incl) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
          if (before < kk && kk <= before + c8[j]) {
            s_less += before;
            s_kk = kk - before;
            s_prefix = prefix | ((lane_ * 8 + j) << shift);
            s_mask = mask | (255u << shift);
          }
          before += c8[j];
        }
      }
    }
    __syncthreads();
    have_kth = s_kk != 0xffffffffu;
  }
  const uint32_t T = have_kth ? s_prefix : 0xffffffffu;  // k-th smallest key
  const uint32_t nless = have_kth ? s_less : 0;          // #keys < T (when have_kth)
  const uint32_t need_ties = have_kth ? k - nless : 0;
  for (uint32_t i = threadIdx.x; i < kpow2; i += 256) sl[i] = ~0ull;
  if (threadIdx.x == 0) {
    s_fill = 0;
    s_tiebase = 0;
  }
  __syncthreads();
  if (!have_kth) {
    // fewer than k admissible rows: gather all of them (count < k <= kpow2)
    for (uint64_t i = threadIdx.x; i < n; i += 256) {
      float v = drow[i];
      if (!admissible(v)) continue;
      uint32_t o = atomicAdd(&s_fill, 1u);
      sl[o] = ((unsigned long long)ordered_key(v) << 32) | (uint32_t)(i + off);
    }
  } else {
    // keys < T: unordered; keys == T: first need_ties in index order (block-ordered compaction)
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint64_t base = 0; base < n; base += 256) {
      uint64_t i = base + threadIdx.x;
      bool tie = false;
      if (i < n) {
        float v = drow[i];
        if (admissible(v)) {
          uint32_t key = ordered_key(v);
          if (key < T) {
            uint32_t o = atomicAdd(&s_fill, 1u);
            sl[o] = ((unsigned long long)key << 32) | (uint32_t)(i + off);
          }
          tie = key == T;
        }
      }
      uint32_t bal = __ballot_sync(0xffffffffu, tie);
      if (lane == 0) s_wsum[w] = __popc(bal);
      __syncthreads();
      uint32_t before = s_tiebase;
      for (uint32_t ww = 0; ww < w; ww++) before += s_wsum[ww];
      uint32_t r = before + __popc(bal & ((1u << lane) - 1u));
      if (tie && r < need_ties) sl[nless + r] = ((unsigned long long)T << 32) | (uint32_t)(i + off);
      __syncthreads();
      if (threadIdx.x == 0) {
        uint32_t tot = 0;
        for (int ww = 0; ww < 8; ww++) tot += s_wsum[ww];
        s_tiebase += tot;
      }
      __syncthreads();
      if (s_tiebase >= need_ties && base + 256 < n) {
        // ties satisfied: finish the (key < T) gather without the ordered machinery
        for (uint64_t j = base + 256 + threadIdx.x; j < n; j += 256) {
          float v = drow[j];
          if (!admissible(v)) continue;
          uint32_t key = ordered_key(v);
          if (key < T) {
            uint32_t o = atomicAdd(&s_fill, 1u);
            sl[o] = ((unsigned long long)key << 32) | (uint32_t)(j + off);
          }
        }
        break;
      }
    }
  }
  __syncthreads();
  const uint32_t cnt = have_kth ? k : s_fill;
  block_bitonic_sort_u64(sl, kpow2);
  for (uint32_t i = threadIdx.x; i < k; i += 256) {
    unsigned long long v = sl[i];
    bool ok = i < cnt;
    out_dist[(size_t)q * k + i] = ok ? key_to_float((uint32_t)(v >> 32)) : __int_as_float(0x7f800000);
    out_idx[(size_t)q * k + i] = ok ? (uint64_t)(uint32_t)v : ~0ull;
  }
  if (threadIdx.x == 0) out_cnt[q] = cnt;
}

}  // namespace lb200

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
// Layout: codes4p = concat over partitions of [M/2][npad] bytes (the reference's transposed 4-bit storage with the row
// count padded to a multiple of 16), rows of a partition sorted by row id at create time so that (distance, local
// index) order == (distance, row id) order.
// Step 3 depends on a row's POSITION in the reference's order (block or tail), so tail4[row] keeps that bit.
// This is the first correct path for 4-bit PQ (SURVEY 8(a) row A5''): distances of a pair are written to an HBM
// scratch row and selected by flat_select_kernel, as the reference's distance_all + sort does; it is not tuned.
#pragma once
#include <cstdint>

#include "coarse.cuh"
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
             const uint64_t* __restrict__ part_off, const uint64_t* __restrict__ pad_off,
             const uint8_t* __restrict__ codes4p, const uint8_t* __restrict__ tail4, float* __restrict__ out,
             size_t out_stride, unsigned long long* __restrict__ stats, const uint64_t* const* __restrict__ bitmask) {
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
  const uint64_t npad = pad_off[p + 1] - pad_off[p];
  const uint8_t* cp = codes4p + pad_off[p] * (M / 2);
  float* orow = out + (size_t)pair * out_stride;
  const uint64_t* bm = bitmask ? bitmask[p] : nullptr;
  if (bm) {
    // prefilter (flat/index.rs:106-122): only the allowed rows, each through PQDistCalculator::distance(id)
    // (pq/storage.rs:602-617) -- the UNQUANTISED f32 table, pair sums added in ascending j
    for (uint64_t i = tid; i < out_stride; i += 256) {
      float v = __int_as_float(0x7fc00000);
      if (i < n && ((bm[i >> 6] >> (i & 63)) & 1)) {
        float sacc = 0.0f;
        for (uint32_t j = 0; j < M / 2; j++) {
          const uint8_t b = cp[(size_t)j * npad + i];
          sacc = __fadd_rn(sacc, __fadd_rn(T[(2 * j) * 16 + (b & 15)], T[(2 * j + 1) * 16 + (b >> 4)]));
        }
        v = sacc;
      }
      orow[i] = v;
    }
    if (stats && tid == 0) {
      atomicAdd(&stats[0], (unsigned long long)n * (M / 2));
      atomicAdd(&stats[1], 1ull);
    }
    return;
  }
  for (uint64_t i = tid; i < out_stride; i += 256) {
    float v = __int_as_float(0x7fc00000);
    if (i < n) {
      uint32_t acc = 0;
      const bool blk = !tail4[off + i];  // was this row in a whole 16-row block of the reference's order?
      for (uint32_t j = 0; j < M / 2; j++) {
        const uint8_t b = cp[(size_t)j * npad + i];
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

// ---- fused variant -------------------------------------------------------------------------------------------
// Same arithmetic, no HBM distance scratch: the pre-dequantisation sum S of a row is a small integer
// (<= (M/2) * 510), so the CTA selects on a shared-memory HISTOGRAM of S instead of on floats.
//   pass 1  S of every admissible row -> hist[S]; after the first 4096 rows the k'-th smallest so far bounds which
//           rows still need counting (rows above it can never be selected), so the atomics stop early;
//   select  k'-th smallest: bin S*; T = f(S*) with f = the reference's dequantisation; because f may map adjacent S
//           to one float, "ties" are all bins with f(S) == T and are taken in row order (= row id order);
//   pass 2  re-reads the codes (L2-resident: one partition), gathers rows below T and the first ties, sorts the
//           <= k' keys and writes (distance, row id, index position).
// Codes: layout `codes4p` = per partition [M/2][npad], npad = n rounded up to 16, base 16-byte aligned, so a thread
// takes 4 consecutive rows of a plane with one 32-bit load and a warp reads 128 contiguous bytes.  The sums of the
// first 4096 rows are kept in shared memory (u16), so partitions up to 4096 rows are scanned ONCE.  Lookups of block rows go through a per-pair 8-bit table
// pt[j][byte] = (qt[2j][lo] + qt[2j+1][hi]) mod 256 built once per CTA (one LDS per code byte instead of two plus the
// nibble split; 256 B per plane keeps bank conflicts at most two-way).
// [lower, upper) is applied on f(S) (exactly the float the reference filters on).
struct Scan4Sel {
  uint32_t s_lt, s_hi, nless, need_ties, have_kth;
};

static __device__ __noinline__ float dequant4(uint32_t S, float range, float qmin) {
  return __fadd_rn(__fdiv_rn(__fmul_rn((float)S, range), 255.0f), qmin);
}

// k-th smallest over hist[0..nbins): fills sel (thread 0), all threads must call. scratch: part[256].
static __device__ __noinline__ void scan4_find_kth(const uint32_t* hist, uint32_t nbins, uint32_t k, float range, float qmin,
                                               uint32_t* part, Scan4Sel* sel) {
  const uint32_t tid = threadIdx.x, chunk = (nbins + 255) / 256;
  const uint32_t b0 = tid * chunk, b1 = min(nbins, b0 + chunk);
  uint32_t c = 0;
  for (uint32_t b = b0; b < b1; b++) c += hist[b];
  part[tid] = c;
  __syncthreads();
  if (tid < 32) {  // warp 0: scan of the 256 chunk sums, 8 per lane; the lane that crosses k finishes the search
    uint32_t loc[8], sum = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      loc[i] = part[tid * 8 + i];
      sum += loc[i];
    }
    uint32_t incl = sum;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (tid >= (uint32_t)o) incl += v;
    }
    uint32_t cum = incl - sum;
    const bool hit = cum < k && incl >= k;
    const uint32_t any = __ballot_sync(0xffffffffu, hit);
    if (!any) {
      if (tid == 31) {
        sel->have_kth = 0;
        sel->s_lt = nbins;   // everything counted is "less"
        sel->s_hi = nbins - 1;
        sel->nless = incl;
        sel->need_ties = 0;
      }
    } else if (hit) {
      uint32_t t = 0;
      for (; t < 8; t++) {
        if (cum + loc[t] >= k) break;
        cum += loc[t];
      }
      uint32_t b = (tid * 8 + t) * chunk;
      for (;; b++) {
        if (cum + hist[b] >= k) break;
        cum += hist[b];
      }
      // bins whose dequantised float equals the threshold's: ties are taken in row order across all of them
      const float T = dequant4(b, range, qmin);
      uint32_t lo = b, hi = b;
      while (lo > 0 && dequant4(lo - 1, range, qmin) == T) {
        lo--;
        cum -= hist[lo];
      }
      while (hi + 1 < nbins && dequant4(hi + 1, range, qmin) == T) hi++;
      sel->have_kth = 1;
      sel->s_lt = lo;
      sel->s_hi = hi;
      sel->nless = cum;
      sel->need_ties = k - cum;
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256, 6)
scan4f_kernel(int metric, const float* __restrict__ keys, const uint32_t* __restrict__ probes, uint32_t P, uint32_t dim,
              uint32_t M, const float* __restrict__ centroids, const float* __restrict__ codebook,
              const uint64_t* __restrict__ part_off, const uint64_t* __restrict__ pad_off /*[nlist+1] padded rows*/,
              const uint8_t* __restrict__ codes4p, const uint8_t* __restrict__ tail4,
              const uint64_t* __restrict__ row_ids, uint32_t kprime, uint32_t kpow2, int has_lower, float lower,
              int has_upper, float upper, float* __restrict__ out_dist, uint64_t* __restrict__ out_rowid,
              uint64_t* __restrict__ out_pos, uint32_t* __restrict__ out_cnt, unsigned long long* __restrict__ stats) {
  extern __shared__ float sm4[];
  const uint32_t nbins = (M / 2) * 510 + 1;
  float* r = sm4;
  float* T = sm4 + dim;
  float* smax = T + (size_t)M * 16;
  unsigned long long* sl = (unsigned long long*)(smax + ((M + 1) & ~1u));  // [kpow2], 8-byte aligned
  uint32_t* hist = (uint32_t*)(sl + kpow2);                                  // [nbins]
  uint8_t* qt = (uint8_t*)(hist + ((nbins + 1) & ~1u));                      // [M][16], 8-byte aligned
  uint8_t* pt = qt + (size_t)M * 16;                                         // [M/2][256] wrapped pair sums
  uint16_t* scache = (uint16_t*)(pt + (size_t)(M / 2) * 256);                // [4096] sums of the first rows
  __shared__ float s_red[8];
  __shared__ float s_qmin, s_qmax;
  __shared__ uint32_t s_part[256], s_wsum[8], s_fill, s_tiebase, s_adm[3];
  __shared__ Scan4Sel s_sel;
  const uint32_t pair = blockIdx.x, q = pair / P, p = probes[pair], tid = threadIdx.x;
  const uint32_t dsub = dim / M;
  const float* key = keys + (size_t)q * dim;
  const float* c = centroids + (size_t)p * dim;
  for (uint32_t i = tid; i < dim; i += 256) r[i] = metric == 2 ? key[i] : __fsub_rn(key[i], c[i]);
  for (uint32_t b = tid; b < nbins; b += 256) hist[b] = 0;
  for (uint32_t i = tid; i < kpow2; i += 256) sl[i] = ~0ull;
  if (tid == 0) {
    s_fill = 0;
    s_tiebase = 0;
  }
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
  const uint64_t poff = pad_off[p], npad = pad_off[p + 1] - poff;
  const uint8_t* cp = codes4p + poff * (M / 2);
  const uint8_t* tl = tail4 + off;
  auto admissible_f = [&](uint32_t S) {
    const float v = dequant4(S, range, qmin);
    bool ok = v == v;
    if (has_lower) ok = ok && v >= lower;
    if (has_upper) ok = ok && v < upper;
    return ok;
  };
  // f is monotone non-decreasing in S when range >= 0 and finite: [lower, upper) on floats becomes a range of S,
  // found once per CTA; no bounds -> every S is admissible (f finite or +inf, never NaN); anything else per row.
  if (tid == 0) {
    const bool fin = (range - range == 0.0f) && (qmin - qmin == 0.0f);
    uint32_t mode = 2, alo = 0, ahi = nbins;
    if (fin && !has_lower && !has_upper) mode = 0;
    else if (fin && range >= 0.0f) {
      mode = 1;
      auto first_ge = [&](float x) {  // first S in [0, nbins] with f(S) >= x
        uint32_t lo = 0, hi = nbins;
        while (lo < hi) {
          const uint32_t mid = (lo + hi) >> 1;
          if (dequant4(mid, range, qmin) >= x) hi = mid; else lo = mid + 1;
        }
        return lo;
      };
      if (has_lower) alo = (lower == lower) ? first_ge(lower) : nbins;
      if (has_upper) ahi = (upper == upper) ? first_ge(upper) : 0;
    }
    s_adm[0] = mode; s_adm[1] = alo; s_adm[2] = ahi;
  }
  __syncthreads();
  const uint32_t amode = s_adm[0], alo = s_adm[1], ahi = s_adm[2];
  auto admissible = [&](uint32_t S) {
    if (amode == 0) return true;
    if (amode == 1) return S >= alo && S < ahi;
    return admissible_f(S);
  };
  // pair table for block rows: pt[j][byte] = (qt[2j][byte & 15] + qt[2j+1][byte >> 4]) mod 256 -- exactly the wrapping
  // u8 lane sum of _mm_add_epi8; one 8-bit lookup per code byte (256 B per plane: at most two words per bank)
  for (uint32_t e = tid; e < (M / 2) * 256; e += 256) {
    const uint32_t j = e >> 8, b = e & 255u;
    pt[e] = (uint8_t)((uint32_t)qt[(2 * j) * 16 + (b & 15)] + (uint32_t)qt[(2 * j + 1) * 16 + (b >> 4)]);
  }
  __syncthreads();
  // S of the 4 rows i0..i0+3 (i0 % 4 == 0, i0 < npad): one 32-bit load per code plane.  Rows of the reference's
  // n % 16 tail add the two looked-up bytes without the wrap (rare: the slow branch).
  auto sums4 = [&](uint64_t i0, uint32_t S[4]) {
    uint32_t tailmask = 0;
#pragma unroll
    for (int rr = 0; rr < 4; rr++) {
      S[rr] = 0;
      if (i0 + rr < n && tl[i0 + rr]) tailmask |= 1u << rr;
    }
    const uint8_t* c0 = cp + i0;
    if (tailmask == 0) {
#pragma unroll 2
      for (uint32_t j = 0; j < M / 2; j++) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(c0 + (size_t)j * npad);
        const uint8_t* t = pt + j * 256;
#pragma unroll
        for (int rr = 0; rr < 4; rr++) S[rr] += t[(w >> (8 * rr)) & 255u];
      }
    } else {
#pragma unroll 1
      for (uint32_t j = 0; j < M / 2; j++) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(c0 + (size_t)j * npad);
        const uint8_t* t0 = qt + (2 * j) * 16;
#pragma unroll
        for (int rr = 0; rr < 4; rr++) {
          const uint32_t b = (w >> (8 * rr)) & 255u;
          const uint32_t v = (uint32_t)t0[b & 15] + (uint32_t)t0[16 + (b >> 4)];
          S[rr] += ((tailmask >> rr) & 1u) ? v : (v & 255u);
        }
      }
    }
  };
  constexpr uint64_t GROUP = 1024, CACHED = 4096;  // rows per CTA step; rows whose sums are kept in shared memory
  const uint32_t lane = tid & 31, wrp = tid >> 5;

  // ---- pass 1: histogram of S; the sums of the first 4096 rows are kept (u16) for pass 2 ---------------------------
  // (loops deliberately NOT unrolled: the first version of this kernel spent most of its issue slots waiting for
  // instruction fetch, ncu "no_instruction" 16.5 per issue)
  uint32_t bound = nbins - 1;
#pragma unroll 1
  for (uint64_t base = 0; base < n; base += GROUP) {
    const uint64_t i0 = base + (uint64_t)tid * 4;
    if (i0 < n) {
      uint32_t S[4];
      sums4(i0, S);
#pragma unroll
      for (int rr = 0; rr < 4; rr++)
        if (i0 + rr < n && S[rr] <= bound && admissible(S[rr])) atomicAdd(&hist[S[rr]], 1u);
      if (i0 < CACHED)
        *reinterpret_cast<uint2*>(scache + i0) = make_uint2(S[0] | (S[1] << 16), S[2] | (S[3] << 16));
    }
    if (base + GROUP == CACHED && n > CACHED) {  // uniform
      __syncthreads();
      scan4_find_kth(hist, nbins, kprime, range, qmin, s_part, &s_sel);
      if (s_sel.have_kth) bound = s_sel.s_hi;
      __syncthreads();
    }
  }
  __syncthreads();
  scan4_find_kth(hist, nbins, kprime, range, qmin, s_part, &s_sel);
  const uint32_t s_lt = s_sel.s_lt, s_hi = s_sel.s_hi, nless = s_sel.nless, need_ties = s_sel.need_ties;
  const bool have_kth = s_sel.have_kth != 0;

  // ---- pass 2: gather (rows below the threshold in any order, ties in row order) ----------------------------------
#pragma unroll 1
  for (uint64_t base = 0; base < n; base += GROUP) {
    const uint64_t i0 = base + (uint64_t)tid * 4;
    uint32_t S[4] = {0, 0, 0, 0};
    uint32_t tmask = 0;  // rows of this thread that tie with the threshold
    if (i0 < n) {
      if (i0 < CACHED) {
        const uint2 c2 = *reinterpret_cast<const uint2*>(scache + i0);
        S[0] = c2.x & 0xffffu; S[1] = c2.x >> 16; S[2] = c2.y & 0xffffu; S[3] = c2.y >> 16;
      } else {
        sums4(i0, S);
      }
#pragma unroll
      for (int rr = 0; rr < 4; rr++) {
        if (i0 + rr >= n || S[rr] > bound || !admissible(S[rr])) continue;
        if (S[rr] < s_lt) {
          const uint32_t o = atomicAdd(&s_fill, 1u);
          sl[o] = ((unsigned long long)ordered_key(dequant4(S[rr], range, qmin)) << 32) | (uint32_t)(i0 + rr);
        } else if (have_kth && S[rr] <= s_hi) {
          tmask |= 1u << rr;
        }
      }
    }
    if (have_kth && s_tiebase < need_ties) {  // uniform: s_tiebase changes only between barriers
      if (!__syncthreads_or(tmask != 0)) continue;
      // rank of this thread's ties in row order: threads hold consecutive rows, warps consecutive threads
      const uint32_t cnt = __popc(tmask);
      uint32_t incl = cnt;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += v;
      }
      if (lane == 31) s_wsum[wrp] = incl;
      __syncthreads();
      uint32_t before = s_tiebase + incl - cnt;
      for (uint32_t ww = 0; ww < wrp; ww++) before += s_wsum[ww];
#pragma unroll
      for (int rr = 0; rr < 4; rr++) {
        if (!((tmask >> rr) & 1u)) continue;
        if (before < need_ties)
          sl[nless + before] = ((unsigned long long)ordered_key(dequant4(S[rr], range, qmin)) << 32) | (uint32_t)(i0 + rr);
        before++;
      }
      __syncthreads();
      if (tid == 0) {
        uint32_t tot = 0;
        for (int ww = 0; ww < 8; ww++) tot += s_wsum[ww];
        s_tiebase += tot;
      }
      __syncthreads();
    }
  }
  __syncthreads();
  const uint32_t cnt = have_kth ? kprime : s_fill;
  if (kpow2 <= 32) {  // one warp sorts in registers: no barriers
    if (tid < 32) {
      unsigned long long v = tid < kpow2 ? sl[tid] : ~0ull;
      for (uint32_t k = 2; k <= 32; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
          const unsigned long long o = __shfl_xor_sync(0xffffffffu, v, j);
          const bool up = (tid & k) == 0, lower_lane = (tid & j) == 0;
          v = (lower_lane == up) ? (v < o ? v : o) : (v < o ? o : v);
        }
      if (tid < kprime) {
        const bool ok = tid < cnt;
        const uint64_t pos = off + (uint32_t)v;
        out_dist[(size_t)pair * kprime + tid] = ok ? key_to_float((uint32_t)(v >> 32)) : __int_as_float(0x7f800000);
        out_rowid[(size_t)pair * kprime + tid] = ok ? row_ids[pos] : ~0ull;
        if (out_pos) out_pos[(size_t)pair * kprime + tid] = ok ? pos : 0;
      }
    }
  } else {
    block_bitonic_sort_u64(sl, kpow2);
    for (uint32_t i = tid; i < kprime; i += 256) {
      const unsigned long long v = sl[i];
      const bool ok = i < cnt;
      const uint64_t pos = off + (uint32_t)v;
      out_dist[(size_t)pair * kprime + i] = ok ? key_to_float((uint32_t)(v >> 32)) : __int_as_float(0x7f800000);
      out_rowid[(size_t)pair * kprime + i] = ok ? row_ids[pos] : ~0ull;
      if (out_pos) out_pos[(size_t)pair * kprime + i] = ok ? pos : 0;
    }
  }
  if (tid == 0) {
    out_cnt[pair] = cnt;
    if (stats) {
      atomicAdd(&stats[0], (unsigned long long)n * (M / 2));
      atomicAdd(&stats[1], 1ull);
    }
  }
}

// ---- IVF_FLAT sub-index ------------------------------------------------------------------------------------------
// FlatDistanceCal::distance_all (rust/lance-index/src/vector/flat/storage.rs:315-375) of one probed partition:
// distance_type.func()(key, row) per row -- l2 / cosine (norm_l2(key), cosine_fast) / dot_distance on the raw f32 rows,
// no residual (FlatQuantizer::use_residual = false, quantizer.rs:46), key already normalised for cosine
// (ivf/v2.rs:370-373).  One CTA per (query, partition) pair; distances go to an HBM scratch row for
// flat_select_kernel.  exact_distance() is the device restatement the refine step already uses.
__global__ void __launch_bounds__(256)
scanflat_kernel(int metric, const float* __restrict__ keys, const uint32_t* __restrict__ probes, uint32_t P, uint32_t dim,
                const uint64_t* __restrict__ part_off, const float* __restrict__ raw, float* __restrict__ out,
                size_t out_stride, unsigned long long* __restrict__ stats, const uint64_t* const* __restrict__ bitmask) {
  extern __shared__ float smf[];  // [dim] key
  __shared__ float s_xn;
  const uint32_t pair = blockIdx.x, q = pair / P, p = probes[pair], tid = threadIdx.x;
  for (uint32_t i = tid; i < dim; i += 256) smf[i] = keys[(size_t)q * dim + i];
  __syncthreads();
  if (tid == 0) s_xn = metric == 1 ? norm_l2(smf, (int)dim) : 0.0f;
  __syncthreads();
  const float xn = s_xn;
  const uint64_t off = part_off[p], n = part_off[p + 1] - off;
  float* orow = out + (size_t)pair * out_stride;
  const uint64_t* bm = bitmask ? bitmask[p] : nullptr;  // prefilter: allowed rows only (flat/index.rs:106-122)
  for (uint64_t i = tid; i < out_stride; i += 256)
    orow[i] = i < n && (!bm || ((bm[i >> 6] >> (i & 63)) & 1)) ? exact_distance(metric, smf, xn, raw + (off + i) * dim, (int)dim)
                                                                 : __int_as_float(0x7fc00000);
  if (stats && tid == 0) {
    atomicAdd(&stats[0], (unsigned long long)n * dim * 4);
    atomicAdd(&stats[1], 1ull);
  }
}

// L2 variant with coalesced row reads: a HALF-WARP per row, lane l = accumulator lane l of l2_scalar::<f32,f32,16>
// (l2.rs:65-99: sums[l] += (x[c*16+l] - y[c*16+l])^2 over the chunks in order), so a chunk is one 64-byte read of
// the half-warp; the 16 lane sums are then added in lane order (shuffles) exactly as `sums.iter().sum()` does, after
// the d % 16 remainder that the reference sums first.  Four rows per half-warp step for load-level parallelism.
__global__ void __launch_bounds__(256)
scanflat_l2_kernel(const float* __restrict__ keys, const uint32_t* __restrict__ probes, uint32_t P, uint32_t dim,
                   const uint64_t* __restrict__ part_off, const float* __restrict__ raw, float* __restrict__ out,
                   size_t out_stride, unsigned long long* __restrict__ stats, const uint64_t* const* __restrict__ bitmask) {
  extern __shared__ float smf[];  // [dim] key
  const uint32_t pair = blockIdx.x, q = pair / P, p = probes[pair], tid = threadIdx.x;
  for (uint32_t i = tid; i < dim; i += 256) smf[i] = keys[(size_t)q * dim + i];
  __syncthreads();
  const uint64_t off = part_off[p], n = part_off[p + 1] - off;
  float* orow = out + (size_t)pair * out_stride;
  const uint32_t hw = tid >> 4, l = tid & 15;
  const uint32_t nchunk = dim >> 4, rem0 = nchunk << 4;
  const uint64_t* bm = bitmask ? bitmask[p] : nullptr;  // prefilter: rows that are not allowed come out as NaN
  for (uint64_t i = n + tid; i < out_stride; i += 256) orow[i] = __int_as_float(0x7fc00000);  // padding of the scratch row
  constexpr int RW = 4;  // rows per half-warp step: RW independent load streams
  for (uint64_t base = 0; base < n; base += 16 * RW) {   // uniform trip count
    uint64_t r[RW];
    const float* y[RW];
    float a[RW], sr[RW];
#pragma unroll
    for (int k = 0; k < RW; k++) {
      r[k] = base + 16 * k + hw;
      y[k] = raw + (off + (r[k] < n ? r[k] : 0)) * dim;
      a[k] = 0.0f;
      sr[k] = 0.0f;
    }
    for (uint32_t i = rem0; i < dim; i++) {   // remainder first (every lane the same value)
#pragma unroll
      for (int k = 0; k < RW; k++) {
        const float d = __fsub_rn(smf[i], y[k][i]);
        sr[k] = __fadd_rn(sr[k], __fmul_rn(d, d));
      }
    }
#pragma unroll 4
    for (uint32_t c = 0; c < nchunk; c++) {
      const float x = smf[c * 16 + l];
#pragma unroll
      for (int k = 0; k < RW; k++) {
        const float d = __fsub_rn(x, y[k][c * 16 + l]);
        a[k] = __fadd_rn(a[k], __fmul_rn(d, d));
      }
    }
#pragma unroll
    for (int k = 0; k < RW; k++) {
      float t = 0.0f;
#pragma unroll
      for (int i = 0; i < 16; i++) t = __fadd_rn(t, __shfl_sync(0xffffffffu, a[k], i, 16));
      if (l == 0 && r[k] < n)
        orow[r[k]] = !bm || ((bm[r[k] >> 6] >> (r[k] & 63)) & 1) ? __fadd_rn(sr[k], t) : __int_as_float(0x7fc00000);
    }
  }
  if (stats && tid == 0) {
    atomicAdd(&stats[0], (unsigned long long)n * dim * 4);
    atomicAdd(&stats[1], 1ull);
  }
}

// L2, several queries per CTA: a work item = (partition, group of up to 4 queries probing it) from the work list of
// the 8-bit path (worklist.cuh), so a partition's rows are read once per query GROUP instead of once per pair (r01:
// one CTA per pair, 0.46 of the roofline with every row crossing L2 once per query).  Arithmetic per (query, row) is
// l2_scalar::<f32,f32,16> as above: thread l of a half-warp owns lane sum l; the 16 lane sums of 16 rows are then
// transposed through shared memory so that thread l adds the sums of row l in lane order (one LDS + FADD per
// value instead of 16 shuffles per row) and the half-warp writes 16 consecutive distances.
// NCH = dim / 16 when the keys fit registers (dim 64 / 128 / 256), else 0: keys from shared memory, any dim.
constexpr int FLATG_QL = 4;
template <int NCH>
__global__ void __launch_bounds__(256)
scanflat_l2_group_kernel(const float* __restrict__ keys, uint32_t P, uint32_t dim, const uint64_t* __restrict__ part_off,
                         const float* __restrict__ raw, const uint4* __restrict__ item_desc,
                         const uint32_t* __restrict__ item_total, const uint32_t* __restrict__ pair_list,
                         const uint32_t* __restrict__ pair_off, float* __restrict__ out, size_t out_stride,
                         unsigned long long* __restrict__ stats, const uint64_t* const* __restrict__ bitmask) {
  constexpr int QL = FLATG_QL;
  extern __shared__ float fg_smem[];
  float* part_sums = fg_smem;                      // [16 half-warps][QL][16 rows][17]
  float* skeys = fg_smem + 16 * QL * 16 * 17;      // [QL][dim] (NCH == 0 only)
  __shared__ uint32_t s_pair[QL];
  const uint32_t tid = threadIdx.x, hw = tid >> 4, l = tid & 15;
  const uint32_t nchunk = NCH > 0 ? (uint32_t)NCH : dim >> 4, rem0 = nchunk << 4;
  float* my = part_sums + (size_t)hw * QL * 16 * 17;
  const uint32_t total = *item_total;
  for (uint32_t item = blockIdx.x; item < total; item += gridDim.x) {
    const uint4 de = item_desc[item];
    const uint32_t part = de.x, g = de.y;
    const uint32_t pbase = pair_off[part] + g * QL;
    const uint32_t nq = min((uint32_t)QL, pair_off[part + 1] - pbase);
    __syncthreads();   // the previous item's readers of s_pair / skeys are done
    if (tid < QL) s_pair[tid] = tid < nq ? pair_list[pbase + tid] : 0u;
    __syncthreads();
    float xq[QL][NCH > 0 ? NCH : 1];
    if (NCH > 0) {
#pragma unroll
      for (int q = 0; q < QL; q++)
#pragma unroll
        for (int c = 0; c < NCH; c++) xq[q][c] = (uint32_t)q < nq ? keys[(size_t)(s_pair[q] / P) * dim + c * 16 + l] : 0.0f;
    } else {
      for (uint32_t i = tid; i < QL * dim; i += 256) {
        const uint32_t q = i / dim;
        skeys[i] = q < nq ? keys[(size_t)(s_pair[q] / P) * dim + (i - q * dim)] : 0.0f;
      }
      __syncthreads();
    }
    const uint64_t off = part_off[part], n = part_off[part + 1] - off;
    const uint64_t* bm = bitmask ? bitmask[part] : nullptr;
    for (uint64_t base = 0; base < n; base += 256) {          // uniform trip count over the CTA
      const uint64_t r0 = base + hw * 16;
      if (NCH > 0) {
        // RW rows per step: all their loads are issued before the first is used (the kernel is latency-bound on the
        // row reads: r02 ncu of the 1-row version, 30 of 35 stall cycles per issue on the loads)
        constexpr int RW = 4;
#pragma unroll 1
        for (uint32_t j = 0; j < 16; j += RW) {
          float yv[RW][NCH > 0 ? NCH : 1];
#pragma unroll
          for (int k = 0; k < RW; k++) {
            const uint64_t r = r0 + j + k;
            const float* y = raw + (off + (r < n ? r : 0)) * dim + l;
#pragma unroll
            for (int c = 0; c < NCH; c++) yv[k][c] = __ldg(y + c * 16);
          }
#pragma unroll
          for (int k = 0; k < RW; k++) {
            float a[QL];
#pragma unroll
            for (int q = 0; q < QL; q++) a[q] = 0.0f;
#pragma unroll
            for (int c = 0; c < NCH; c++)
#pragma unroll
              for (int q = 0; q < QL; q++) {
                const float d = __fsub_rn(xq[q][c], yv[k][c]);
                a[q] = __fadd_rn(a[q], __fmul_rn(d, d));
              }
#pragma unroll
            for (int q = 0; q < QL; q++) my[(q * 16 + j + k) * 17 + l] = a[q];
          }
        }
      } else {
#pragma unroll 2
        for (uint32_t j = 0; j < 16; j++) {
          const uint64_t r = r0 + j;
          float a[QL];
#pragma unroll
          for (int q = 0; q < QL; q++) a[q] = 0.0f;
          if (r < n) {
            const float* y = raw + (off + r) * dim + l;
#pragma unroll 8
            for (uint32_t c = 0; c < nchunk; c++) {
              const float yv = __ldg(y + c * 16);
#pragma unroll
              for (int q = 0; q < QL; q++) {
                const float d = __fsub_rn(skeys[q * dim + c * 16 + l], yv);
                a[q] = __fadd_rn(a[q], __fmul_rn(d, d));
              }
            }
          }
#pragma unroll
          for (int q = 0; q < QL; q++) my[(q * 16 + j) * 17 + l] = a[q];
        }
      }
      __syncwarp();
      // thread l finishes row r0 + l of every query: remainder first (l2.rs:73-83), then the lanes in order
      const uint64_t r = r0 + l;
      if (r < n) {
        const bool allowed = !bm || ((bm[r >> 6] >> (r & 63)) & 1);
#pragma unroll
        for (int q = 0; q < QL; q++) {
          if ((uint32_t)q >= nq) break;
          float v = __int_as_float(0x7fc00000);
          if (allowed) {
            float sr = 0.0f;
            if (NCH == 0) {
              const float* y = raw + (off + r) * dim;
              for (uint32_t i = rem0; i < dim; i++) {
                const float d = __fsub_rn(skeys[q * dim + i], y[i]);
                sr = __fadd_rn(sr, __fmul_rn(d, d));
              }
            }
            float t = 0.0f;
#pragma unroll
            for (int i = 0; i < 16; i++) t = __fadd_rn(t, my[(q * 16 + l) * 17 + i]);
            v = __fadd_rn(sr, t);
          }
          out[(size_t)s_pair[q] * out_stride + r] = v;
        }
      }
      __syncwarp();
    }
    // padding of the scratch rows (flat_select reads out_stride entries)
    for (uint32_t q = 0; q < nq; q++)
      for (uint64_t i = n + tid; i < out_stride; i += 256) out[(size_t)s_pair[q] * out_stride + i] = __int_as_float(0x7fc00000);
    if (stats && tid == 0) {
      atomicAdd(&stats[0], (unsigned long long)n * dim * 4 * nq);
      atomicAdd(&stats[1], (unsigned long long)nq);
    }
  }
}

// pairs that probe an EMPTY partition have no work item: their scratch rows are all padding
__global__ void flat_empty_pairs_kernel(const uint32_t* __restrict__ probes, uint32_t npairs, const uint64_t* __restrict__ part_off,
                                        float* __restrict__ out, size_t out_stride) {
  const uint32_t e = blockIdx.x;
  if (e >= npairs) return;
  const uint32_t p = probes[e];
  if (part_off[p + 1] != part_off[p]) return;
  for (uint64_t i = threadIdx.x; i < out_stride; i += blockDim.x) out[(size_t)e * out_stride + i] = __int_as_float(0x7fc00000);
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

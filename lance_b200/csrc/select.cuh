// select.cuh -- warp-level top-k primitives over small candidate buffers in shared memory.
//
// Replaces the `res.sort_unstable(); take(k)` of FlatIndex::search
// (rust/lance-index/src/vector/flat/index.rs:124-130): candidates are (ordered distance key,
// local row position); ties on the key are ordered by row id (the reference's order is
// unstable there, the oracle pins (distance,row id)).
#pragma once
#include "dist.cuh"


namespace lb200 {

constexpr uint32_t FULL = 0xffffffffu;

// (key,rowid) strict-less with lazy row-id fetch (only on equal keys).
__device__ __forceinline__ bool cand_less(uint32_t ka, uint32_t pa, uint32_t kb, uint32_t pb,
                                          const uint64_t* __restrict__ rowids) {
  if (ka != kb) return ka < kb;
  uint64_t ra = rowids[pa], rb = rowids[pb];
  if (ra != rb) return ra < rb;
  return pa < pb;
}

// Keep the k smallest of (key[i],pos[i]), i<n, n>=k>=1, by (key,rowid); they end up in
// [0,k) in unspecified order.  Returns the k-th smallest key.  All 32 lanes must call.
static __device__ __noinline__ uint32_t warp_select_k(uint32_t* key, uint32_t* pos, uint32_t n,
                                               uint32_t k, const uint64_t* __restrict__ rowids) {
  const uint32_t lane = threadIdx.x & 31;
  // bitwise binary search for T = k-th smallest key: largest T with #(key < T) < k
  uint32_t T = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit) {
    uint32_t cand = T | (1u << bit);
    uint32_t c = 0;
    for (uint32_t i = lane; i < n; i += 32) c += (key[i] < cand) ? 1u : 0u;
    c = __reduce_add_sync(FULL, c);
    if (c < k) T = cand;
  }
  uint32_t nl = 0, ne = 0;
  for (uint32_t i = lane; i < n; i += 32) {
    uint32_t kk = key[i];
    nl += (kk < T) ? 1u : 0u;
    ne += (kk == T) ? 1u : 0u;
  }
  nl = __reduce_add_sync(FULL, nl);
  ne = __reduce_add_sync(FULL, ne);
  const uint32_t need = k - nl;  // ties to keep, 1 <= need <= ne
  const bool tie_split = ne > need;
  if (tie_split) {
    // Rare: the k-th key is shared by more rows than fit.  Rank the ties by (rowid,pos) on the
    // unmodified array, then knock the losers out (key := max) before compacting.  n <= 1024.
    uint32_t rej = 0, idx = 0;
    for (uint32_t i = lane; i < n; i += 32, idx++) {
      if (key[i] == T) {
        uint32_t pp = pos[i], r = 0;
        for (uint32_t j = 0; j < n; j++)
          if (j != i && key[j] == T && cand_less(T, pos[j], T, pp, rowids)) r++;
        if (r >= need) rej |= 1u << idx;
      }
    }
    __syncwarp();
    idx = 0;
    for (uint32_t i = lane; i < n; i += 32, idx++)
      if ((rej >> idx) & 1u) key[i] = 0xffffffffu;
    __syncwarp();
  }
  // in-place stable compaction, 32 entries per round (dst index <= src index always)
  uint32_t outn = 0;
  for (uint32_t base = 0; base < n; base += 32) {
    uint32_t i = base + lane;
    uint32_t kk = 0xffffffffu, pp = 0;
    bool keep = false;
    if (i < n) {
      kk = key[i];
      pp = pos[i];
      keep = kk <= T && !(tie_split && kk == 0xffffffffu);
    }
    uint32_t m = __ballot_sync(FULL, keep);
    __syncwarp();
    if (keep) {
      uint32_t dst = outn + __popc(m & ((1u << lane) - 1u));
      key[dst] = kk;
      pos[dst] = pp;
    }
    outn += __popc(m);
    __syncwarp();
  }
  return T;
}

// Sort (key,pos)[0..n) ascending by (key,rowid) via rank counting and write the result to
// global memory.  n is small (<= 2k').  All 32 lanes must call.
static __device__ __noinline__ void warp_sort_emit(const uint32_t* key, const uint32_t* pos, uint32_t n,
                                            const uint64_t* __restrict__ rowids /*partition base*/,
                                            uint64_t part_row_base, float* __restrict__ out_dist,
                                            uint64_t* __restrict__ out_pos,
                                            uint64_t* __restrict__ out_rowid) {
  const uint32_t lane = threadIdx.x & 31;
  for (uint32_t i = lane; i < n; i += 32) {
    uint32_t ki = key[i], pi = pos[i];
    uint32_t r = 0;
    for (uint32_t j = 0; j < n; j++) {
      uint32_t kj = key[j];
      if (kj < ki)
        r++;
      else if (kj == ki && j != i && cand_less(kj, pos[j], ki, pi, rowids))
        r++;
    }
    out_dist[r] = key_to_float(ki);
    out_pos[r] = part_row_base + pi;
    out_rowid[r] = rowids[pi];
  }
}

}  // namespace lb200

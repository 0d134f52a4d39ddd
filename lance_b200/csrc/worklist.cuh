// worklist.cuh -- device-side scheduling of (query, partition) pairs, K5 merge and K6 refine.
#pragma once
#include "select.cuh"

namespace lb200 {

// Repack the reference layout (concat over p of transposed [M][n_p], pq/storage.rs:357-377)
// into tile16 layout: partition p starts at tile tile_off[p]; tile t = [M][16] bytes.
__global__ void repack_codes_kernel(const uint8_t* __restrict__ src, const uint64_t* __restrict__ part_off,
                                    const uint64_t* __restrict__ tile_off, uint32_t nlist, uint32_t M,
                                    uint8_t* __restrict__ dst) {
  // one block per partition-slice: blockIdx.x = partition, blockIdx.y strides over tiles
  uint32_t p = blockIdx.x;
  uint64_t o = part_off[p], n = part_off[p + 1] - o;
  uint64_t ntile = (n + 15) >> 4;
  const uint8_t* s = src + o * M;
  uint8_t* d = dst + tile_off[p] * 16 * M;
  uint64_t total = ntile * 16 * M;
  for (uint64_t i = (uint64_t)blockIdx.y * blockDim.x + threadIdx.x; i < total;
       i += (uint64_t)gridDim.y * blockDim.x) {
    uint64_t t = i / (16 * M);
    uint32_t r = (uint32_t)(i - t * 16 * M);
    uint32_t m = r >> 4, k = r & 15;
    uint64_t row = t * 16 + k;
    d[i] = row < n ? s[(uint64_t)m * n + row] : 0;
  }
}

// Reference layout (concat over p of transposed [M][n_p]) -> skew streams.
// grid = (nlist, Y); one thread per 16-byte unit-lane.
__global__ void repack_skew_kernel(const uint8_t* __restrict__ src, const uint64_t* __restrict__ part_off,
                                   const uint64_t* __restrict__ sb_off, uint32_t M, uint32_t L, uint32_t D,
                                   uint8_t* __restrict__ dst) {
  const uint32_t p = blockIdx.x;
  const uint64_t o = part_off[p], n = part_off[p + 1] - o;
  const uint64_t sb0 = sb_off[p], nsb = sb_off[p + 1] - sb0;
  const uint32_t units = L * (M / 16) + 1;
  const uint8_t* s = src + o * M;
  const uint64_t total = nsb * units * 32;  // 16-byte records
  for (uint64_t rec = (uint64_t)blockIdx.y * blockDim.x + threadIdx.x; rec < total;
       rec += (uint64_t)gridDim.y * blockDim.x) {
    const uint32_t lane = (uint32_t)(rec & 31);
    const uint64_t ug = rec >> 5;
    const uint64_t sb = ug / units;
    const uint32_t g = (uint32_t)(ug - sb * units);
    const uint32_t d = lane % D;
    uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
    for (int b = 0; b < 16; b++) {
      int64_t t = (int64_t)g * 16 + b - (int64_t)d;
      if (t >= 0 && t < (int64_t)L * M) {
        uint32_t ri = (uint32_t)(t / M), m = (uint32_t)(t % M);
        uint64_t row = sb * 32 * L + (uint64_t)ri * 32 + lane;
        if (row < n) w[b >> 2] |= (uint32_t)s[(uint64_t)m * n + row] << (8 * (b & 3));
      }
    }
    reinterpret_cast<uint4*>(dst)[(sb0 * units) * 32 + rec] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__global__ void fill_u32_kernel(uint32_t* p, uint32_t v, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// Step 0: the partition each query meets FIRST in scheduling order (work items run in ascending
// partition id): the smallest probed, non-empty partition.  Its pair is the query's "cold" pair
// (no running bound yet); step 1/3 pack the cold pairs of a partition into the same work items so
// that few items pay the cold start (with G shards a query has only P/G pairs per shard: without
// the packing 1-(1-G/P)^QL of all items would hold a cold query).
// `queue_rank` (or null: the partition id itself) = position of a partition in the order of the item queue.
__global__ void first_part_kernel(const uint32_t* __restrict__ probes, uint32_t Q, uint32_t P,
                                  const uint64_t* __restrict__ part_off, uint32_t* __restrict__ minp,
                                  const uint32_t* __restrict__ queue_rank) {
  uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  uint32_t m = 0xffffffffu, mr = 0xffffffffu;
  for (uint32_t r = 0; r < P; r++) {
    uint32_t p = probes[(size_t)q * P + r];
    if (part_off[p + 1] > part_off[p]) {
      const uint32_t rk = queue_rank ? queue_rank[p] : p;
      if (rk < mr) { mr = rk; m = p; }
    }
  }
  minp[q] = m;
}

// Seeding launch (api_search.inl): the pair a query's first bound comes from = its NEAREST probed partition that holds
// rows here (probe lists are in ascending centroid distance).  No such partition -> the first probe (empty: no item).
// first_only (partition shards): only the query's globally nearest partition counts -- the shard that holds it seeds the
// query and the bound reaches the other shards through the peer tables, instead of every shard seeding every query.
__global__ void seed_probes_kernel(const uint32_t* __restrict__ probes, uint32_t Q, uint32_t P,
                                   const uint64_t* __restrict__ part_off, uint32_t* __restrict__ seed, int first_only) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  uint32_t best = probes[(size_t)q * P];
  for (uint32_t r = 0; r < (first_only ? 1u : P); r++) {
    const uint32_t p = probes[(size_t)q * P + r];
    if (part_off[p + 1] > part_off[p]) { best = p; break; }
  }
  seed[q] = best;
}

// Step 1: histogram of probed partitions (cold pairs and the others counted apart);
// rank[e] = arrival order of pair e in its class, bit 31 = cold.
__global__ void count_pairs_kernel(const uint32_t* __restrict__ probes, uint32_t npairs, uint32_t P,
                                   const uint32_t* __restrict__ minp, uint32_t* __restrict__ cnt_cold,
                                   uint32_t* __restrict__ cnt_other, uint32_t* __restrict__ rank) {
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= npairs) return;
  const uint32_t p = probes[e];
  if (p == minp[e / P])
    rank[e] = atomicAdd(&cnt_cold[p], 1u) | 0x80000000u;
  else
    rank[e] = atomicAdd(&cnt_other[p], 1u);
}

// Exclusive prefix of one value per thread over a 1024-thread block, for two sequences at once: warp scans by
// shuffle, the 32 warp totals scanned by warp 0.  Returns the exclusive prefixes in (ex_a, ex_b) and the block totals
// in (tot_a, tot_b).  Two barriers (the first block scan here was a 10-step Hillis-Steele over shared memory per 1024
// partitions: 80 barriers and 48 us for nlist = 4096 -- 3 % of an 8-GPU batch, twice per batch with the seeding launch).
__device__ __forceinline__ void block_scan2_1024(uint32_t a, uint32_t b, uint32_t* wa /*shared [33]*/, uint32_t* wb /*shared [33]*/,
                                                 uint32_t& ex_a, uint32_t& ex_b, uint32_t& tot_a, uint32_t& tot_b) {
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t ia = a, ib = b;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t va = __shfl_up_sync(0xffffffffu, ia, o), vb = __shfl_up_sync(0xffffffffu, ib, o);
    if (lane >= (uint32_t)o) { ia += va; ib += vb; }
  }
  if (lane == 31) { wa[w] = ia; wb[w] = ib; }
  __syncthreads();
  if (w == 0) {
    const uint32_t xa = wa[lane], xb = wb[lane];
    uint32_t ja = xa, jb = xb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t va = __shfl_up_sync(0xffffffffu, ja, o), vb = __shfl_up_sync(0xffffffffu, jb, o);
      if (lane >= (uint32_t)o) { ja += va; jb += vb; }
    }
    wa[lane] = ja - xa;  // exclusive prefix of the warp totals
    wb[lane] = jb - xb;
    if (lane == 31) { wa[32] = ja; wb[32] = jb; }
  }
  __syncthreads();
  ex_a = wa[w] + ia - a;
  ex_b = wb[w] + ib - b;
  tot_a = wa[32];
  tot_b = wb[32];
}

// Step 2 (single block of 1024 threads): exclusive scans over partitions -> pair offsets and item offsets.
// items(p) = ceil(cnt[p]/QL) * ceil(n_p/chunk_vecs).  Also accumulates the algorithmic scan
// bytes sum_pairs n_p * M and the pair count into stats[0..1].
// Each thread owns a run of ceil(nlist / 1024) consecutive partitions: local sums, one block scan, second pass.
__global__ void __launch_bounds__(1024)
offsets_kernel(const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ cnt2,
               const uint64_t* __restrict__ part_off, uint32_t nlist,
               uint32_t QL, uint32_t chunk_vecs, uint32_t M, uint32_t* __restrict__ pair_off,
               uint32_t* __restrict__ item_off, unsigned long long* __restrict__ stats) {
  __shared__ uint32_t wa[33], wb[33];
  __shared__ unsigned long long w_bytes[32];
  const uint32_t t = threadIdx.x, lane = t & 31, w = t >> 5;
  const uint32_t per = (nlist + 1023) / 1024;
  const uint32_t p0 = min(nlist, t * per), p1 = min(nlist, p0 + per);
  uint32_t sp = 0, si = 0;
  unsigned long long bytes = 0;
  for (uint32_t p = p0; p < p1; p++) {
    const uint32_t c = cnt[p] + cnt2[p];
    const uint64_t n = part_off[p + 1] - part_off[p];
    const uint32_t nch = (uint32_t)((n + chunk_vecs - 1) / chunk_vecs);
    sp += c;
    si += ((c + QL - 1) / QL) * nch;
    bytes += (unsigned long long)c * n * M;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bytes += __shfl_down_sync(0xffffffffu, bytes, o);
  if (lane == 0) w_bytes[w] = bytes;
  uint32_t ep, ei, tp, ti;
  block_scan2_1024(sp, si, wa, wb, ep, ei, tp, ti);  // (its barriers also publish w_bytes)
  for (uint32_t p = p0; p < p1; p++) {
    const uint32_t c = cnt[p] + cnt2[p];
    const uint64_t n = part_off[p + 1] - part_off[p];
    const uint32_t nch = (uint32_t)((n + chunk_vecs - 1) / chunk_vecs);
    pair_off[p] = ep;
    item_off[p] = ei;
    ep += c;
    ei += ((c + QL - 1) / QL) * nch;
  }
  if (t == 0) {
    unsigned long long tb = 0;
    for (int i = 0; i < 32; i++) tb += w_bytes[i];
    pair_off[nlist] = tp;
    item_off[nlist] = ti;
    stats[0] = tb;
    stats[1] = tp;
  }
}

// Step 2b: one descriptor per work item -- (partition, query group, chunk) -- so that the scan kernel takes an
// item with ONE load (prefetched while the previous item runs) instead of a binary search over item_off.
// One warp per partition.
// `queue_off` (or null: item_off) says where a partition's items sit in the QUEUE the scan CTAs draw from.
__global__ void expand_items_kernel(const uint32_t* __restrict__ item_off, const uint64_t* __restrict__ part_off,
                                    uint32_t nlist, uint32_t chunk_vecs, uint4* __restrict__ desc,
                                    const uint32_t* __restrict__ queue_off) {
  const uint32_t p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= nlist) return;
  const uint32_t i0 = item_off[p], i1 = item_off[p + 1];
  if (i1 == i0) return;
  const uint64_t n = part_off[p + 1] - part_off[p];
  const uint32_t nch = (uint32_t)((n + chunk_vecs - 1) / chunk_vecs);
  const uint32_t q0 = queue_off ? queue_off[p] : i0;
  for (uint32_t local = lane; local < i1 - i0; local += 32) desc[q0 + local] = make_uint4(p, local / nch, local % nch, 0u);
}

// Longest items first: the queue holds the partitions in descending size (`order`, fixed at index creation), so the
// last items a CTA draws are the short ones and the CTAs finish together.  With 1/8 of the items per GPU (partition
// shards) the ragged end of a partition-id-ordered queue was ~10 % of the scan.  queue_off[p] = items of the
// partitions before p in that order.  One block.
__global__ void __launch_bounds__(1024)
queue_offsets_kernel(const uint32_t* __restrict__ item_off, const uint32_t* __restrict__ order, uint32_t nlist,
                     uint32_t* __restrict__ queue_off) {
  __shared__ uint32_t wa[33], wb[33];
  const uint32_t t = threadIdx.x;
  const uint32_t per = (nlist + 1023) / 1024;
  const uint32_t i0 = min(nlist, t * per), i1 = min(nlist, i0 + per);
  uint32_t sc = 0;
  for (uint32_t i = i0; i < i1; i++) {
    const uint32_t p = order[i];
    sc += item_off[p + 1] - item_off[p];
  }
  uint32_t ex, ex2, tot, tot2;
  block_scan2_1024(sc, 0u, wa, wb, ex, ex2, tot, tot2);
  for (uint32_t i = i0; i < i1; i++) {
    const uint32_t p = order[i];
    queue_off[p] = ex;
    ex += item_off[p + 1] - item_off[p];
  }
}

// Step 3: scatter pair ids into partition-grouped order, cold pairs first.
__global__ void scatter_pairs_kernel(const uint32_t* __restrict__ probes, const uint32_t* __restrict__ rank,
                                     const uint32_t* __restrict__ pair_off, const uint32_t* __restrict__ cnt_cold,
                                     uint32_t npairs, uint32_t* __restrict__ pair_list) {
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= npairs) return;
  const uint32_t p = probes[e], r = rank[e];
  const uint32_t pos = (r & 0x80000000u) ? (r & 0x7fffffffu) : cnt_cold[p] + r;
  pair_list[pair_off[p] + pos] = e;
}

// ---------------------------------------------------------------------------------------
// K5 merge: for each output group g, the `L` candidate lists [g*L, (g+1)*L) (each <= stride
// valid entries, count in cnt[]) are merged to the `kout` smallest by (distance,row id).
// == concat_batches + sort_to_indices(limit) of ivf/v2.rs:383-400 (and the DataFusion
// SortExec fetch=k' of scanner.rs:2147-2157).  One block per group; block-wide bitonic sort of
// (key<<32 | src index) with a row-id tie-break inside the comparator.
__device__ __forceinline__ bool merge_less(unsigned long long a, unsigned long long b,
                                           const uint64_t* __restrict__ rowid) {
  uint32_t ka = (uint32_t)(a >> 32), kb = (uint32_t)(b >> 32);
  if (ka != kb) return ka < kb;
  if (ka == 0xffffffffu) return (uint32_t)a < (uint32_t)b;  // padding
  uint64_t ra = rowid[(uint32_t)a], rb = rowid[(uint32_t)b];
  if (ra != rb) return ra < rb;
  return (uint32_t)a < (uint32_t)b;
}

// Addressing: entry r of list l of group g lives at g*group_stride + l*list_stride + r;
// its count at g*cnt_gs + l*cnt_ls (covers both [group][list][r] and [list][group][r] inputs).
__global__ void __launch_bounds__(1024)
merge_lists_kernel(const float* __restrict__ in_dist, const uint64_t* __restrict__ in_rowid,
                   const uint64_t* __restrict__ in_pos /*or null*/, const float* __restrict__ in_aux /*or null*/,
                   const uint32_t* __restrict__ in_cnt, uint32_t L, uint32_t stride,
                   size_t group_stride, size_t list_stride, size_t cnt_gs, size_t cnt_ls,
                   uint32_t npow2, uint32_t kout, uint32_t out_stride, float* __restrict__ out_dist,
                   uint64_t* __restrict__ out_rowid, uint64_t* __restrict__ out_pos /*or null*/,
                   float* __restrict__ out_aux /*or null*/, uint32_t* __restrict__ out_cnt,
                   uint32_t sub_groups, uint32_t lists_total) {
  extern __shared__ unsigned long long sk[];
  __shared__ uint32_t s_n;
  const uint32_t g = blockIdx.x;
  // sub_groups > 1 (first level of a two-level merge): block g = (group g / sub_groups, slice g % sub_groups)
  // merges lists [slice*L, min(lists_total, slice*L + L)) of its group
  size_t base = (size_t)g * group_stride, cbase = (size_t)g * cnt_gs;
  if (sub_groups > 1) {
    const uint32_t gg = g / sub_groups, sl = g - gg * sub_groups;
    base = (size_t)gg * group_stride + (size_t)sl * L * list_stride;
    cbase = (size_t)gg * cnt_gs + (size_t)sl * L * cnt_ls;
    L = min(L, lists_total - sl * L);
  }
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  // the low word of a sort key is the entry's offset from `base` (must fit 32 bits).
  // Only the valid entries are gathered (most per-partition lists are short once the running
  // threshold has tightened), so the bitonic network runs on next_pow2(valid), not on L*stride.
  for (uint32_t i = threadIdx.x; i < L * stride; i += blockDim.x) {
    uint32_t l = i / stride, r = i - l * stride;
    if (r < in_cnt[cbase + (size_t)l * cnt_ls]) {
      size_t off = (size_t)l * list_stride + r;
      sk[atomicAdd(&s_n, 1u)] = ((unsigned long long)ordered_key(in_dist[base + off]) << 32) | (uint32_t)off;
    }
  }
  __syncthreads();
  uint32_t n2 = 2;
  while (n2 < s_n) n2 <<= 1;  // <= npow2
  for (uint32_t i = s_n + threadIdx.x; i < n2; i += blockDim.x) sk[i] = ~0ull;
  __syncthreads();
  const uint64_t* rid = in_rowid + base;
  for (uint32_t k = 2; k <= n2; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t i = threadIdx.x; i < n2; i += blockDim.x) {
        uint32_t ixj = i ^ j;
        if (ixj > i) {
          unsigned long long a = sk[i], b = sk[ixj];
          bool up = (i & k) == 0;
          bool a_gt_b = merge_less(b, a, rid);
          if (a_gt_b == up) {
            sk[i] = b;
            sk[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  const uint32_t n = min(s_n, kout);
  for (uint32_t i = threadIdx.x; i < kout; i += blockDim.x) {
    size_t o = (size_t)g * out_stride + i;
    if (i < n) {
      uint32_t src = (uint32_t)sk[i];
      out_dist[o] = in_dist[base + src];
      out_rowid[o] = in_rowid[base + src];
      if (out_pos) out_pos[o] = in_pos[base + src];
      if (out_aux) out_aux[o] = in_aux[base + src];
    } else {
      out_dist[o] = __int_as_float(0x7f800000);
      out_rowid[o] = ~0ull;
      if (out_pos) out_pos[o] = ~0ull;
      if (out_aux) out_aux[o] = __int_as_float(0x7fc00000);
    }
  }
  if (threadIdx.x == 0) out_cnt[g] = n;
}

// ---------------------------------------------------------------------------------------
// K6 refine, part 1: exact distance of every merged ANN candidate on the raw vectors with the
// ORIGINAL query and the INDEX metric (scanner.rs:1706-1721 -> flat_knn -> compute_distance,
// flat.rs:28-86).  One thread per candidate keeps the reference's lane structure exactly.
__global__ void refine_distance_kernel(int metric, const float* __restrict__ queries /*original*/,
                                       const float* __restrict__ raw, uint32_t dim,
                                       const uint64_t* __restrict__ pos, const uint32_t* __restrict__ cnt,
                                       uint32_t Q, uint32_t kprime, float* __restrict__ exact) {
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= Q * kprime) return;
  uint32_t q = e / kprime, r = e - q * kprime;
  float out = __int_as_float(0x7fc00000);
  if (r < cnt[q]) {
    const float* x = queries + (size_t)q * dim;
    float xn = metric == 1 ? norm_l2(x, (int)dim) : 0.0f;
    out = exact_distance(metric, x, xn, raw + (size_t)pos[e] * dim, (int)dim);
  }
  exact[e] = out;
}

// K6 part 2 (and the re-rank half of the multi-GPU merge): among the first cnt[q] candidates
// keep those with lower <= exact < upper, order by (exact,row id), emit the first k.
// == FilterExec(bounds) + SortExec(fetch=k) + NOT NULL of scanner.rs:2066-2136.
// One warp per query; kprime <= 1024.
__global__ void rerank_kernel(const float* __restrict__ exact, const uint64_t* __restrict__ rowid,
                              const uint32_t* __restrict__ cnt, uint32_t Q, uint32_t kprime,
                              uint32_t k, int has_lower, float lower, int has_upper, float upper,
                              float* __restrict__ out_dist, uint64_t* __restrict__ out_rowid,
                              uint32_t* __restrict__ out_cnt) {
  uint32_t q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (q >= Q) return;
  const float* ex = exact + (size_t)q * kprime;
  const uint64_t* rid = rowid + (size_t)q * kprime;
  uint32_t n = cnt[q];
  uint32_t nvalid = 0;
  for (uint32_t i = lane; i < k; i += 32) {
    out_dist[(size_t)q * k + i] = __int_as_float(0x7f800000);
    out_rowid[(size_t)q * k + i] = ~0ull;
  }
  __syncwarp();
  for (uint32_t i = lane; i < ((n + 31) & ~31u); i += 32) {
    bool ok = false;
    float di = 0.0f;
    uint64_t ri = 0;
    if (i < n) {
      di = ex[i];
      ri = rid[i];
      ok = di == di;  // IS NOT NULL
      if (has_lower) ok = ok && (di >= lower);
      if (has_upper) ok = ok && (di < upper);
    }
    if (ok) {
      uint32_t ki = ordered_key(di), r = 0;
      for (uint32_t j = 0; j < n; j++) {
        float dj = ex[j];
        bool okj = dj == dj;
        if (has_lower) okj = okj && (dj >= lower);
        if (has_upper) okj = okj && (dj < upper);
        if (!okj || j == i) continue;
        uint32_t kj = ordered_key(dj);
        if (kj < ki || (kj == ki && (rid[j] < ri || (rid[j] == ri && j < i)))) r++;
      }
      if (r < k) {
        out_dist[(size_t)q * k + r] = di;
        out_rowid[(size_t)q * k + r] = ri;
      }
    }
    nvalid += __popc(__ballot_sync(FULL, ok));
  }
  if (lane == 0) out_cnt[q] = min(nvalid, k);
}

}  // namespace lb200

// scan.cuh -- K2 (ADC lookup table in shared memory) + K3 (PQ code scan) + K4 (per-pair top-k').
//
// One persistent CTA per SM pulls work items (partition, group of <=QL queries probing it,
// chunk of the partition's rows) from a device-side queue.  Per item it
//   A3  forms the residual  r = q - centroid[p]            (ivf/v2.rs:285-301; L2|Cosine only)
//   A4  builds the ADC tables T[m][j] in shared memory     (pq/distance.rs:17-85)
//   A5  streams the partition's u8 codes and accumulates   (pq/distance.rs:102-135)
//         dist[i] = ((0 + T[0][c0]) + T[1][c1]) + ... in ASCENDING m order (bit-exact)
//   A6  keeps the k' smallest (distance,row id) per (query,partition,chunk)
//                                                          (flat/index.rs:76-137)
// Paths relative to /root/reference/rust/lance-index/src/vector/ unless noted.
//
// Device code layout ("tile16"): a partition's codes are stored as consecutive 16-row tiles,
// each tile = [M][16] bytes, so one 128-bit load fetches sub-vector m of 16 consecutive rows.
//
// VARIANT 1 (generic, any M): one query per CTA, table [M][256]; lane = one tile16.
// VARIANT 2 (bank-skewed):    QL queries per CTA, VG = 32/QL row groups per warp.
//   Table layout [M/VG][256][VG][QL] f32: the 32 words of row (mg,j) are one 128-byte line.
//   Lane (v,q) handles query slot q and tile16 v of the warp's block; at step u it looks up
//   sub-vector m = u - v, so the VG groups of a warp always sit on VG different m (mod VG),
//   i.e. on disjoint bank groups: every LDS is conflict-free although the code bytes are
//   random, and each row still sums its m in ascending order.
#pragma once
#include "select.cuh"

namespace lb200 {

struct ScanParams {
  // index (device)
  const uint8_t* codes;           // tile16 layout
  const uint64_t* tile_off;       // [nlist+1] first tile16 of each partition
  const uint64_t* part_off;       // [nlist+1] first row of each partition
  const uint64_t* row_ids;        // [N]
  const float* centroids;         // [nlist][dim]
  const float* codebook;          // [M][256][dsub]
  uint32_t dim, M, dsub, nlist;
  int32_t residual;               // 1: L2/cosine (residual + l2 table), 0: dot
  // queries
  const float* queries;           // [Q][dim], already normalised for cosine
  uint32_t Q, P, kprime;
  // work list (device, built by worklist kernels)
  const uint32_t* pair_list;      // pair ids (q*P + r) grouped by partition
  const uint32_t* part_pair_off;  // [nlist+1]
  const uint32_t* item_off;       // [nlist+1] prefix sum of items per partition
  uint32_t* item_counter;         // work queue head
  uint32_t chunk_vecs;            // rows per chunk (multiple of 16)
  uint32_t max_chunks;            // output slots per pair
  uint32_t capw;                  // candidate buffer entries per (warp, query slot)
  // bounds / prefilter
  float lower, upper;
  int32_t has_lower, has_upper;
  const uint64_t* const* bitmask; // [nlist] device pointers or null
  // outputs, slot = pair * max_chunks + chunk
  float* out_dist;                // [slots][kprime]
  uint64_t* out_pos;              // [slots][kprime] global row position
  uint64_t* out_rowid;            // [slots][kprime]
  uint32_t* out_count;            // [slots]
};

__device__ __forceinline__ float lds_f32(uint32_t saddr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(saddr));
  return v;
}
__device__ __forceinline__ uint4 ldg_u4(const uint8_t* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

// One ADC table element: l2_distance_batch / dot_distance_batch of a query sub-vector against
// one codebook centroid (pq/distance.rs:39-48,75-84).
template <int DSUB>
__device__ __forceinline__ float lut_elem(bool residual, const float* x, const float* y,
                                          uint32_t dsub) {
  if (DSUB == 8) return residual ? l2_once8(x, y) : dot_distance(x, y, 8);
  if (DSUB == 16) return residual ? l2_once16(x, y) : dot_distance(x, y, 16);
  return residual ? l2_batch_elem(x, y, (int)dsub) : dot_distance(x, y, (int)dsub);
}

template <int QL>
__device__ __forceinline__ uint32_t same_query_mask(int qs) {
  uint32_t base = QL == 1 ? 0xffffffffu : QL == 2 ? 0x55555555u : QL == 4 ? 0x11111111u : 0x01010101u;
  return base << qs;
}

// Append the lanes' passing candidates (one per lane at most) to their (warp, query-slot)
// buffers; prunes a buffer to k' first when it would overflow.  Warp-uniform call.
template <int QL, int NW>
__device__ __forceinline__ void append_candidates(bool pass, float dist, uint32_t pos, int qs,
                                                  uint32_t qmask, uint32_t w, uint32_t lane,
                                                  uint32_t capw, uint32_t kprime, uint32_t* ckey,
                                                  uint32_t* cpos, uint32_t* ccnt, uint32_t* sthr,
                                                  const uint64_t* prow) {
  uint32_t mask = __ballot_sync(FULL, pass);
  if (mask == 0) return;
  uint32_t mq = mask & qmask;
  uint32_t np = __popc(mq);
  uint32_t cnt = ccnt[w * QL + qs];
  bool overflow = np > 0 && cnt + np > capw;
  uint32_t ovm = __ballot_sync(FULL, overflow);
  if (ovm) {
#pragma unroll 1
    for (int qq = 0; qq < QL; qq++) {
      if (ovm & same_query_mask<QL>(qq)) {
        uint32_t n = ccnt[w * QL + qq];
        uint32_t* kb = ckey + (size_t)(w * QL + qq) * capw;
        uint32_t* pb = cpos + (size_t)(w * QL + qq) * capw;
        uint32_t T = warp_select_k(kb, pb, n, kprime, prow);
        if (lane == 0) {
          ccnt[w * QL + qq] = kprime;
          atomicMin(&sthr[qq], T);
        }
      }
    }
    __syncwarp();
    pass = pass && ordered_key(dist) <= sthr[qs];
    mask = __ballot_sync(FULL, pass);
    mq = mask & qmask;
    np = __popc(mq);
    cnt = ccnt[w * QL + qs];
  }
  if (pass) {
    uint32_t r = __popc(mq & ((1u << lane) - 1u));
    size_t o = (size_t)(w * QL + qs) * capw + cnt + r;
    ckey[o] = ordered_key(dist);
    cpos[o] = pos;
    if (r == 0) ccnt[w * QL + qs] = cnt + np;
  }
  __syncwarp();
}

template <int VARIANT, int QL, int NW, int DSUB>
__global__ void __launch_bounds__(NW * 32, 1) scan_kernel(const ScanParams p,
                                                          const uint32_t* __restrict__ total_items_p) {
  constexpr int VG = 32 / QL;
  constexpr int NT = NW * 32;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* lut = reinterpret_cast<float*>(smem_raw);
  const uint32_t MG = VARIANT == 1 ? 0 : p.M / VG;
  const uint32_t lut_floats = VARIANT == 1 ? p.M * 256 : MG * 256 * 32;
  float* resid = lut + lut_floats;
  uint32_t* ckey = reinterpret_cast<uint32_t*>(resid + QL * p.dim);
  uint32_t* cpos = ckey + (size_t)NW * QL * p.capw;
  uint32_t* ccnt = cpos + (size_t)NW * QL * p.capw;
  uint32_t* sthr = ccnt + NW * QL;
  uint32_t* spair = sthr + QL;
  uint32_t* sitem = spair + QL;

  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int qs = VARIANT == 1 ? 0 : (int)(lane % QL);
  const int vg = VARIANT == 1 ? 0 : (int)(lane / QL);
  const uint32_t qmask = same_query_mask<QL>(qs);
  const uint32_t lut_saddr = (uint32_t)__cvta_generic_to_shared(lut);
  const uint32_t M = p.M, dsub = p.dsub, dim = p.dim, kprime = p.kprime, capw = p.capw;
  const uint32_t total_items = *total_items_p;

  // VARIANT 2: per-lane byte offset of word (mo,q) for each phase b = u mod VG, where
  // mo = (b - v) mod VG; lanes with b < v are still on the previous table row group.
  int32_t Lb[VG];
#pragma unroll
  for (int b = 0; b < VG; b++) {
    int mo = (b - vg + VG) % VG;
    Lb[b] = (mo * QL + qs) * 4 - (b < vg ? 32768 : 0);
  }

  for (;;) {
    __syncthreads();
    if (tid == 0) *sitem = atomicAdd(p.item_counter, 1u);
    __syncthreads();
    const uint32_t item = *sitem;
    if (item >= total_items) break;
    uint32_t lo = 0, hi = p.nlist;
    while (hi - lo > 1) {
      uint32_t mid = (lo + hi) >> 1;
      if (p.item_off[mid] <= item) lo = mid; else hi = mid;
    }
    const uint32_t part = lo;
    const uint64_t row_base = p.part_off[part];
    const uint32_t n_p = (uint32_t)(p.part_off[part + 1] - row_base);
    const uint32_t nch = (n_p + p.chunk_vecs - 1) / p.chunk_vecs;
    const uint32_t local = item - p.item_off[part];
    const uint32_t g = local / nch, c = local % nch;
    const uint32_t pbase = p.part_pair_off[part] + g * QL;
    const uint32_t nq = min((uint32_t)QL, p.part_pair_off[part + 1] - pbase);
    const uint64_t* prow = p.row_ids + row_base;
    const uint64_t* bm = p.bitmask ? p.bitmask[part] : nullptr;

    if (tid < QL) {
      spair[tid] = tid < nq ? p.pair_list[pbase + tid] : 0xffffffffu;
      sthr[tid] = 0xff800000u;  // ordered_key(+inf)
    }
    if (tid < NW * QL) ccnt[tid] = 0;
    for (uint32_t i = tid; i < QL * dim; i += NT) {
      uint32_t qq = i / dim, e = i - qq * dim;
      float val = 0.0f;
      if (qq < nq) {
        uint32_t qidx = p.pair_list[pbase + qq] / p.P;
        float x = p.queries[(size_t)qidx * dim + e];
        val = p.residual ? __fsub_rn(x, p.centroids[(size_t)part * dim + e]) : x;
      }
      resid[i] = val;
    }
    __syncthreads();

    // ---- A4: ADC tables ------------------------------------------------------------
    if (VARIANT == 1) {
      for (uint32_t idx = tid; idx < M * 256; idx += NT) {
        uint32_t m = idx >> 8;
        const float* y = p.codebook + (size_t)idx * dsub;
        float yr[16];
        const float* yp = y;
        if (DSUB == 8 || DSUB == 16) {
#pragma unroll
          for (int i = 0; i < DSUB / 4; i++) {
            float4 v4 = __ldg(reinterpret_cast<const float4*>(y) + i);
            yr[4 * i] = v4.x; yr[4 * i + 1] = v4.y; yr[4 * i + 2] = v4.z; yr[4 * i + 3] = v4.w;
          }
          yp = yr;
        }
        lut[idx] = lut_elem<DSUB>(p.residual != 0, resid + m * dsub, yp, dsub);
      }
    } else {
#pragma unroll 1
      for (uint32_t mg = 0; mg < MG; mg++) {
        const uint32_t m = mg * VG + vg;
        float xr[16];
        const float* xp = resid + qs * dim + m * dsub;
        if (DSUB == 8 || DSUB == 16) {
#pragma unroll
          for (int i = 0; i < DSUB; i++) xr[i] = xp[i];
          xp = xr;
        }
        for (uint32_t j = w; j < 256; j += NW) {
          const float* y = p.codebook + ((size_t)m * 256 + j) * dsub;
          float yr[16];
          const float* yp = y;
          if (DSUB == 8 || DSUB == 16) {
#pragma unroll
            for (int i = 0; i < DSUB / 4; i++) {
              float4 v4 = __ldg(reinterpret_cast<const float4*>(y) + i);
              yr[4 * i] = v4.x; yr[4 * i + 1] = v4.y; yr[4 * i + 2] = v4.z; yr[4 * i + 3] = v4.w;
            }
            yp = yr;
          }
          lut[((size_t)mg * 256 + j) * 32 + lane] = lut_elem<DSUB>(p.residual != 0, xp, yp, dsub);
        }
      }
    }
    __syncthreads();

    // ---- A5 + A6: scan the chunk ------------------------------------------------------
    const uint32_t v0 = c * p.chunk_vecs;
    const uint32_t vend = min(n_p, v0 + p.chunk_vecs);
    const uint32_t t0 = v0 >> 4, tend = (vend + 15) >> 4;
    const uint8_t* cbase = p.codes + (size_t)p.tile_off[part] * 16 * M;
    const float lower = p.lower, upper = p.upper;
    const bool has_lower = p.has_lower != 0, has_upper = p.has_upper != 0;

    const uint32_t ntile = tend - t0;
    const uint32_t nblk = VARIANT == 1 ? (ntile + NT - 1) / NT : (ntile + VG - 1) / VG;
#pragma unroll 1
    for (uint32_t blk = VARIANT == 1 ? 0 : w; blk < nblk; blk += VARIANT == 1 ? 1 : NW) {
      const uint32_t t = VARIANT == 1 ? t0 + blk * NT + tid : t0 + blk * VG + vg;
      const bool tile_valid = t < tend && (VARIANT == 1 || qs < (int)nq);
      const uint32_t tc = min(t, tend - 1);
      const uint8_t* tp = cbase + (size_t)tc * 16 * M;
      float acc[16];
#pragma unroll
      for (int k = 0; k < 16; k++) acc[k] = 0.0f;

      if (VARIANT == 1) {
#pragma unroll 2
        for (uint32_t m = 0; m < M; m++) {
          uint4 W = ldg_u4(tp + m * 16);
          const uint32_t cb = lut_saddr + m * 1024;
          const uint32_t ww[4] = {W.x, W.y, W.z, W.w};
#pragma unroll
          for (int k = 0; k < 16; k++) {
            uint32_t a = __dp4a(ww[k >> 2], 4u << (8 * (k & 3)), cb);
            acc[k] = __fadd_rn(acc[k], lds_f32(a));
          }
        }
      } else {
        // step u = a*VG + b handles m = u - v on this lane; active iff 0 <= u - v < M.
        const uint8_t* tpl = tp - vg * 16;
#pragma unroll 1
        for (uint32_t a = 0; a <= MG; a++) {
          const uint32_t abase = lut_saddr + a * 32768u;
#pragma unroll
          for (int b = 0; b < VG; b++) {
            const bool active = (a > 0 || vg <= b) && (a < MG || vg > b);
            if (a == MG && b == VG - 1) continue;  // u = M+VG-1: nobody is active
            if (active) {
              uint4 W = ldg_u4(tpl + (a * VG + b) * 16);
              const uint32_t cb = abase + (uint32_t)Lb[b];
              const uint32_t ww[4] = {W.x, W.y, W.z, W.w};
#pragma unroll
              for (int k = 0; k < 16; k++) {
                uint32_t ad = __dp4a(ww[k >> 2], 128u << (8 * (k & 3)), cb);
                acc[k] = __fadd_rn(acc[k], lds_f32(ad));
              }
            }
          }
        }
      }

      // ---- candidate filter: dist <= threshold (k'-th best so far), bounds, prefilter ----
      const float thr = key_to_float(sthr[qs]);
      uint32_t pm = 0;
#pragma unroll
      for (int k = 0; k < 16; k++) {
        bool ok = acc[k] <= thr;
        if (has_lower) ok = ok && (lower <= acc[k]);
        if (has_upper) ok = ok && (acc[k] < upper);
        pm |= ok ? (1u << k) : 0u;
      }
      const uint32_t i0 = t * 16;
      if (!tile_valid) pm = 0;
      else if (i0 + 16 > vend) pm &= (1u << (vend - i0)) - 1u;
      if (bm && pm) {
        uint32_t bits = (uint32_t)(bm[i0 >> 6] >> (i0 & 63)) & 0xffffu;
        pm &= bits;
      }
      if (__any_sync(FULL, pm != 0)) {
#pragma unroll
        for (int k = 0; k < 16; k++)
          append_candidates<QL, NW>((pm >> k) & 1u, acc[k], i0 + k, qs, qmask, w, lane, capw,
                                    kprime, ckey, cpos, ccnt, sthr, prow);
      }
    }
    __syncthreads();

    // ---- finalize: per-warp prune, cross-warp merge, sorted emit ----------------------
#pragma unroll 1
    for (uint32_t qq = 0; qq < nq; qq++) {
      uint32_t n = ccnt[w * QL + qq];
      if (n > kprime) {
        warp_select_k(ckey + (size_t)(w * QL + qq) * capw, cpos + (size_t)(w * QL + qq) * capw, n,
                      kprime, prow);
        if (lane == 0) ccnt[w * QL + qq] = kprime;
      }
    }
    __syncthreads();
#pragma unroll 1
    for (uint32_t qq = w; qq < nq; qq += NW) {
      uint32_t* dk = ckey + (size_t)(0 * QL + qq) * capw;
      uint32_t* dp = cpos + (size_t)(0 * QL + qq) * capw;
      uint32_t n = ccnt[0 * QL + qq];
      for (uint32_t w2 = 1; w2 < NW; w2++) {
        uint32_t n2 = ccnt[w2 * QL + qq];
        if (n2 == 0) continue;
        if (n + n2 > capw) {  // capw >= 2k': after pruning to k' there is room again
          warp_select_k(dk, dp, n, kprime, prow);
          n = kprime;
        }
        const uint32_t* sk = ckey + (size_t)(w2 * QL + qq) * capw;
        const uint32_t* sp = cpos + (size_t)(w2 * QL + qq) * capw;
        for (uint32_t i = lane; i < n2; i += 32) {
          dk[n + i] = sk[i];
          dp[n + i] = sp[i];
        }
        n += n2;
        __syncwarp();
      }
      if (n > kprime) {
        warp_select_k(dk, dp, n, kprime, prow);
        n = kprime;
      }
      const size_t slot = (size_t)spair[qq] * p.max_chunks + c;
      warp_sort_emit(dk, dp, n, prow, row_base, p.out_dist + slot * kprime,
                     p.out_pos + slot * kprime, p.out_rowid + slot * kprime);
      if (lane == 0) p.out_count[slot] = n;
    }
  }
}

}  // namespace lb200

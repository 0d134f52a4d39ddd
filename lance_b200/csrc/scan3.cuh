// scan3.cuh -- VARIANT 3 of K2+K3+K4: lane = row, QL queries per CTA, skewed code streams.
//
// Same arithmetic as scan.cuh (A3 residual, A4 ADC table, A5 ascending-m f32 accumulation, A6
// per-pair top-k'); what changes is how the work is laid on the SM so that the shared-memory
// gather -- the real limiter of the scan (SURVEY 7 "hard parts") -- runs conflict-free at one
// wide LDS per code byte:
//
//  * The ADC tables of the CTA's QL queries are interleaved: T[chunk][code][slot][q], slot = m%16,
//    chunk = m/16, so ONE LDS.128 (QL=4) / LDS.64 (QL=2) returns T_q[m][code] for all QL queries.
//    The word address is chunk*CB + code*(64*QL) + slot*(4*QL) bytes: the bank group of an access
//    depends only on the slot, never on the (random) code.
//  * A wide LDS is served in phases of 8 (128-bit) / 16 (64-bit) consecutive lanes.  Lane l runs
//    d = l % D bytes BEHIND in its own code stream, so at any step the lanes of a phase sit on D
//    consecutive sub-vectors m -> D distinct slots -> distinct bank groups: no conflicts, although
//    every lane still adds its row's M table entries in ASCENDING m order (bit-exact with
//    pq/distance.rs:102-135).
//  * The skew is baked into the device layout ("skew stream"): a partition is cut into
//    superblocks of 32*L rows; lane l of superblock s owns rows s*32L + i*32 + l (i < L), its
//    stream is the concatenation of those rows' M code bytes delayed by d zero bytes, stored as
//    16-byte units interleaved over the 32 lanes (unit g of lane l at (g*32 + l)*16): every
//    LDG.128 of a warp is one fully coalesced 512-byte line group.  L*M/16 + 1 units per
//    superblock (the last one drains the d delayed bytes).
//  * A lane finishes row i-1 and starts row i at step d of macro-iteration i ("switch").  The
//    finished sum is captured and the accumulator cleared either with selects (SAFE) or, when
//    every table entry is finite and small enough that no partial sum can overflow, with exact
//    FMAs by 0.0/1.0 on the FMA pipe (acc*1+v == acc+v, acc*0+v == v bit for bit).
//  * Candidates are filtered against a running k'-th-best threshold that is shared by the warps
//    of the CTA and -- for merged searches -- by all work items of the same query (global
//    atomicMin), so that after a query's first partition almost nothing reaches the buffers.
//  * Life of a work item in a CTA: take the descriptor (fetched during the previous item) -> tables (built here from
//    the lane-order codebook image, or copied from the pre-pass: bulk copies of a group image / 4-byte gathers of
//    per-query tables, in the wide CTAs both ISSUED during the previous item's select) -> row loop -> barrier ->
//    next item's table fetch starts -> one warp per query selects k' (in registers for <= 128 candidates) and emits.
#pragma once
#include "scan.cuh"

#ifndef LB200_OPT_TABLE
#define LB200_OPT_TABLE 1
#endif
#ifndef LB200_CODE_LOAD
#define LB200_CODE_LOAD 0
#endif

#ifndef LB200_FINALIZE_SMALL
#define LB200_FINALIZE_SMALL 1
#endif
#ifndef LB200_FINALIZE_SMALL_INLINE
#define LB200_FINALIZE_SMALL_INLINE 0   // the call-free (BOOT = false) text: as a call 2-4 % slower there
#endif
#ifndef LB200_CB_DEPTH
#define LB200_CB_DEPTH 1
#endif

namespace lb200 {

struct Scan3Params {
  // index (device)
  const uint8_t* codes;           // skew-stream layout
  const uint64_t* sb_off;         // [nlist+1] first superblock of each partition
  const uint64_t* part_off;       // [nlist+1] first row of each partition
  const uint64_t* row_ids;        // [N]
  const float* centroids;         // [nlist][dim]
  const float* codebook;          // DSUB == 0 instantiations: [M][256][dsub]; DSUB > 0: the lane-order image (cb_image_f4)
  uint32_t dim, M, dsub, nlist;
  uint32_t L;                     // rows per lane per superblock
  int32_t residual;               // 1: L2/cosine, 0: dot
  // queries
  const float* queries;           // [Q][dim], already normalised for cosine
  uint32_t Q, P, kprime;
  // work list
  const uint32_t* pair_list;
  const uint32_t* part_pair_off;
  const uint32_t* item_off;
  uint32_t* item_counter;
  const uint4* item_desc;         // [items] (partition, query group, chunk, 0), see expand_items_kernel
  uint32_t chunk_vecs;            // rows per chunk (multiple of 32*L)
  uint32_t max_chunks;
  uint32_t capw;                  // candidate buffer entries per (warp, query slot)
  const float* lut_src;           // ADC tables built by lut_prepass_kernel: [unit][M/16][256][16] f32, or null
  const uint32_t* lut_max;        // [unit] 0, or the largest |entry| bit pattern if some entry is >= 2^118
  int32_t lut_by_query;           // unit = query index (dot: no residual) instead of pair id
  int32_t lut_image;              // tables of a query group are stored as ONE interleaved image (= the shared-memory
                                  // table of its work items), group (part, g) at lut_image_base(); lut_max by pair-list position
  uint32_t* qthr;                 // [Q] ordered keys: per-query global threshold, or null
  uint32_t* const* peers;         // [n_peers] the same table on the peer GPUs (NVLink peer memory), or null
  uint32_t n_peers;
  int32_t no_boot;                // diagnostic: disable the pooled bootstrap
  int32_t boot2;                  // second pooled prune at the end of each warp's first superblock
  // bounds / prefilter
  float lower, upper;
  int32_t has_lower, has_upper;
  const uint64_t* const* bitmask;
  // outputs, slot = pair * max_chunks + chunk
  float* out_dist;
  uint64_t* out_pos;
  uint64_t* out_rowid;
  uint32_t* out_count;
};

constexpr uint32_t KEY_PINF = 0xff800000u;  // ordered_key(+inf)

// (packed FP32 helpers -- f32x2_t, pk2, upk2, add2, sub2, mul2, fma2 -- live in dist.cuh)
// l2_once<f32x8,8> / l2_once<f32x16,16> (l2.rs:207-216) with the reduce_sum trees of simd/f32.rs,
// on packed pairs X[i] = (x[2i], x[2i+1]).  Same operations, same order per value as dist.cuh.
template <int DSUB>
__device__ __forceinline__ float l2_once_packed(const f32x2_t* X, const f32x2_t* Y) {
  float sq[DSUB];
#pragma unroll
  for (int i = 0; i < DSUB / 2; i++) {
    const f32x2_t d = sub2(X[i], Y[i]);
    upk2(mul2(d, d), sq[2 * i], sq[2 * i + 1]);
  }
  if (DSUB == 16) {
    float y8[8];
#pragma unroll
    for (int i = 0; i < 8; i++) y8[i] = __fadd_rn(sq[i], sq[i + 8]);       // reduce_sum16: halves first
    const f32x2_t a01 = add2(pk2(y8[0], y8[1]), pk2(y8[4], y8[5]));        // reduce_sum8: (a0, a1)
    const f32x2_t a23 = add2(pk2(y8[2], y8[3]), pk2(y8[6], y8[7]));        //              (a2, a3)
    float lo, hi;
    upk2(add2(a01, a23), lo, hi);                                           // (a0+a2, a1+a3)
    return __fadd_rn(lo, hi);
  } else {
    float a[4];
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = __fadd_rn(sq[i], sq[i + 4]);
    float lo, hi;
    upk2(add2(pk2(a[0], a[1]), pk2(a[2], a[3])), lo, hi);
    return __fadd_rn(lo, hi);
  }
}

// ---------------------------------------------------------------------------------------------
// The same trees with the products formed by fma.rn.f32x2(d, d, -0.0): x * x + (-0) rounds once, exactly like
// mul.rn (the sum of a product and -0 is the product, +0 stays +0), and a packed FMA feeding a packed ADD cannot be
// contracted any further -- so the first tree level runs packed too instead of unpacking every product (the
// contraction problem described above).  `nz` = two packed -0.0f that ptxas cannot see through (a kernel parameter):
// with a visible constant it would turn the fma back into a mul and contract it into the add.
template <int DSUB>
__device__ __forceinline__ float l2_once_packed_nz(const f32x2_t* X, const f32x2_t* Y, f32x2_t nz) {
  f32x2_t P[DSUB / 2];
#pragma unroll
  for (int i = 0; i < DSUB / 2; i++) {
    const f32x2_t d = sub2(X[i], Y[i]);
    P[i] = fma2(d, d, nz);
  }
  float lo, hi;
  if (DSUB == 16) {
    // reduce_sum16: y8[i] = sq[i] + sq[i + 8]  ->  (y8[2i], y8[2i+1]) = P[i] + P[i + 4]
    const f32x2_t y01 = add2(P[0], P[4]), y23 = add2(P[1], P[5]), y45 = add2(P[2], P[6]), y67 = add2(P[3], P[7]);
    const f32x2_t a01 = add2(y01, y45);   // reduce_sum8: (a0, a1) = (y0 + y4, y1 + y5)
    const f32x2_t a23 = add2(y23, y67);   //              (a2, a3)
    upk2(add2(a01, a23), lo, hi);         // (a0 + a2, a1 + a3)
  } else {
    const f32x2_t a01 = add2(P[0], P[2]), a23 = add2(P[1], P[3]);   // a[i] = sq[i] + sq[i + 4]
    upk2(add2(a01, a23), lo, hi);
  }
  return __fadd_rn(lo, hi);
}

// Codebook image for the in-kernel table build.  A build thread is (slot = tid % 16 -> sub-vector, code = tid / 16 + step):
// in the [M][256][dsub] array the 32 lanes of one load instruction sit 32 bytes or 8 KB apart -- 32 cache lines per
// instruction, and the build's loads alone cost the LSU pipe as many wavefronts as the whole row loop of a 4k-row
// partition (r02_bt ncu, cfg 1: the build was 36 % of the kernel's warp time).  The image stores, per 16-sub-vector
// chunk and per PAIR of codes, the i-th float4 of the 32 (code parity, slot) rows side by side: one load instruction
// = 512 contiguous bytes.  Offset in float4 units; built once per handle by cb_image_kernel.
__host__ __device__ __forceinline__ size_t cb_image_f4(uint32_t ch, uint32_t code, uint32_t slot, uint32_t i, uint32_t dsub4) {
  return (((size_t)ch * 128 + (code >> 1)) * dsub4 + i) * 32 + (code & 1u) * 16 + slot;
}
static __global__ void cb_image_kernel(const float* __restrict__ codebook, uint32_t M, uint32_t dsub, float4* __restrict__ image) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x, dsub4 = dsub / 4;
  if (e >= M * 256 * dsub4) return;
  const uint32_t i = e % dsub4, row = e / dsub4, code = row & 255u, m = row >> 8;
  image[cb_image_f4(m >> 4, code, m & 15u, i, dsub4)] = reinterpret_cast<const float4*>(codebook)[e];
}

// ADC tables as a pre-pass (A3 + A4 out of the scan kernel).  The scan kernel builds its tables per
// work item, re-reading the whole codebook (M*256*dsub*4 bytes) for QL queries; when tables are
// expensive relative to the rows they serve (small partitions, many sub-vectors) or do not depend
// on the partition at all (dot: no residual -> one table per QUERY, not per pair), they are built
// here instead, with each thread's codebook rows held in registers across 64 units, and the scan
// kernel only copies them into shared memory (lut_fetch).  Same arithmetic as the in-kernel build:
// residual = q - centroid (v2.rs:285-301), entries by lut_elem (pq/distance.rs:17-85).
// grid.x = (M/16) * LUTPRE_SPLIT: 16 sub-vectors x a slice of the 256 codes; grid.y = unit tiles.
// Output layout per unit: [M/16][256][16] f32 == the scan kernel's table order for one query.
// first table slot of query group g of partition `part` in image mode: partition p's groups start at
// pair_off[p] + p * (ql - 1), which leaves room for its last, partly filled group
__host__ __device__ __forceinline__ uint64_t lut_image_base(uint32_t pair_off_p, uint32_t part, uint32_t g, uint32_t ql) {
  return (uint64_t)pair_off_p + (uint64_t)part * (ql - 1) + (uint64_t)g * ql;
}
constexpr uint32_t LUTPRE_UNITS = 64;
#ifndef LUTPRE_STAGE_BATCH
#define LUTPRE_STAGE_BATCH 16
#endif
constexpr int LUTPRE_CPT = 4;                              // codes per thread
constexpr uint32_t LUTPRE_SPLIT = 256 / (16 * LUTPRE_CPT);  // blocks per 16-sub-vector chunk
template <int DSUB>
__global__ void __launch_bounds__(256)
lut_prepass_kernel(const float* __restrict__ queries, const float* __restrict__ centroids,
                   const float* __restrict__ codebook, const uint32_t* __restrict__ probes /*[Q][P]; null: by query*/,
                   const uint64_t* __restrict__ part_off, uint32_t n_units, uint32_t P, uint32_t dim, uint32_t M,
                   float* __restrict__ out, uint32_t* __restrict__ out_max /*[n_units], zeroed*/,
                   const uint32_t* __restrict__ pair_list /*image mode: unit = position in the pair list; else null*/,
                   const uint32_t* __restrict__ pair_off, uint32_t ql, unsigned long long negzero2 /*0x8000000080000000*/) {
  constexpr int PITCH = DSUB + 4;  // floats; 8 lanes of an LDS.128 phase land on 8 distinct bank groups
  constexpr int UB = 16;           // units staged at a time
  __shared__ __align__(16) float xs[UB][16 * PITCH];
  __shared__ uint32_t spart[LUTPRE_UNITS];  // partition of each unit of the tile; 0xffffffff: no table needed
  __shared__ uint32_t squery[LUTPRE_UNITS]; // its query
  __shared__ uint64_t sobase[LUTPRE_UNITS]; // where its table starts in `out` (floats)
  const uint32_t ch = blockIdx.x / LUTPRE_SPLIT, quarter = blockIdx.x % LUTPRE_SPLIT;
  const uint32_t t = threadIdx.x, slot = t & 15, cq = t >> 4;
  const uint32_t m = ch * 16 + slot;
  const bool residual = probes != nullptr;
  const uint32_t u0 = blockIdx.y * LUTPRE_UNITS, u1 = min(n_units, u0 + LUTPRE_UNITS);
  // image mode: the tables of the (up to ql) queries of a group are interleaved -- entry e of slot s at
  // (e * ql + s) -- which is the layout the scan kernel's shared-memory table has, so it fetches a group's tables
  // with one bulk copy (lut_image_issue) instead of 4-byte gathers.
  const uint32_t qs = pair_list ? ql : 1u;
  if (t < LUTPRE_UNITS) {
    uint32_t part = 0xffffffffu, query = 0;
    uint64_t obase = (uint64_t)(u0 + t) * M * 256;
    if (u0 + t < u1) {
      part = 0;
      query = u0 + t;
      if (residual) {  // a shard holds all nlist partitions, most of them empty: no items, no table
        const uint32_t pair = pair_list ? pair_list[u0 + t] : u0 + t;
        part = probes[pair];
        query = pair / P;
        if (pair_list) {
          const uint32_t local = u0 + t - pair_off[part];
          obase = lut_image_base(pair_off[part], part, local / ql, ql) * M * 256 + local % ql;
        }
        if (part_off[part + 1] == part_off[part]) part = 0xffffffffu;
      }
    }
    spart[t] = part;
    squery[t] = query;
    sobase[t] = obase;
  }
  float y[LUTPRE_CPT][DSUB];
#pragma unroll
  for (int i = 0; i < LUTPRE_CPT; i++) {
    const uint32_t code = quarter * (16 * LUTPRE_CPT) + cq + 16 * i;
    const float4* src = reinterpret_cast<const float4*>(codebook + ((size_t)m * 256 + code) * DSUB);
#pragma unroll
    for (int j = 0; j < DSUB / 4; j++) {
      const float4 v = __ldg(src + j);
      y[i][4 * j] = v.x; y[i][4 * j + 1] = v.y; y[i][4 * j + 2] = v.z; y[i][4 * j + 3] = v.w;
    }
  }
#pragma unroll 1
  for (uint32_t ub = u0; ub < u1; ub += UB) {
    __syncthreads();
    // stage the residual (or query) chunk of UB units.  The loads are issued in batches of SB per thread BEFORE any of
    // them is used: written as load / subtract / store per element the compiler kept them in program order and every
    // element paid its own L2 / DRAM latency (r02_bv ncu: 30 % of the kernel's warp time sat on these subtractions).
    constexpr int NIT = UB * 16 * DSUB / 256, SB = LUTPRE_STAGE_BATCH < NIT ? LUTPRE_STAGE_BATCH : NIT;
#pragma unroll
    for (int it0 = 0; it0 < NIT; it0 += SB) {
      float qv[SB], cv[SB];
#pragma unroll
      for (int it = 0; it < SB; it++) {
        const uint32_t idx = t + (uint32_t)(it0 + it) * 256u;
        const uint32_t ui = idx / (16 * DSUB), rem = idx % (16 * DSUB), u = ub + ui;
        const uint32_t part = u < u1 ? spart[u - u0] : 0xffffffffu;
        qv[it] = 0.0f; cv[it] = 0.0f;
        if (part != 0xffffffffu) {
          const uint32_t e = ch * 16 * DSUB + rem;
          qv[it] = __ldg(queries + (size_t)squery[u - u0] * dim + e);
          if (residual) cv[it] = __ldg(centroids + (size_t)part * dim + e);
        }
      }
#pragma unroll
      for (int it = 0; it < SB; it++) {
        const uint32_t idx = t + (uint32_t)(it0 + it) * 256u;
        const uint32_t ui = idx / (16 * DSUB), rem = idx % (16 * DSUB);
        xs[ui][(rem / DSUB) * PITCH + rem % DSUB] = residual ? __fsub_rn(qv[it], cv[it]) : qv[it];
      }
    }
    __syncthreads();
#pragma unroll 2
    for (uint32_t ui = 0; ui < UB; ui++) {
      if (ub + ui >= u1 || spart[ub + ui - u0] == 0xffffffffu) continue;
      float x[DSUB];
#pragma unroll
      for (int j = 0; j < DSUB / 4; j++) {
        const float4 v = *reinterpret_cast<const float4*>(&xs[ui][slot * PITCH + 4 * j]);
        x[4 * j] = v.x; x[4 * j + 1] = v.y; x[4 * j + 2] = v.z; x[4 * j + 3] = v.w;
      }
      float* o = out + sobase[ub + ui - u0] + ((size_t)ch * 4096 + slot) * qs;
      uint32_t mb = 0;
      f32x2_t X[DSUB / 2];
#pragma unroll
      for (int j = 0; j < DSUB / 2; j++) X[j] = pk2(x[2 * j], x[2 * j + 1]);
#pragma unroll
      for (int i = 0; i < LUTPRE_CPT; i++) {
        float v;
        if (residual) {
          f32x2_t Y[DSUB / 2];
#pragma unroll
          for (int j = 0; j < DSUB / 2; j++) Y[j] = pk2(y[i][2 * j], y[i][2 * j + 1]);
          v = l2_once_packed_nz<DSUB>(X, Y, negzero2);
        } else {
          v = lut_elem<DSUB>(false, x, y[i], DSUB);
        }
        o[(size_t)(quarter * (16 * LUTPRE_CPT) + cq + 16 * i) * 16 * qs] = v;
        mb = max(mb, __float_as_uint(v) & 0x7fffffffu);
      }
      // only magnitudes that rule out the exact-FMA row switch (>= 2^118) are recorded: (almost) never
      if (mb >= 0x7a800000u) atomicMax(out_max + ub + ui, mb);
    }
  }
}

// Copy the pre-built tables of the item's queries into the interleaved shared-memory table
// (entry e of query q at byte (e*QL + q)*4) with 4-byte cp.async: lane -> (q = lane % QL,
// e = base + lane / QL), so one warp instruction reads QL runs of 32/QL consecutive floats (whole
// 32-byte sectors) and writes 128 consecutive shared bytes (conflict-free); nothing is staged in
// registers, so every thread keeps its whole share of the table in flight at once -- the copy is
// bandwidth- not latency-bound.  Returns the largest |entry| bit pattern of the item's tables (from
// the pre-pass) for the exact-FMA row switch test.  Out of line: runs once per item.
// NWARP (the number of issuing warps) is a template parameter so that the copies of an unrolled block differ by
// compile-time offsets: one LDGSTS per copy plus two pointer bumps per eight (the first version, with a run-time
// stride and a bound check per copy, spent 14 instructions per 4-byte copy: r02_cb ncu, 18 % of cfg 5's instructions).
template <int QL, int NWARP>
static __device__ __noinline__ void lut_fetch_issue(unsigned char* lut, const float* __restrict__ src, size_t unit_stride,
                                                    const uint32_t* units /*shared [QL]*/, uint32_t nq, uint32_t nent,
                                                    uint32_t w /*rank among the issuing warps*/) {
  constexpr uint32_t EPW = 32 / QL, STEP = NWARP * EPW;  // entries per warp instruction / per round of the issuing warps
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t unit = units[min(lane % QL, nq - 1)];  // unused slots replicate a valid table (never read back)
  const float* s = src + (size_t)unit * unit_stride + lane / QL + w * EPW;
  uint32_t dst = (uint32_t)__cvta_generic_to_shared(lut) + lane * 4 + w * EPW * (4 * QL);
  uint32_t nit = nent > w * EPW ? (nent - w * EPW + STEP - 1) / STEP : 0u;
#pragma unroll 1
  for (; nit >= 8; nit -= 8) {
#define LB200_CP4(K) \
    asm volatile("cp.async.ca.shared.global [%0 + %2], [%1 + %3], 4;" ::"r"(dst), "l"(s), "n"((K) * STEP * 4 * QL), "n"((K) * STEP * 4) : "memory")
    LB200_CP4(0); LB200_CP4(1); LB200_CP4(2); LB200_CP4(3); LB200_CP4(4); LB200_CP4(5); LB200_CP4(6); LB200_CP4(7);
#undef LB200_CP4
    dst += 8 * STEP * 4 * QL;
    s += 8 * STEP;
  }
#pragma unroll 1
  for (; nit; nit--) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(s) : "memory");
    dst += STEP * 4 * QL;
    s += STEP;
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int QL>
static __device__ __noinline__ uint32_t lut_fetch_wait(const uint32_t* __restrict__ lut_max, const uint32_t* units, uint32_t nq) {
  const uint32_t mb = lut_max[units[min((threadIdx.x & 31) % QL, nq - 1)]];
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  return mb;
}
template <int QL, int NWARP>
static __device__ __forceinline__ uint32_t lut_fetch(unsigned char* lut, const float* __restrict__ src,
                                                     const uint32_t* __restrict__ lut_max, size_t unit_stride,
                                                     const uint32_t* units /*shared [QL]*/, uint32_t nq, uint32_t nent) {
  lut_fetch_issue<QL, NWARP>(lut, src, unit_stride, units, nq, nent, threadIdx.x >> 5);
  return lut_fetch_wait<QL>(lut_max, units, nq);
}

// Image mode: the group's interleaved tables are one contiguous block -> bulk copies (TMA 1-D) issued by one thread,
// completion on an mbarrier every thread waits on.  No LSU instructions, no registers.  Issue and wait are separate
// calls: the copy for the NEXT item is issued as soon as the current item's row loop is over (the end-of-item select
// does not read the table), so the fetch overlaps the select instead of following it (r02_ae ncu, cfg 3: 13 % of
// the warp time sat in the wait, 25 % behind the select).
static __device__ __noinline__ void lut_image_issue(unsigned char* lut, const float* __restrict__ src, uint32_t bytes,
                                                    uint64_t* bar) {
  const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(bytes) : "memory");
  const uint32_t dst = (uint32_t)__cvta_generic_to_shared(lut);
  for (uint32_t o = 0; o < bytes; o += 32768u) {
    const uint32_t n = min(32768u, bytes - o);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst + o), "l"(reinterpret_cast<const unsigned char*>(src) + o), "r"(n), "r"(bar_a) : "memory");
  }
}
static __device__ __noinline__ uint32_t lut_image_wait(const uint32_t* __restrict__ lut_max, uint32_t pbase, uint32_t nq,
                                                       uint64_t* bar, const uint32_t* phase_s) {
  const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar);
  const uint32_t phase = *phase_s;   // flipped by thread 0 after the barrier that follows this call
  uint32_t mb = 0;
  if (threadIdx.x < nq) mb = lut_max[pbase + threadIdx.x];
  uint32_t ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar_a), "r"(phase) : "memory");
  } while (!ok);
  return mb;
}

// One 16-byte unit of a lane's code stream.  The stream is read exactly once per work item by this CTA
// (other CTAs scanning the same partition find it in L2), so it need not be kept in L1.
__device__ __forceinline__ uint4 load_codes(const uint4* p) {
#if LB200_CODE_LOAD == 1
  return __ldcg(p);
#elif LB200_CODE_LOAD == 2
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
#elif LB200_CODE_LOAD == 3
  return __ldcs(p);
#else
  return __ldg(p);
#endif
}

template <int QL>
struct LutVec;
template <>
struct LutVec<4> { using T = float4; };
template <>
struct LutVec<2> { using T = float2; };

template <int QL>
__device__ __forceinline__ void lut_load(const unsigned char* lut, uint32_t off, float (&v)[QL]) {
  if (QL == 4) {
    float4 t = *reinterpret_cast<const float4*>(lut + off);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  } else {
    float2 t = *reinterpret_cast<const float2*>(lut + off);
    v[0] = t.x; v[1] = t.y;
  }
}

// byte k of `word` times the table row pitch (64*QL bytes), merged with / added to `b`.
template <int QL>
__device__ __forceinline__ uint32_t code_addr(uint32_t word, int k, uint32_t b) {
  if (QL == 4) {
    // pitch 256: result = {b.byte0, word.byte_k, b.byte2, b.byte3}; b.byte1 is 0 by construction
    return __byte_perm(word, b, 0x7604u | ((uint32_t)k << 4));
  } else {
    return __dp4a(word, 128u << (8 * k), b);
  }
}

// One 16-byte unit of the lane's stream = 16 lookups for each of the QL queries.
// SWITCH: this is unit 0 of a row (steps 0..15 of the macro-iteration): lanes with d == j finish
// the previous row at step j.  NSTEP < 16 only for the drain unit.
template <int QL, int D, bool SWITCH, bool SAFE, int NSTEP>
__device__ __forceinline__ void scan_unit(const uint4 W, const unsigned char* lut_u /*lut + (u-1)*CB*/,
                                          const uint32_t (&Bp)[16], const uint32_t (&B0)[D],
                                          const float (&keepf)[D], const float (&swf)[D], uint32_t d,
                                          float (&acc)[QL], float (&done)[QL]) {
  const uint32_t ww[4] = {W.x, W.y, W.z, W.w};
#pragma unroll
  for (int j = 0; j < NSTEP; j++) {
    const uint32_t b = (SWITCH && j < D) ? B0[j < D ? j : 0] : Bp[j];
    // B0 is relative to the table start, Bp to (chunk u-1): the caller passes lut_u accordingly.
    float v[QL];
    if (SWITCH && j < D) {
      lut_load<QL>(lut_u + (size_t)(256 * 64 * QL), code_addr<QL>(ww[j >> 2], j & 3, b), v);
    } else {
      lut_load<QL>(lut_u, code_addr<QL>(ww[j >> 2], j & 3, b), v);
    }
    if (SWITCH && j < D) {
      if (SAFE) {
        const bool sw = d == (uint32_t)j;
#pragma unroll
        for (int q = 0; q < QL; q++) {
          done[q] = sw ? acc[q] : done[q];
          acc[q] = __fadd_rn(sw ? 0.0f : acc[q], v[q]);
        }
      } else {
#pragma unroll
        for (int q = 0; q < QL; q++) {
          done[q] = j == 0 ? __fmul_rn(acc[q], swf[0]) : __fmaf_rn(acc[q], swf[j < D ? j : 0], done[q]);
          acc[q] = __fmaf_rn(acc[q], keepf[j < D ? j : 0], v[q]);
        }
      }
    } else {
#pragma unroll
      for (int q = 0; q < QL; q++) acc[q] = __fadd_rn(acc[q], v[q]);
    }
  }
}

// Cheap bound for pruning: the k-th smallest of the 32 lanes' values `m` (0xffffffff = none) is the
// k-th smallest of a SUBSET of the candidates, hence >= the true k-th best: a valid threshold
// (tightness ~ k-th of all when each lane passes the minimum of its share).  Needs k <= 32.
// Returns 0xffffffff when fewer than k lanes hold a value.  ~100 instructions, no memory traffic.
static __device__ __noinline__ uint32_t warp_bound_k(uint32_t m, uint32_t k, uint32_t lane) {
  uint32_t r = 0;
#pragma unroll
  for (int j = 0; j < 32; j++) {
    const uint32_t o = __shfl_sync(FULL, m, j);
    r += (o < m || (o == m && (uint32_t)j < lane)) ? 1u : 0u;
  }
  const uint32_t sel = __ballot_sync(FULL, r == k - 1);
  return __shfl_sync(FULL, m, __ffs(sel) - 1);  // exactly one lane has rank k-1
}

// Keep the entries with key <= T of a warp's buffer (stable, in place).  Returns the new count.
static __device__ __noinline__ uint32_t warp_compact_le(uint32_t* kb, uint32_t* pb, uint32_t cnt, uint32_t T,
                                                    uint32_t lane) {
  uint32_t outn = 0;
  for (uint32_t base = 0; base < cnt; base += 32) {
    const uint32_t i = base + lane;
    uint32_t kk = 0xffffffffu, pp = 0;
    if (i < cnt) { kk = kb[i]; pp = pb[i]; }
    const bool keep = i < cnt && kk <= T;
    const uint32_t mk = __ballot_sync(FULL, keep);
    __syncwarp();
    if (keep) {
      const uint32_t dst = outn + __popc(mk & ((1u << lane) - 1u));
      kb[dst] = kk;
      pb[dst] = pp;
    }
    outn += __popc(mk);
    __syncwarp();
  }
  return outn;
}

// Append the warp's passing rows for query slot q to its (query, warp) buffer; prune first when
// it would overflow.  Warp-uniform call.
__device__ __forceinline__ void append_rows(bool pass, float dist, uint32_t pos, uint32_t lane, uint32_t capw,
                                            uint32_t kprime, uint32_t* kb, uint32_t* pb, uint32_t* cnt_p,
                                            uint32_t* sthr_q, float* sthrf_q, uint32_t* gthr_q,
                                            const uint64_t* prow) {
  uint32_t mask = __ballot_sync(FULL, pass);
  if (mask == 0) return;
  uint32_t np = __popc(mask);
  uint32_t cnt = *cnt_p;
  if (cnt + np > capw) {  // capw >= k'+32, so cnt > k' here
    uint32_t T = warp_select_k(kb, pb, cnt, kprime, prow);
    cnt = kprime;
    if (lane == 0) {
      uint32_t old = atomicMin(sthr_q, T);
      *sthrf_q = key_to_float(min(old, T));
      if (gthr_q) atomicMin(gthr_q, T);
    }
    __syncwarp();
    pass = pass && ordered_key(dist) <= *sthr_q;
    mask = __ballot_sync(FULL, pass);
    np = __popc(mask);
  }
  if (pass) {
    uint32_t r = __popc(mask & ((1u << lane) - 1u));
    kb[cnt + r] = ordered_key(dist);
    pb[cnt + r] = pos;
  }
  if (lane == 0) *cnt_p = cnt + np;
  __syncwarp();
}

// The thresholds change under the scanning warps' feet (other warps tighten them): volatile.
template <int QL>
__device__ __forceinline__ void lds_volatile(uint32_t saddr, float (&v)[QL]) {
  if (QL == 4)
    asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(saddr));
  else
    asm volatile("ld.volatile.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v[0]), "=f"(v[1]) : "r"(saddr));
}

// Per-item state the COLD paths need, kept in shared memory so that those paths can live in
// out-of-line functions: the hot loop must stay small (r01: with the cold code inlined the kernel
// grew from 4.9k to 21k instructions and the scan lost 10% to instruction-cache misses and spills).
struct ItemCtx {
  uint32_t *ckey, *cpos, *ccnt, *sthr, *sqidx, *spair, *qthr;
  uint32_t* const* peers;
  uint32_t n_peers;
  float* sthrf;
  const uint64_t* prow;
  const uint64_t* bm;
  float *out_dist;
  uint64_t *out_pos, *out_rowid;
  uint32_t* out_count;
  uint64_t row_base;
  uint32_t capw, kprime, nq, vend, nw, max_chunks, chunk;
  float lower, upper;
  int32_t has_lower, has_upper;
};

// Shared-memory footprint of scan3_kernel (host + device agree through this one function).
__host__ __device__ inline size_t scan3_smem_bytes(int QL, int NW, uint32_t M, uint32_t dim, uint32_t capw) {
  size_t lut = (size_t)(M / 16) * 256 * 64 * QL;
  return lut + (size_t)QL * dim * 4 + (size_t)QL * NW * capw * 8 + (size_t)QL * NW * 4 + (size_t)QL * 16 + 32 + sizeof(ItemCtx) + 16;  // 32: next item (16) + its descriptor (16)
}

// Publish a query's bound (one thread): into this GPU's table and -- partition shards on several
// GPUs -- into every peer's table with a fire-and-forget atomic-min over NVLink peer memory.  The
// peers get the best value known here, so bounds found by the in-loop prunes travel too.
static __device__ __noinline__ void publish_bound(const ItemCtx* c, uint32_t qidx, uint32_t T) {
  const uint32_t old = atomicMin(c->qthr + qidx, T);
  const uint32_t best = min(old, T);
  for (uint32_t j = 0; j < c->n_peers; j++) atomicMin(c->peers[j] + qidx, best);
}

// Pull the NW per-warp candidate segments of one query (segment w at w*capw, cnt[w] entries) together
// at the start of the region.  dst index <= src index always, so the forward copy is safe.
// Returns the total.  One warp.
static __device__ __noinline__ uint32_t gather_segments(uint32_t* dk, uint32_t* dp, const uint32_t* cnt, uint32_t nw,
                                                    uint32_t capw, uint32_t lane) {
  uint32_t n = cnt[0];
#pragma unroll 1
  for (uint32_t w2 = 1; w2 < nw; w2++) {
    const uint32_t n2 = cnt[w2];
    const uint32_t* sk = dk + (size_t)w2 * capw;
    const uint32_t* sp = dp + (size_t)w2 * capw;
    for (uint32_t i0 = 0; i0 < n2; i0 += 32) {
      uint32_t i = i0 + lane, kk = 0, pp = 0;
      if (i < n2) { kk = sk[i]; pp = sp[i]; }
      __syncwarp();
      if (i < n2) { dk[n + i] = kk; dp[n + i] = pp; }
      __syncwarp();
    }
    n += n2;
  }
  return n;
}

// Pooled prune of a cold item (see the kernel).  Called by ALL threads of the CTA.
static __device__ __noinline__ void bootstrap_fn(const ItemCtx* c) {
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t NW = c->nw, capw = c->capw, kprime = c->kprime, nq = c->nq;
  __syncthreads();
  if (kprime <= 32) {
    // bound = k'-th smallest of the lanes' minima over the pooled NW segments of the query
    for (uint32_t qq = w; qq < nq; qq += NW) {
      uint32_t m = 0xffffffffu;
      for (uint32_t w2 = 0; w2 < NW; w2++) {
        const uint32_t n2 = c->ccnt[qq * NW + w2];
        const uint32_t* sk = c->ckey + (size_t)(qq * NW + w2) * capw;
        for (uint32_t i = lane; i < n2; i += 32) m = min(m, sk[i]);
      }
      const uint32_t T = warp_bound_k(m, kprime, lane);
      if (T != 0xffffffffu && lane == 0) {
        uint32_t old = atomicMin(c->sthr + qq, T);
        c->sthrf[qq] = key_to_float(min(old, T));
        if (c->qthr) publish_bound(c, c->sqidx[qq], T);
      }
    }
    __syncthreads();
    // every warp drops what the new bounds exclude from its own buffers
    for (uint32_t qq = 0; qq < nq; qq++) {
      const uint32_t b = qq * NW + w;
      const uint32_t n0 = c->ccnt[b];
      if (n0) {
        const uint32_t n1 = warp_compact_le(c->ckey + (size_t)b * capw, c->cpos + (size_t)b * capw, n0, c->sthr[qq], lane);
        if (lane == 0) c->ccnt[b] = n1;
      }
    }
    __syncwarp();
    return;
  }
  for (uint32_t qq = w; qq < nq; qq += NW) {
    uint32_t* dk = c->ckey + (size_t)qq * NW * capw;
    uint32_t* dp = c->cpos + (size_t)qq * NW * capw;
    uint32_t n = gather_segments(dk, dp, c->ccnt + qq * NW, NW, capw, lane);
    if (n > kprime) {
      uint32_t T = warp_select_k(dk, dp, n, kprime, c->prow);
      n = kprime;
      if (lane == 0) {
        uint32_t old = atomicMin(c->sthr + qq, T);
        c->sthrf[qq] = key_to_float(min(old, T));
        if (c->qthr) publish_bound(c, c->sqidx[qq], T);
      }
    }
    __syncwarp();
    if (lane < NW) c->ccnt[qq * NW + lane] = lane == 0 ? n : 0u;
  }
  __syncthreads();
}

// End of an item for one query, small candidate counts (n <= 128, the usual case once the bounds are warm): the
// candidates go from the NW per-warp segments straight into registers (4 per lane, located through a prefix scan of
// the segment counts), every candidate is ranked against every other by (key, row id, position) with shuffles, and
// the first k' are written to their rank.  Same order as warp_select_k + warp_sort_emit (select.cuh), without the
// segment-by-segment gather (16 dependent shared-memory round trips) and the 32-step bisection over shared memory
// (r02_bt ncu, cfg 5: the select was 16 % of the warp time with nq of 16 warps busy).  Returns false when n > 128.
// Returns bit 63: handled, bit 62: the query has k' candidates and bits 0-31 hold the k'-th best key (the bound to
// publish), bits 32-47: rows written.
static __device__ __noinline__ uint64_t rank_emit_small(const uint32_t* dk, const uint32_t* dp, const uint32_t* cnt, uint32_t NW,
                                                        uint32_t capw, uint32_t kprime, const uint64_t* __restrict__ prow,
                                                        uint64_t row_base, float* __restrict__ out_dist,
                                                        uint64_t* __restrict__ out_pos, uint64_t* __restrict__ out_rowid,
                                                        uint32_t lane) {
  const uint32_t cn = lane < NW ? cnt[lane] : 0u;
  uint32_t inc = cn;  // inclusive prefix over the segments
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(FULL, inc, o);
    if (lane >= (uint32_t)o) inc += v;
  }
  const uint32_t n = __shfl_sync(FULL, inc, 31);
  if (n > 128) return 0ull;
  uint32_t seg[4], base[4];
#pragma unroll
  for (int r = 0; r < 4; r++) seg[r] = base[r] = 0;
  for (uint32_t l = 0; l < NW; l++) {  // segment of element idx = number of segments that end at or before it
    const uint32_t e = __shfl_sync(FULL, inc, l);
#pragma unroll
    for (int r = 0; r < 4; r++)
      if (e <= lane + 32u * r) { seg[r] = l + 1; base[r] = e; }
  }
  uint32_t key[4], pos[4], rank[4];
#pragma unroll
  for (int r = 0; r < 4; r++) {
    const uint32_t idx = lane + 32u * r;
    key[r] = 0xffffffffu; pos[r] = 0; rank[r] = 0;
    if (idx < n) {
      key[r] = dk[(size_t)seg[r] * capw + idx - base[r]];
      pos[r] = dp[(size_t)seg[r] * capw + idx - base[r]];
    }
  }
#pragma unroll
  for (int rj = 0; rj < 4; rj++) {
    const uint32_t nj = n > 32u * rj ? min(32u, n - 32u * rj) : 0u;
    for (uint32_t jl = 0; jl < nj; jl++) {
      const uint32_t kj = __shfl_sync(FULL, key[rj], jl), pj = __shfl_sync(FULL, pos[rj], jl);
#pragma unroll
      for (int r = 0; r < 4; r++) {
        if (kj < key[r]) rank[r]++;
        else if (kj == key[r] && !(rj == r && jl == lane) && lane + 32u * r < n && cand_less(kj, pj, key[r], pos[r], prow)) rank[r]++;
      }
    }
  }
  const uint32_t nout = min(n, kprime);
  uint32_t myT = 0;
  bool mine = false;
#pragma unroll
  for (int r = 0; r < 4; r++) {
    if (lane + 32u * r < n && rank[r] < nout) {
      out_dist[rank[r]] = key_to_float(key[r]);
      out_pos[rank[r]] = row_base + pos[r];
      out_rowid[rank[r]] = prow[pos[r]];
      if (rank[r] == kprime - 1) { myT = key[r]; mine = true; }  // the k'-th best
    }
  }
  const uint32_t who = __ballot_sync(FULL, mine);
  uint64_t ret = (1ull << 63) | ((uint64_t)nout << 32);
  if (who) ret |= (1ull << 62) | __shfl_sync(FULL, myT, __ffs(who) - 1);
  return ret;
}
static __device__ __forceinline__ bool finalize_small(const ItemCtx* c, uint32_t qq, uint32_t lane) {
  const size_t slot = (size_t)c->spair[qq] * c->max_chunks + c->chunk;
  const uint64_t r = rank_emit_small(c->ckey + (size_t)qq * c->nw * c->capw, c->cpos + (size_t)qq * c->nw * c->capw,
                                     c->ccnt + qq * c->nw, c->nw, c->capw, c->kprime, c->prow, c->row_base,
                                     c->out_dist + slot * c->kprime, c->out_pos + slot * c->kprime,
                                     c->out_rowid + slot * c->kprime, lane);
  if (!(r >> 63)) return false;
  if (lane == 0) {
    if (((r >> 62) & 1ull) && c->qthr) publish_bound(c, c->sqidx[qq], (uint32_t)r);
    c->out_count[slot] = (uint32_t)(r >> 32) & 0xffffu;
  }
  return true;
}

// End of an item: one warp per query gathers the warps' buffers, selects k', emits sorted.
static __device__ __noinline__ void finalize_fn(const ItemCtx* c) {
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t NW = c->nw, capw = c->capw, kprime = c->kprime;
  for (uint32_t qq = w; qq < c->nq; qq += NW) {
    if (LB200_FINALIZE_SMALL && finalize_small(c, qq, lane)) continue;
    uint32_t* dk = c->ckey + (size_t)qq * NW * capw;
    uint32_t* dp = c->cpos + (size_t)qq * NW * capw;
    uint32_t n = gather_segments(dk, dp, c->ccnt + qq * NW, NW, capw, lane);
    if (n > kprime) {
      uint32_t T = warp_select_k(dk, dp, n, kprime, c->prow);
      n = kprime;
      if (c->qthr && lane == 0) publish_bound(c, c->sqidx[qq], T);
    } else if (n == kprime && c->qthr) {
      uint32_t mx = 0;
      for (uint32_t i = lane; i < n; i += 32) mx = max(mx, dk[i]);
      mx = __reduce_max_sync(FULL, mx);
      if (lane == 0) publish_bound(c, c->sqidx[qq], mx);
    }
    const size_t slot = (size_t)c->spair[qq] * c->max_chunks + c->chunk;
    warp_sort_emit(dk, dp, n, c->prow, c->row_base, c->out_dist + slot * kprime, c->out_pos + slot * kprime,
                   c->out_rowid + slot * kprime);
    if (lane == 0) c->out_count[slot] = n;
  }
}

// One kernel, two instantiations per shape.  BOOT = true compiles the pooled-bootstrap calls in (shards, few
// probes per query, per-pair outputs: most work items are cold) and keeps the end-of-item code out of line;
// BOOT = false is the call-free text for batches whose items are mostly warm (merged search, nprobes >= 16):
// no ItemCtx, end-of-item code inline.  (r01 kept two verbatim copies of this function because the row loop is
// sensitive to what else lives in it; `if constexpr` gives the same two texts from one source.)
template <int QL, int D, int NW, int MINB, int DSUB, bool BOOT>
__global__ void __launch_bounds__(NW * 32, MINB) scan3_kernel(const Scan3Params p,
                                                              const uint32_t* __restrict__ total_items_p) {
  constexpr int NT = NW * 32;
  constexpr uint32_t CB = 256u * 64u * QL;  // bytes per table chunk (16 sub-vectors)
  constexpr uint32_t SLOTB = 4u * QL;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char* lut = smem_raw;
  const uint32_t C16 = p.M >> 4;
  float* resid = reinterpret_cast<float*>(smem_raw + (size_t)C16 * CB);
  uint32_t* ckey = reinterpret_cast<uint32_t*>(resid + (p.lut_src ? 0u : QL * p.dim));  // no residuals with pre-built tables
  uint32_t* cpos = ckey + (size_t)QL * NW * p.capw;
  uint32_t* ccnt = cpos + (size_t)QL * NW * p.capw;
  float* sthrf = reinterpret_cast<float*>(ccnt + QL * NW);  // 16-byte aligned (all sizes are multiples of 16)
  uint32_t* sthr = reinterpret_cast<uint32_t*>(sthrf + QL);
  uint32_t* spair = sthr + QL;
  uint32_t* sqidx = spair + QL;
  uint32_t* sitem = sqidx + QL;                            // [0]: the item to run next
  uint4* sdesc = reinterpret_cast<uint4*>(sitem + 4);      // its descriptor (partition, query group, chunk)
  ItemCtx* ictx = reinterpret_cast<ItemCtx*>(sitem + 8);   // 16-byte aligned
  const uint32_t sthrf_saddr = (uint32_t)__cvta_generic_to_shared(sthrf);
  __shared__ __align__(8) uint64_t lut_bar;   // completion of the bulk table fetch (image mode)
  __shared__ uint32_t lut_phase;              // its parity (kept out of the registers: the row loop is at the 128 cap)
  __shared__ uint32_t sunit_n[4];             // wide CTAs, pre-built tables: the NEXT item's table units, its query count,
  __shared__ uint32_t snq_n, spref;           // and whether the item now starting had its tables fetched ahead

  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (MINB == 1 && p.lut_image && tid == 0) {
    lut_phase = 0;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&lut_bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const uint32_t d = lane % D;
  const uint32_t M = p.M, dsub = p.dsub, dim = p.dim, kprime = p.kprime, capw = p.capw, L = p.L;
  const uint32_t units = L * C16 + 1;
  const uint32_t total_items = *total_items_p;

  // per-lane table offsets for static step j (see header): slot = (j-d) mod 16;
  //   Bp[j]: relative to chunk (u-1): + CB when the lane is already on chunk u (j >= d)
  //   B0[j]: unit 0 of a row, relative to the table start: lanes with j < d are on the LAST chunk
  uint32_t Bp[16], B0[D];
  float keepf[D], swf[D];
#pragma unroll
  for (int j = 0; j < 16; j++) Bp[j] = (((uint32_t)j - d) & 15u) * SLOTB + ((uint32_t)j >= d ? CB : 0u);
#pragma unroll
  for (int j = 0; j < D; j++) {
    B0[j] = (((uint32_t)j - d) & 15u) * SLOTB + ((uint32_t)j >= d ? 0u : (C16 - 1) * CB);
    keepf[j] = d == (uint32_t)j ? 0.0f : 1.0f;
    swf[j] = d == (uint32_t)j ? 1.0f : 0.0f;
  }
  // Keep these per-lane constants in registers: left alone, ptxas rematerialises them inside the
  // hot loop (ISETP+FSEL per switch step; r01 profile) which costs more issue slots than the FMAs.
#pragma unroll
  for (int j = 0; j < 16; j++) asm volatile("" : "+r"(Bp[j]));
#pragma unroll
  for (int j = 0; j < D; j++) {
    asm volatile("" : "+r"(B0[j]));
    asm volatile("" : "+f"(keepf[j]));
    asm volatile("" : "+f"(swf[j]));
  }

  // Work items come from a device queue.  Taking one costs an atomic and a descriptor load, ~2 us of latency
  // that r02's ncu put at 9.5 % of the warp time when it sat between items (atomicAdd -> barrier -> binary
  // search over item_off).  Now the NEXT item is taken while the current one runs: the atomic is issued at the
  // top of an item and its result is first used after the table build; the descriptor then travels global ->
  // shared with cp.async during the row scan.
  if (tid == 0) {
    spref = 0;
    const uint32_t it = atomicAdd(p.item_counter, 1u);
    sitem[0] = it;
    if (it < total_items) {
      const uint4 d0 = p.item_desc[it];
      *sdesc = d0;
      if (MINB == 1 && p.lut_image)
        lut_image_issue(lut, p.lut_src + lut_image_base(p.part_pair_off[d0.x], d0.x, d0.y, QL) * M * 256, C16 * CB, &lut_bar);
    }
  }
  for (;;) {
    __syncthreads();
    const uint32_t item = sitem[0];
    if (item >= total_items) break;
    const uint4 de = *sdesc;
    uint32_t nxt = 0;
    if (tid == 0) nxt = atomicAdd(p.item_counter, 1u);
    const uint32_t part = de.x, g = de.y, c = de.z;
    const uint64_t row_base = p.part_off[part];
    const uint32_t n_p = (uint32_t)(p.part_off[part + 1] - row_base);
    const uint32_t pbase = p.part_pair_off[part] + g * QL;
    const uint32_t nq = min((uint32_t)QL, p.part_pair_off[part + 1] - pbase);
    const uint64_t* prow = p.row_ids + row_base;
    const uint64_t* bm = p.bitmask ? p.bitmask[part] : nullptr;

    bool mycold = false;
    if (tid < QL) {
      uint32_t pr = tid < nq ? p.pair_list[pbase + tid] : 0xffffffffu;
      spair[tid] = pr;
      const uint32_t qi = tid < nq ? pr / p.P : 0u;
      sqidx[tid] = qi;
      uint32_t t0 = KEY_PINF;
      if (p.qthr && tid < nq) t0 = __ldcg(p.qthr + qi);
      sthr[tid] = t0;
      // unused query slots never pass the fast test
      sthrf[tid] = tid < nq ? key_to_float(t0) : __int_as_float(0xff800000);
      mycold = tid < nq && t0 == KEY_PINF;
    }
    if (tid < NW * QL) ccnt[tid] = 0;
    if (BOOT && tid == 0) {
      ItemCtx ic;
      ic.ckey = ckey; ic.cpos = cpos; ic.ccnt = ccnt; ic.sthr = sthr; ic.sqidx = sqidx; ic.spair = spair;
      ic.qthr = p.qthr; ic.peers = p.peers; ic.n_peers = p.n_peers; ic.sthrf = sthrf; ic.prow = prow; ic.bm = bm;
      ic.out_dist = p.out_dist; ic.out_pos = p.out_pos; ic.out_rowid = p.out_rowid; ic.out_count = p.out_count;
      ic.row_base = row_base; ic.capw = capw; ic.kprime = kprime; ic.nq = nq;
      ic.vend = min(n_p, c * p.chunk_vecs + p.chunk_vecs); ic.nw = NW; ic.max_chunks = p.max_chunks; ic.chunk = c;
      ic.lower = p.lower; ic.upper = p.upper; ic.has_lower = p.has_lower; ic.has_upper = p.has_upper;
      *ictx = ic;
    }
    if (!p.lut_src)
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
    bool cold_item = false;
    if (BOOT) cold_item = __syncthreads_or(mycold) != 0;  // some query of the item has no bound yet
    else __syncthreads();

    // ---- A4: ADC tables, interleaved over the QL queries --------------------------------
    // thread -> (slot = tid%16, code = tid/16 + k*NT/16): a quarter/half warp writes consecutive
    // slots of one code row -> conflict-free wide stores.
    uint32_t maxbits = 0;
    if (MINB == 1 && p.lut_image) {  // (one-CTA-per-SM instantiations only: the others sit at the register cap)
      maxbits = lut_image_wait(p.lut_max, pbase, nq, &lut_bar, &lut_phase);  // issued at the end of the previous item
    } else if (p.lut_src) {
      if (NW == 16 && spref)  // issued by the idle warps at the end of the previous item (below)
        maxbits = lut_fetch_wait<QL>(p.lut_max, p.lut_by_query ? sqidx : spair, nq);
      else
        maxbits = lut_fetch<QL, NW>(lut, p.lut_src, p.lut_max, (size_t)M * 256, p.lut_by_query ? sqidx : spair, nq, C16 * 4096u);
    } else {
      const uint32_t slot = tid & 15;
#pragma unroll 1
      for (uint32_t ch = 0; ch < C16; ch++) {
        const uint32_t m = ch * 16 + slot;
        float xr[QL][DSUB > 0 ? DSUB : 1];
        if (DSUB > 0) {
#pragma unroll
          for (int q = 0; q < QL; q++)
#pragma unroll
            for (int i = 0; i < DSUB; i++) xr[q][i] = resid[q * dim + m * DSUB + i];
        }
        // the codebook rows come from L2 (128 KB+ per item, more than L1 keeps): the next row is loaded while the
        // current one is worked on (r02 ncu: 5.6 % of the kernel's warp time sat on the first use of this load)
        constexpr int CBD = LB200_CB_DEPTH;                       // loads in flight per thread (code steps ahead)
        constexpr uint32_t CBSTEP = (NT / 32) * (DSUB > 0 ? DSUB / 4 : 1) * 32;  // float4s between code steps
        float4 nx[CBD][DSUB > 0 ? DSUB / 4 : 1];
        // lane-order image: this warp's 32 (code parity, slot) rows of one load are 512 contiguous bytes
        const float4* cb4 = reinterpret_cast<const float4*>(p.codebook) + cb_image_f4(ch, tid >> 4, slot, 0, DSUB > 0 ? DSUB / 4 : 1);
        if (DSUB > 0) {
#pragma unroll
          for (int d = 0; d < CBD; d++)
#pragma unroll
            for (int i = 0; i < DSUB / 4; i++) nx[d][i] = __ldg(cb4 + d * CBSTEP + i * 32);
        }
#pragma unroll 1
        for (uint32_t code0 = tid >> 4; code0 < 256; code0 += CBD * (NT / 16)) {
#pragma unroll
        for (int d = 0; d < CBD; d++) {
          const uint32_t code = code0 + d * (NT / 16);
          const float* y = p.codebook + ((size_t)m * 256 + code) * dsub;
          float yr[DSUB > 0 ? DSUB : 1];
          if (DSUB > 0) {
#pragma unroll
            for (int i = 0; i < DSUB / 4; i++) {
              yr[4 * i] = nx[d][i].x; yr[4 * i + 1] = nx[d][i].y; yr[4 * i + 2] = nx[d][i].z; yr[4 * i + 3] = nx[d][i].w;
            }
            if (code + CBD * (NT / 16) < 256) {
#pragma unroll
              for (int i = 0; i < DSUB / 4; i++) nx[d][i] = __ldg(cb4 + (d + CBD) * CBSTEP + i * 32);
            }
          }
          float val[QL];
          if (LB200_OPT_TABLE && DSUB > 0 && p.residual) {  // l2_once on packed pairs: FADD2 / FMUL2, bit-identical per half
            f32x2_t Y[DSUB > 0 ? DSUB / 2 : 1];
#pragma unroll
            for (int i = 0; i < DSUB / 2; i++) Y[i] = pk2(yr[2 * i], yr[2 * i + 1]);
#pragma unroll
            for (int q = 0; q < QL; q++) {
              f32x2_t X[DSUB > 0 ? DSUB / 2 : 1];
#pragma unroll
              for (int i = 0; i < DSUB / 2; i++) X[i] = pk2(xr[q][2 * i], xr[q][2 * i + 1]);
              val[q] = l2_once_packed<(DSUB > 0 ? DSUB : 8)>(X, Y);
              maxbits = max(maxbits, __float_as_uint(val[q]) & 0x7fffffffu);
            }
          } else {
#pragma unroll
          for (int q = 0; q < QL; q++) {
            if (DSUB > 0)
              val[q] = lut_elem<DSUB>(p.residual != 0, xr[q], yr, dsub);
            else
              val[q] = lut_elem<0>(p.residual != 0, resid + q * dim + m * dsub, y, dsub);
            maxbits = max(maxbits, __float_as_uint(val[q]) & 0x7fffffffu);
          }
          }
          unsigned char* dst = lut + (size_t)ch * CB + code * (64 * QL) + slot * SLOTB;
          if (QL == 4)
            *reinterpret_cast<float4*>(dst) = make_float4(val[0], val[1], val[2], val[3]);
          else
            *reinterpret_cast<float2*>(dst) = make_float2(val[0], val[1]);
        }
        cb4 += CBD * CBSTEP;
        }
      }
    }
    // exact-FMA switch only if no partial sum of <=256 entries can overflow: |T| < 2^118
    const bool unsafe = __syncthreads_or(maxbits >= 0x7a800000u) != 0;
    if (tid == 0) {  // everyone has read sitem / sdesc (two barriers ago): hand over the next item
      if (MINB == 1 && p.lut_image) lut_phase ^= 1u;
      sitem[0] = nxt;
      if (nxt < total_items)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(sdesc)),
                     "l"(p.item_desc + nxt) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
    }

    // ---- A5 + A6: scan the chunk's rows -----------------------------------------------------
    const uint32_t v0 = c * p.chunk_vecs;
    const uint32_t vend = min(n_p, v0 + p.chunk_vecs);
    const uint32_t sb_lo = v0 / (32 * L);
    const float lower = p.lower, upper = p.upper;
    const bool has_lower = p.has_lower != 0, has_upper = p.has_upper != 0;
    const uint8_t* cbase = p.codes + (size_t)p.sb_off[part] * units * 512 + lane * 16;

    // Pooled bootstrap of a cold item (some query has no bound yet): the candidates the NW warps
    // buffered in their first rows are pooled and pruned once (#1 after ~96 rows per warp, #2 at the
    // end of each warp's first slice).  The calls sit BETWEEN segments of the row loop, never
    // inside it: the loop body must stay free of calls and of inlined cold code (r01 bisect: either
    // costs the scan ~10% through register pressure / instruction-cache misses).
    const bool cold = BOOT && cold_item && !p.no_boot;
    bool need1 = cold, need2 = cold && p.boot2 != 0;
    const uint32_t boot_i = 3u;
    // The chunk's rows are dealt to the NW warps in row groups (32 rows = one row per lane) so that all warps
    // finish an item together: whole superblocks per warp left 15 % of the warp time in the end-of-item barrier
    // (r01 ncu: ceil(nsb / NW) rounds for nsb ~ 24 superblocks).  A warp's share [rg, rg_end) is cut into slices
    // (sb, rows i0..i1 of each lane) at superblock borders.  A slice may start and end mid-stream: its first
    // switch output (the tail of row i0-1) is dropped, and the tail of its last row is read from the first D
    // bytes of the NEXT unit -- row i1's first unit, or the superblock's drain unit when i1 == L.
    const uint32_t Tv = (vend - v0 + 31) / 32;
    uint32_t rg = (uint32_t)(((uint64_t)Tv * w) / NW);
    const uint32_t rg_end = (uint32_t)(((uint64_t)Tv * (w + 1)) / NW);
#pragma unroll 1
    while (rg < rg_end) {
      const uint32_t sbr = rg / L, i0 = rg - sbr * L, i1 = min(L, i0 + (rg_end - rg));
      const uint32_t sb = sb_lo + sbr;
      rg += i1 - i0;
      const uint4* up = reinterpret_cast<const uint4*>(cbase + ((size_t)sb * units + (size_t)i0 * C16) * 512);
      uint4 cur = load_codes(up), n1 = load_codes(up + 32);
      const uint4* pf = up + 64;  // prefetch pointer, 2 units ahead (the buffer is padded by 2 units)
      float acc[QL], done[QL];
#pragma unroll
      for (int q = 0; q < QL; q++) acc[q] = done[q] = 0.0f;
      const uint32_t row0 = sb * 32 * L + lane;
      // pick up bounds found meanwhile by other work items of the same queries
      if (p.qthr) {
        if (lane < nq) {
          uint32_t t = __ldcg(p.qthr + sqidx[lane]);
          if (t < sthr[lane]) {
            uint32_t old = atomicMin(sthr + lane, t);
            sthrf[lane] = key_to_float(min(old, t));
          }
        }
        __syncwarp();
      }
      uint32_t i = i0;
      uint32_t stop = (BOOT && need1) ? min(i1, i0 + boot_i) + 1 : i1 + 1;
#pragma unroll 1
      for (;;) {
#pragma unroll 1
      for (; i < stop; i++) {
        const uint4 n2 = load_codes(pf);
        pf += 32;
        if (i < i1) {
          if (unsafe) scan_unit<QL, D, true, true, 16>(cur, lut - CB, Bp, B0, keepf, swf, d, acc, done);
          else scan_unit<QL, D, true, false, 16>(cur, lut - CB, Bp, B0, keepf, swf, d, acc, done);
        } else {
          if (unsafe) scan_unit<QL, D, true, true, D>(cur, lut - CB, Bp, B0, keepf, swf, d, acc, done);
          else scan_unit<QL, D, true, false, D>(cur, lut - CB, Bp, B0, keepf, swf, d, acc, done);
        }
        cur = n1; n1 = n2;
        if (i > i0) {
          // `done` = distance of row i-1 of this lane.  Fast test: running threshold only; bounds,
          // tail rows and the prefilter are applied on the (rare) slow path.
          float thr[QL];
          lds_volatile<QL>(sthrf_saddr, thr);
          bool anyp = done[0] <= thr[0];
#pragma unroll
          for (int q = 1; q < QL; q++) anyp = anyp || (done[q] <= thr[q]);
          if (__any_sync(FULL, anyp)) {
            const uint32_t r = row0 + (i - 1) * 32;
            uint32_t pm = 0;
#pragma unroll
            for (int q = 0; q < QL; q++) {
              bool ok = done[q] <= thr[q];
              if (has_lower) ok = ok && (lower <= done[q]);
              if (has_upper) ok = ok && (done[q] < upper);
              pm |= ok ? (1u << q) : 0u;
            }
            if (r >= vend) pm = 0;
            pm &= (1u << nq) - 1u;
            if (bm && pm) {
              if (!((bm[r >> 6] >> (r & 63)) & 1ull)) pm = 0;
            }
            if (__any_sync(FULL, pm != 0)) {
#pragma unroll
              for (int q = 0; q < QL; q++) {
                const uint32_t b = (uint32_t)q * NW + w;
                append_rows((pm >> q) & 1u, done[q], r, lane, capw, kprime, ckey + (size_t)b * capw,
                            cpos + (size_t)b * capw, ccnt + b, sthr + q, sthrf + q,
                            (p.qthr && (uint32_t)q < nq) ? p.qthr + sqidx[q] : nullptr, prow);
              }
            }
          }
        }
        if (i == i1) continue;  // drain iteration: no further units (and the last of the loop)
#pragma unroll 1
        for (uint32_t u = 1; u < C16; u++) {
          const uint4 n3 = load_codes(pf);
          pf += 32;
          scan_unit<QL, D, false, true, 16>(cur, lut + (size_t)(u - 1) * CB, Bp, B0, keepf, swf, d, acc, done);
          cur = n1; n1 = n3;
        }
      }
        if (BOOT && need1) {
          bootstrap_fn(ictx);
          need1 = false;
          stop = i1 + 1;
        }
        if (!BOOT || i > i1) break;
      }
      if (BOOT && need2) {
        bootstrap_fn(ictx);
        need2 = false;
      }
    }
    if (BOOT && need1) bootstrap_fn(ictx);  // warps that drew no rows still owe the barriers
    if (BOOT && need2) bootstrap_fn(ictx);
    if (tid == 0) asm volatile("cp.async.wait_group 0;" ::: "memory");  // the next item's descriptor has landed
    __syncthreads();
    // image mode: nobody reads this item's table any more -> start the next item's fetch now, from a warp that has
    // no query to select for (nq <= QL < NW), so that it runs under the end-of-item select
    if (MINB == 1 && p.lut_image && tid == NT - 32 && sitem[0] < total_items) {
      const uint4 dn = *sdesc;
      lut_image_issue(lut, p.lut_src + lut_image_base(p.part_pair_off[dn.x], dn.x, dn.y, QL) * M * 256, C16 * CB, &lut_bar);
    }
    // wide CTAs with pre-built tables that are gathered per query (lut_fetch): the end-of-item select keeps nq <= QL
    // warps busy and the rest waited for it, then everyone issued the next item's 4-byte copies (r02_bt ncu, cfg 5:
    // 16 % of the warp time behind the select, 17 % of the instructions in the fetch).  The warps without a query
    // now work out the next item's units and issue its fetch while the select runs.
    if constexpr (NW == 16 && BOOT) {
      if (p.lut_src && !p.lut_image) {
        const bool have = sitem[0] < total_items;
        if (tid == 0) spref = have ? 1u : 0u;
        if (have && w >= QL) {
          const uint32_t i = tid - QL * 32;
          if (i < QL) {
            const uint4 dn = *sdesc;
            const uint32_t pb = p.part_pair_off[dn.x] + dn.y * QL;
            const uint32_t nqn = min((uint32_t)QL, p.part_pair_off[dn.x + 1] - pb);
            uint32_t u = 0;
            if (i < nqn) {
              u = p.pair_list[pb + i];
              if (p.lut_by_query) u /= p.P;
            }
            sunit_n[i] = u;
            if (i == 0) snq_n = nqn;
          }
          asm volatile("bar.sync 1, %0;" ::"n"((NW - QL) * 32) : "memory");
          lut_fetch_issue<QL, NW - QL>(lut, p.lut_src, (size_t)M * 256, sunit_n, snq_n, C16 * 4096u, w - QL);
        }
      }
    }

    if constexpr (BOOT) {
      finalize_fn(ictx);
    } else {
      // ---- end of item, inline: one warp per query gathers the warps' buffers, selects k', emits sorted ----
#pragma unroll 1
      for (uint32_t qq = w; qq < nq; qq += NW) {
        uint32_t* dk = ckey + (size_t)qq * NW * capw;
        uint32_t* dp = cpos + (size_t)qq * NW * capw;
        if (LB200_FINALIZE_SMALL_INLINE) {
          const size_t slot = (size_t)spair[qq] * p.max_chunks + c;
          const uint64_t r = rank_emit_small(dk, dp, ccnt + qq * NW, NW, capw, kprime, prow, row_base, p.out_dist + slot * kprime,
                                             p.out_pos + slot * kprime, p.out_rowid + slot * kprime, lane);
          if (r >> 63) {
            if (lane == 0) {
              if (((r >> 62) & 1ull) && p.qthr) atomicMin(p.qthr + sqidx[qq], (uint32_t)r);
              p.out_count[slot] = (uint32_t)(r >> 32) & 0xffffu;
            }
            continue;
          }
        }
        uint32_t n = ccnt[qq * NW];
#pragma unroll 1
        for (uint32_t w2 = 1; w2 < NW; w2++) {
          const uint32_t n2 = ccnt[qq * NW + w2];
          const uint32_t* sk = dk + (size_t)w2 * capw;
          const uint32_t* sp = dp + (size_t)w2 * capw;
          for (uint32_t j0 = 0; j0 < n2; j0 += 32) {
            uint32_t j = j0 + lane, kk = 0, pp = 0;
            if (j < n2) { kk = sk[j]; pp = sp[j]; }
            __syncwarp();
            if (j < n2) { dk[n + j] = kk; dp[n + j] = pp; }
            __syncwarp();
          }
          n += n2;
        }
        if (n > kprime) {
          uint32_t T = warp_select_k(dk, dp, n, kprime, prow);
          n = kprime;
          if (p.qthr && lane == 0) atomicMin(p.qthr + sqidx[qq], T);
        } else if (n == kprime && p.qthr) {
          uint32_t mx = 0;
          for (uint32_t j = lane; j < n; j += 32) mx = max(mx, dk[j]);
          mx = __reduce_max_sync(FULL, mx);
          if (lane == 0) atomicMin(p.qthr + sqidx[qq], mx);
        }
        const size_t slot = (size_t)spair[qq] * p.max_chunks + c;
        warp_sort_emit(dk, dp, n, prow, row_base, p.out_dist + slot * kprime, p.out_pos + slot * kprime,
                       p.out_rowid + slot * kprime);
        if (lane == 0) p.out_count[slot] = n;
      }
    }
  }
}

}  // namespace lb200

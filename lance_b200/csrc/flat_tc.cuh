// flat_tc.cuh -- tensor-core CANDIDATE FILTER of the flat kNN path (K7), hand-written for sm_100a:
// TMA (cp.async.bulk.tensor) stages bf16 tiles in shared memory, one elected thread issues tcgen05.mma
// (cta_group::1, kind::f16, bf16 x bf16 -> f32) into TMEM accumulators, four epilogue warps read them back with
// tcgen05.ld and keep the rows that CAN be among a query's k nearest.
//
// Why a filter and not the distance itself: the flat path must return the reference's f32 distances bit for bit
// (l2_scalar<16> lane order, flat.rs:28-86), and no GEMM formulation keeps that arithmetic.  "Bound and verify" does:
//   approx(q, y) = |q|^2 + |y|^2 - 2 <bf16(q), bf16(y)>          (this kernel)
//   |approx - exact_f32| <= E(q, y) = e1 |q| |y| + e2 (|q|^2 + |y|^2)   (proved bound, api_column.inl)
//   T_q  = exact k-th distance of q over a SAMPLE of the rows    (exact kernels; T_q >= the true k-th distance)
//   keep row y  iff  approx - E <= T_q                            (a superset of the true top k, ties included)
// and the kept rows (a few hundred of a million) are re-scored by the EXACT fp32 kernels and selected by
// (distance, row index) as before -- identical results by construction, 1/1000 of the exact arithmetic.
//
// Tile shapes: UMMA M = 128 queries (TMEM lanes) x N = 128 rows (TMEM columns) x K = 16 per instruction;
// operands K-major in 128-byte-swizzled [rows][64 bf16] blocks (one TMA box each), dim = 64 * KB, KB <= 4.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "dist.cuh"

namespace lb200 {

constexpr uint32_t FTC_M = 128;        // queries per CTA (TMEM lanes)
constexpr uint32_t FTC_N = 128;        // rows per tile (TMEM columns per accumulator stage)
constexpr uint32_t FTC_KBLK = 64;      // bf16 elements per swizzle row (128 bytes)
constexpr uint32_t FTC_THREADS = 384;  // warp 0 TMA, warp 1 MMA, warp 2 TMEM alloc, warps 4-11 epilogue
constexpr uint32_t FTC_TILE_BYTES = FTC_N * FTC_KBLK * 2;  // one [128][64] bf16 block = 16 KB

__device__ __forceinline__ uint32_t ftc_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ftc_mbar_init(uint64_t* b, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ftc_smem(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void ftc_mbar_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ftc_smem(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ftc_mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ftc_smem(b)) : "memory");
}
__device__ __forceinline__ void ftc_mbar_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(ftc_smem(b)), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void ftc_tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(ftc_smem(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(ftc_smem(bar)), "r"(c0), "r"(c1) : "memory");
}
// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor): start address
// >> 4 | LBO (= 1, unused for swizzled K-major) << 16 | SBO (8 rows x 128 B = 1024 B >> 4 = 64) << 32 | version 1 << 46
// | layout type 2 (SWIZZLE_128B) << 61.  The tile base is 1024-byte aligned (base_offset 0).
__device__ __forceinline__ uint64_t ftc_desc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3ffffu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, f32 accumulate; issued by ONE thread for the CTA
__device__ __forceinline__ void ftc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
               ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void ftc_commit(uint64_t* bar) {  // arrives on `bar` when all MMAs issued so far have finished
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(ftc_smem(bar)) : "memory");
}

struct FlatTcParams {
  const float2* row_norm;   // [ntiles * 128] (|y|^2 (1 - e2'), |y|) from the f32 rows; (+inf, 0) past n
  const float2* q_norm;     // [Qpad] (|q|^2, |q|)
  const float* q_thr;       // [Qpad] T_q (exact k-th distance over the sample; -inf for padding queries)
  uint32_t* cand;           // [Qpad][cap] row indices
  uint32_t* cand_cnt;       // [Qpad] (may exceed cap: overflow, the host falls back to the exact path)
  uint32_t cap;
  uint32_t n;               // rows
  uint32_t ntiles;          // ceil(n / FTC_N)
  uint32_t kb;              // dim / 64
  uint32_t stages;          // B pipeline depth
  float e1, e2;             // error bound coefficients, each + 2^-21 (see the epilogue)
  float* dump;              // DUMP mode: [nq][n] upper bounds of the exact distances (approx + E), instead of candidates
  uint32_t nq;              // DUMP mode: real queries (rows of `dump`)
};

// MET: 0 L2, 1 cosine (map_v = rows normalised to unit length before the bf16 rounding), 2 dot.
// DUMP: no filtering -- every (query, row) gets approx + E, an UPPER bound of its exact distance, written to p.dump.
// Run over the column's sample it yields the thresholds without any exact distance: the k-th smallest upper bound
// over a subset of the rows is >= the k-th smallest exact distance over all rows.  (row_norm then holds
// (|y|^2 (1 + e2), |y|) for L2.)
template <int MET, bool DUMP>
__global__ void __launch_bounds__(FTC_THREADS, 1)
flat_tc_filter_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_v, const FlatTcParams p) {
  extern __shared__ __align__(1024) unsigned char ftc_smem_raw[];
  // [A: kb blocks][B: stages x kb blocks][barriers]
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(ftc_smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* smem_a = base;
  unsigned char* smem_b = base + (size_t)p.kb * FTC_TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + (size_t)p.stages * p.kb * FTC_TILE_BYTES);
  uint64_t* full = bars;                 // [stages] TMA -> MMA
  uint64_t* empty = bars + 8;            // [stages] MMA -> TMA
  uint64_t* tfull = bars + 16;           // [2] MMA -> epilogue
  uint64_t* tempty = bars + 18;          // [2] epilogue -> MMA
  uint64_t* afull = bars + 20;           // [1] A tile landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 21);
  float2* norm_buf = reinterpret_cast<float2*>(bars + 32);   // [8 epilogue warps][64] row norms of the current tile

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t qblk = blockIdx.y;
  const uint32_t S = p.stages, KB = p.kb;

  if (warp == 1 && lane == 0) {
    for (uint32_t i = 0; i < S; i++) { ftc_mbar_init(full + i, 1); ftc_mbar_init(empty + i, 1); }
    for (uint32_t i = 0; i < 2; i++) { ftc_mbar_init(tfull + i, 1); ftc_mbar_init(tempty + i, 8); }
    ftc_mbar_init(afull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {  // 2 accumulator stages x 128 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(ftc_smem(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  // tiles of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...
  const uint32_t my_tiles = p.ntiles > blockIdx.x ? (p.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    ftc_mbar_expect_tx(afull, KB * FTC_TILE_BYTES);
    for (uint32_t kb = 0; kb < KB; kb++)
      ftc_tma_load_2d(smem_a + (size_t)kb * FTC_TILE_BYTES, &map_q, afull, (int32_t)(kb * FTC_KBLK), (int32_t)(qblk * FTC_M));
    for (uint32_t t = 0; t < my_tiles; t++) {
      const uint32_t s = t % S, ph = (t / S) & 1;
      ftc_mbar_wait(empty + s, ph ^ 1);
      ftc_mbar_expect_tx(full + s, KB * FTC_TILE_BYTES);
      const uint32_t row0 = (blockIdx.x + t * gridDim.x) * FTC_N;
      for (uint32_t kb = 0; kb < KB; kb++)
        ftc_tma_load_2d(smem_b + ((size_t)s * KB + kb) * FTC_TILE_BYTES, &map_v, full + s, (int32_t)(kb * FTC_KBLK), (int32_t)row0);
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    // instruction descriptor (InstrDescriptor): c_format F32 (1 << 4), a/b format BF16 (1 << 7, 1 << 10), K-major both,
    // N >> 3 at bit 17, M >> 4 at bit 24
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((FTC_N >> 3) << 17) | ((FTC_M >> 4) << 24);
    ftc_mbar_wait(afull, 0);
    for (uint32_t t = 0; t < my_tiles; t++) {
      const uint32_t s = t % S, ph = (t / S) & 1, acc = t & 1, aph = (t >> 1) & 1;
      ftc_mbar_wait(tempty + acc, aph ^ 1);
      ftc_mbar_wait(full + s, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_tmem = tmem_base + acc * FTC_N;
      for (uint32_t kb = 0; kb < KB; kb++) {
        const uint64_t ad = ftc_desc(ftc_smem(smem_a + (size_t)kb * FTC_TILE_BYTES));
        const uint64_t bd = ftc_desc(ftc_smem(smem_b + ((size_t)s * KB + kb) * FTC_TILE_BYTES));
#pragma unroll
        for (uint32_t k = 0; k < FTC_KBLK / 16; k++)  // +32 bytes (2 x 16 B units) per K = 16 step inside the swizzle row
          ftc_mma(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
      }
      ftc_commit(empty + s);     // the stage's shared memory may be refilled once these MMAs are done
      ftc_commit(tfull + acc);   // and the accumulator may be read
    }
  } else if (warp >= 4) {
    // ===== epilogue: 8 warps.  A warp may only read the TMEM lanes of its quarter (warp % 4): thread = one query;
    // the two warps of a quarter take 64 of the tile's 128 rows (TMEM columns) each.
    // keep row y  iff  NY (1 - e2) - 2 P - e1 |q| |y| <= T - NQ (1 - e2)   (the bound of api_column.inl, terms of the
    // row on the left, terms of the query on the right; every coefficient is enlarged by 2^-21 for the f32 roundings
    // of this form).  row_norm holds (NY (1 - e2'), |y|) and is padded with (+inf, 0) to whole tiles.
    const uint32_t quarter = warp & 3, half = (warp - 4) >> 2;
    const uint32_t lq = quarter * 32 + lane;
    const uint32_t qi = qblk * FTC_M + lq;
    const float2 qn = p.q_norm[qi];
    const float thr = p.q_thr[qi];
    // dot:     keep iff  -P - e1 |q||y| <= T - 1 + e0            (distance 1 - q.y, E = e1 |q||y| + e0; e0 in p.e2)
    // cosine:  keep iff  -P' <= (T - 1 + E) |q|                  (rows pre-normalised: P' ~ q.y / |y|; E in p.e1)
    // each right-hand side gets 2^-21 (|T| + 1) (|q| for cosine) of slack for its own roundings
    const float ne1q = -(p.e1 * qn.y);
    float cq;
    if (thr == __int_as_float(0xff800000)) cq = thr;
    else if (MET == 0) cq = fmaf(p.e2, qn.x, thr - qn.x) + 4.76837158203125e-07f * (fabsf(thr) + qn.x);
    else if (MET == 2) cq = (thr - 1.0f) + p.e2 + 4.76837158203125e-07f * (fabsf(thr) + 1.0f);
    else cq = ((thr - 1.0f) + p.e1 + 4.76837158203125e-07f * (fabsf(thr) + 1.0f)) * qn.y;
    float2* nb = norm_buf + (warp - 4) * 64;
    uint32_t row0 = blockIdx.x * FTC_N + half * 64;
    float2 n0 = make_float2(0.f, 0.f), n1 = n0;
    if (MET != 1 && my_tiles) { n0 = __ldg(p.row_norm + row0 + lane); n1 = __ldg(p.row_norm + row0 + 32 + lane); }
    for (uint32_t t = 0; t < my_tiles; t++) {
      const uint32_t acc = t & 1, aph = (t >> 1) & 1;
      __syncwarp();
      nb[lane] = n0;
      nb[32 + lane] = n1;
      __syncwarp();
      const uint32_t row_cur = row0;
      row0 += gridDim.x * FTC_N;
      if (MET != 1 && t + 1 < my_tiles) { n0 = __ldg(p.row_norm + row0 + lane); n1 = __ldg(p.row_norm + row0 + 32 + lane); }
      ftc_mbar_wait(tfull + acc, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
      for (uint32_t c0 = 0; c0 < 64; c0 += 32) {
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((quarter * 32u) << 16) + acc * FTC_N + half * 64 + c0;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
              "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
              "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const uint32_t nb_addr = ftc_smem(nb + c0);
        if (DUMP) {
          // upper bounds: L2  |q|^2 (1 + e2) + |y|^2 (1 + e2) - 2 P + e1 |q||y|;  dot  1 - P + e1 |q||y| + e0;
          // cosine  1 - P' / |q| + E
          if (qi < p.nq) {
            float* drow = p.dump + (size_t)qi * p.n + row_cur + c0;
            const float inv_q = MET == 1 ? __fdiv_rn(1.0f, qn.y) : 0.0f;
            const float base = MET == 0 ? fmaf(p.e2, qn.x, qn.x) : MET == 2 ? 1.0f + p.e2 : 1.0f + p.e1;
#pragma unroll
            for (int j = 0; j < 32; j++) {
              float2 yn = make_float2(0.f, 0.f);
              if (MET != 1) asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(yn.x), "=f"(yn.y) : "r"(nb_addr + j * 8));
              const float P = __uint_as_float(v[j]);
              float ub;
              if (MET == 0) ub = fmaf(-ne1q, yn.y, fmaf(-2.0f, P, yn.x)) + base;
              else if (MET == 2) ub = fmaf(-ne1q, yn.y, -P) + base;
              else ub = fmaf(-P, inv_q, base);
              if (row_cur + c0 + j < p.n) drow[j] = ub;
            }
          }
          continue;
        }
        // one bit per row that passes; a few rows of a thousand do, so nearly every WARP has some lane with a bit
        // set in a group of 32 rows -- the test must not branch per row
        uint32_t mask = 0;
#pragma unroll
        for (int j = 0; j < 32; j += 2) {   // one LDS.128 (a broadcast) brings the norms of two rows
          float4 y2 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (MET != 1)
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(y2.x), "=f"(y2.y), "=f"(y2.z), "=f"(y2.w) : "r"(nb_addr + j * 8));
          float g0, g1;
          if (MET == 0) {
            g0 = fmaf(ne1q, y2.y, fmaf(-2.0f, __uint_as_float(v[j]), y2.x));
            g1 = fmaf(ne1q, y2.w, fmaf(-2.0f, __uint_as_float(v[j + 1]), y2.z));
          } else if (MET == 2) {
            g0 = fmaf(ne1q, y2.y, -__uint_as_float(v[j]));
            g1 = fmaf(ne1q, y2.w, -__uint_as_float(v[j + 1]));
          } else {
            g0 = -__uint_as_float(v[j]);
            g1 = -__uint_as_float(v[j + 1]);
          }
          mask |= g0 <= cq ? 1u << j : 0u;
          mask |= g1 <= cq ? 2u << j : 0u;
        }
        while (mask) {
          const uint32_t j = __ffs(mask) - 1;
          mask &= mask - 1;
          if (MET != 0 && row_cur + c0 + j >= p.n) continue;   // zero-filled rows past the end (L2: excluded by their +inf norm)
          const uint32_t pos = atomicAdd(p.cand_cnt + qi, 1u);
          if (pos < p.cap) p.cand[(size_t)qi * p.cap + pos] = row_cur + c0 + j;
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) ftc_mbar_arrive(tempty + acc);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem_base) : "memory");
}

// f32 rows -> bf16 rows (round to nearest even) + (scale |y|^2, |y|) from the f32 values; rows n <= r < n_pad get
// (+inf, 0).  A row whose norm is not finite (NaN / inf values, overflow) raises *nonfinite: the caller then leaves the
// filter out, its bound assumes ordinary values.  One warp per row.
__global__ void to_bf16_norm_kernel(const float* __restrict__ x, uint64_t n, uint64_t n_pad, uint32_t dim, float scale,
                                    __nv_bfloat16* __restrict__ xb, float2* __restrict__ norms, uint32_t* __restrict__ nonfinite) {
  const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (r >= n_pad) return;
  if (r >= n) {
    if (lane == 0) norms[r] = make_float2(__int_as_float(0x7f800000), 0.0f);
    return;
  }
  float s = 0.0f;
  for (uint32_t j = lane; j < dim; j += 32) {
    const float v = x[r * dim + j];
    xb[r * dim + j] = __float2bfloat16_rn(v);
    s = fmaf(v, v, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) {
    norms[r] = make_float2(s * scale, sqrtf(s));
    if (!(s < __int_as_float(0x7f800000))) atomicExch(nonfinite, 1u);
  }
}

// cosine: f32 rows -> unit-length rows rounded to bf16 (y / |y| per element; a zero row becomes NaN and can never pass
// the filter, like its NaN cosine distance can never be returned).  One warp per row.
__global__ void to_bf16_unit_kernel(const float* __restrict__ x, uint64_t n, uint32_t dim, __nv_bfloat16* __restrict__ xb) {
  const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (r >= n) return;
  float s = 0.0f;
  for (uint32_t j = lane; j < dim; j += 32) {
    const float v = x[r * dim + j];
    s = fmaf(v, v, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float ny = sqrtf(s);
  for (uint32_t j = lane; j < dim; j += 32) xb[r * dim + j] = __float2bfloat16_rn(__fdiv_rn(x[r * dim + j], ny));
}

// kept rows as the u64 positions the generic exact kernels take (cosine / dot re-score); ~0 past the count
__global__ void ftc_widen_kernel(const uint32_t* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t Q, uint32_t cap,
                                 uint64_t* __restrict__ pos) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (uint64_t)Q * cap) return;
  const uint32_t q = (uint32_t)(i / cap), j = (uint32_t)(i - (uint64_t)q * cap);
  pos[i] = j < min(cnt[q], cap) ? (uint64_t)cand[i] : ~0ull;
}

// rows [i * stride] of x, i < ns -> contiguous sample
__global__ void gather_rows_kernel(const float* __restrict__ x, uint64_t stride_rows, uint32_t ns, uint32_t dim, float* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (uint64_t)ns * dim) return;
  const uint64_t r = i / dim;
  out[i] = x[r * stride_rows * dim + (i - r * dim)];
}

// per query: threshold T_q = the exact k-th smallest of its ns sample distances (radix select over the order-preserving
// u32 keys of dist.cuh, 8 bits per pass, values held in registers); padding queries get -inf.  Also clears the
// candidate counters.  One block of 256 threads per query, ns = 256 * PER.
template <int PER>
__global__ void __launch_bounds__(256)
ftc_kth_kernel(const float* __restrict__ D /*[Q][stride]*/, size_t stride, uint32_t Q, uint32_t k, float* __restrict__ thr,
               uint32_t* __restrict__ cnt) {
  __shared__ uint32_t hist[8][256];
  __shared__ uint32_t s_digit, s_k;
  const uint32_t q = blockIdx.x, t = threadIdx.x, w = t >> 5, lane = t & 31;
  if (t == 0) cnt[q] = 0;
  if (q >= Q) {
    if (t == 0) thr[q] = __int_as_float(0xff800000);
    return;
  }
  uint32_t key[PER];
#pragma unroll
  for (int i = 0; i < PER; i++) key[i] = ordered_key(D[(size_t)q * stride + (size_t)i * 256 + t]);
  uint32_t prefix = 0, pmask = 0, kk = k;
#pragma unroll 1
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (uint32_t i = t; i < 8 * 256; i += 256) (&hist[0][0])[i] = 0;
    __syncthreads();
    // runs of equal digits are counted in registers first: the top bits of a query's distances are nearly all alike
    uint32_t run_d = 0xffffffffu, run_c = 0;
#pragma unroll
    for (int i = 0; i < PER; i++) {
      if ((key[i] & pmask) == prefix) {
        const uint32_t dg = (key[i] >> shift) & 255u;
        if (dg == run_d) run_c++;
        else {
          if (run_c) atomicAdd(&hist[w][run_d], run_c);
          run_d = dg;
          run_c = 1;
        }
      }
    }
    if (run_c) atomicAdd(&hist[w][run_d], run_c);
    __syncthreads();
    uint32_t tot = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) tot += hist[j][t];
    __syncthreads();
    hist[0][t] = tot;
    __syncthreads();
    if (w == 0) {  // lane owns bins 8 lane .. 8 lane + 7
      uint32_t c8[8], sum = 0;
#pragma unroll
      for (int j = 0; j < 8; j++) { c8[j] = hist[0][lane * 8 + j]; sum += c8[j]; }
      uint32_t incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += v;
      }
      uint32_t before = incl - sum;
      if (before < kk && kk <= incl) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
          if (before < kk && kk <= before + c8[j]) { s_digit = lane * 8 + j; s_k = kk - before; }
          before += c8[j];
        }
      }
    }
    __syncthreads();
    prefix |= s_digit << shift;
    pmask |= 255u << shift;
    kk = s_k;
    __syncthreads();
  }
  if (t == 0) thr[q] = key_to_float(prefix);
}

// exact L2 of the kept rows: l2_scalar::<f32,f32,16> (distance/l2.rs:65-99) with its 16 lane sums on 16 THREADS -- thread
// i of a half-warp owns sums[i] (sub, mul, add per 16-element chunk, in chunk order), then the lanes are added in lane
// order.  The half-warp reads 64 contiguous bytes of the row per chunk.  dim % 16 == 0, dim <= 256.
// grid (blocks per query, Q), 256 threads = 16 half-warps striding over the query's kept rows.
__global__ void __launch_bounds__(256)
ftc_rescore_l2_kernel(const float* __restrict__ queries, const float* __restrict__ rows, uint32_t dim, const uint32_t* __restrict__ cand,
                      const uint32_t* __restrict__ cnt, uint32_t cap, float* __restrict__ exact /*[Q][cap]*/,
                      uint64_t* __restrict__ pos /*[Q][cap]*/) {
  const uint32_t q = blockIdx.y, hl = threadIdx.x & 15, hw = threadIdx.x >> 4;
  const uint32_t n = min(cnt[q], cap), nchunk = dim >> 4;
  float xq[16];
#pragma unroll
  for (int c = 0; c < 16; c++) xq[c] = (uint32_t)c < nchunk ? queries[(size_t)q * dim + c * 16 + hl] : 0.0f;
  for (uint32_t i0 = blockIdx.x * 16; i0 < n; i0 += gridDim.x * 16) {   // n is uniform over the block
    const uint32_t i = i0 + hw;
    const bool live = i < n;
    const uint32_t r = live ? cand[(size_t)q * cap + i] : 0;
    const float* y = rows + (size_t)r * dim + hl;
    float yv[16];
#pragma unroll
    for (int c = 0; c < 16; c++) yv[c] = live && (uint32_t)c < nchunk ? __ldg(y + c * 16) : 0.0f;
    float sum = 0.0f;
#pragma unroll
    for (int c = 0; c < 16; c++)
      if ((uint32_t)c < nchunk) {
        const float df = __fsub_rn(xq[c], yv[c]);
        sum = __fadd_rn(sum, __fmul_rn(df, df));
      }
    float tt = 0.0f;
#pragma unroll
    for (int l = 0; l < 16; l++) tt = __fadd_rn(tt, __shfl_sync(0xffffffffu, sum, l, 16));
    if (hl == 0 && live) {
      exact[(size_t)q * cap + i] = __fadd_rn(0.0f, tt);   // l2_scalar's `s + t` with an empty remainder
      pos[(size_t)q * cap + i] = (uint64_t)r;
    }
  }
}

// before the final (distance, row) sort: drop the re-scored rows that cannot be among the k smallest.  Radix select of
// the k-th smallest exact distance of the query (keys re-read from L2 per pass), then the rows with distance <= that
// (the k smallest and every tie of the k-th) are packed to the front of the query's list, in place through shared
// memory.  A query with more than PRUNE_MAX survivors (a sea of equal distances) keeps its whole list.
constexpr uint32_t PRUNE_MAX = 2048;
__global__ void __launch_bounds__(256)
ftc_prune_kernel(float* __restrict__ exact /*[Q][cap]*/, uint64_t* __restrict__ pos, uint32_t* __restrict__ cnt, uint32_t cap, uint32_t k) {
  __shared__ uint32_t hist[8][256];
  __shared__ uint32_t s_digit, s_k, s_m;
  __shared__ float keep_d[PRUNE_MAX];
  __shared__ uint64_t keep_p[PRUNE_MAX];
  const uint32_t q = blockIdx.x, t = threadIdx.x, w = t >> 5, lane = t & 31;
  const uint32_t n = cnt[q];
  if (n <= k || n <= 32) return;
  float* ex = exact + (size_t)q * cap;
  uint64_t* ps = pos + (size_t)q * cap;
  uint32_t prefix = 0, pmask = 0, kk = k;
#pragma unroll 1
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (uint32_t i = t; i < 8 * 256; i += 256) (&hist[0][0])[i] = 0;
    __syncthreads();
    uint32_t run_d = 0xffffffffu, run_c = 0;
    for (uint32_t i = t; i < n; i += 256) {
      const uint32_t key = ordered_key(ex[i]);
      if ((key & pmask) == prefix) {
        const uint32_t dg = (key >> shift) & 255u;
        if (dg == run_d) run_c++;
        else {
          if (run_c) atomicAdd(&hist[w][run_d], run_c);
          run_d = dg;
          run_c = 1;
        }
      }
    }
    if (run_c) atomicAdd(&hist[w][run_d], run_c);
    __syncthreads();
    uint32_t tot = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) tot += hist[j][t];
    __syncthreads();
    hist[0][t] = tot;
    __syncthreads();
    if (w == 0) {
      uint32_t c8[8], sum = 0;
#pragma unroll
      for (int j = 0; j < 8; j++) { c8[j] = hist[0][lane * 8 + j]; sum += c8[j]; }
      uint32_t incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += v;
      }
      uint32_t before = incl - sum;
      if (before < kk && kk <= incl) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
          if (before < kk && kk <= before + c8[j]) { s_digit = lane * 8 + j; s_k = kk - before; }
          before += c8[j];
        }
      }
    }
    __syncthreads();
    prefix |= s_digit << shift;
    pmask |= 255u << shift;
    kk = s_k;
    __syncthreads();
  }
  if (t == 0) s_m = 0;
  __syncthreads();
  for (uint32_t i = t; i < n; i += 256) {
    const float d = ex[i];
    if (ordered_key(d) <= prefix) {
      const uint32_t o = atomicAdd(&s_m, 1u);
      if (o < PRUNE_MAX) { keep_d[o] = d; keep_p[o] = ps[i]; }
    }
  }
  __syncthreads();
  const uint32_t m = s_m;
  if (m > PRUNE_MAX) return;
  for (uint32_t i = t; i < m; i += 256) { ex[i] = keep_d[i]; ps[i] = keep_p[i]; }
  if (t == 0) cnt[q] = m;
}

// clamp the candidate counts to cap and flag overflow
__global__ void ftc_clamp_kernel(uint32_t* __restrict__ cnt, uint32_t Q, uint32_t cap, uint32_t* __restrict__ raw_cnt,
                                 uint32_t* __restrict__ overflow) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  const uint32_t c = cnt[q];
  raw_cnt[q] = c;
  if (c > cap) atomicExch(overflow, 1u);
  cnt[q] = min(c, cap);
}

}  // namespace lb200

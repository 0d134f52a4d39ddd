// api_column.inl -- a vector column kept RESIDENT on the device for flat kNN (KNNVectorDistanceExec path, A9 / K7;
// included by api.cu).  lb200_flat_topk gets a fresh batch per call like compute_distance (flat.rs:28-86) and so
// re-uploads the vectors every time (r01: 512 MB of PCIe per call at cfg 2); a column handle uploads once, and it
// also holds what the tensor-core candidate filter needs (flat_tc.cuh): a bf16 copy of the rows, their norms, and
// a strided sample of the rows for the per-query thresholds.
//
// Results are bit-identical to lb200_flat_topk (the filter only decides WHICH rows the exact fp32 kernels score):
//   exact  R = fl32(sum (q_i - y_i)^2), any summation order          |R - D| <= (d + 4) u32 D,  u32 = 2^-24,  D = |q - y|^2
//   approx a = NQ + NY - 2 P,  P = tensor-core sum of bf16(q_i) bf16(y_i) accumulated in f32
//     |bf16(x) - x| <= 2^-9 |x|  ->  |sum bf16(q_i)bf16(y_i) - q.y| <= (2^-8 + 2^-18) sum |q_i y_i| <= (2^-8 + 2^-18) |q||y|
//     f32 accumulation inside the tensor core: <= d 2^-22 sum |bf16(q_i) bf16(y_i)| (a generous bound for a non-IEEE adder tree)
//     NQ, NY: f32 sums of squares, relative error <= (d + 1) u32; final f32 combination <= 2^-22 (NQ + NY)
//   => |a - R| <= 2 c |q||y| + ((3 d + 12) 2^-24 + 2^-22) (|q|^2 + |y|^2),  c = 2^-8 (1 + 2^-10) + d 2^-21   (using D <= 2(|q|^2 + |y|^2))
//   E = 1.02 x that (the norms used are themselves f32 values).  A row of the true top k has R <= R_k <= T_q (T_q is the
//   exact k-th distance over a subset of the rows), so a - E <= R <= T_q: it is kept.  Ties at the k-th distance are kept too.
// Not covered (the filter is skipped): validity bitmaps, distance bounds, metrics other than L2, dim % 64 != 0 or > 256,
// small columns.  Values are assumed to be of ordinary magnitude (no f32 overflow / underflow in the squares).

#include "flat_tc.cuh"

struct lb200_column {
  int device = 0;
  uint32_t dim = 0;
  uint64_t n = 0;
  float* d_v = nullptr;
  uint8_t* d_valid = nullptr;
  __nv_bfloat16* d_vb = nullptr;
  __nv_bfloat16* d_vb_unit = nullptr;   // cosine: unit-length rows in bf16, built on the first cosine call
  CUtensorMap map_v_unit;
  // the sample in the filter's own formats: thresholds come from upper bounds computed by the tensor-core kernel
  __nv_bfloat16 *d_sb = nullptr, *d_sb_unit = nullptr;
  float2* d_snorm = nullptr;            // (|y|^2 (1 + e2), |y|) of the sample rows
  CUtensorMap map_s, map_s_unit;
  float2* d_norm = nullptr;
  float* d_sample = nullptr;
  uint32_t ns = 0;
  bool tc_ok = false;
  CUtensorMap map_v;
  std::mutex mu;          // calls on one column serialise (shared scratch)
  Arena scratch;
  std::mutex io_mu;       // host-buffer calls: device copies of the queries / results
  Arena io;
  cudaStream_t stream = nullptr;
  int sm_count = 148;
  // last call
  uint32_t cap_boost = 1;
  uint32_t last_tc = 0, last_fallback = 0;
  uint64_t last_candidates = 0;
};

namespace {

typedef CUresult (*lb200_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                          const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                          CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static lb200_encode_tiled_fn encode_tiled_fn() {
  static lb200_encode_tiled_fn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess)
      p = nullptr;
    cudaGetLastError();
    return reinterpret_cast<lb200_encode_tiled_fn>(p);
  }();
  return fn;
}

// bf16 [rows][dim] row-major, boxes of [128 rows][64 elements], 128-byte swizzle, zero fill past the end
static int make_tile_map(CUtensorMap* map, const void* base, uint64_t rows, uint32_t dim) {
  lb200_encode_tiled_fn fn = encode_tiled_fn();
  if (!fn) return set_err(LB200_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t gdim[2] = {dim, rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)dim * 2};
  const cuuint32_t box[2] = {FTC_KBLK, FTC_N};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_err(LB200_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return 0;
}

constexpr uint32_t COL_SAMPLE = 16384;
constexpr uint32_t COL_QCHUNK = 2048;

// coefficients of E (header comment), each enlarged by 2^-21: the kernel evaluates the test as
// NY (1 - e2) - 2 P - e1 |q||y| <= T - NQ (1 - e2), four more f32 roundings than the proof's form
static void ftc_coeffs(uint32_t dim, float* e1, float* e2) {
  const double cc = 0.00390625 * (1.0 + 1.0 / 1024.0) + (double)dim / 2097152.0;
  *e1 = (float)(2.0 * cc * 1.02 + 1.0 / 2097152.0);
  *e2 = (float)((((3.0 * dim + 12.0) / 16777216.0) + 1.0 / 4194304.0) * 1.02 + 1.0 / 2097152.0);
}
// dot: exact R = fl(1 - fl_dot(q, y)), |R - (1 - q.y)| <= (d + 2) u |q||y| + u (u = 2^-24); approx 1 - P with
// |P - q.y| <= cc |q||y|  =>  E = e1 |q||y| + e0,  e1 = 1.02 (cc + (d + 2) u) + 2^-21 (roundings of the test), e0 = 2^-22
static void ftc_coeffs_dot(uint32_t dim, float* e1, float* e0) {
  const double cc = 0.00390625 * (1.0 + 1.0 / 1024.0) + (double)dim / 2097152.0;
  *e1 = (float)(1.02 * (cc + (dim + 2.0) / 16777216.0) + 1.0 / 2097152.0);
  *e0 = (float)(1.0 / 4194304.0);
}
// cosine: exact R = cosine_fast, |R - (1 - cos)| <= (2 d + 7) u; approx 1 - P' / |q| on rows normalised before the bf16
// rounding, |P' / |q| - cos| <= cc + (d + 7) u  =>  E = 1.02 (cc + (3 d + 14) u) + 2^-20, independent of the norms
static void ftc_coeffs_cos(uint32_t dim, float* e) {
  const double cc = 0.00390625 * (1.0 + 1.0 / 1024.0) + (double)dim / 2097152.0;
  *e = (float)(1.02 * (cc + (3.0 * dim + 14.0) / 16777216.0) + 1.0 / 1048576.0);
}

static int column_topk(lb200_column* c, int metric, const float* d_q, uint32_t Q, uint32_t k, bool has_lower, float lower, bool has_upper,
                       float upper, float* d_od, uint64_t* d_oi, uint32_t* d_oc, cudaStream_t st) {
  c->last_tc = 0;
  c->last_fallback = 0;
  c->last_candidates = 0;
  if (Q > COL_QCHUNK) {  // bounds the scratch (the sample distances are [Q][16384] f32)
    uint32_t tc = 1, fb = 0;
    uint64_t cand = 0;
    for (uint32_t q0 = 0; q0 < Q; q0 += COL_QCHUNK) {
      const uint32_t qn = std::min(COL_QCHUNK, Q - q0);
      int rc = column_topk(c, metric, d_q + (size_t)q0 * c->dim, qn, k, has_lower, lower, has_upper, upper, d_od + (size_t)q0 * k,
                           d_oi + (size_t)q0 * k, d_oc + q0, st);
      if (rc) return rc;
      tc &= c->last_tc; fb |= c->last_fallback; cand += c->last_candidates;
    }
    c->last_tc = tc; c->last_fallback = fb; c->last_candidates = cand;
    return 0;
  }
  const uint64_t n = c->n;
  const uint32_t dim = c->dim;
  // expected candidates per query ~ k n / ns (rows below the sample's k-th distance); 4x head-room
  const uint64_t expect = c->ns ? (uint64_t)k * n / c->ns : ~0ull;
  // thresholds: exact k-th distance over the sample for L2 (the slab kernel makes that 0.1 ms), k-th smallest UPPER
  // bound from the tensor-core kernel for cosine / dot (their exact kernels would cost 2 ms on the sample); the looser
  // threshold keeps about twice the rows, hence the larger candidate buffers.  c->cap_boost doubles after an overflow.
  const bool tc_thr = getenv("LB200_FLAT_TC_THR") ? atoi(getenv("LB200_FLAT_TC_THR")) != 0 : metric != LB200_L2;
  const uint64_t want_cap = ((tc_thr ? 8 : 4) * expect + 256) * c->cap_boost;
  uint32_t cap = 1024;
  while (cap < want_cap && cap < 16384) cap <<= 1;
  const bool use_tc = c->tc_ok && (metric == LB200_L2 || metric == LB200_COSINE || metric == LB200_DOT) && !has_lower && !has_upper &&
                      k <= c->ns && (tc_thr ? 8 : 4) * expect + 256 <= 16384 && Q > 0 && !getenv("LB200_FLAT_NO_TC");
  if (!use_tc)
    return flat_topk_device_impl(metric, dim, d_q, Q, c->d_v, c->d_valid, n, k, has_lower, lower, has_upper, upper, d_od, d_oi, d_oc, st);
  if (metric == LB200_COSINE && !c->d_vb_unit) {   // first cosine call: the unit-length bf16 copy (c->mu is held)
    cudaError_t e = cudaMalloc(&c->d_vb_unit, (size_t)n * dim * 2);
    if (e != cudaSuccess) return set_err(LB200_ENOMEM, "cosine copy of the column: %s", cudaGetErrorString(e));
    to_bf16_unit_kernel<<<(uint32_t)((n * 32 + 255) / 256), 256, 0, st>>>(c->d_v, n, dim, c->d_vb_unit);
    CK(cudaGetLastError());
    int rcm = make_tile_map(&c->map_v_unit, c->d_vb_unit, n, dim);
    if (rcm) return rcm;
    e = cudaMalloc(&c->d_sb_unit, (size_t)c->ns * dim * 2);
    if (e != cudaSuccess) return set_err(LB200_ENOMEM, "cosine copy of the sample: %s", cudaGetErrorString(e));
    to_bf16_unit_kernel<<<(uint32_t)(((uint64_t)c->ns * 32 + 255) / 256), 256, 0, st>>>(c->d_sample, c->ns, dim, c->d_sb_unit);
    CK(cudaGetLastError());
    if ((rcm = make_tile_map(&c->map_s_unit, c->d_sb_unit, c->ns, dim))) return rcm;
  }
  const uint32_t Qpad = (Q + FTC_M - 1) / FTC_M * FTC_M;
  const uint32_t ns = c->ns;
  size_t need = 0;
  auto add = [&](size_t b) { need += Arena::pad(b) + 256; };
  add((size_t)Qpad * dim * 2); add((size_t)Qpad * 8); add((size_t)Qpad * 4); add((size_t)Qpad * cap * 4); add((size_t)Qpad * 4);
  add((size_t)Q * cap * 8); add((size_t)Q * cap * 4); add(64); add((size_t)Q * ns * 4); add((size_t)Q * 4);
  int rc = c->scratch.reserve(need);
  if (rc) return rc;
  c->scratch.reset();
  __nv_bfloat16* d_qb = c->scratch.take<__nv_bfloat16>((size_t)Qpad * dim);
  float2* d_qn = c->scratch.take<float2>(Qpad);
  float* d_thr = c->scratch.take<float>(Qpad);
  uint32_t* d_cand = c->scratch.take<uint32_t>((size_t)Qpad * cap);
  uint32_t* d_cnt = c->scratch.take<uint32_t>(Qpad);
  uint64_t* d_pos = c->scratch.take<uint64_t>((size_t)Q * cap);
  float* d_exact = c->scratch.take<float>((size_t)Q * cap);
  uint32_t* d_flag = c->scratch.take<uint32_t>(16);
  float* d_sd = c->scratch.take<float>((size_t)Q * ns);
  uint32_t* d_raw = c->scratch.take<uint32_t>(Q);
  // (1) queries -> bf16 + norms (padding rows zero)
  CK(cudaMemsetAsync(d_qb, 0, (size_t)Qpad * dim * 2, st));
  CK(cudaMemsetAsync(d_qn, 0, (size_t)Qpad * 8, st));
  CK(cudaMemsetAsync(d_flag, 0, 64, st));
  to_bf16_norm_kernel<<<(uint32_t)(((uint64_t)Q * 32 + 255) / 256), 256, 0, st>>>(d_q, Q, Q, dim, 1.0f, d_qb, d_qn, d_flag);  // a non-finite query -> exact path
  CK(cudaGetLastError());
  CUtensorMap map_q;
  if ((rc = make_tile_map(&map_q, d_qb, Qpad, dim))) return rc;
  FlatTcParams p = {};
  p.q_norm = d_qn; p.q_thr = d_thr; p.cand = d_cand; p.cand_cnt = d_cnt; p.cap = cap; p.kb = dim / FTC_KBLK;
  const size_t blk = (size_t)p.kb * FTC_TILE_BYTES;
  p.stages = (uint32_t)std::min<size_t>(8, (200 * 1024 - blk) / blk);
  const size_t smem = 1024 + blk + (size_t)p.stages * blk + 256 + 8 * 64 * sizeof(float2);
  const uint32_t qblocks = Qpad / FTC_M;
  if (metric == LB200_L2) ftc_coeffs(dim, &p.e1, &p.e2);
  else if (metric == LB200_DOT) ftc_coeffs_dot(dim, &p.e1, &p.e2);
  else ftc_coeffs_cos(dim, &p.e1);
#define LB200_FTC_LAUNCH(MET, DUMPF, MAPV)                                                                                  \
  do {                                                                                                                      \
    CK(cudaFuncSetAttribute(flat_tc_filter_kernel<MET, DUMPF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));    \
    flat_tc_filter_kernel<MET, DUMPF><<<grid, FTC_THREADS, smem, st>>>(map_q, MAPV, p);                                     \
  } while (0)
  // (2) thresholds T_q >= the true k-th distance: the k-th smallest over the sample of
  //     (a) an UPPER bound of each exact distance, approx + E, from the tensor-core kernel itself (default), or
  //     (b) the exact distances (L2; LB200_FLAT_TC_THR=0 forces it for cosine / dot, which then spend 2 ms here)
  if (tc_thr) {
    p.row_norm = c->d_snorm; p.n = ns; p.ntiles = ns / FTC_N; p.dump = d_sd; p.nq = Q;
    dim3 grid(std::max(1u, std::min(p.ntiles, (uint32_t)c->sm_count / std::min(qblocks, (uint32_t)c->sm_count))), qblocks);
    if (metric == LB200_L2) LB200_FTC_LAUNCH(0, true, c->map_s);
    else if (metric == LB200_DOT) LB200_FTC_LAUNCH(2, true, c->map_s);
    else LB200_FTC_LAUNCH(1, true, c->map_s_unit);
    CK(cudaGetLastError());
  } else if ((rc = flat_dist_launch(metric, d_q, Q, c->d_sample, nullptr, ns, dim, d_sd, ns, st))) {
    return rc;
  }
  static_assert(COL_SAMPLE == 256 * 64 && COL_SAMPLE % FTC_N == 0, "ftc_kth_kernel<64> holds the sample in 64 registers per thread");
  ftc_kth_kernel<64><<<Qpad, 256, 0, st>>>(d_sd, ns, Q, k, d_thr, d_cnt);
  // (3) tensor-core filter over the whole column
  {
    p.row_norm = c->d_norm; p.n = (uint32_t)n; p.ntiles = (uint32_t)((n + FTC_N - 1) / FTC_N); p.dump = nullptr; p.nq = Q;
    dim3 grid(std::max(1u, std::min(p.ntiles, (uint32_t)c->sm_count / std::min(qblocks, (uint32_t)c->sm_count))), qblocks);
    if (metric == LB200_L2) LB200_FTC_LAUNCH(0, false, c->map_v);
    else if (metric == LB200_DOT) LB200_FTC_LAUNCH(2, false, c->map_v);
    else LB200_FTC_LAUNCH(1, false, c->map_v_unit);
    CK(cudaGetLastError());
  }
#undef LB200_FTC_LAUNCH
  // (4) exact re-score of the kept rows, top k by (distance, row index)
  ftc_clamp_kernel<<<(Q + 255) / 256, 256, 0, st>>>(d_cnt, Q, cap, d_raw, d_flag);
  std::vector<uint32_t> h_cnt(Q);
  uint32_t h_flag = 0;
  CK(cudaMemcpyAsync(h_cnt.data(), d_raw, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&h_flag, d_flag, 4, cudaMemcpyDeviceToHost, st));
  dim3 rgrid(std::max(1u, std::min(cap / 16, (uint32_t)(8 * c->sm_count + Q - 1) / Q)), Q);
  if (metric == LB200_L2) {
    ftc_rescore_l2_kernel<<<rgrid, 256, 0, st>>>(d_q, c->d_v, dim, d_cand, d_cnt, cap, d_exact, d_pos);
  } else {   // cosine / dot: the generic exact kernel of the refine step (cosine_fast, dot_distance; dist.cuh)
    ftc_widen_kernel<<<(uint32_t)(((uint64_t)Q * cap + 255) / 256), 256, 0, st>>>(d_cand, d_cnt, Q, cap, d_pos);
    refine_distance_kernel<<<(uint32_t)(((size_t)Q * cap + 127) / 128), 128, 0, st>>>(metric, d_q, c->d_v, dim, d_pos, d_cnt, Q, cap, d_exact);
  }
  ftc_prune_kernel<<<Q, 256, 0, st>>>(d_exact, d_pos, d_cnt, cap, k);
  CK(cudaGetLastError());
  rc = launch_merge(d_exact, d_pos, nullptr, nullptr, d_cnt, Q, 1, cap, cap, cap, 1, 1, k, k, d_od, d_oi, nullptr, nullptr, d_oc, st);
  if (rc) return rc;
  CK(cudaStreamSynchronize(st));
  for (uint32_t q = 0; q < Q; q++) c->last_candidates += h_cnt[q];
  c->last_tc = 1;
  if (h_flag) {  // some query kept more rows than the buffer holds: the exact path answers the whole batch
    c->last_fallback = 1;
    if (c->cap_boost < 8) c->cap_boost *= 2;   // and the next call gets larger buffers
    return flat_topk_device_impl(metric, dim, d_q, Q, c->d_v, c->d_valid, n, k, false, 0.f, false, 0.f, d_od, d_oi, d_oc, st);
  }
  return 0;
}

}  // namespace

extern "C" void lb200_column_destroy(lb200_column* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  cudaFree(c->d_v);
  cudaFree(c->d_valid);
  cudaFree(c->d_vb);
  cudaFree(c->d_vb_unit);
  cudaFree(c->d_sb);
  cudaFree(c->d_sb_unit);
  cudaFree(c->d_snorm);
  cudaFree(c->d_norm);
  cudaFree(c->d_sample);
  c->scratch.release();
  c->io.release();
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

extern "C" int32_t lb200_column_create(int32_t device_id, uint32_t dim, const float* vectors, const uint8_t* validity_or_null, uint64_t n,
                                       lb200_column** out) {
  if (!out) return set_err(LB200_EINVAL, "out is null");
  *out = nullptr;
  if (!dim) return set_err(LB200_EINVAL, "dim must be > 0");
  if (n && !vectors) return set_err(LB200_EINVAL, "null vectors");
  if (n >= (1ull << 32)) return set_err(LB200_EUNSUPPORTED, "n >= 2^32 rows");
  int rc = select_device(device_id);
  if (rc) return rc;
  lb200_column* c = new lb200_column();
  c->device = device_id;
  c->dim = dim;
  c->n = n;
  auto fail = [&](int code) { lb200_column_destroy(c); return code; };
#define CKC(call)                                                                                                   \
  do {                                                                                                              \
    cudaError_t e_ = (call);                                                                                        \
    if (e_ != cudaSuccess)                                                                                          \
      return fail(set_err(e_ == cudaErrorMemoryAllocation ? LB200_ENOMEM : LB200_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e_))); \
  } while (0)
  cudaDeviceProp prop;
  CKC(cudaGetDeviceProperties(&prop, device_id));
  c->sm_count = prop.multiProcessorCount;
  CKC(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CKC(cudaMalloc(&c->d_v, std::max<size_t>((size_t)n * dim * 4, 16)));
  if (n) CKC(cudaMemcpyAsync(c->d_v, vectors, (size_t)n * dim * 4, cudaMemcpyHostToDevice, c->stream));
  if (validity_or_null) {
    CKC(cudaMalloc(&c->d_valid, (size_t)(n + 7) / 8 + 1));
    CKC(cudaMemcpyAsync(c->d_valid, validity_or_null, (size_t)(n + 7) / 8, cudaMemcpyHostToDevice, c->stream));
  }
  // what the tensor-core filter needs; only for shapes it serves
  if (!validity_or_null && dim % FTC_KBLK == 0 && dim <= 256 && n >= 4 * (uint64_t)COL_SAMPLE && prop.major == 10 && encode_tiled_fn()) {
    CKC(cudaMalloc(&c->d_vb, (size_t)n * dim * 2));
    const uint64_t n_pad = (n + FTC_N - 1) / FTC_N * FTC_N;
    float e1, e2;
    ftc_coeffs(dim, &e1, &e2);
    uint32_t* d_bad = nullptr;
    uint32_t h_bad = 0;
    CKC(cudaMalloc(&c->d_norm, (size_t)n_pad * 8));
    CKC(cudaMalloc(&d_bad, 4));
    CKC(cudaMemsetAsync(d_bad, 0, 4, c->stream));
    c->ns = COL_SAMPLE;
    CKC(cudaMalloc(&c->d_sample, (size_t)c->ns * dim * 4));
    to_bf16_norm_kernel<<<(uint32_t)((n_pad * 32 + 255) / 256), 256, 0, c->stream>>>(c->d_v, n, n_pad, dim, 1.0f - e2, c->d_vb, c->d_norm, d_bad);
    gather_rows_kernel<<<(uint32_t)(((uint64_t)c->ns * dim + 255) / 256), 256, 0, c->stream>>>(c->d_v, n / c->ns, c->ns, dim, c->d_sample);
    CKC(cudaGetLastError());
    CKC(cudaMalloc(&c->d_sb, (size_t)c->ns * dim * 2));
    CKC(cudaMalloc(&c->d_snorm, (size_t)c->ns * 8));
    to_bf16_norm_kernel<<<(uint32_t)(((uint64_t)c->ns * 32 + 255) / 256), 256, 0, c->stream>>>(c->d_sample, c->ns, c->ns, dim, 1.0f + e2, c->d_sb,
                                                                                           c->d_snorm, d_bad);
    CKC(cudaGetLastError());
    if ((rc = make_tile_map(&c->map_s, c->d_sb, c->ns, dim))) return fail(rc);
    cudaError_t eb = cudaMemcpyAsync(&h_bad, d_bad, 4, cudaMemcpyDeviceToHost, c->stream);
    if (eb == cudaSuccess) eb = cudaStreamSynchronize(c->stream);
    cudaFree(d_bad);
    CKC(eb);
    if ((rc = make_tile_map(&c->map_v, c->d_vb, n, dim))) return fail(rc);
    c->tc_ok = h_bad == 0;   // NaN / inf values (or squares that overflow): the bound does not hold, exact path only
  }
  CKC(cudaStreamSynchronize(c->stream));
#undef CKC
  *out = c;
  return 0;
}

extern "C" int32_t lb200_column_flat_topk_device(lb200_column* c, int32_t metric, const float* d_queries, uint32_t Q, uint32_t k,
                                                 const float* lower_or_null, const float* upper_or_null, float* d_out_dist,
                                                 uint64_t* d_out_idx, uint32_t* d_out_count, void* stream) {
  if (!c) return set_err(LB200_EINVAL, "null column");
  int rc = flat_check(metric, c->dim, d_queries, Q, c->d_v, c->n);
  if (rc) return rc;
  if (Q && (!d_out_dist || !d_out_idx || !d_out_count)) return set_err(LB200_EINVAL, "null output");
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (k > 1024) return set_err(LB200_EUNSUPPORTED, "k=%u > 1024", k);
  if (!Q) return 0;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  return column_topk(c, metric, d_queries, Q, k, lower_or_null != nullptr, lower_or_null ? *lower_or_null : 0.f, upper_or_null != nullptr,
                     upper_or_null ? *upper_or_null : 0.f, d_out_dist, d_out_idx, d_out_count, stream ? (cudaStream_t)stream : c->stream);
}

extern "C" int32_t lb200_column_flat_topk(lb200_column* c, int32_t metric, const float* queries, uint32_t Q, uint32_t k,
                                          const float* lower_or_null, const float* upper_or_null, float* out_dist, uint64_t* out_idx,
                                          uint32_t* out_count) {
  if (!c) return set_err(LB200_EINVAL, "null column");
  int rc = flat_check(metric, c->dim, queries, Q, c->d_v, c->n);
  if (rc) return rc;
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (!Q) return 0;
  if (!out_dist || !out_idx || !out_count) return set_err(LB200_EINVAL, "null output");
  CK(cudaSetDevice(c->device));
  std::lock_guard<std::mutex> lk(c->io_mu);
  const size_t bq = Arena::pad((size_t)Q * c->dim * 4), bod = Arena::pad((size_t)Q * k * 4), boi = Arena::pad((size_t)Q * k * 8);
  if ((rc = c->io.reserve(bq + bod + boi + Arena::pad((size_t)Q * 4) + 1024))) return rc;
  c->io.reset();
  float* d_q = c->io.take<float>((size_t)Q * c->dim);
  float* d_od = c->io.take<float>((size_t)Q * k);
  uint64_t* d_oi = c->io.take<uint64_t>((size_t)Q * k);
  uint32_t* d_oc = c->io.take<uint32_t>(Q);
  CK(cudaMemcpy(d_q, queries, (size_t)Q * c->dim * 4, cudaMemcpyHostToDevice));
  rc = lb200_column_flat_topk_device(c, metric, d_q, Q, k, lower_or_null, upper_or_null, d_od, d_oi, d_oc, nullptr);
  if (rc) return rc;
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaMemcpy(out_dist, d_od, (size_t)Q * k * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(out_idx, d_oi, (size_t)Q * k * 8, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(out_count, d_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost));
  return 0;
}

// what the last call on this column did: used_tc = the tensor-core filter ran, candidates = rows it kept (sum over the
// queries), fell_back = a candidate buffer overflowed and the exact path answered instead
extern "C" int32_t lb200_column_last_call(const lb200_column* c, uint32_t* used_tc, uint64_t* candidates, uint32_t* fell_back) {
  if (!c) return set_err(LB200_EINVAL, "null column");
  if (used_tc) *used_tc = c->last_tc;
  if (candidates) *candidates = c->last_candidates;
  if (fell_back) *fell_back = c->last_fallback;
  return 0;
}

// api_pq4.inl -- 4-bit PQ indexes and IVF_FLAT indexes (bits == 32: raw f32 rows instead of codes) behind the same
// entry points (included by api.cu): create, and the search pipeline search_pipeline() hands over to when
// ix->bits != 8.

namespace {

static int search4_pipeline(lb200_index* ix, CallCtx& cx, const SearchArgs& a, cudaStream_t st) {
  const uint32_t Q = a.Q, dim = ix->dim, nlist = ix->nlist, M = ix->M;
  const uint32_t P = a.d_part_ids ? a.P : std::min(a.nprobes, nlist);
  const uint32_t kprime = a.k * std::max(a.refine_factor, 1u);
  if (a.k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (kprime > 512) return set_err(LB200_EUNSUPPORTED, "k*refine_factor=%u > 512", kprime);
  const bool flat = ix->bits == 32;
  if (P == 0 || Q == 0) return 0;
  const bool refine = a.refine_factor > 0 && ix->d_raw != nullptr && !a.per_pair_output;
  if (a.refine_factor > 0 && !ix->d_raw && !a.per_pair_output && !(a.shard_mode && ix->N == 0))
    return set_err(LB200_EINVAL, "refine_factor>0 needs raw vectors in the index handle");
  if (!a.per_pair_output && (uint64_t)P * kprime > 16384)
    return set_err(LB200_EUNSUPPORTED, "4-bit PQ: nprobes*k'=%llu > 16384", (unsigned long long)P * kprime);
  const size_t stride = (size_t)std::max<uint64_t>(ix->max_part, 1);
  // queries per tile: the [pairs][stride] distance scratch stays below 512 MB
  const uint32_t tileQ = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(Q, (512ull << 20) / (stride * 4 * P)));
  const size_t tpairs = (size_t)tileQ * P, npairs = (size_t)Q * P;
  // fused histogram-select kernel (scan4.cuh) unless its shared memory does not fit or LB200_PQ4_SIMPLE=1 asks for
  // the distance-scratch path (kept as the A/B check of the fused select)
  const uint32_t kp2 = std::max(2u, next_pow2(kprime));
  const uint32_t nbins = (M / 2) * 510 + 1;
  const size_t fsmem = ((size_t)dim + (size_t)M * 16 + ((M + 1) & ~1u)) * 4 + (size_t)kp2 * 8 + (size_t)((nbins + 1) & ~1u) * 4 + (size_t)M * 16 +
                       (size_t)(M / 2) * 256 + 4096 * 2;
  // (a prefilter switches the reference to per-row distances on the unquantised table: scan4_kernel's masked path)
  const bool fused = !flat && fsmem <= ix->smem_optin && !getenv("LB200_PQ4_SIMPLE") && !a.d_bitmask;
  size_t need = 0;
  auto add = [&](size_t b) { need += Arena::pad(b) + 256; };
  add((size_t)Q * dim * 4); add((size_t)Q * nlist * 4); add(npairs * 4); add(64);
  add(fused ? 256 : tpairs * stride * 4);
  add(npairs * kprime * 4); add(npairs * kprime * 8); add(npairs * kprime * 8); add(npairs * kprime * 8); add(npairs * 4);
  add((size_t)Q * kprime * 4); add((size_t)Q * kprime * 8); add((size_t)Q * kprime * 8); add((size_t)Q * 4);
  add((size_t)Q * kprime * 4);
  // grouped IVF_FLAT L2 scan (several queries share a partition's rows): work list of one query tile
  const bool grouped = flat && ix->metric != LB200_DOT && ix->metric != LB200_COSINE && !getenv("LB200_FLAT_SIMPLE") &&
                       !getenv("LB200_FLAT_PAIRWISE");
  if (grouped) {
    add((size_t)nlist * 8); add((size_t)tileQ * 4); add(tpairs * 4); add((size_t)(nlist + 1) * 8); add(tpairs * 4);
    add(tpairs * 16); add(64);
  }
  int rc = cx.arena.reserve(need);
  if (rc) return rc;
  cx.arena.reset();
  float* d_keys = cx.arena.take<float>((size_t)Q * dim);
  float* d_D = cx.arena.take<float>((size_t)Q * nlist);
  uint32_t* d_probes = cx.arena.take<uint32_t>(npairs);
  unsigned long long* d_stats = cx.arena.take<unsigned long long>(8);
  float* d_dist = cx.arena.take<float>(fused ? 64 : tpairs * stride);
  float* d_sdist = cx.arena.take<float>(npairs * kprime);
  uint64_t* d_sidx = cx.arena.take<uint64_t>(npairs * kprime);
  uint64_t* d_srow = cx.arena.take<uint64_t>(npairs * kprime);
  uint64_t* d_spos = cx.arena.take<uint64_t>(npairs * kprime);
  uint32_t* d_scnt = cx.arena.take<uint32_t>(npairs);
  float* d_mdist = cx.arena.take<float>((size_t)Q * kprime);
  uint64_t* d_mrow = cx.arena.take<uint64_t>((size_t)Q * kprime);
  uint64_t* d_mpos = cx.arena.take<uint64_t>((size_t)Q * kprime);
  uint32_t* d_mcnt = cx.arena.take<uint32_t>(Q);
  float* d_exact = cx.arena.take<float>((size_t)Q * kprime);
  uint32_t *g_cnt = nullptr, *g_minp = nullptr, *g_rank = nullptr, *g_pair_off = nullptr, *g_item_off = nullptr, *g_pair_list = nullptr;
  uint4* g_desc = nullptr;
  unsigned long long* g_stats = nullptr;
  if (grouped) {
    g_cnt = cx.arena.take<uint32_t>((size_t)nlist * 2);
    g_minp = cx.arena.take<uint32_t>(tileQ);
    g_rank = cx.arena.take<uint32_t>(tpairs);
    g_pair_off = cx.arena.take<uint32_t>(nlist + 1);
    g_item_off = cx.arena.take<uint32_t>(nlist + 1);
    g_pair_list = cx.arena.take<uint32_t>(tpairs);
    g_desc = cx.arena.take<uint4>(tpairs);
    g_stats = cx.arena.take<unsigned long long>(8);
  }

  uint32_t launches = 0;
  lb200_stats stats = {};
  const bool prof = ix->profiling;
  if (prof) CK(cudaEventRecord(cx.ev[0], st));
  const float* keys = a.d_queries;
  if (a.normalize && ix->metric == LB200_COSINE) {
    normalize_kernel<<<(Q * 32 + 255) / 256, 256, 0, st>>>(a.d_queries, d_keys, Q, dim);
    CK(cudaGetLastError());
    launches++;
    keys = d_keys;
  }
  const uint32_t* probes = a.d_part_ids;
  if (!probes) {
    rc = coarse_device(ix, keys, Q, P, d_D, d_probes, nullptr, st, &launches);
    if (rc) return rc;
    probes = d_probes;
  }
  if (prof) CK(cudaEventRecord(cx.ev[1], st));
  CK(cudaMemsetAsync(d_stats, 0, 32, st));
  const size_t smem = ((size_t)dim + (size_t)M * 16 + M) * 4 + (size_t)M * 16;
  if (smem > 48 * 1024) CK(cudaFuncSetAttribute(scan4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // the per-pair cut: FlatIndex::search keeps [lower, upper) and the first k' (flat/index.rs:97-126)
  float* pd = a.per_pair_output ? a.d_out_dist : d_sdist;
  uint32_t* pc = a.per_pair_output ? a.d_out_count : d_scnt;
  uint64_t* prow = a.per_pair_output ? a.d_out_rowid : d_srow;
  if (prof) CK(cudaEventRecord(cx.ev[2], st));
  if (fused) {
    if (fsmem > 48 * 1024) CK(cudaFuncSetAttribute(scan4f_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
    scan4f_kernel<<<(uint32_t)npairs, 256, fsmem, st>>>(ix->metric, keys, probes, P, dim, M, ix->d_centroids,
                                                        ix->d_codebook, ix->d_part_off, ix->d_pad_off, ix->d_codes4,
                                                        ix->d_tail4, ix->d_row_ids, kprime, kp2, a.has_lower, a.lower,
                                                        a.has_upper, a.upper, pd, prow,
                                                        a.per_pair_output ? nullptr : d_spos, pc, d_stats);
    CK(cudaGetLastError());
    launches++;
  }
  for (uint32_t q0 = 0; q0 < Q && !fused; q0 += tileQ) {
    const uint32_t qn = std::min(tileQ, Q - q0);
    const size_t p0 = (size_t)q0 * P, pn = (size_t)qn * P;
    if (flat) {
      const size_t fs = (size_t)dim * 4;
      if (fs > 48 * 1024) CK(cudaFuncSetAttribute(scanflat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs));
      if (grouped) {
        // work list of the tile: pairs grouped by partition, items = (partition, group of FLATG_QL queries)
        CK(cudaMemsetAsync(g_cnt, 0, (size_t)nlist * 8, st));
        CK(cudaMemsetAsync(g_minp, 0xff, (size_t)qn * 4, st));   // no cold packing here
        count_pairs_kernel<<<(uint32_t)((pn + 255) / 256), 256, 0, st>>>(probes + p0, (uint32_t)pn, P, g_minp, g_cnt, g_cnt + nlist, g_rank);
        offsets_kernel<<<1, 1024, 0, st>>>(g_cnt, g_cnt + nlist, ix->d_part_off, nlist, (uint32_t)FLATG_QL,
                                           (uint32_t)std::max<uint64_t>(ix->max_part, 1), dim * 4, g_pair_off, g_item_off, g_stats);
        scatter_pairs_kernel<<<(uint32_t)((pn + 255) / 256), 256, 0, st>>>(probes + p0, g_rank, g_pair_off, g_cnt, (uint32_t)pn, g_pair_list);
        expand_items_kernel<<<(nlist * 32 + 255) / 256, 256, 0, st>>>(g_item_off, ix->d_part_off, nlist,
                                                                      (uint32_t)std::max<uint64_t>(ix->max_part, 1), g_desc, nullptr);
        flat_empty_pairs_kernel<<<(uint32_t)pn, 128, 0, st>>>(probes + p0, (uint32_t)pn, ix->d_part_off, d_dist, stride);
        const int nch = dim % 16 == 0 && (dim == 64 || dim == 128 || dim == 256) ? (int)(dim / 16) : 0;
        const size_t gs = ((size_t)16 * FLATG_QL * 16 * 17 + (nch ? 0 : (size_t)FLATG_QL * dim)) * 4;
        const uint32_t gg = (uint32_t)std::min<size_t>(pn, (size_t)ix->sm_count * 3);
#define LB200_FLATG(N)                                                                                                     \
  do {                                                                                                                     \
    CK(cudaFuncSetAttribute(scanflat_l2_group_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gs));           \
    scanflat_l2_group_kernel<N><<<gg, 256, gs, st>>>(keys + (size_t)q0 * dim, P, dim, ix->d_part_off, ix->d_raw, g_desc,   \
                                                     g_item_off + nlist, g_pair_list, g_pair_off, d_dist, stride, d_stats, \
                                                     a.d_bitmask);                                                         \
  } while (0)
        if (nch == 4) LB200_FLATG(4);
        else if (nch == 8) LB200_FLATG(8);
        else if (nch == 16) LB200_FLATG(16);
        else LB200_FLATG(0);
#undef LB200_FLATG
        launches += 5;
      } else if (ix->metric == LB200_L2 && !getenv("LB200_FLAT_SIMPLE")) {
        if (fs > 48 * 1024) CK(cudaFuncSetAttribute(scanflat_l2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs));
        scanflat_l2_kernel<<<(uint32_t)pn, 256, fs, st>>>(keys + (size_t)q0 * dim, probes + p0, P, dim, ix->d_part_off,
                                                           ix->d_raw, d_dist, stride, d_stats, a.d_bitmask);
      } else
      scanflat_kernel<<<(uint32_t)pn, 256, fs, st>>>(ix->metric, keys + (size_t)q0 * dim, probes + p0, P, dim,
                                                      ix->d_part_off, ix->d_raw, d_dist, stride, d_stats, a.d_bitmask);
    } else {
      scan4_kernel<<<(uint32_t)pn, 256, smem, st>>>(ix->metric, keys + (size_t)q0 * dim, probes + p0, P, dim, M,
                                                     ix->d_centroids, ix->d_codebook, ix->d_part_off, ix->d_pad_off,
                                                     ix->d_codes4, ix->d_tail4, d_dist, stride, d_stats, a.d_bitmask);
    }
    CK(cudaGetLastError());
    flat_select_kernel<<<(uint32_t)pn, 256, (size_t)kp2 * 8, st>>>(d_dist, stride, stride, 1, stride, kprime, kp2,
                                                                   a.has_lower, a.lower, a.has_upper, a.upper,
                                                                   pd + p0 * kprime, d_sidx + p0 * kprime, pc + p0);
    CK(cudaGetLastError());
    launches += 2;
  }
  if (prof) CK(cudaEventRecord(cx.ev[3], st));
  if (!fused) {
    gather4_kernel<<<(uint32_t)((npairs * kprime + 255) / 256), 256, 0, st>>>(
        probes, ix->d_part_off, ix->d_row_ids, d_sidx, pc, npairs, kprime, prow, a.per_pair_output ? nullptr : d_spos);
    CK(cudaGetLastError());
    launches++;
  }
  if (!a.per_pair_output) {
    const bool direct = !refine && !a.shard_mode;
    rc = launch_merge(d_sdist, d_srow, d_spos, nullptr, d_scnt, Q, P, kprime, (size_t)P * kprime, kprime, P, 1, kprime,
                      kprime, direct ? a.d_out_dist : d_mdist, direct ? a.d_out_rowid : d_mrow, d_mpos, nullptr,
                      direct ? a.d_out_count : d_mcnt, st);
    if (rc) return rc;
    launches++;
    if (a.shard_mode) {
      // a shard's answer: its merged k' list before the final cut, plus the exact distances when it refines
      // (same contract as the 8-bit pipeline; the outputs may be peer memory of the root shard)
      if (refine) {
        refine_distance_kernel<<<(uint32_t)(((size_t)Q * kprime + 127) / 128), 128, 0, st>>>(
            ix->metric, a.d_queries, ix->d_raw, dim, d_mpos, d_mcnt, Q, kprime, a.d_out_exact);
        CK(cudaGetLastError());
        launches++;
      } else {
        CK(cudaMemsetAsync(a.d_out_exact, 0xff, (size_t)Q * kprime * 4, st));  // NaN
      }
      CK(cudaMemcpyAsync(a.d_out_dist, d_mdist, (size_t)Q * kprime * 4, cudaMemcpyDefault, st));
      CK(cudaMemcpyAsync(a.d_out_rowid, d_mrow, (size_t)Q * kprime * 8, cudaMemcpyDefault, st));
      CK(cudaMemcpyAsync(a.d_out_count, d_mcnt, (size_t)Q * 4, cudaMemcpyDefault, st));
    } else if (refine) {
      refine_distance_kernel<<<(uint32_t)(((size_t)Q * kprime + 127) / 128), 128, 0, st>>>(
          ix->metric, a.d_queries, ix->d_raw, dim, d_mpos, d_mcnt, Q, kprime, d_exact);
      CK(cudaGetLastError());
      rerank_kernel<<<(Q * 32 + 127) / 128, 128, 0, st>>>(d_exact, d_mrow, d_mcnt, Q, kprime, a.k, a.has_lower, a.lower,
                                                        a.has_upper, a.upper, a.d_out_dist, a.d_out_rowid,
                                                        a.d_out_count);
      CK(cudaGetLastError());
      launches += 2;
    }
  }
  if (prof) {
    CK(cudaEventRecord(cx.ev[4], st));
    CK(cudaStreamSynchronize(st));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, cx.ev[0], cx.ev[1])); stats.coarse_ms = ms;
    CK(cudaEventElapsedTime(&ms, cx.ev[2], cx.ev[3])); stats.scan_ms = ms;
    CK(cudaEventElapsedTime(&ms, cx.ev[0], cx.ev[4])); stats.total_ms = ms;
    unsigned long long hs[2];
    CK(cudaMemcpy(hs, d_stats, 16, cudaMemcpyDeviceToHost));
    stats.scan_code_bytes = hs[0];
    stats.pairs = hs[1];
    stats.scan_items = (uint32_t)npairs;
  }
  stats.kernel_launches = launches;
  stats.scan_variant = flat ? 32 : 4;
  stats.scan_items = (uint32_t)npairs;
  stats.scan_ql = 1;
  {
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->stats = stats;
  }
  return 0;
}

// lb200_index_create for num_bits == 4: codebook [M][16][dim/M], codes concat over p of [M/2][n_p].
int create_index4(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist, const float* centroids,
                         uint32_t M, const float* codebook, const uint64_t* part_offsets, const uint8_t* codes,
                         const uint64_t* row_ids, const float* raw, lb200_index** out) {
  if (metric != LB200_L2 && metric != LB200_COSINE && metric != LB200_DOT) return set_err(LB200_EINVAL, "unknown metric %d", metric);
  if (!dim || !nlist || !M || dim % M) return set_err(LB200_EINVAL, "bad shape dim=%u nlist=%u M=%u", dim, nlist, M);
  if (M % 2)  // pq.rs:131-138
    return set_err(LB200_EINVAL, "PQ: num_sub_vectors must be divisible by 2 for num_bits=4, but got %u", M);
  if (M > 512) return set_err(LB200_EUNSUPPORTED, "num_sub_vectors=%u > 512", M);
  if (!centroids || !codebook || !part_offsets) return set_err(LB200_EINVAL, "null index buffers");
  if (part_offsets[0] != 0) return set_err(LB200_EINVAL, "part_offsets[0] must be 0");
  for (uint32_t p = 0; p < nlist; p++)
    if (part_offsets[p + 1] < part_offsets[p]) return set_err(LB200_EINVAL, "part_offsets not monotone at %u", p);
  const uint64_t N = part_offsets[nlist];
  if (N >= (1ull << 32)) return set_err(LB200_EUNSUPPORTED, "N=%llu rows >= 2^32", (unsigned long long)N);
  if (N && (!codes || !row_ids)) return set_err(LB200_EINVAL, "null codes/row_ids");
  int rc = select_device(device_id);
  if (rc) return rc;
  lb200_index* ix = new lb200_index();
  ix->device = device_id; ix->metric = metric; ix->dim = dim; ix->nlist = nlist; ix->M = M; ix->dsub = dim / M;
  ix->N = N; ix->bits = 4;
  ix->h_part_off.assign(part_offsets, part_offsets + nlist + 1);
  auto fail = [&](int code) { lb200_index_destroy(ix); return code; };
#define CK4(call)                                                                                          \
  do {                                                                                                     \
    cudaError_t e_ = (call);                                                                               \
    if (e_ != cudaSuccess)                                                                                 \
      return fail(set_err(e_ == cudaErrorMemoryAllocation ? LB200_ENOMEM : LB200_ECUDA, "%s failed: %s (%s:%d)", \
                          #call, cudaGetErrorString(e_), __FILE__, __LINE__));                             \
  } while (0)
  // rows of a partition sorted by row id: (distance, local index) order == (distance, row id) order (scan4.cuh)
  const uint32_t cb = M / 2;
  std::vector<uint64_t> pad_off(nlist + 1, 0);
  for (uint32_t p = 0; p < nlist; p++) pad_off[p + 1] = pad_off[p] + ((part_offsets[p + 1] - part_offsets[p] + 15) & ~15ull);
  const uint64_t NP = pad_off[nlist];
  std::vector<uint8_t> h_codes((size_t)NP * cb, 0);
  std::vector<uint64_t> h_rid((size_t)N);
  std::vector<uint8_t> h_tail((size_t)N);
  std::vector<float> h_raw(raw ? (size_t)N * dim : 0);
  std::vector<uint32_t> perm;
  ix->h_perm.resize((size_t)N);  // sorted position -> the caller's position in its partition (prefilter bit order)
  uint32_t nonempty = 0;
  for (uint32_t p = 0; p < nlist; p++) {
    const uint64_t o = part_offsets[p], n = part_offsets[p + 1] - o;
    ix->max_part = std::max(ix->max_part, n);
    nonempty += n > 0;
    perm.resize(n);
    for (uint64_t i = 0; i < n; i++) perm[i] = (uint32_t)i;
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t x, uint32_t y) { return row_ids[o + x] < row_ids[o + y]; });
    const uint8_t* src = codes + o * cb;
    const uint64_t npad = pad_off[p + 1] - pad_off[p];
    uint8_t* dst = h_codes.data() + pad_off[p] * cb;
    const uint64_t blocks = n >= 16 ? n - (n & 15) : 0;  // rows the reference runs through u8x16 (distance.rs:158)
    for (uint64_t i = 0; i < n; i++) {
      ix->h_perm[o + i] = perm[i];
      h_rid[o + i] = row_ids[o + perm[i]];
      h_tail[o + i] = perm[i] >= blocks;
      for (uint32_t j = 0; j < cb; j++) dst[(size_t)j * npad + i] = src[(size_t)j * n + perm[i]];
      if (raw) memcpy(h_raw.data() + (o + i) * dim, raw + (o + perm[i]) * dim, (size_t)dim * 4);
    }
  }
  ix->avg_np = (double)N / std::max(1u, nonempty);
  cudaDeviceProp prop;
  CK4(cudaGetDeviceProperties(&prop, device_id));
  ix->sm_count = prop.multiProcessorCount;
  ix->smem_optin = prop.sharedMemPerBlockOptin;
  ix->smem_per_sm = prop.sharedMemPerMultiprocessor;
  CK4(cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking));
  const size_t cent_b = (size_t)nlist * dim * 4, cb_b = (size_t)M * 16 * ix->dsub * 4;
  CK4(cudaMalloc(&ix->d_centroids, cent_b));
  CK4(cudaMalloc(&ix->d_codebook, cb_b));
  CK4(cudaMalloc(&ix->d_part_off, (nlist + 1) * 8));
  CK4(cudaMalloc(&ix->d_row_ids, std::max<uint64_t>(N, 1) * 8));
  CK4(cudaMalloc(&ix->d_codes4, std::max<size_t>((size_t)NP * cb, 16)));
  CK4(cudaMalloc(&ix->d_pad_off, (nlist + 1) * 8));
  CK4(cudaMemcpyAsync(ix->d_pad_off, pad_off.data(), (nlist + 1) * 8, cudaMemcpyHostToDevice, ix->stream));
  CK4(cudaMalloc(&ix->d_tail4, std::max<uint64_t>(N, 1)));
  CK4(cudaMemcpyAsync(ix->d_centroids, centroids, cent_b, cudaMemcpyHostToDevice, ix->stream));
  CK4(cudaMemcpyAsync(ix->d_codebook, codebook, cb_b, cudaMemcpyHostToDevice, ix->stream));
  CK4(cudaMemcpyAsync(ix->d_part_off, part_offsets, (nlist + 1) * 8, cudaMemcpyHostToDevice, ix->stream));
  if (N) {
    CK4(cudaMemcpyAsync(ix->d_row_ids, h_rid.data(), N * 8, cudaMemcpyHostToDevice, ix->stream));
    CK4(cudaMemcpyAsync(ix->d_codes4, h_codes.data(), (size_t)NP * cb, cudaMemcpyHostToDevice, ix->stream));
    CK4(cudaMemcpyAsync(ix->d_tail4, h_tail.data(), (size_t)N, cudaMemcpyHostToDevice, ix->stream));
    if (raw) {
      CK4(cudaMalloc(&ix->d_raw, (size_t)N * dim * 4));
      CK4(cudaMemcpyAsync(ix->d_raw, h_raw.data(), (size_t)N * dim * 4, cudaMemcpyHostToDevice, ix->stream));
    }
  }
  CK4(cudaStreamSynchronize(ix->stream));
#undef CK4
  *out = ix;
  return 0;
}


// lb200_index_create for num_bits == 32: IVF_FLAT -- no codebook, no codes; raw f32 rows per partition.
int create_index_flat(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist, const float* centroids,
                      const uint64_t* part_offsets, const uint64_t* row_ids, const float* raw, lb200_index** out) {
  if (metric != LB200_L2 && metric != LB200_COSINE && metric != LB200_DOT) return set_err(LB200_EINVAL, "unknown metric %d", metric);
  if (!dim || !nlist) return set_err(LB200_EINVAL, "bad shape dim=%u nlist=%u", dim, nlist);
  if (!centroids || !part_offsets) return set_err(LB200_EINVAL, "null index buffers");
  if (part_offsets[0] != 0) return set_err(LB200_EINVAL, "part_offsets[0] must be 0");
  for (uint32_t p = 0; p < nlist; p++)
    if (part_offsets[p + 1] < part_offsets[p]) return set_err(LB200_EINVAL, "part_offsets not monotone at %u", p);
  const uint64_t N = part_offsets[nlist];
  if (N >= (1ull << 32)) return set_err(LB200_EUNSUPPORTED, "N=%llu rows >= 2^32", (unsigned long long)N);
  if (N && (!raw || !row_ids)) return set_err(LB200_EINVAL, "IVF_FLAT needs the raw vectors and row ids");
  int rc = select_device(device_id);
  if (rc) return rc;
  lb200_index* ix = new lb200_index();
  ix->device = device_id; ix->metric = metric; ix->dim = dim; ix->nlist = nlist; ix->M = 1; ix->dsub = dim;
  ix->N = N; ix->bits = 32;
  ix->h_part_off.assign(part_offsets, part_offsets + nlist + 1);
  auto fail = [&](int code) { lb200_index_destroy(ix); return code; };
#define CKX(call)                                                                                          \
  do {                                                                                                     \
    cudaError_t e_ = (call);                                                                               \
    if (e_ != cudaSuccess)                                                                                 \
      return fail(set_err(e_ == cudaErrorMemoryAllocation ? LB200_ENOMEM : LB200_ECUDA, "%s failed: %s (%s:%d)", \
                          #call, cudaGetErrorString(e_), __FILE__, __LINE__));                             \
  } while (0)
  // rows of a partition sorted by row id: (distance, local index) order == (distance, row id) order
  std::vector<uint64_t> h_rid((size_t)N);
  std::vector<float> h_raw((size_t)N * dim);
  std::vector<uint32_t> perm;
  ix->h_perm.resize((size_t)N);  // sorted position -> the caller's position in its partition (prefilter bit order)
  uint32_t nonempty = 0;
  for (uint32_t p = 0; p < nlist; p++) {
    const uint64_t o = part_offsets[p], n = part_offsets[p + 1] - o;
    ix->max_part = std::max(ix->max_part, n);
    nonempty += n > 0;
    perm.resize(n);
    for (uint64_t i = 0; i < n; i++) perm[i] = (uint32_t)i;
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t x, uint32_t y) { return row_ids[o + x] < row_ids[o + y]; });
    for (uint64_t i = 0; i < n; i++) {
      ix->h_perm[o + i] = perm[i];
      h_rid[o + i] = row_ids[o + perm[i]];
      memcpy(h_raw.data() + (o + i) * dim, raw + (o + perm[i]) * dim, (size_t)dim * 4);
    }
  }
  ix->avg_np = (double)N / std::max(1u, nonempty);
  cudaDeviceProp prop;
  CKX(cudaGetDeviceProperties(&prop, device_id));
  ix->sm_count = prop.multiProcessorCount;
  ix->smem_optin = prop.sharedMemPerBlockOptin;
  ix->smem_per_sm = prop.sharedMemPerMultiprocessor;
  CKX(cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking));
  CKX(cudaMalloc(&ix->d_centroids, (size_t)nlist * dim * 4));
  CKX(cudaMalloc(&ix->d_part_off, (nlist + 1) * 8));
  CKX(cudaMalloc(&ix->d_row_ids, std::max<uint64_t>(N, 1) * 8));
  CKX(cudaMalloc(&ix->d_raw, std::max<size_t>((size_t)N * dim * 4, 16)));
  CKX(cudaMemcpyAsync(ix->d_centroids, centroids, (size_t)nlist * dim * 4, cudaMemcpyHostToDevice, ix->stream));
  CKX(cudaMemcpyAsync(ix->d_part_off, part_offsets, (nlist + 1) * 8, cudaMemcpyHostToDevice, ix->stream));
  if (N) {
    CKX(cudaMemcpyAsync(ix->d_row_ids, h_rid.data(), N * 8, cudaMemcpyHostToDevice, ix->stream));
    CKX(cudaMemcpyAsync(ix->d_raw, h_raw.data(), (size_t)N * dim * 4, cudaMemcpyHostToDevice, ix->stream));
  }
  CKX(cudaStreamSynchronize(ix->stream));
#undef CKX
  *out = ix;
  return 0;
}

}  // namespace

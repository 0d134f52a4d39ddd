// api_search.inl -- the batched IVF_PQ query pipeline (included by api.cu).

namespace {

static size_t scan_smem_bytes(int variant, int ql, uint32_t M, uint32_t dim, uint32_t capw) {
  int vg = 32 / ql;
  size_t lut = variant == 1 ? (size_t)M * 256 * 4 : (size_t)(M / vg) * 256 * 32 * 4;
  return lut + (size_t)ql * dim * 4 + (size_t)SCAN_NW * ql * capw * 8 + (size_t)SCAN_NW * ql * 4 +
         (size_t)ql * 8 + 16;
}

// Choose the scan kernel: the bank-skewed multi-query variant when the table layout allows it
// (M % (32/QL) == 0) and the batch brings enough queries per partition, else the generic one.
static int choose_scan(const lb200_index* ix, uint32_t kprime, double avg_q_per_part, ScanConfig* out) {
  const uint32_t need_cap = std::max(2 * kprime, kprime + 32);
  const size_t limit = ix->smem_optin;
  auto fit = [&](int variant, int ql, ScanConfig* c) {
    if (variant == 2 && ix->M % (32 / ql)) return false;
    size_t fixed = scan_smem_bytes(variant, ql, ix->M, ix->dim, 0);
    if (fixed >= limit) return false;
    size_t capw = (limit - fixed) / ((size_t)SCAN_NW * ql * 8);
    capw = std::min<size_t>(capw, 1024);
    if (capw < need_cap) return false;
    if (capw > 4 * (size_t)need_cap) capw = 4 * (size_t)need_cap;  // enough; keep L1 for the codes
    c->variant = variant;
    c->ql = ql;
    c->capw = (uint32_t)capw;
    c->smem = scan_smem_bytes(variant, ql, ix->M, ix->dim, (uint32_t)capw);
    c->dsub_t = ix->dsub == 8 ? 8 : ix->dsub == 16 ? 16 : 0;
    return true;
  };
  out->minb = 1;
  if (ix->d_codes3 && (ix->force_variant == 0 || ix->force_variant == 3)) {
    // variant 3: QL is tied to the layout's lane-delay period (8 -> LDS.128/QL 4, 16 -> LDS.64/QL 2)
    const int ql = ix->skew_D == 8 ? 4 : 2;
    uint32_t capw = std::max(64u, (kprime + 32 + 3) & ~3u);
    // (f16 handles always fetch pre-built tables: no residual staging area, see search_pipeline)
    size_t smem = scan3_smem_bytes(ql, SCAN_NW, ix->M, ix->f16 ? 0 : ix->dim, capw);
    if (smem <= limit) {
      out->variant = 3;
      out->ql = ql;
      out->capw = capw;
      out->smem = smem;
      out->dsub_t = ix->dsub == 8 ? 8 : ix->dsub == 16 ? 16 : 0;
      out->minb = (2 * (smem + 1024) <= ix->smem_per_sm) ? 2 : 1;
      return 0;
    }
    if (ix->force_variant == 3 || !ix->d_codes)
      return set_err(LB200_EUNSUPPORTED, "k'=%u does not fit the shared-memory candidate buffers (M=%u, dim=%u)",
                     kprime, ix->M, ix->dim);
  }
  if (!ix->d_codes) return set_err(LB200_EUNSUPPORTED, "tile16 layout missing (LB200_LAYOUTS=all)");
  if (ix->force_variant != 1) {
    int want = avg_q_per_part >= 6 ? 8 : avg_q_per_part >= 3 ? 4 : avg_q_per_part >= 1.5 ? 2 : 1;
    if (ix->force_variant == 2 && want < 2) want = 8;
    for (int ql = want; ql >= 1; ql >>= 1)
      if ((ql > 1 || ix->force_variant == 2 || ix->M % 32 == 0) && fit(2, ql, out)) return 0;
    if (ix->force_variant == 2)
      for (int ql = 8; ql > want; ql >>= 1)
        if (fit(2, ql, out)) return 0;
  }
  if (fit(1, 1, out)) return 0;
  return set_err(LB200_EUNSUPPORTED, "k'=%u does not fit the shared-memory candidate buffers (M=%u, dim=%u)",
                 kprime, ix->M, ix->dim);
}

struct SearchArgs {
  const float* d_queries;      // [Q][dim] device, as given by the caller
  uint32_t Q;
  uint32_t nprobes;            // used when d_part_ids == nullptr
  const uint32_t* d_part_ids;  // [Q][P] device (search_partitions) or nullptr (search)
  uint32_t P;                  // partitions per query (search_partitions)
  uint32_t k, refine_factor;
  bool normalize;              // cosine + VectorIndex::search entry
  bool per_pair_output;        // search_partitions: [Q][P][k'] instead of merged [Q][k']
  bool shard_mode;             // stop before the final cut, also return exact distances
  bool has_lower, has_upper;
  float lower, upper;
  const uint64_t* const* d_bitmask;  // device array [nlist] of device pointers, or nullptr
  float* d_out_dist;
  uint64_t* d_out_rowid;
  uint32_t* d_out_count;
  float* d_out_exact;          // shard mode only
  const struct PrebuiltWork* pw = nullptr;  // work list built on the host (small search_partitions calls)
  const float* d_keys_override = nullptr;   // f16 handles: keys already normalised in f16 arithmetic (f16pq.cuh)
};

// Scan plan of one call: kernel configuration and how partitions are cut into chunks.
struct ScanPlan {
  ScanConfig sc;
  uint32_t chunk_vecs, max_chunks;
};

// Work list of a small call, built on the host and already uploaded: search_in_partition calls of the
// drop-in path carry one (query, partition) pair each, and the four work-list kernels + three memsets of
// the batched path would be most of their latency.
struct PrebuiltWork {
  ScanPlan plan;
  const uint32_t *d_pair_list, *d_pair_off, *d_item_off;
  const uint4* d_item_desc;
  uint32_t* d_counter;  // zeroed
  uint64_t code_bytes, pairs;
  uint32_t items;
};

// find_partitions on device: distances [Q][nlist] + per-query sort.  Output [Q][P].
static int coarse_device(lb200_index* ix, const float* d_keys, uint32_t Q, uint32_t P, float* d_D,
                         uint32_t* d_probes, float* d_pdists, cudaStream_t st, uint32_t* launches) {
  if (ix->nlist > 8192)
    return set_err(LB200_EUNSUPPORTED, "nlist=%u > 8192 not supported by the coarse select kernel yet", ix->nlist);
  dim3 grid((ix->nlist + COARSE_THREADS - 1) / COARSE_THREADS, (Q + COARSE_TQ - 1) / COARSE_TQ);
  size_t smem = (size_t)COARSE_TQ * ix->dim * 4;
  const size_t tsm = ((size_t)ix->dim * CQ_PITCH + (size_t)CQ_QB * ix->dim) * 4;
  const size_t ssm2 = (2 * (size_t)CD_SLAB + (size_t)CD_QB * ix->dim) * 4;
  if (ix->f16 && (ix->metric == LB200_DOT || ix->dim <= 16)) {
    // f16::dot is dot_scalar::<f16,f32,32>, f16::l2 has no dimension-8/16 special case: not the f32 kernels' order
    const uint64_t tot = (uint64_t)Q * ix->nlist;
    coarse_dist_f16order_kernel<<<(uint32_t)((tot + 255) / 256), 256, 0, st>>>(ix->metric == LB200_DOT, d_keys, ix->d_centroids, Q,
                                                                             ix->nlist, ix->dim, d_D);
  } else if (ix->metric != LB200_DOT && ix->dim % 64 == 0 && ssm2 <= ix->smem_optin && !getenv("LB200_COARSE_SIMPLE") &&
      !getenv("LB200_COARSE_TILED")) {
    CK(cudaFuncSetAttribute(coarse_dist_l2_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm2));
    dim3 g2((ix->nlist + 127) / 128, (Q + CD_QB - 1) / CD_QB);
    coarse_dist_l2_slab_kernel<<<g2, 512, ssm2, st>>>(d_keys, ix->d_centroids, Q, ix->nlist, ix->dim, d_D, ix->nlist,
                                                      nullptr);
  } else if (ix->metric != LB200_DOT && ix->dim % 16 == 0 && ix->dim >= 32 && tsm <= ix->smem_optin &&
      !getenv("LB200_COARSE_SIMPLE")) {
    CK(cudaFuncSetAttribute(coarse_dist_l2_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm));
    dim3 g2((ix->nlist + 127) / 128, (Q + CQ_QB - 1) / CQ_QB);
    coarse_dist_l2_tiled_kernel<<<g2, 128, tsm, st>>>(d_keys, ix->d_centroids, Q, ix->nlist, ix->dim, d_D, ix->nlist,
                                                      nullptr);
  } else if (ix->metric == LB200_DOT && ix->dim % 64 == 0 &&
             (2 * (size_t)CD_SLAB + (size_t)CD_QB * ix->dim) * 4 <= ix->smem_optin && !getenv("LB200_COARSE_SIMPLE")) {
    const size_t dsm = (2 * (size_t)CD_SLAB + (size_t)CD_QB * ix->dim) * 4;
    CK(cudaFuncSetAttribute(coarse_dist_dot_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsm));
    dim3 g2((ix->nlist + 127) / 128, (Q + CD_QB - 1) / CD_QB);
    coarse_dist_dot_slab_kernel<<<g2, 512, dsm, st>>>(d_keys, ix->d_centroids, Q, ix->nlist, ix->dim, d_D, ix->nlist);
  } else if (ix->metric == LB200_DOT)
    coarse_dist_kernel<false><<<grid, COARSE_THREADS, smem, st>>>(d_keys, ix->d_centroids, Q, ix->nlist, ix->dim, d_D);
  else
    coarse_dist_kernel<true><<<grid, COARSE_THREADS, smem, st>>>(d_keys, ix->d_centroids, Q, ix->nlist, ix->dim, d_D);
  CK(cudaGetLastError());
  if (P <= 256 && !getenv("LB200_COARSE_FULLSORT")) {
    coarse_select2_kernel<<<Q, 256, 0, st>>>(d_D, ix->nlist, P, d_probes, d_pdists);
    CK(cudaGetLastError());
    *launches += 2;
    return 0;
  }
  uint32_t np2 = next_pow2(ix->nlist);
  size_t ssm = (size_t)np2 * 8;
  if (ssm > 48 * 1024)
    CK(cudaFuncSetAttribute(coarse_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
  coarse_select_kernel<<<Q, 256, ssm, st>>>(d_D, ix->nlist, np2, P, d_probes, d_pdists);
  CK(cudaGetLastError());
  *launches += 2;
  return 0;
}

static int launch_merge(const float* in_dist, const uint64_t* in_rowid, const uint64_t* in_pos, const float* in_aux,
                        const uint32_t* in_cnt, uint32_t groups, uint32_t L, uint32_t stride, size_t group_stride,
                        size_t list_stride, size_t cnt_gs, size_t cnt_ls, uint32_t kout, uint32_t out_stride,
                        float* out_dist, uint64_t* out_rowid, uint64_t* out_pos, float* out_aux, uint32_t* out_cnt,
                        cudaStream_t st) {
  uint64_t tot = (uint64_t)L * stride;
  if (tot > 16384) return set_err(LB200_EUNSUPPORTED, "merge of %llu candidates per query exceeds 16384", (unsigned long long)tot);
  uint32_t np2 = std::max(2u, next_pow2((uint32_t)tot));
  size_t smem = (size_t)np2 * 8;
  if (smem > 48 * 1024)
    CK(cudaFuncSetAttribute(merge_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // few groups with many candidates (tiny batches of chunked partitions): a wide block per group
  const uint32_t mthreads = tot > 2048 ? 1024 : 256;
  merge_lists_kernel<<<groups, mthreads, smem, st>>>(in_dist, in_rowid, in_pos, in_aux, in_cnt, L, stride, group_stride,
                                                list_stride, cnt_gs, cnt_ls, np2, kout, out_stride, out_dist,
                                                out_rowid, out_pos, out_aux, out_cnt, 1, L);
  CK(cudaGetLastError());
  return 0;
}

// Lists per slice of the first level of a two-level merge, or 0 when one level is enough.
static uint32_t merge_slice_lists(uint64_t L, uint32_t stride) {
  return L * stride > 16384 ? std::max(1u, 16384u / stride) : 0u;
}
// Two-level merge for L*stride > 16384 candidates per group (many probes x large k'): slices of F lists
// are merged to `stride` survivors each into t_* ([groups][slices][stride]), then the slices are merged.
// The survivors of a slice are its exact top-`stride` by (distance, row id), so the result is the same.
static int launch_merge2(const float* in_dist, const uint64_t* in_rowid, const uint64_t* in_pos,
                         const uint32_t* in_cnt, uint32_t groups, uint32_t L, uint32_t stride, float* t_dist,
                         uint64_t* t_rowid, uint64_t* t_pos, uint32_t* t_cnt, float* out_dist, uint64_t* out_rowid,
                         uint64_t* out_pos, uint32_t* out_cnt, cudaStream_t st) {
  const uint32_t F = merge_slice_lists(L, stride), S = (L + F - 1) / F;
  if ((uint64_t)S * stride > 16384)
    return set_err(LB200_EUNSUPPORTED, "merge of %u lists of %u per query exceeds two merge levels", L, stride);
  uint32_t np2 = std::max(2u, next_pow2(F * stride));
  size_t smem = (size_t)np2 * 8;
  CK(cudaFuncSetAttribute(merge_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem, 48 * 1024)));
  merge_lists_kernel<<<groups * S, 1024, smem, st>>>(in_dist, in_rowid, in_pos, nullptr, in_cnt, F, stride,
                                                    (size_t)L * stride, stride, L, 1, np2, stride, stride, t_dist,
                                                    t_rowid, t_pos, nullptr, t_cnt, S, L);
  CK(cudaGetLastError());
  return launch_merge(t_dist, t_rowid, t_pos, nullptr, t_cnt, groups, S, stride, (size_t)S * stride, stride, S, 1,
                      stride, stride, out_dist, out_rowid, out_pos, nullptr, out_cnt, st);
}

// Kernel choice and chunking for a batch of Q queries x P partitions.
static int plan_scan(const lb200_index* ix, uint32_t Q, uint32_t P, uint32_t kprime, ScanPlan* out) {
  const uint32_t nlist = ix->nlist;
  int rc;
  ScanConfig sc;
  double avg_q = (double)Q * P / std::max(1u, nlist);
  rc = choose_scan(ix, kprime, avg_q, &sc);
  if (rc) return rc;
  // chunking: enough work items to fill the SMs even for tiny batches
  uint64_t est_items = std::max<uint64_t>(1, (uint64_t)Q * P / sc.ql);
  uint64_t want_items = 4ull * ix->sm_count;
  uint32_t chunk_vecs = (uint32_t)std::max<uint64_t>(16, (ix->max_part + 15) / 16 * 16);
  if (est_items < want_items) {
    uint64_t nch = (want_items + est_items - 1) / est_items;
    uint64_t cv = (ix->max_part + nch - 1) / nch;
    cv = std::max<uint64_t>(2048, (cv + 511) / 512 * 512);
    chunk_vecs = (uint32_t)std::min<uint64_t>(chunk_vecs, cv);
  }
  if (const char* e = getenv("LB200_CHUNK_ROWS")) {  // experiment: bound the rows of one work item
    const uint32_t cr = (uint32_t)atoi(e);
    if (cr >= 512) chunk_vecs = std::min(chunk_vecs, cr);
  }
  {
    // the per-query merge sorts at most 16384 candidates: P * chunks * k' must stay below that
    const uint64_t allowed = std::max<uint64_t>(1, 16384ull / ((uint64_t)P * kprime));
    const uint64_t cv_min = (ix->max_part + allowed - 1) / allowed;
    if (chunk_vecs < cv_min) chunk_vecs = (uint32_t)((cv_min + 15) / 16 * 16);
  }
  if (sc.variant == 3) {  // chunks are whole superblocks
    const uint32_t sbr = 32 * ix->skew_L;
    chunk_vecs = (chunk_vecs + sbr - 1) / sbr * sbr;
  }
  const uint32_t max_chunks = (uint32_t)std::max<uint64_t>(1, (ix->max_part + chunk_vecs - 1) / chunk_vecs);

  out->sc = sc;
  out->chunk_vecs = chunk_vecs;
  out->max_chunks = max_chunks;
  return 0;
}

// The whole device pipeline for one batch of queries; asynchronous on `st`.
static int search4_pipeline(lb200_index* ix, CallCtx& cx, const SearchArgs& a, cudaStream_t st);  // api_pq4.inl
static int search_pipeline(lb200_index* ix, CallCtx& cx, const SearchArgs& a, cudaStream_t st) {
  if (ix->bits != 8) return search4_pipeline(ix, cx, a, st);  // 4-bit PQ, IVF_FLAT
  if (ix->f16 && !a.d_keys_override)
    return set_err(LB200_EINVAL, "a Float16 handle takes Float16 keys: use lb200_search_f16");
  const uint32_t Q = a.Q, dim = ix->dim, nlist = ix->nlist;
  const uint32_t P = a.d_part_ids ? a.P : std::min(a.nprobes, nlist);
  const uint32_t kprime = a.k * std::max(a.refine_factor, 1u);
  if (a.k == 0) return set_err(LB200_EINVAL, "k must be > 0");       // scanner.rs:642 validation
  if (kprime > 512) return set_err(LB200_EUNSUPPORTED, "k*refine_factor=%u > 512", kprime);
  if (P == 0 || Q == 0) return 0;
  const bool has_raw = ix->d_raw != nullptr || ix->d_raw16 != nullptr;
  const bool refine = a.refine_factor > 0 && has_raw && !a.per_pair_output;
  // (a shard that owns no rows has no raw vectors and nothing to refine; any other handle must hold them)
  if (a.refine_factor > 0 && !has_raw && !a.per_pair_output && !(a.shard_mode && ix->N == 0))
    return set_err(LB200_EINVAL, "refine_factor>0 needs raw vectors in the index handle");
  uint32_t launches = 0;
  lb200_stats stats = {};
  int rc;

  ScanPlan plan;
  if (a.pw) plan = a.pw->plan;
  else if ((rc = plan_scan(ix, Q, P, kprime, &plan))) return rc;
  ScanConfig sc = plan.sc;
  const uint32_t chunk_vecs = plan.chunk_vecs, max_chunks = plan.max_chunks;

  // ---- workspace -----------------------------------------------------------------------
  const size_t npairs = (size_t)Q * P, nslots = npairs * max_chunks;
  if (npairs >= (1ull << 31)) return set_err(LB200_EUNSUPPORTED, "Q*P too large");
  size_t need = 0;
  auto add = [&](size_t b) { need += Arena::pad(b) + 256; };
  add((size_t)Q * dim * 4);         // normalised keys
  add((size_t)Q * nlist * 4);       // coarse distances
  add(npairs * 4);                  // probes
  add((size_t)nlist * 4 * 2);       // cnt (cold, other)
  add((size_t)Q * 4);               // first partition per query
  add(npairs * 4);                  // rank
  add((size_t)nlist * 4);           // queue offsets (size-ordered item queue)
  // seeding launch (below): its own small work list and throw-away outputs, one pair per query
  add((size_t)Q * 4 * 4); add((size_t)nlist * 4 * 2); add((size_t)(nlist + 1) * 4 * 2); add((size_t)Q * 16); add(64);
  add((size_t)Q * kprime * 20); add((size_t)Q * 4);
  add((size_t)(nlist + 1) * 4 * 2); // pair_off, item_off
  add(npairs * 4);                  // pair_list
  add(nslots * 16);                 // item descriptors (items <= pairs * chunks)
  add(64);                          // counter + stats
  add(nslots * kprime * 4); add(nslots * kprime * 8); add(nslots * kprime * 8); add(nslots * 4);
  add((size_t)Q * kprime * 4); add((size_t)Q * kprime * 8); add((size_t)Q * kprime * 8); add((size_t)Q * 4);
  add((size_t)Q * kprime * 4);      // exact
  add((size_t)Q * 4);               // per-query thresholds
  const uint32_t mslice = a.per_pair_output ? 0 : merge_slice_lists((uint64_t)P * max_chunks, kprime);
  const uint32_t mslices = mslice ? (P * max_chunks + mslice - 1) / mslice : 0;
  if (mslice) { add((size_t)Q * mslices * kprime * 4); add((size_t)Q * mslices * kprime * 8); add((size_t)Q * mslices * kprime * 8); add((size_t)Q * mslices * 4); }
  // ADC tables as a pre-pass (scan3.cuh, lut_prepass_kernel): always for dot (one table per query serves all
  // its probes), for residual metrics when a pair's table costs more than its rows (256*dsub flops per
  // sub-vector against n_p lookups)
  bool use_pre = false, pre_by_query = false;
  size_t pre_units = 0;
  if (sc.variant == 3 && (ix->dsub == 8 || ix->dsub == 16)) {
    pre_by_query = ix->metric == LB200_DOT;
    pre_units = pre_by_query ? (size_t)Q : npairs;
    bool want = pre_by_query || 256.0 * ix->dsub > ix->avg_np;
    if (const char* e = getenv("LB200_LUT_PREPASS")) want = atoi(e) != 0;
    double cap_gb = 24.0;
    if (const char* e = getenv("LB200_LUT_SCRATCH_GB")) cap_gb = atof(e);
    use_pre = want && (double)pre_units * ix->M * 1024.0 <= cap_gb * 1e9 && pre_units <= 65535ull * LUTPRE_UNITS;
  }
  if (ix->f16) {
    // f16 handles: the table arithmetic differs from the f32 kernels' (f16pq.cuh), so tables always come from the
    // f16 pre-pass and the scan only fetches them
    if (sc.variant != 3)
      return set_err(LB200_EUNSUPPORTED, "Float16 IVF_PQ: k*refine_factor=%u does not fit the skew-stream scan's shared memory (M=%u)",
                     kprime, ix->M);
    pre_by_query = ix->metric == LB200_DOT;
    pre_units = pre_by_query ? (size_t)Q : npairs;
    if ((double)pre_units * ix->M * 1024.0 > 24e9)
      return set_err(LB200_EUNSUPPORTED, "Float16 IVF_PQ: %zu ADC tables in one batch exceed the table scratch", pre_units);
    use_pre = true;
  }
  // per-pair tables stored as one interleaved image per query group (scan3.cuh lut_image_issue / lut_image_wait): a few more table
  // slots (the last group of a partition may be partly filled), positions in the pair list as unit numbers
  const bool pre_image = use_pre && !pre_by_query && !ix->f16 && sc.minb == 1 && !getenv("LB200_NO_LUT_IMAGE");
  const size_t pre_slots = pre_image ? pre_units + (size_t)nlist * (sc.ql - 1) : pre_units;
  if (use_pre) { add(pre_slots * ix->M * 1024); add(pre_units * 4); }
  // Tables that leave room for one CTA per SM only (M >= 48): with the tables pre-built the residual
  // staging area is not needed, which makes room for the candidate buffers of 16 warps instead of 8.
  if (use_pre && sc.minb == 1 && !getenv("LB200_NO_WIDE")) {
    const size_t smem16 = scan3_smem_bytes(sc.ql, 16, ix->M, 0, sc.capw);
    if (smem16 <= ix->smem_optin) {
      sc.nw = 16;
      sc.smem = smem16;
    }
  }
  rc = cx.arena.reserve(need);
  if (rc) return rc;
  cx.arena.reset();
  float* d_keys = cx.arena.take<float>((size_t)Q * dim);
  float* d_D = cx.arena.take<float>((size_t)Q * nlist);
  uint32_t* d_probes = cx.arena.take<uint32_t>(npairs);
  uint32_t* d_cnt = cx.arena.take<uint32_t>((size_t)nlist * 2);
  uint32_t* d_cnt2 = d_cnt + nlist;
  uint32_t* d_minp = cx.arena.take<uint32_t>(Q);
  uint32_t* d_rank = cx.arena.take<uint32_t>(npairs);
  uint32_t* d_qoff = cx.arena.take<uint32_t>(nlist);
  uint32_t* s_probe = cx.arena.take<uint32_t>(Q);
  uint32_t* s_minp = cx.arena.take<uint32_t>(Q);
  uint32_t* s_rank = cx.arena.take<uint32_t>(Q);
  uint32_t* s_pair_list = cx.arena.take<uint32_t>(Q);
  uint32_t* s_cnt = cx.arena.take<uint32_t>((size_t)nlist * 2);
  uint32_t* s_pair_off = cx.arena.take<uint32_t>(nlist + 1);
  uint32_t* s_item_off = cx.arena.take<uint32_t>(nlist + 1);
  uint4* s_desc = cx.arena.take<uint4>(Q);
  unsigned long long* s_stats = cx.arena.take<unsigned long long>(4);
  uint32_t* s_counter = reinterpret_cast<uint32_t*>(s_stats + 2);
  float* s_odist = cx.arena.take<float>((size_t)Q * kprime);
  uint64_t* s_opos = cx.arena.take<uint64_t>((size_t)Q * kprime);
  uint64_t* s_orow = cx.arena.take<uint64_t>((size_t)Q * kprime);
  uint32_t* s_ocnt = cx.arena.take<uint32_t>(Q);
  uint32_t* d_pair_off = cx.arena.take<uint32_t>(nlist + 1);
  uint32_t* d_item_off = cx.arena.take<uint32_t>(nlist + 1);
  uint32_t* d_pair_list = cx.arena.take<uint32_t>(npairs);
  uint4* d_item_desc = cx.arena.take<uint4>(nslots);
  unsigned long long* d_stats = cx.arena.take<unsigned long long>(4);
  uint32_t* d_counter = reinterpret_cast<uint32_t*>(d_stats + 2);
  float* d_sdist = cx.arena.take<float>(nslots * kprime);
  uint64_t* d_spos = cx.arena.take<uint64_t>(nslots * kprime);
  uint64_t* d_srow = cx.arena.take<uint64_t>(nslots * kprime);
  uint32_t* d_scnt = cx.arena.take<uint32_t>(nslots);
  float* d_mdist = cx.arena.take<float>((size_t)Q * kprime);
  uint64_t* d_mrow = cx.arena.take<uint64_t>((size_t)Q * kprime);
  uint64_t* d_mpos = cx.arena.take<uint64_t>((size_t)Q * kprime);
  uint32_t* d_mcnt = cx.arena.take<uint32_t>(Q);
  float* d_exact = cx.arena.take<float>((size_t)Q * kprime);
  uint32_t* d_qthr = cx.arena.take<uint32_t>(Q);
  float* t_mdist = mslice ? cx.arena.take<float>((size_t)Q * mslices * kprime) : nullptr;
  uint64_t* t_mrow = mslice ? cx.arena.take<uint64_t>((size_t)Q * mslices * kprime) : nullptr;
  uint64_t* t_mpos = mslice ? cx.arena.take<uint64_t>((size_t)Q * mslices * kprime) : nullptr;
  uint32_t* t_mcnt = mslice ? cx.arena.take<uint32_t>((size_t)Q * mslices) : nullptr;
  float* d_lut = use_pre ? cx.arena.take<float>(pre_slots * ix->M * 256) : nullptr;
  uint32_t* d_lut_max = use_pre ? cx.arena.take<uint32_t>(pre_units) : nullptr;

  const bool prof = ix->profiling;
  if (prof) CK(cudaEventRecord(cx.ev[0], st));

  // ---- A1 normalise (VectorIndex::search on a cosine index, ivf/v2.rs:370-373) ----------
  const float* keys = a.d_queries;
  if (a.d_keys_override) {
    keys = a.d_keys_override;
  } else if (a.normalize && ix->metric == LB200_COSINE) {
    normalize_kernel<<<(Q * 32 + 255) / 256, 256, 0, st>>>(a.d_queries, d_keys, Q, dim);
    CK(cudaGetLastError());
    launches++;
    keys = d_keys;
  }
  // ---- A2 find_partitions ---------------------------------------------------------------
  const uint32_t* probes = a.d_part_ids;
  if (!probes) {
    rc = coarse_device(ix, keys, Q, P, d_D, d_probes, nullptr, st, &launches);
    if (rc) return rc;
    probes = d_probes;
  }
  if (prof) CK(cudaEventRecord(cx.ev[1], st));
  // ---- work list ------------------------------------------------------------------------
  // Item queue in descending partition size (longest first) when a CTA draws few items: the ragged end of the queue is
  // then a visible part of the scan (partition shards: 1/G of the items per GPU; cfg 4 at N = 8: +3.3 %).  With
  // hundreds of items per CTA the id-ordered queue is the faster one (cfg 4 at N = 1: -1 %, cfg 1: -4 %; same-box A/B
  // profiles/r02_ai, r02_aj).
  // (a shard holds rows in a fraction of the partitions only: that fraction of the pairs becomes work items)
  bool lpt = ix->d_size_order != nullptr &&
             (double)Q * P / sc.ql * ((double)ix->nonempty / std::max(1u, nlist)) < 64.0 * ix->sm_count * std::max(1, sc.minb);
  if (const char* e = getenv("LB200_LPT")) lpt = ix->d_size_order != nullptr && atoi(e) != 0;
  CK(cudaMemsetAsync(d_scnt, 0, nslots * 4, st));
  const uint32_t* total_items_p = d_item_off + nlist;
  if (a.pw) {  // built on the host and uploaded with the queries (small search_partitions calls)
    d_pair_list = const_cast<uint32_t*>(a.pw->d_pair_list);
    d_pair_off = const_cast<uint32_t*>(a.pw->d_pair_off);
    d_item_off = const_cast<uint32_t*>(a.pw->d_item_off);
    d_counter = a.pw->d_counter;
    d_item_desc = const_cast<uint4*>(a.pw->d_item_desc);
    total_items_p = d_item_off + nlist;
  } else {
  CK(cudaMemsetAsync(d_cnt, 0, (size_t)nlist * 4 * 2, st));
  CK(cudaMemsetAsync(d_stats, 0, 32, st));
  if (getenv("LB200_NO_COLDPACK"))
    CK(cudaMemsetAsync(d_minp, 0xff, (size_t)Q * 4, st));
  else
    first_part_kernel<<<(Q + 255) / 256, 256, 0, st>>>(probes, Q, P, ix->d_part_off, d_minp,
                                                       lpt ? ix->d_size_order + nlist : nullptr);
  count_pairs_kernel<<<(uint32_t)((npairs + 255) / 256), 256, 0, st>>>(probes, (uint32_t)npairs, P, d_minp, d_cnt,
                                                                      d_cnt2, d_rank);
  offsets_kernel<<<1, 1024, 0, st>>>(d_cnt, d_cnt2, ix->d_part_off, nlist, (uint32_t)sc.ql, chunk_vecs, ix->M,
                                     d_pair_off, d_item_off, d_stats);
  scatter_pairs_kernel<<<(uint32_t)((npairs + 255) / 256), 256, 0, st>>>(probes, d_rank, d_pair_off, d_cnt,
                                                                        (uint32_t)npairs, d_pair_list);
  if (sc.variant == 3) {
    const uint32_t* qoff = nullptr;
    if (lpt) {
      queue_offsets_kernel<<<1, 1024, 0, st>>>(d_item_off, ix->d_size_order, nlist, d_qoff);
      qoff = d_qoff;
      launches++;
    }
    expand_items_kernel<<<(nlist * 32 + 255) / 256, 256, 0, st>>>(d_item_off, ix->d_part_off, nlist, chunk_vecs, d_item_desc, qoff);
    launches++;
  }
  CK(cudaGetLastError());
  launches += 4;
  }
  // ---- A3-A6 scan -----------------------------------------------------------------------
  if (sc.variant == 3) {
    Scan3Params sp = {};
    sp.codes = ix->d_codes3; sp.sb_off = ix->d_sb_off; sp.part_off = ix->d_part_off; sp.row_ids = ix->d_row_ids;
    sp.centroids = ix->d_centroids; sp.codebook = sc.dsub_t ? ix->d_codebook_img : ix->d_codebook;  // in-kernel build: DSUB > 0 reads the image
    sp.dim = dim; sp.M = ix->M; sp.dsub = ix->dsub; sp.nlist = nlist; sp.L = ix->skew_L;
    sp.residual = ix->metric != LB200_DOT;
    sp.queries = keys; sp.Q = Q; sp.P = P; sp.kprime = kprime;
    sp.pair_list = d_pair_list; sp.part_pair_off = d_pair_off; sp.item_off = d_item_off; sp.item_counter = d_counter;
    sp.item_desc = d_item_desc;
    sp.chunk_vecs = chunk_vecs; sp.max_chunks = max_chunks; sp.capw = sc.capw;
    // merged searches share one running k'-th-best bound per query across all its partitions;
    // search_in_partition semantics (one full list per pair) must not.
    sp.qthr = (a.per_pair_output || getenv("LB200_NO_QTHR")) ? nullptr : d_qthr;
    sp.no_boot = getenv("LB200_NO_BOOT") != nullptr;
    sp.boot2 = getenv("LB200_NO_BOOT2") == nullptr;
    sp.lower = a.lower; sp.upper = a.upper; sp.has_lower = a.has_lower; sp.has_upper = a.has_upper;
    sp.bitmask = a.d_bitmask;
    sp.out_dist = d_sdist; sp.out_pos = d_spos; sp.out_rowid = d_srow; sp.out_count = d_scnt;
    uint32_t* reset = d_qthr;
    uint32_t nreset = Q;
    if (sp.qthr && a.shard_mode && ix->pthr_peers && Q <= ix->pthr_cap && !getenv("LB200_NO_PEER_THR")) {
      // bounds shared with the peer shards: this batch uses table (epoch & 1) -- reset one batch ago, peers
      // may already have published into it -- and clears the other one for the next batch
      std::lock_guard<std::mutex> lk(ix->mu);
      const uint32_t par = (uint32_t)(ix->pthr_epoch & 1);
      sp.qthr = ix->pthr_local + (size_t)par * ix->pthr_cap;
      sp.peers = ix->d_pthr_tbl + (size_t)par * LB200_MAX_PEERS;
      sp.n_peers = ix->pthr_peers;
      reset = ix->pthr_local + (size_t)(par ^ 1) * ix->pthr_cap;
      nreset = ix->pthr_cap;
      ix->pthr_epoch++;
    }
    static int keep_calls = 0;   // experiment (LB200_KEEP_QTHR): after the first batch keep the bounds -- valid only for a repeated batch
    if (sp.qthr && !(getenv("LB200_KEEP_QTHR") && keep_calls++ > 0)) {
      fill_u32_kernel<<<(nreset + 255) / 256, 256, 0, st>>>(reset, KEY_PINF, nreset);
      CK(cudaGetLastError());
      launches++;
    }
    if (prof) CK(cudaEventRecord(cx.ev[2], st));
    if (use_pre) {  // inside the scan's timed span: the tables are part of the scan's cost
      CK(cudaMemsetAsync(d_lut_max, 0, pre_units * 4, st));
      dim3 g((ix->M / 16) * LUTPRE_SPLIT, (uint32_t)((pre_units + LUTPRE_UNITS - 1) / LUTPRE_UNITS));
      const uint32_t* pr = pre_by_query ? nullptr : probes;
      if (ix->f16)
        lut_prepass_f16_kernel<<<(uint32_t)pre_units, 256, (size_t)dim * 4, st>>>(keys, ix->d_centroids, ix->d_codebook, pr,
                                                                                 ix->d_part_off, P, dim, ix->M, ix->dsub, d_lut,
                                                                                 d_lut_max);
      else if (ix->dsub == 8)
        lut_prepass_kernel<8><<<g, 256, 0, st>>>(keys, ix->d_centroids, ix->d_codebook, pr, ix->d_part_off,
                                                 (uint32_t)pre_units, P, dim, ix->M, d_lut, d_lut_max,
                                                 pre_image ? d_pair_list : nullptr, d_pair_off, (uint32_t)sc.ql, 0x8000000080000000ull);
      else
        lut_prepass_kernel<16><<<g, 256, 0, st>>>(keys, ix->d_centroids, ix->d_codebook, pr, ix->d_part_off,
                                                  (uint32_t)pre_units, P, dim, ix->M, d_lut, d_lut_max,
                                                  pre_image ? d_pair_list : nullptr, d_pair_off, (uint32_t)sc.ql, 0x8000000080000000ull);
      CK(cudaGetLastError());
      launches++;
      sp.lut_src = d_lut;
      sp.lut_max = d_lut_max;
      sp.lut_by_query = pre_by_query;
      sp.lut_image = pre_image;
    }
    // ---- seeding launch: every query gets a first bound before the main scan -----------------------------------
    // Without it a query's work items run with no bound (cold) or a loose one until its near partitions have been
    // scanned; the same batch with the final bounds known from the start scans 15 % faster at N = 1 and 39 % faster on a
    // shard of 8 (profiles/r02_bj_*: LB200_KEEP_QTHR).  So the first 2048 rows of each query's NEAREST probed partition
    // are scanned first, by the same kernel on a one-pair-per-query work list; its finalize publishes the k'-th best
    // of those rows (a valid bound: k' real rows at or below it; to the peer shards too).  Outputs are thrown away --
    // the main launch scans those rows again.  Only with in-kernel tables: per-pair pre-pass tables (cfg 3) sit in slots
    // that follow the main work list, and with per-query tables (cfg 5, dot) the extra launch cost more than it saved
    // (same-box A/B: 8.39 vs 7.98 ms).
    bool seed = sp.qthr != nullptr && !use_pre && !a.pw && !a.d_bitmask && P >= 8 && npairs >= 4096 && kprime <= 64;
    if (const char* e = getenv("LB200_SEED")) seed = seed && atoi(e) != 0;
    if (seed) {
      CK(cudaMemsetAsync(s_cnt, 0, (size_t)nlist * 8, st));
      CK(cudaMemsetAsync(s_minp, 0xff, (size_t)Q * 4, st));
      CK(cudaMemsetAsync(s_stats, 0, 32, st));
      CK(cudaMemsetAsync(s_ocnt, 0, (size_t)Q * 4, st));
      const int first_only = a.shard_mode && sp.n_peers > 0 && !getenv("LB200_SEED_ALL");
      seed_probes_kernel<<<(Q + 255) / 256, 256, 0, st>>>(probes, Q, P, ix->d_part_off, s_probe, first_only);
      count_pairs_kernel<<<(Q + 255) / 256, 256, 0, st>>>(s_probe, Q, 1, s_minp, s_cnt, s_cnt + nlist, s_rank);
      const uint32_t whole = (uint32_t)std::max<uint64_t>(16, (ix->max_part + 15) / 16 * 16);   // one item per query group
      offsets_kernel<<<1, 1024, 0, st>>>(s_cnt, s_cnt + nlist, ix->d_part_off, nlist, (uint32_t)sc.ql, whole, ix->M, s_pair_off,
                                         s_item_off, s_stats);
      scatter_pairs_kernel<<<(Q + 255) / 256, 256, 0, st>>>(s_probe, s_rank, s_pair_off, s_cnt, Q, s_pair_list);
      expand_items_kernel<<<(nlist * 32 + 255) / 256, 256, 0, st>>>(s_item_off, ix->d_part_off, nlist, whole, s_desc, nullptr);
      CK(cudaGetLastError());
      Scan3Params ss = sp;
      ss.P = 1;                                   // pair id == query
      ss.pair_list = s_pair_list; ss.part_pair_off = s_pair_off; ss.item_off = s_item_off; ss.item_counter = s_counter;
      ss.item_desc = s_desc;
      uint32_t seed_rows = 2048;
      if (const char* e = getenv("LB200_SEED_ROWS")) seed_rows = std::max(1024u, (uint32_t)atoi(e) / 1024u * 1024u);
      ss.chunk_vecs = seed_rows; ss.max_chunks = 1;    // chunk 0 only: rows [0, seed_rows) of the partition
      ss.no_boot = 0; ss.boot2 = 1;
      ss.lut_src = nullptr; ss.lut_max = nullptr; ss.lut_image = 0;
      ss.out_dist = s_odist; ss.out_pos = s_opos; ss.out_rowid = s_orow; ss.out_count = s_ocnt;
      ScanConfig scs = sc;
      CK(launch_scan3(scs, true, ix->sm_count, st, ss, s_item_off + nlist));
      launches += 7;
    }
    // pooled bootstrap when most work items are cold: a shard (few of a query's probes are local), few
    // probes per query, or per-pair outputs (no cross-item bound at all)
    bool boot = (a.shard_mode && !seed) || a.per_pair_output || P < 16;
    if (const char* eb = getenv("LB200_BOOT")) boot = atoi(eb) != 0;
    if (!boot) sp.no_boot = 1;
    // wide CTAs (16 warps, pre-built tables) exist as the BOOT instantiation only; its calls are off via no_boot
    CK(launch_scan3(sc, boot || sc.nw == 16, ix->sm_count, st, sp, total_items_p));
    launches++;
  } else {
  ScanParams sp = {};
  sp.codes = ix->d_codes; sp.tile_off = ix->d_tile_off; sp.part_off = ix->d_part_off; sp.row_ids = ix->d_row_ids;
  sp.centroids = ix->d_centroids; sp.codebook = ix->d_codebook;
  sp.dim = dim; sp.M = ix->M; sp.dsub = ix->dsub; sp.nlist = nlist;
  sp.residual = ix->metric != LB200_DOT;
  sp.queries = keys; sp.Q = Q; sp.P = P; sp.kprime = kprime;
  sp.pair_list = d_pair_list; sp.part_pair_off = d_pair_off; sp.item_off = d_item_off; sp.item_counter = d_counter;
  sp.chunk_vecs = chunk_vecs; sp.max_chunks = max_chunks; sp.capw = sc.capw;
  sp.lower = a.lower; sp.upper = a.upper; sp.has_lower = a.has_lower; sp.has_upper = a.has_upper;
  sp.bitmask = a.d_bitmask;
  sp.out_dist = d_sdist; sp.out_pos = d_spos; sp.out_rowid = d_srow; sp.out_count = d_scnt;
  if (prof) CK(cudaEventRecord(cx.ev[2], st));
  CK(launch_scan(sc, ix->sm_count, st, sp, total_items_p));
  launches++;
  }
  if (prof) CK(cudaEventRecord(cx.ev[3], st));
  // ---- A7 merge ---------------------------------------------------------------------------
  if (a.per_pair_output) {
    // search_in_partition semantics: one list per (query, partition); merge only the chunks
    rc = launch_merge(d_sdist, d_srow, nullptr, nullptr, d_scnt, (uint32_t)npairs, max_chunks, kprime,
                      (size_t)max_chunks * kprime, kprime, max_chunks, 1, kprime, kprime, a.d_out_dist,
                      a.d_out_rowid, nullptr, nullptr, a.d_out_count, st);
    if (rc) return rc;
    launches++;
  } else {
    const uint32_t L = P * max_chunks;
    const bool direct = !refine && !a.shard_mode;  // merged list is the final answer
    if (mslice) {
      rc = launch_merge2(d_sdist, d_srow, d_spos, d_scnt, Q, L, kprime, t_mdist, t_mrow, t_mpos, t_mcnt,
                         direct ? a.d_out_dist : d_mdist, direct ? a.d_out_rowid : d_mrow, d_mpos,
                         direct ? a.d_out_count : d_mcnt, st);
      launches++;
    } else
    rc = launch_merge(d_sdist, d_srow, d_spos, nullptr, d_scnt, Q, L, kprime, (size_t)L * kprime, kprime, L, 1,
                      kprime, kprime, direct ? a.d_out_dist : d_mdist, direct ? a.d_out_rowid : d_mrow, d_mpos,
                      nullptr, direct ? a.d_out_count : d_mcnt, st);
    if (rc) return rc;
    launches++;
    if (!direct) {
      // ---- A8 refine ------------------------------------------------------------------------
      float* exact = a.shard_mode ? a.d_out_exact : d_exact;
      if (a.refine_factor > 0 && ix->d_raw16) {
        refine_distance_f16_kernel<<<(uint32_t)(((size_t)Q * kprime + 127) / 128), 128, 0, st>>>(
            ix->metric, a.d_queries, ix->d_raw16, dim, d_mpos, d_mcnt, Q, kprime, exact);
        CK(cudaGetLastError());
        launches++;
      } else if (a.refine_factor > 0 && ix->d_raw) {
        refine_distance_kernel<<<(uint32_t)(((size_t)Q * kprime + 127) / 128), 128, 0, st>>>(
            ix->metric, a.d_queries, ix->d_raw, dim, d_mpos, d_mcnt, Q, kprime, exact);
        CK(cudaGetLastError());
        launches++;
      } else if (a.shard_mode) {
        CK(cudaMemsetAsync(exact, 0xff, (size_t)Q * kprime * 4, st));  // NaN
      }
      if (a.shard_mode) {
        // (the outputs may live in the ROOT shard's gather buffer on another GPU: peer memory, api_group.inl)
        CK(cudaMemcpyAsync(a.d_out_dist, d_mdist, (size_t)Q * kprime * 4, cudaMemcpyDefault, st));
        CK(cudaMemcpyAsync(a.d_out_rowid, d_mrow, (size_t)Q * kprime * 8, cudaMemcpyDefault, st));
        CK(cudaMemcpyAsync(a.d_out_count, d_mcnt, (size_t)Q * 4, cudaMemcpyDefault, st));
      } else {
        rerank_kernel<<<(Q * 32 + 127) / 128, 128, 0, st>>>(exact, d_mrow, d_mcnt, Q, kprime, a.k, a.has_lower, a.lower,
                                                          a.has_upper, a.upper, a.d_out_dist, a.d_out_rowid,
                                                          a.d_out_count);
        CK(cudaGetLastError());
        launches++;
      }
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
    uint32_t items = 0;
    if (a.pw) {
      hs[0] = a.pw->code_bytes; hs[1] = a.pw->pairs; items = a.pw->items;
    } else {
      CK(cudaMemcpy(hs, d_stats, 16, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(&items, d_item_off + nlist, 4, cudaMemcpyDeviceToHost));
    }
    stats.scan_code_bytes = hs[0];
    stats.pairs = hs[1];
    stats.scan_items = items;
  }
  stats.kernel_launches = launches;
  stats.scan_variant = sc.variant;
  stats.scan_ql = sc.ql;
  {
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->stats = stats;
  }
  return 0;
}

static int check_queries(const lb200_index* ix, const void* q, uint32_t Q) {
  if (!ix) return set_err(LB200_EINVAL, "null index");
  if (Q && !q) return set_err(LB200_EINVAL, "null queries");
  return 0;
}

}  // namespace

namespace {
// One-shot device scratch of the handle-less host-pointer entry points (merge_topk, rerank, flat, f16).
struct HostStage {
  void* p = nullptr;
  ~HostStage() { if (p) cudaFree(p); }
};
}  // namespace

namespace {
// Capture-once / replay of the device side of a small host-pointer call (see CallCtx::graphs).
struct GraphRun {
  CallCtx& cx;
  cudaStream_t st;
  std::vector<uint64_t> key;
  bool capturing = false;
  static uint64_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
  // 1: replayed (the device work is enqueued); 0: the caller enqueues it (eagerly, or into a capture); < 0: error
  int begin(bool eligible) {
    if (!eligible) return 0;
    if (cudaGraphExec_t exec = cx.find_graph(key)) {
      CK(cudaGraphLaunch(exec, st));
      return 1;
    }
    if (!cx.seen_before(key)) return 0;  // first call of a shape runs eagerly: it sizes the context's buffers
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    capturing = true;
    return 0;
  }
  int end(int rc) {
    if (!capturing) return rc;
    capturing = false;
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture(st, &g);
    if (rc) {
      if (g) cudaGraphDestroy(g);
      cudaGetLastError();
      return rc;
    }
    if (e != cudaSuccess || !g) return set_err(LB200_ECUDA, "graph capture failed: %s", cudaGetErrorString(e));
    cudaGraphExec_t exec = nullptr;
    e = cudaGraphInstantiate(&exec, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) return set_err(LB200_ECUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e));
    cx.store_graph(key, exec);
    CK(cudaGraphLaunch(exec, st));
    return 0;
  }
};
static bool graphs_enabled(const lb200_index* ix) { return !ix->profiling && !getenv("LB200_NO_GRAPHS"); }
}  // namespace

// Host<->device staging of the host-pointer entry points.  Small payloads go through the context's pinned
// buffer (one async copy, no driver-side staging of pageable memory); large ones are copied straight from /
// to the caller's memory (pinned by the caller in the benchmarks).
constexpr size_t LB200_PIN_BYTES = 256 * 1024;

extern "C" int32_t lb200_find_partitions(lb200_index* ix, const float* queries, uint32_t Q, uint32_t nprobes,
                                         uint32_t* out_part_ids, float* out_part_dists_or_null) {
  int rc = check_queries(ix, queries, Q);
  if (rc) return rc;
  if (!out_part_ids) return set_err(LB200_EINVAL, "null output");
  const uint32_t P = std::min(nprobes, ix->nlist);
  if (!Q || !P) return 0;
  CtxLease ls;
  if ((rc = ls.acquire(ix, nullptr))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;
  const size_t nq = (size_t)Q * ix->dim * 4, np = (size_t)Q * P * 4;
  size_t need = Arena::pad(nq) + Arena::pad((size_t)Q * ix->nlist * 4) + Arena::pad(np) * 2 + 2048;
  if ((rc = cx.arena.reserve(need))) return rc;
  cx.arena.reset();
  float* d_q = cx.arena.take<float>((size_t)Q * ix->dim);
  float* d_D = cx.arena.take<float>((size_t)Q * ix->nlist);
  uint32_t* d_p = cx.arena.take<uint32_t>((size_t)Q * P);   // d_p and d_pd are adjacent (pad(np) apart)
  float* d_pd = cx.arena.take<float>((size_t)Q * P);
  const bool pin = nq + 2 * Arena::pad(np) <= LB200_PIN_BYTES;
  uint32_t launches = 0;
  if (pin) {
    if ((rc = cx.pin_reserve(nq + 2 * Arena::pad(np)))) return rc;
    memcpy(cx.h_pin, queries, nq);
    const size_t span = out_part_dists_or_null ? Arena::pad(np) + np : np;
    GraphRun gr{cx, st, {1, Q, P, span, (uint64_t)(uintptr_t)cx.arena.base, (uint64_t)(uintptr_t)cx.h_pin}};
    int r = gr.begin(graphs_enabled(ix));
    if (r < 0) return r;
    if (r == 0) {
      auto enqueue = [&]() -> int {
        CK(cudaMemcpyAsync(d_q, cx.h_pin, nq, cudaMemcpyHostToDevice, st));
        int rc2 = coarse_device(ix, d_q, Q, P, d_D, d_p, d_pd, st, &launches);
        if (rc2) return rc2;
        CK(cudaMemcpyAsync(cx.h_pin, d_p, span, cudaMemcpyDeviceToHost, st));
        return 0;
      };
      if ((rc = gr.end(enqueue()))) return rc;
    }
    CK(cudaStreamSynchronize(st));
    memcpy(out_part_ids, cx.h_pin, np);
    if (out_part_dists_or_null) memcpy(out_part_dists_or_null, cx.h_pin + Arena::pad(np), np);
    return 0;
  }
  CK(cudaMemcpyAsync(d_q, queries, nq, cudaMemcpyHostToDevice, st));
  rc = coarse_device(ix, d_q, Q, P, d_D, d_p, d_pd, st, &launches);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out_part_ids, d_p, np, cudaMemcpyDeviceToHost, st));
  if (out_part_dists_or_null) CK(cudaMemcpyAsync(out_part_dists_or_null, d_pd, np, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

// ---- prefilter objects ----------------------------------------------------------------------------
namespace {
// the allow bits of partition p in the handle's row order.  8-bit handles keep the caller's order; 4-bit and
// IVF_FLAT handles sort a partition's rows by row id at create time (h_perm: sorted position -> caller's position)
static void put_mask_words(const lb200_index* ix, uint32_t p, const uint64_t* src, uint64_t* dst) {
  const uint64_t o = ix->h_part_off[p], n = ix->h_part_off[p + 1] - o;
  const size_t nw = (size_t)((n + 63) / 64);
  if (ix->h_perm.empty()) {
    memcpy(dst, src, nw * 8);
    return;
  }
  memset(dst, 0, nw * 8);
  for (uint64_t i = 0; i < n; i++) {
    const uint32_t s = ix->h_perm[o + i];
    if ((src[s >> 6] >> (s & 63)) & 1) dst[i >> 6] |= 1ull << (i & 63);
  }
}
}  // namespace

extern "C" int32_t lb200_prefilter_create(lb200_index* ix, const uint64_t* const* part_bitmask, lb200_prefilter** out) {
  if (!out) return set_err(LB200_EINVAL, "null output");
  *out = nullptr;
  if (!ix || !part_bitmask) return set_err(LB200_EINVAL, "null argument");
  CK(cudaSetDevice(ix->device));
  const uint32_t nlist = ix->nlist;
  size_t words = 0;
  for (uint32_t p = 0; p < nlist; p++)
    if (part_bitmask[p]) words += (size_t)((ix->h_part_off[p + 1] - ix->h_part_off[p] + 63) / 64) + 1;
  const size_t tbl = Arena::pad((size_t)nlist * 8), total = tbl + words * 8;
  std::vector<uint64_t> h;
  try {
    h.assign(total / 8, 0);
  } catch (...) {
    return set_err(LB200_ENOMEM, "out of host memory for %zu prefilter words", words);
  }
  lb200_prefilter* pf = new lb200_prefilter();
  pf->ix = ix;
  cudaError_t e = cudaMalloc((void**)&pf->base, std::max<size_t>(total, 8));
  if (e != cudaSuccess) {
    delete pf;
    return set_err(LB200_ENOMEM, "cudaMalloc(%zu) failed: %s", total, cudaGetErrorString(e));
  }
  size_t w = tbl / 8;
  for (uint32_t p = 0; p < nlist; p++) {
    if (!part_bitmask[p]) continue;
    const size_t nw = (size_t)((ix->h_part_off[p + 1] - ix->h_part_off[p] + 63) / 64);
    h[p] = (uint64_t)(uintptr_t)(pf->base + w * 8);
    put_mask_words(ix, p, part_bitmask[p], h.data() + w);  // + one zero word: the scan may test one row past a ragged end
    w += nw + 1;
  }
  e = cudaMemcpy(pf->base, h.data(), total, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cudaFree(pf->base);
    delete pf;
    return set_err(LB200_ECUDA, "prefilter upload failed: %s", cudaGetErrorString(e));
  }
  pf->d_ptrs = reinterpret_cast<const uint64_t**>(pf->base);
  *out = pf;
  return 0;
}

extern "C" void lb200_prefilter_destroy(lb200_prefilter* pf) {
  if (!pf) return;
  if (pf->ix) cudaSetDevice(pf->ix->device);
  cudaFree(pf->base);
  delete pf;
}

namespace {

// search_in_partition for explicit pairs; masks either as host pointers (uploaded with this call, only the
// partitions the call probes) or as a device-resident lb200_prefilter.
static int search_partitions_impl(lb200_index* ix, const float* queries, uint32_t Q, const uint32_t* part_ids, uint32_t P,
                                  uint32_t k_prime, const float* lower_or_null, const float* upper_or_null,
                                  const uint64_t* const* host_masks, const lb200_prefilter* pf, float* out_dist,
                                  uint64_t* out_rowid, uint32_t* out_count) {
  int rc = check_queries(ix, queries, Q);
  if (rc) return rc;
  if (Q && P && (!part_ids || !out_dist || !out_rowid || !out_count)) return set_err(LB200_EINVAL, "null argument");
  if (k_prime == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (k_prime > 512) return set_err(LB200_EUNSUPPORTED, "k=%u > 512", k_prime);
  if (!Q || !P) return 0;
  if (pf && pf->ix != ix) return set_err(LB200_EINVAL, "prefilter belongs to another index handle");
  const uint32_t nlist = ix->nlist;
  const size_t npairs = (size_t)Q * P;
  if (npairs >= (1ull << 31)) return set_err(LB200_EUNSUPPORTED, "Q*P too large");
  for (size_t i = 0; i < npairs; i++)
    if (part_ids[i] >= nlist)  // ivf/v2.rs:288-293 "partition centroid {} does not exist"
      return set_err(LB200_EINVAL, "partition centroid %u does not exist", part_ids[i]);
  CtxLease ls;
  if ((rc = ls.acquire(ix, nullptr))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;

  // ---- plan; small calls get their work list from the host ----
  ScanPlan plan = {};
  bool small = ix->bits == 8 && npairs <= 2048 && !getenv("LB200_NO_HOST_WORKLIST");
  if (ix->bits == 8) {
    if ((rc = plan_scan(ix, Q, P, k_prime, &plan))) return rc;
    small = small && plan.sc.variant == 3;
  }
  // ---- staging layout (device `stage` mirrors the pinned host buffer byte for byte on the way in) ----
  const size_t nq = (size_t)Q * ix->dim * 4, np = npairs * 4, no = npairs * k_prime;
  size_t in_bytes = 0;
  auto put = [&](size_t b) { size_t o = in_bytes; in_bytes += Arena::pad(b); return o; };
  const size_t o_q = put(nq), o_p = put(np);
  size_t o_pl = 0, o_po = 0, o_io = 0, o_ctr = 0, o_tbl = 0, o_masks = 0, mask_words = 0;
  size_t o_desc = 0;
  if (small) {
    o_pl = put(np); o_po = put((size_t)(nlist + 1) * 4); o_io = put((size_t)(nlist + 1) * 4); o_ctr = put(16);
    o_desc = put(npairs * plan.max_chunks * 16);  // items <= pairs * chunks
  }
  std::vector<uint32_t> mparts;  // probed partitions that carry a host mask
  if (host_masks) {
    for (size_t i = 0; i < npairs; i++)
      if (host_masks[part_ids[i]]) mparts.push_back(part_ids[i]);
    std::sort(mparts.begin(), mparts.end());
    mparts.erase(std::unique(mparts.begin(), mparts.end()), mparts.end());
    for (uint32_t p : mparts) mask_words += (size_t)((ix->h_part_off[p + 1] - ix->h_part_off[p] + 63) / 64) + 1;
    o_tbl = put((size_t)nlist * 8);
    o_masks = put(mask_words * 8);
  }
  const size_t o_od = in_bytes, o_or = o_od + Arena::pad(no * 4), o_oc = o_or + Arena::pad(no * 8);
  const size_t out_bytes = o_oc + Arena::pad(np) - o_od;
  if ((rc = cx.stage.reserve(in_bytes + out_bytes))) return rc;
  char* dbase = cx.stage.base;
  const bool pin_in = in_bytes <= 4 * LB200_PIN_BYTES, pin_out = out_bytes <= LB200_PIN_BYTES;
  if ((rc = cx.pin_reserve((pin_in ? in_bytes : 0) + (pin_out ? out_bytes : 0)))) return rc;
  char* hin = cx.h_pin;
  char* hout = cx.h_pin + (pin_in ? in_bytes : 0);
  std::vector<char> pageable;
  if (!pin_in) {  // large call: build the small tables in a pageable buffer, copy queries straight from the caller
    pageable.resize(in_bytes - Arena::pad(nq));
    hin = pageable.data() - Arena::pad(nq);
  }
  if (pin_in) memcpy(hin + o_q, queries, nq);
  memcpy(hin + o_p, part_ids, np);

  PrebuiltWork pw = {};
  if (small) {
    uint32_t* pl = reinterpret_cast<uint32_t*>(hin + o_pl);
    uint32_t* po = reinterpret_cast<uint32_t*>(hin + o_po);
    uint32_t* io = reinterpret_cast<uint32_t*>(hin + o_io);
    memset(po, 0, (size_t)(nlist + 1) * 4);
    for (size_t i = 0; i < npairs; i++) po[part_ids[i]]++;
    uint32_t pacc = 0, iacc = 0;
    uint64_t bytes = 0;
    const uint32_t ql = (uint32_t)plan.sc.ql;
    uint32_t* desc = reinterpret_cast<uint32_t*>(hin + o_desc);
    for (uint32_t p = 0; p < nlist; p++) {  // == offsets_kernel + expand_items_kernel
      const uint32_t c = po[p];
      const uint64_t n = ix->h_part_off[p + 1] - ix->h_part_off[p];
      po[p] = pacc;
      io[p] = iacc;
      if (c) {
        const uint32_t nch = (uint32_t)((n + plan.chunk_vecs - 1) / plan.chunk_vecs), ni = ((c + ql - 1) / ql) * nch;
        for (uint32_t local = 0; local < ni; local++) {
          uint32_t* de = desc + (size_t)(iacc + local) * 4;
          de[0] = p; de[1] = local / nch; de[2] = local % nch; de[3] = 0;
        }
        pacc += c;
        iacc += ni;
        bytes += (uint64_t)c * n * ix->M;
      }
    }
    po[nlist] = pacc;
    io[nlist] = iacc;
    cx.h_work.assign(nlist, 0);  // fill position per partition
    for (size_t i = 0; i < npairs; i++) pl[po[part_ids[i]] + cx.h_work[part_ids[i]]++] = (uint32_t)i;  // == scatter_pairs_kernel
    memset(hin + o_ctr, 0, 16);
    pw.plan = plan;
    pw.d_pair_list = reinterpret_cast<uint32_t*>(dbase + o_pl);
    pw.d_pair_off = reinterpret_cast<uint32_t*>(dbase + o_po);
    pw.d_item_off = reinterpret_cast<uint32_t*>(dbase + o_io);
    pw.d_counter = reinterpret_cast<uint32_t*>(dbase + o_ctr);
    pw.d_item_desc = reinterpret_cast<const uint4*>(dbase + o_desc);
    pw.code_bytes = bytes;
    pw.pairs = pacc;
    pw.items = iacc;
  }
  if (host_masks) {
    uint64_t* tbl = reinterpret_cast<uint64_t*>(hin + o_tbl);
    uint64_t* mw = reinterpret_cast<uint64_t*>(hin + o_masks);
    memset(tbl, 0, (size_t)nlist * 8);
    size_t w = 0;
    for (uint32_t p : mparts) {
      const size_t nw = (size_t)((ix->h_part_off[p + 1] - ix->h_part_off[p] + 63) / 64);
      tbl[p] = (uint64_t)(uintptr_t)(dbase + o_masks + w * 8);
      put_mask_words(ix, p, host_masks[p], mw + w);
      mw[w + nw] = 0;
      w += nw + 1;
    }
  }
  SearchArgs a = {};
  a.d_queries = reinterpret_cast<const float*>(dbase + o_q); a.Q = Q;
  a.d_part_ids = reinterpret_cast<const uint32_t*>(dbase + o_p); a.P = P; a.k = k_prime; a.refine_factor = 0;
  a.normalize = false; a.per_pair_output = true;
  a.has_lower = lower_or_null != nullptr; a.has_upper = upper_or_null != nullptr;
  a.lower = lower_or_null ? *lower_or_null : 0.f; a.upper = upper_or_null ? *upper_or_null : 0.f;
  a.d_bitmask = host_masks ? reinterpret_cast<const uint64_t* const*>(dbase + o_tbl) : pf ? pf->d_ptrs : nullptr;
  a.d_out_dist = reinterpret_cast<float*>(dbase + o_od);
  a.d_out_rowid = reinterpret_cast<uint64_t*>(dbase + o_or);
  a.d_out_count = reinterpret_cast<uint32_t*>(dbase + o_oc);
  a.pw = small ? &pw : nullptr;
  if (pin_in && pin_out) {
    // fixed device-side sequence for this shape and these buffers: captured once, replayed afterwards
    GraphRun gr{cx, st, {2, Q, P, k_prime, (uint64_t)a.has_lower, GraphRun::fbits(a.lower), (uint64_t)a.has_upper,
                         GraphRun::fbits(a.upper), (uint64_t)(uintptr_t)(pf ? pf->d_ptrs : nullptr), (uint64_t)small,
                         in_bytes, (uint64_t)(uintptr_t)cx.arena.base, (uint64_t)(uintptr_t)dbase,
                         (uint64_t)(uintptr_t)cx.h_pin}};
    // which partitions are probed reaches the kernels only through device memory (the uploaded work list); the
    // chunk plan and every launch dimension depend on (Q, P, k') alone.  Host masks change the upload size per
    // call: not captured (lb200_prefilter objects are).
    int r = gr.begin(graphs_enabled(ix) && small && !host_masks);
    if (r < 0) return r;
    if (r == 0) {
      auto enqueue = [&]() -> int {
        CK(cudaMemcpyAsync(dbase, hin, in_bytes, cudaMemcpyHostToDevice, st));
        int rc2 = search_pipeline(ix, cx, a, st);
        if (rc2) return rc2;
        CK(cudaMemcpyAsync(hout, dbase + o_od, out_bytes, cudaMemcpyDeviceToHost, st));
        return 0;
      };
      if ((rc = gr.end(enqueue()))) return rc;
    }
    CK(cudaStreamSynchronize(st));
    memcpy(out_dist, hout, no * 4);
    memcpy(out_rowid, hout + (o_or - o_od), no * 8);
    memcpy(out_count, hout + (o_oc - o_od), np);
    return 0;
  }
  if (pin_in) {
    CK(cudaMemcpyAsync(dbase, hin, in_bytes, cudaMemcpyHostToDevice, st));
  } else {
    CK(cudaMemcpyAsync(dbase + o_q, queries, nq, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(dbase + o_p, hin + o_p, in_bytes - Arena::pad(nq), cudaMemcpyHostToDevice, st));
  }
  rc = search_pipeline(ix, cx, a, st);
  if (rc) return rc;
  if (pin_out) {
    CK(cudaMemcpyAsync(hout, dbase + o_od, out_bytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    memcpy(out_dist, hout, no * 4);
    memcpy(out_rowid, hout + (o_or - o_od), no * 8);
    memcpy(out_count, hout + (o_oc - o_od), np);
    return 0;
  }
  CK(cudaMemcpyAsync(out_dist, dbase + o_od, no * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_rowid, dbase + o_or, no * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_count, dbase + o_oc, np, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

}  // namespace

extern "C" int32_t lb200_search_partitions(lb200_index* ix, const float* queries, uint32_t Q,
                                           const uint32_t* part_ids, uint32_t P, uint32_t k_prime,
                                           const float* lower_or_null, const float* upper_or_null,
                                           const uint64_t* const* part_bitmask_or_null, float* out_dist,
                                           uint64_t* out_rowid, uint32_t* out_count) {
  return search_partitions_impl(ix, queries, Q, part_ids, P, k_prime, lower_or_null, upper_or_null,
                                part_bitmask_or_null, nullptr, out_dist, out_rowid, out_count);
}

extern "C" int32_t lb200_search_partitions_filtered(lb200_index* ix, const float* queries, uint32_t Q,
                                                    const uint32_t* part_ids, uint32_t P, uint32_t k_prime,
                                                    const float* lower_or_null, const float* upper_or_null,
                                                    const lb200_prefilter* prefilter_or_null, float* out_dist,
                                                    uint64_t* out_rowid, uint32_t* out_count) {
  return search_partitions_impl(ix, queries, Q, part_ids, P, k_prime, lower_or_null, upper_or_null, nullptr,
                                prefilter_or_null, out_dist, out_rowid, out_count);
}

namespace {
static int search_device_impl(lb200_index* ix, const float* d_queries, uint32_t Q, uint32_t nprobes,
                              const uint32_t* d_part_ids, uint32_t P, uint32_t k, uint32_t refine_factor,
                              const float* lower_or_null, const float* upper_or_null, float* d_out_dist,
                              uint64_t* d_out_rowid, uint32_t* d_out_count, float* d_out_exact_or_null, void* stream) {
  CtxLease ls;
  int rc = ls.acquire(ix, (cudaStream_t)stream);
  if (rc) return rc;
  SearchArgs a = {};
  a.d_queries = d_queries; a.Q = Q; a.nprobes = nprobes; a.d_part_ids = d_part_ids; a.P = P; a.k = k;
  a.refine_factor = refine_factor;
  a.normalize = true; a.shard_mode = d_out_exact_or_null != nullptr;
  a.has_lower = lower_or_null != nullptr; a.has_upper = upper_or_null != nullptr;
  a.lower = lower_or_null ? *lower_or_null : 0.f; a.upper = upper_or_null ? *upper_or_null : 0.f;
  a.d_out_dist = d_out_dist; a.d_out_rowid = d_out_rowid; a.d_out_count = d_out_count;
  a.d_out_exact = d_out_exact_or_null;
  rc = search_pipeline(ix, *ls.cx, a, ls.st);
  if (rc) return rc;
  if (!stream) CK(cudaStreamSynchronize(ls.st));
  return 0;
}
}  // namespace

extern "C" int32_t lb200_search_device(lb200_index* ix, const float* d_queries, uint32_t Q, uint32_t nprobes,
                                       uint32_t k, uint32_t refine_factor, const float* lower_or_null,
                                       const float* upper_or_null, float* d_out_dist, uint64_t* d_out_rowid,
                                       uint32_t* d_out_count, float* d_out_exact_or_null, void* stream) {
  int rc = check_queries(ix, d_queries, Q);
  if (rc) return rc;
  if (Q && (!d_out_dist || !d_out_rowid || !d_out_count)) return set_err(LB200_EINVAL, "null output");
  if (nprobes == 0) return set_err(LB200_EINVAL, "nprobes must be > 0");
  return search_device_impl(ix, d_queries, Q, nprobes, nullptr, 0, k, refine_factor, lower_or_null, upper_or_null,
                            d_out_dist, d_out_rowid, d_out_count, d_out_exact_or_null, stream);
}

extern "C" int32_t lb200_find_partitions_device(lb200_index* ix, const float* d_queries, uint32_t Q, uint32_t nprobes,
                                                int32_t normalize_if_cosine, uint32_t* d_out_part_ids,
                                                void* stream) {
  int rc = check_queries(ix, d_queries, Q);
  if (rc) return rc;
  if (!d_out_part_ids) return set_err(LB200_EINVAL, "null output");
  if (nprobes == 0) return set_err(LB200_EINVAL, "nprobes must be > 0");
  const uint32_t P = std::min(nprobes, ix->nlist);
  if (!Q) return 0;
  CtxLease ls;
  if ((rc = ls.acquire(ix, (cudaStream_t)stream))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;
  size_t need = Arena::pad((size_t)Q * ix->dim * 4) + Arena::pad((size_t)Q * ix->nlist * 4) + 1024;
  rc = cx.arena.reserve(need);
  if (rc) return rc;
  cx.arena.reset();
  float* d_keys = cx.arena.take<float>((size_t)Q * ix->dim);
  float* d_D = cx.arena.take<float>((size_t)Q * ix->nlist);
  const float* keys = d_queries;
  uint32_t launches = 0;
  if (normalize_if_cosine && ix->metric == LB200_COSINE) {
    normalize_kernel<<<(Q * 32 + 255) / 256, 256, 0, st>>>(d_queries, d_keys, Q, ix->dim);
    CK(cudaGetLastError());
    keys = d_keys;
  }
  rc = coarse_device(ix, keys, Q, P, d_D, d_out_part_ids, nullptr, st, &launches);
  if (rc) return rc;
  if (!stream) CK(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int32_t lb200_search_probed_device(lb200_index* ix, const float* d_queries, uint32_t Q,
                                              const uint32_t* d_part_ids, uint32_t P, uint32_t k,
                                              uint32_t refine_factor, const float* lower_or_null,
                                              const float* upper_or_null, float* d_out_dist, uint64_t* d_out_rowid,
                                              uint32_t* d_out_count, float* d_out_exact_or_null, void* stream) {
  int rc = check_queries(ix, d_queries, Q);
  if (rc) return rc;
  if (Q && (!d_out_dist || !d_out_rowid || !d_out_count || !d_part_ids)) return set_err(LB200_EINVAL, "null argument");
  if (P == 0 || P > ix->nlist) return set_err(LB200_EINVAL, "P must be in [1, nlist]");
  return search_device_impl(ix, d_queries, Q, 0, d_part_ids, P, k, refine_factor, lower_or_null, upper_or_null,
                            d_out_dist, d_out_rowid, d_out_count, d_out_exact_or_null, stream);
}

extern "C" int32_t lb200_search(lb200_index* ix, const float* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                                uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                                float* out_dist, uint64_t* out_rowid, uint32_t* out_count) {
  int rc = check_queries(ix, queries, Q);
  if (rc) return rc;
  if (Q && (!out_dist || !out_rowid || !out_count)) return set_err(LB200_EINVAL, "null output");
  if (k == 0) return set_err(LB200_EINVAL, "k must be > 0");
  if (nprobes == 0) return set_err(LB200_EINVAL, "nprobes must be > 0");
  if (!Q) return 0;
  const bool refined = refine_factor > 0 && ix->d_raw;
  if (refine_factor > 0 && !ix->d_raw) return set_err(LB200_EINVAL, "refine_factor>0 needs raw vectors in the index handle");
  const uint32_t ostride = refined ? k : k * std::max(refine_factor, 1u);
  CtxLease ls;
  if ((rc = ls.acquire(ix, nullptr))) return rc;
  CallCtx& cx = *ls.cx;
  cudaStream_t st = ls.st;
  // staging for queries/results: grow-only buffers owned by the call context (no allocation per call)
  const size_t nq = (size_t)Q * ix->dim * 4, no = (size_t)Q * ostride;
  const size_t o_od = Arena::pad(nq), o_or = o_od + Arena::pad(no * 4), o_oc = o_or + Arena::pad(no * 8);
  const size_t total = o_oc + Arena::pad((size_t)Q * 4);
  if ((rc = cx.stage.reserve(total))) return rc;
  char* b = cx.stage.base;
  const bool pin = total <= LB200_PIN_BYTES;
  SearchArgs a = {};
  a.d_queries = reinterpret_cast<const float*>(b); a.Q = Q; a.nprobes = nprobes; a.k = k; a.refine_factor = refine_factor;
  a.normalize = true;
  a.has_lower = lower_or_null != nullptr; a.has_upper = upper_or_null != nullptr;
  a.lower = lower_or_null ? *lower_or_null : 0.f; a.upper = upper_or_null ? *upper_or_null : 0.f;
  a.d_out_dist = reinterpret_cast<float*>(b + o_od);
  a.d_out_rowid = reinterpret_cast<uint64_t*>(b + o_or);
  a.d_out_count = reinterpret_cast<uint32_t*>(b + o_oc);
  if (pin) {
    if ((rc = cx.pin_reserve(total))) return rc;
    memcpy(cx.h_pin, queries, nq);
    GraphRun gr{cx, st, {3, Q, nprobes, k, refine_factor, (uint64_t)a.has_lower, GraphRun::fbits(a.lower),
                         (uint64_t)a.has_upper, GraphRun::fbits(a.upper), (uint64_t)(uintptr_t)cx.arena.base,
                         (uint64_t)(uintptr_t)b, (uint64_t)(uintptr_t)cx.h_pin}};
    int r = gr.begin(graphs_enabled(ix) && ix->bits == 8);
    if (r < 0) return r;
    if (r == 0) {
      auto enqueue = [&]() -> int {
        CK(cudaMemcpyAsync(b, cx.h_pin, nq, cudaMemcpyHostToDevice, st));
        int rc2 = search_pipeline(ix, cx, a, st);
        if (rc2) return rc2;
        CK(cudaMemcpyAsync(cx.h_pin + o_od, b + o_od, total - o_od, cudaMemcpyDeviceToHost, st));
        return 0;
      };
      if ((rc = gr.end(enqueue()))) return rc;
    }
    CK(cudaStreamSynchronize(st));
    memcpy(out_dist, cx.h_pin + o_od, no * 4);
    memcpy(out_rowid, cx.h_pin + o_or, no * 8);
    memcpy(out_count, cx.h_pin + o_oc, (size_t)Q * 4);
    return 0;
  }
  CK(cudaMemcpyAsync(b, queries, nq, cudaMemcpyHostToDevice, st));
  rc = search_pipeline(ix, cx, a, st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out_dist, b + o_od, no * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_rowid, b + o_or, no * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_count, b + o_oc, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int32_t lb200_merge_topk_device(int32_t device_id, const float* d_dist, const uint64_t* d_rowid,
                                           const float* d_exact_or_null, const uint32_t* d_count, uint32_t G,
                                           uint32_t Q, uint32_t stride, uint32_t kprime, uint32_t k,
                                           const float* lower_or_null, const float* upper_or_null,
                                           float* d_out_dist, uint64_t* d_out_rowid, uint32_t* d_out_count,
                                           void* stream) {
  int rc = select_device(device_id);
  if (rc) return rc;
  if (!G || !Q) return 0;
  if (!d_dist || !d_rowid || !d_count || !d_out_dist || !d_out_rowid || !d_out_count)
    return set_err(LB200_EINVAL, "null argument");
  if (kprime == 0 || k == 0 || kprime > stride * G) kprime = std::min(std::max(kprime, 1u), stride * G);
  cudaStream_t st = (cudaStream_t)stream;
  if (!d_exact_or_null) {
    uint32_t kout = std::min(k, kprime);
    rc = launch_merge(d_dist, d_rowid, nullptr, nullptr, d_count, Q, G, stride, stride, (size_t)Q * stride, 1, Q,
                      kout, k, d_out_dist, d_out_rowid, nullptr, nullptr, d_out_count, st);
    if (rc) return rc;
  } else {
    float *t_dist = nullptr, *t_exact = nullptr;
    uint64_t* t_row = nullptr;
    uint32_t* t_cnt = nullptr;
    HostStage hs;
    size_t n = (size_t)Q * kprime;
    CK(cudaMalloc(&hs.p, Arena::pad(n * 4) * 2 + Arena::pad(n * 8) + Arena::pad((size_t)Q * 4)));
    char* b = (char*)hs.p;
    t_dist = (float*)b; b += Arena::pad(n * 4);
    t_exact = (float*)b; b += Arena::pad(n * 4);
    t_row = (uint64_t*)b; b += Arena::pad(n * 8);
    t_cnt = (uint32_t*)b;
    rc = launch_merge(d_dist, d_rowid, nullptr, d_exact_or_null, d_count, Q, G, stride, stride, (size_t)Q * stride, 1,
                      Q, kprime, kprime, t_dist, t_row, nullptr, t_exact, t_cnt, st);
    if (rc) return rc;
    rerank_kernel<<<(Q * 32 + 127) / 128, 128, 0, st>>>(t_exact, t_row, t_cnt, Q, kprime, k, lower_or_null != nullptr,
                                                      lower_or_null ? *lower_or_null : 0.f,
                                                      upper_or_null != nullptr, upper_or_null ? *upper_or_null : 0.f,
                                                      d_out_dist, d_out_rowid, d_out_count);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));  // scratch is freed on return
  }
  return 0;
}

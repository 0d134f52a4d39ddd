// api.cu -- host side of liblance_b200.so: the device-resident index object and the C ABI.
//
// `B200IvfPq` mirrors IVFIndex<FlatIndex, ProductQuantizer> (rust/lance/src/index/vector/ivf/v2.rs:76)
// behind `trait VectorIndex` (rust/lance-index/src/vector.rs:117-217): find_partitions,
// search_in_partition (batched as search_partitions) and search.  No CPU fallback exists:
// every entry point fails with LB200_ECUDA when no CUDA device is usable.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/lance_b200.h"
#include "coarse.cuh"
#include "f16pq.cuh"
#include "flat.cuh"
#include "launch.h"
#include "scan4.cuh"
#include "worklist.cuh"

using namespace lb200;

static thread_local std::string g_err;
static int set_err(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return set_err(e_ == cudaErrorMemoryAllocation ? LB200_ENOMEM : LB200_ECUDA,            \
                     "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

extern "C" const char* lb200_last_error(void) { return g_err.c_str(); }
extern "C" const char* lb200_version(void) { return "lance_b200 0.1 (sm_100a)"; }

static inline uint32_t next_pow2(uint32_t v) {
  uint32_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

// Grow-only device arena; one per handle, carved per call.
struct Arena {
  char* base = nullptr;
  size_t cap = 0, used = 0;
  // Outgrown buffers are retired, not freed: cudaFree synchronises the whole device, and on a shard whose stream
  // holds a kernel waiting for a peer that this very host thread has yet to launch (one process driving several
  // GPUs, api_group.inl) that would dead-lock.  Growth is geometric, so the retired buffers stay below the live one.
  std::vector<char*> retired;
  void release() {
    for (char* p : retired) cudaFree(p);
    retired.clear();
    cudaFree(base);
    base = nullptr;
    cap = 0;
  }
  int reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (base) retired.push_back(base);
    base = nullptr;
    cap = 0;
    size_t want = bytes + (bytes >> 1);
    cudaError_t e = cudaMalloc((void**)&base, want);
    if (e != cudaSuccess) {
      cudaGetLastError();
      e = cudaMalloc((void**)&base, bytes);
      want = bytes;
    }
    if (e != cudaSuccess) return set_err(LB200_ENOMEM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    cap = want;
    return 0;
  }
  void reset() { used = 0; }
  template <typename T>
  T* take(size_t n) {
    size_t off = (used + 255) & ~(size_t)255;
    used = off + n * sizeof(T);
    return reinterpret_cast<T*>(base + off);
  }
  static size_t pad(size_t b) { return (b + 255) & ~(size_t)255; }
};

// Per-call context: stream, device scratch, host<->device staging (device side + pinned host side) and
// events.  A handle owns a small pool of them so that concurrent calls -- the reference fires up to
// get_num_compute_intensive_cpus() search_in_partition calls per query (rust/lance/src/io/exec/knn.rs:606-634)
// -- run side by side on their own streams instead of serialising on one arena.  Grow-only: no allocation
// in steady state.
struct CallCtx {
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;  // recorded after the last GPU work that touches arena / stage
  cudaEvent_t ev[6] = {};      // profiling
  Arena arena;                 // pipeline scratch
  Arena stage;                 // device side of the host-pointer entry points
  char* h_pin = nullptr;       // pinned host side
  size_t h_pin_cap = 0;
  bool busy = false;
  std::vector<uint32_t> h_work;  // host-built work list of small calls
  // CUDA graphs of small calls: the whole device side of a call (H2D from the pinned buffer, kernels, D2H into
  // it) is fixed by the call's shape and this context's buffer addresses, so it is captured the second time a
  // shape is seen and replayed afterwards -- ONE driver call per search_in_partition instead of ~6, which is
  // what bounds the drop-in pattern (nprobes calls per query from a pool of host threads all contend on the
  // driver's per-context lock).
  struct GraphEntry {
    std::vector<uint64_t> key;
    cudaGraphExec_t exec = nullptr;
    uint64_t last_use = 0;
  };
  std::vector<GraphEntry> graphs;
  std::vector<std::vector<uint64_t>> seen;  // shapes seen once (first call runs eagerly and sizes the buffers)
  uint64_t tick = 0;
  cudaGraphExec_t find_graph(const std::vector<uint64_t>& key) {
    for (auto& g : graphs)
      if (g.key == key) { g.last_use = ++tick; return g.exec; }
    return nullptr;
  }
  bool seen_before(const std::vector<uint64_t>& key) {
    for (auto& k : seen)
      if (k == key) return true;
    if (seen.size() >= 16) seen.erase(seen.begin());
    seen.push_back(key);
    return false;
  }
  void store_graph(const std::vector<uint64_t>& key, cudaGraphExec_t exec) {
    if (graphs.size() >= 8) {  // evict the least recently used
      size_t lru = 0;
      for (size_t i = 1; i < graphs.size(); i++)
        if (graphs[i].last_use < graphs[lru].last_use) lru = i;
      cudaGraphExecDestroy(graphs[lru].exec);
      graphs.erase(graphs.begin() + lru);
    }
    GraphEntry e;
    e.key = key;
    e.exec = exec;
    e.last_use = ++tick;
    graphs.push_back(e);
  }
  int pin_reserve(size_t bytes) {
    if (bytes <= h_pin_cap) return 0;
    if (h_pin) cudaFreeHost(h_pin);
    h_pin = nullptr;
    h_pin_cap = 0;
    const size_t want = std::max<size_t>(bytes + (bytes >> 2), 64 * 1024);
    cudaError_t e = cudaHostAlloc((void**)&h_pin, want, cudaHostAllocDefault);
    if (e != cudaSuccess) return set_err(LB200_ENOMEM, "cudaHostAlloc(%zu) failed: %s", want, cudaGetErrorString(e));
    h_pin_cap = want;
    return 0;
  }
};

// Device-resident prefilter (lb200_prefilter_create): the per-partition row bitmasks of one RowIdMask,
// uploaded once and shared by every search_in_partition call of a query (Arc<dyn PreFilter>).
struct lb200_prefilter {
  lb200_index* ix = nullptr;
  char* base = nullptr;            // [nlist] device pointers, then the mask words
  const uint64_t** d_ptrs = nullptr;
};

constexpr uint32_t LB200_MAX_PEERS = 15;
struct ShardComm;  // api_group.inl: the exchange block of a partition shard (multi-GPU)
struct lb200_index {
  ShardComm* comm = nullptr;
  int device = 0, metric = 0;
  uint32_t dim = 0, nlist = 0, M = 0, dsub = 0;
  uint64_t N = 0, ntiles = 0, max_part = 0;
  double avg_np = 0;  // mean rows per NON-EMPTY partition
  uint32_t nonempty = 0;  // partitions that hold rows
  float *d_centroids = nullptr, *d_codebook = nullptr, *d_raw = nullptr;
  float* d_codebook_img = nullptr;     // lane-order image of the codebook for the in-kernel table build (scan3.cuh cb_image_f4)
  bool f16 = false;                    // Float16 handle (api_f16pq.inl): centroids / codebook are exact widenings
  __half* d_raw16 = nullptr;           // its raw vectors for the refine step
  uint64_t *d_part_off = nullptr, *d_tile_off = nullptr, *d_row_ids = nullptr;
  uint32_t* d_size_order = nullptr;    // partitions in descending size: order of the scan's item queue (worklist.cuh)
  uint8_t* d_codes = nullptr;          // tile16 layout (variants 1/2), or null
  uint8_t* d_codes3 = nullptr;         // skew-stream layout (variant 3), or null
  uint8_t* d_codes4 = nullptr;         // 4-bit PQ: concat over p of [M/2][n_p], rows sorted by row id
  uint64_t* d_pad_off = nullptr;       // 4-bit PQ: [nlist+1] prefix sum of partition rows padded to 16
  uint8_t* d_tail4 = nullptr;          // 4-bit PQ: [N] 1 = row sat in the n % 16 tail of the reference's order
  uint32_t bits = 8;                   // PQ code width
  uint64_t* d_sb_off = nullptr;        // [nlist+1] first superblock per partition
  uint32_t skew_L = 0, skew_D = 0;     // superblock = 32*skew_L rows; lane delay period
  std::vector<uint64_t> h_part_off;
  std::vector<uint32_t> h_perm;   // 4-bit / IVF_FLAT handles: sorted row position -> the caller's position in its partition
  cudaStream_t stream = nullptr;  // create-time uploads and the peer-table set-up
  std::mutex mu;                  // guards stats, pthr_* and the admin calls (never held across a search)
  std::vector<CallCtx*> ctxs;     // pool of per-call contexts (LB200_CTX_POOL, default 16)
  std::mutex pool_mu;
  std::condition_variable pool_cv;
  uint32_t max_ctx = 16;
  bool profiling = false;
  int force_variant = 0;
  lb200_stats stats = {};
  int sm_count = 148;
  size_t smem_optin = 0, smem_per_sm = 0;
  // shared pruning bounds between shards (lb200_peer_thresholds_*): two tables of pthr_cap keys
  // (batch parity), the peers' addresses of both, and which of them were opened through CUDA IPC
  uint32_t* pthr_local = nullptr;
  uint32_t pthr_cap = 0, pthr_peers = 0;
  uint64_t pthr_epoch = 0;
  uint32_t** d_pthr_tbl = nullptr;  // device [2][LB200_MAX_PEERS]
  void* pthr_ipc[LB200_MAX_PEERS] = {};
};

// Lease of one CallCtx for the duration of a call.  `user` = the caller's stream for the *_device entry
// points (the work is only enqueued; `done` orders the next user of the context behind it), or null for
// the context's own stream.
struct CtxLease {
  lb200_index* ix = nullptr;
  CallCtx* cx = nullptr;
  cudaStream_t st = nullptr;
  int acquire(lb200_index* ix_, cudaStream_t user) {
    ix = ix_;
    cudaError_t e = cudaSetDevice(ix->device);
    if (e != cudaSuccess) return set_err(LB200_ECUDA, "cudaSetDevice(%d) failed: %s", ix->device, cudaGetErrorString(e));
    {
      std::unique_lock<std::mutex> lk(ix->pool_mu);
      for (;;) {
        for (CallCtx* c : ix->ctxs)
          if (!c->busy) { cx = c; break; }
        if (cx) break;
        if (ix->ctxs.size() < ix->max_ctx) {
          CallCtx* c = new CallCtx();
          cudaError_t e1 = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
          if (e1 == cudaSuccess) e1 = cudaEventCreateWithFlags(&c->done, cudaEventDisableTiming);
          for (auto& ev : c->ev)
            if (e1 == cudaSuccess) e1 = cudaEventCreate(&ev);
          if (e1 != cudaSuccess) {
            if (c->stream) cudaStreamDestroy(c->stream);
            if (c->done) cudaEventDestroy(c->done);
            for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
            delete c;
            return set_err(LB200_ECUDA, "cannot create a call context: %s", cudaGetErrorString(e1));
          }
          ix->ctxs.push_back(c);
          cx = c;
          break;
        }
        ix->pool_cv.wait(lk);
      }
      cx->busy = true;
    }
    st = user ? user : cx->stream;
    e = cudaStreamWaitEvent(st, cx->done, 0);  // earlier work on this context's scratch (maybe on another stream)
    if (e != cudaSuccess) return set_err(LB200_ECUDA, "cudaStreamWaitEvent failed: %s", cudaGetErrorString(e));
    return 0;
  }
  ~CtxLease() {
    if (!cx) return;
    cudaEventRecord(cx->done, st);
    {
      std::lock_guard<std::mutex> lk(ix->pool_mu);
      cx->busy = false;
    }
    ix->pool_cv.notify_one();
  }
};

static int select_device(int device_id) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return set_err(LB200_ECUDA, "no CUDA device available (%s); lance_b200 has no CPU fallback",
                   e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device_id < 0 || device_id >= n) return set_err(LB200_EINVAL, "device_id %d out of range [0,%d)", device_id, n);
  CK(cudaSetDevice(device_id));
  return 0;
}

namespace {
int create_index4(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist, const float* centroids, uint32_t M,
                  const float* codebook, const uint64_t* part_offsets, const uint8_t* codes, const uint64_t* row_ids,
                  const float* raw, lb200_index** out);
int create_index_flat(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist, const float* centroids,
                      const uint64_t* part_offsets, const uint64_t* row_ids, const float* raw, lb200_index** out);
}
extern "C" int32_t lb200_index_create(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist,
                                      const float* centroids, uint32_t M, uint32_t num_bits,
                                      const float* codebook, const uint64_t* part_offsets,
                                      const uint8_t* codes, const uint64_t* row_ids,
                                      const float* raw_vectors_or_null, lb200_index** out) {
  if (!out) return set_err(LB200_EINVAL, "out is null");
  *out = nullptr;
  if (num_bits == 4)
    return create_index4(device_id, metric, dim, nlist, centroids, M, codebook, part_offsets, codes, row_ids,
                         raw_vectors_or_null, out);
  if (num_bits == 32)  // IVF_FLAT: raw f32 rows, no quantizer
    return create_index_flat(device_id, metric, dim, nlist, centroids, part_offsets, row_ids, raw_vectors_or_null, out);
  if (num_bits != 8)
    return set_err(LB200_EUNSUPPORTED, "num_bits=%u: only 8-bit and 4-bit PQ are supported", num_bits);
  if (metric != LB200_L2 && metric != LB200_COSINE && metric != LB200_DOT)
    return set_err(LB200_EINVAL, "unknown metric %d", metric);
  if (!dim || !nlist || !M || dim % M) return set_err(LB200_EINVAL, "bad shape dim=%u nlist=%u M=%u", dim, nlist, M);
  if (!centroids || !codebook || !part_offsets) return set_err(LB200_EINVAL, "null index buffers");
  if (M > 192) return set_err(LB200_EUNSUPPORTED, "num_sub_vectors=%u > 192 (ADC table would exceed shared memory)", M);
  for (uint32_t p = 0; p < nlist; p++)
    if (part_offsets[p + 1] < part_offsets[p]) return set_err(LB200_EINVAL, "part_offsets not monotone at %u", p);
  if (part_offsets[0] != 0) return set_err(LB200_EINVAL, "part_offsets[0] must be 0");
  uint64_t N = part_offsets[nlist];
  if (N >= (1ull << 32)) return set_err(LB200_EUNSUPPORTED, "N=%llu rows >= 2^32", (unsigned long long)N);
  if (N && (!codes || !row_ids)) return set_err(LB200_EINVAL, "null codes/row_ids");
  int rc = select_device(device_id);
  if (rc) return rc;

  lb200_index* ix = new lb200_index();
  ix->device = device_id;
  ix->metric = metric;
  ix->dim = dim;
  ix->nlist = nlist;
  ix->M = M;
  ix->dsub = dim / M;
  ix->N = N;
  ix->h_part_off.assign(part_offsets, part_offsets + nlist + 1);
  std::vector<uint32_t> size_order(nlist);
  for (uint32_t p = 0; p < nlist; p++) size_order[p] = p;
  std::stable_sort(size_order.begin(), size_order.end(), [&](uint32_t a, uint32_t b) {
    return part_offsets[a + 1] - part_offsets[a] > part_offsets[b + 1] - part_offsets[b];
  });
  {
    uint32_t ne = 0;
    for (uint32_t p = 0; p < nlist; p++) ne += part_offsets[p + 1] > part_offsets[p];
    ix->avg_np = (double)N / std::max(1u, ne);
    ix->nonempty = ne;
  }
  std::vector<uint64_t> tile_off(nlist + 1);
  uint64_t t = 0;
  for (uint32_t p = 0; p < nlist; p++) {
    tile_off[p] = t;
    uint64_t n = part_offsets[p + 1] - part_offsets[p];
    ix->max_part = std::max(ix->max_part, n);
    if (n >= (1ull << 31)) { delete ix; return set_err(LB200_EUNSUPPORTED, "partition %u too large", p); }
    t += (n + 15) >> 4;
  }
  tile_off[nlist] = t;
  ix->ntiles = t;
  cudaDeviceProp prop;
  auto fail = [&](int code) {
    lb200_index_destroy(ix);
    return code;
  };
#define CKF(call)                                                                                  \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return fail(set_err(e_ == cudaErrorMemoryAllocation ? LB200_ENOMEM : LB200_ECUDA,            \
                          "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__)); \
  } while (0)
  CKF(cudaGetDeviceProperties(&prop, device_id));
  ix->sm_count = prop.multiProcessorCount;
  ix->smem_optin = prop.sharedMemPerBlockOptin;
  ix->smem_per_sm = prop.sharedMemPerMultiprocessor;
  CKF(cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking));
  if (const char* e = getenv("LB200_CTX_POOL")) ix->max_ctx = (uint32_t)std::min(64, std::max(1, atoi(e)));
  size_t cent_b = (size_t)nlist * dim * 4, cb_b = (size_t)M * 256 * ix->dsub * 4;
  CKF(cudaMalloc(&ix->d_centroids, cent_b));
  CKF(cudaMalloc(&ix->d_codebook, cb_b));
  CKF(cudaMalloc(&ix->d_part_off, (nlist + 1) * 8));
  size_order.resize(2 * (size_t)nlist);  // [nlist] order, then [nlist] its inverse (rank of a partition in the queue)
  for (uint32_t i = 0; i < nlist; i++) size_order[nlist + size_order[i]] = i;
  CKF(cudaMalloc(&ix->d_size_order, (size_t)nlist * 8));
  CKF(cudaMemcpy(ix->d_size_order, size_order.data(), (size_t)nlist * 8, cudaMemcpyHostToDevice));
  CKF(cudaMalloc(&ix->d_tile_off, (nlist + 1) * 8));
  CKF(cudaMalloc(&ix->d_row_ids, std::max<uint64_t>(N, 1) * 8));
  // Device code layouts.  Variant 3 (skew streams) serves every M that is a multiple of 16; the
  // tile16 layout of variants 1/2 is built otherwise, or when LB200_LAYOUTS=all asks for both
  // (A/B runs of the scan variants).
  const char* lay = getenv("LB200_LAYOUTS");
  const bool want_all = lay && !strcmp(lay, "all");
  // QL=4 tables (M*4 KB) fit up to M=48; beyond that QL=2 tables (M*2 KB) up to M=96.
  const bool want_skew = (M % 16 == 0) && M <= 96 && !(lay && !strcmp(lay, "tile16"));
  const bool want_tile = !want_skew || want_all;
  std::vector<uint64_t> sb_off(nlist + 1, 0);
  if (want_skew) {
    // rows per lane per superblock.  The scan deals ROW GROUPS (32 rows) to its warps, not superblocks, so the
    // depth no longer has to follow the partition size: the deepest one always wins (one drain unit per
    // L * M / 16 units; a partition's last superblock may be mostly padding, which the scan never reads).
    uint32_t L = 32;
    if (const char* el = getenv("LB200_SKEW_L")) {  // test hook: force the superblock depth
      int v = atoi(el);
      if (v >= 1 && v <= 256) L = (uint32_t)v;
    }
    ix->skew_L = L;
    ix->skew_D = M <= 48 ? 8 : 16;
    uint64_t sb = 0;
    for (uint32_t p = 0; p < nlist; p++) {
      sb_off[p] = sb;
      sb += (part_offsets[p + 1] - part_offsets[p] + 32 * L - 1) / (32 * L);
    }
    sb_off[nlist] = sb;
    CKF(cudaMalloc(&ix->d_sb_off, (nlist + 1) * 8));
    CKF(cudaMemcpyAsync(ix->d_sb_off, sb_off.data(), (nlist + 1) * 8, cudaMemcpyHostToDevice, ix->stream));
    size_t units = (size_t)L * (M / 16) + 1;
    size_t b3 = (size_t)sb * units * 512 + 2 * 512;  // + 2 units: the scan prefetches 2 units ahead
    CKF(cudaMalloc(&ix->d_codes3, b3));
    CKF(cudaMemsetAsync(ix->d_codes3 + (b3 - 1024), 0, 1024, ix->stream));
  }
  size_t codes_b = (size_t)(t + 64) * 16 * M;
  if (want_tile) {
    // tile16 codes, padded by one extra block of tiles so clamped loads stay in bounds
    CKF(cudaMalloc(&ix->d_codes, codes_b));
    CKF(cudaMemsetAsync(ix->d_codes, 0, codes_b, ix->stream));
  }
  CKF(cudaMemcpyAsync(ix->d_centroids, centroids, cent_b, cudaMemcpyHostToDevice, ix->stream));
  CKF(cudaMemcpyAsync(ix->d_codebook, codebook, cb_b, cudaMemcpyHostToDevice, ix->stream));
  if (M % 16 == 0 && (ix->dsub == 8 || ix->dsub == 16)) {
    CKF(cudaMalloc(&ix->d_codebook_img, cb_b));
    const uint32_t n4 = M * 256 * (ix->dsub / 4);
    cb_image_kernel<<<(n4 + 255) / 256, 256, 0, ix->stream>>>(ix->d_codebook, M, ix->dsub, reinterpret_cast<float4*>(ix->d_codebook_img));
    CKF(cudaGetLastError());
  }
  CKF(cudaMemcpyAsync(ix->d_part_off, part_offsets, (nlist + 1) * 8, cudaMemcpyHostToDevice, ix->stream));
  CKF(cudaMemcpyAsync(ix->d_tile_off, tile_off.data(), (nlist + 1) * 8, cudaMemcpyHostToDevice, ix->stream));
  if (N) {
    CKF(cudaMemcpyAsync(ix->d_row_ids, row_ids, N * 8, cudaMemcpyHostToDevice, ix->stream));
    uint8_t* d_src = nullptr;
    CKF(cudaMalloc(&d_src, (size_t)N * M));
    cudaError_t e1 = cudaMemcpyAsync(d_src, codes, (size_t)N * M, cudaMemcpyHostToDevice, ix->stream);
    if (e1 == cudaSuccess && want_tile) {
      dim3 grid(nlist, 8);
      repack_codes_kernel<<<grid, 256, 0, ix->stream>>>(d_src, ix->d_part_off, ix->d_tile_off, nlist, M, ix->d_codes);
      e1 = cudaGetLastError();
    }
    if (e1 == cudaSuccess && want_skew) {
      dim3 grid(nlist, 8);
      repack_skew_kernel<<<grid, 256, 0, ix->stream>>>(d_src, ix->d_part_off, ix->d_sb_off, M, ix->skew_L,
                                                       ix->skew_D, ix->d_codes3);
      e1 = cudaGetLastError();
    }
    if (e1 == cudaSuccess) e1 = cudaStreamSynchronize(ix->stream);
    cudaFree(d_src);
    CKF(e1);
    if (raw_vectors_or_null) {
      CKF(cudaMalloc(&ix->d_raw, (size_t)N * dim * 4));
      CKF(cudaMemcpyAsync(ix->d_raw, raw_vectors_or_null, (size_t)N * dim * 4, cudaMemcpyHostToDevice, ix->stream));
    }
  }
  CKF(cudaStreamSynchronize(ix->stream));
#undef CKF
  *out = ix;
  return 0;
}

static void shard_comm_free(lb200_index* ix);  // api_group.inl
extern "C" void lb200_index_destroy(lb200_index* ix) {
  if (!ix) return;
  cudaSetDevice(ix->device);
  shard_comm_free(ix);
  cudaFree(ix->d_centroids);
  cudaFree(ix->d_codebook);
  cudaFree(ix->d_codebook_img);
  cudaFree(ix->d_raw);
  cudaFree(ix->d_raw16);
  cudaFree(ix->d_part_off);
  cudaFree(ix->d_size_order);
  cudaFree(ix->d_tile_off);
  cudaFree(ix->d_row_ids);
  cudaFree(ix->d_codes);
  cudaFree(ix->d_codes3);
  cudaFree(ix->d_codes4);
  cudaFree(ix->d_tail4);
  cudaFree(ix->d_pad_off);
  cudaFree(ix->d_sb_off);
  for (CallCtx* c : ix->ctxs) {
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (auto& g : c->graphs) cudaGraphExecDestroy(g.exec);
    c->arena.release();
    c->stage.release();
    if (c->h_pin) cudaFreeHost(c->h_pin);
    if (c->done) cudaEventDestroy(c->done);
    for (auto& e : c->ev)
      if (e) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
  }
  ix->ctxs.clear();
  lb200_peer_thresholds_detach(ix);
  cudaFree(ix->pthr_local);
  cudaFree(ix->d_pthr_tbl);
  if (ix->stream) cudaStreamDestroy(ix->stream);
  delete ix;
}

// ---- shared pruning bounds between shards (include/lance_b200.h) ---------------------------------
extern "C" int32_t lb200_peer_thresholds_export(lb200_index* ix, uint32_t max_queries, void* handle_out_64,
                                                uint64_t* dev_ptr_out) {
  if (!ix || !max_queries) return set_err(LB200_EINVAL, "null index or max_queries == 0");
  std::lock_guard<std::mutex> lk(ix->mu);
  CK(cudaSetDevice(ix->device));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (ix->pthr_peers) return set_err(LB200_EINVAL, "peer tables are attached; detach first");
  if (ix->pthr_local && ix->pthr_cap < max_queries) {
    CK(cudaFree(ix->pthr_local));
    ix->pthr_local = nullptr;
  }
  if (!ix->pthr_local) {
    // plain cudaMalloc: pooled / stream-ordered allocations cannot be exported over IPC
    CK(cudaMalloc(&ix->pthr_local, (size_t)2 * max_queries * 4));
    ix->pthr_cap = max_queries;
  }
  fill_u32_kernel<<<(2 * ix->pthr_cap + 255) / 256, 256, 0, ix->stream>>>(ix->pthr_local, KEY_PINF,
                                                                          (size_t)2 * ix->pthr_cap);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(ix->stream));
  if (handle_out_64) {
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, ix->pthr_local));
    memcpy(handle_out_64, &h, 64);
  }
  if (dev_ptr_out) *dev_ptr_out = (uint64_t)(uintptr_t)ix->pthr_local;
  return 0;
}

extern "C" int32_t lb200_peer_thresholds_detach(lb200_index* ix) {
  if (!ix) return set_err(LB200_EINVAL, "null index");
  cudaSetDevice(ix->device);
  for (auto& p : ix->pthr_ipc)
    if (p) {
      cudaIpcCloseMemHandle(p);
      p = nullptr;
    }
  ix->pthr_peers = 0;
  ix->pthr_epoch = 0;
  return 0;
}

extern "C" int32_t lb200_peer_thresholds_attach(lb200_index* ix, uint32_t world, uint32_t self,
                                                const void* handles_or_null,
                                                const uint64_t* same_process_ptrs_or_null) {
  if (!ix) return set_err(LB200_EINVAL, "null index");
  if (!ix->pthr_local) return set_err(LB200_EINVAL, "call lb200_peer_thresholds_export first");
  if (world < 1 || world > LB200_MAX_PEERS + 1 || self >= world)
    return set_err(LB200_EINVAL, "world=%u self=%u out of range (at most %u shards)", world, self, LB200_MAX_PEERS + 1);
  if (!handles_or_null && !same_process_ptrs_or_null) return set_err(LB200_EINVAL, "no peer handles given");
  lb200_peer_thresholds_detach(ix);
  std::lock_guard<std::mutex> lk(ix->mu);
  CK(cudaSetDevice(ix->device));
  uint32_t* tbl[2 * LB200_MAX_PEERS] = {};
  uint32_t n = 0;
  for (uint32_t r = 0; r < world; r++) {
    if (r == self) continue;
    uint32_t* base = nullptr;
    if (same_process_ptrs_or_null) {
      base = reinterpret_cast<uint32_t*>((uintptr_t)same_process_ptrs_or_null[r]);
      if (!base) return set_err(LB200_EINVAL, "peer %u: null table address", r);
    } else {
      cudaIpcMemHandle_t h;
      memcpy(&h, (const char*)handles_or_null + (size_t)r * 64, 64);
      void* p = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        for (uint32_t j = 0; j < n; j++) {
          cudaIpcCloseMemHandle(ix->pthr_ipc[j]);
          ix->pthr_ipc[j] = nullptr;
        }
        return set_err(LB200_ECUDA, "cudaIpcOpenMemHandle(peer %u) failed: %s", r, cudaGetErrorString(e));
      }
      ix->pthr_ipc[n] = p;
      base = (uint32_t*)p;
    }
    // every shard exported the same max_queries: parity b of peer r lives at base + b*cap
    tbl[n] = base;
    tbl[LB200_MAX_PEERS + n] = base + ix->pthr_cap;
    n++;
  }
  if (!ix->d_pthr_tbl) CK(cudaMalloc(&ix->d_pthr_tbl, sizeof(tbl)));
  CK(cudaMemcpy(ix->d_pthr_tbl, tbl, sizeof(tbl), cudaMemcpyHostToDevice));
  ix->pthr_peers = n;
  ix->pthr_epoch = 0;
  return 0;
}

extern "C" int32_t lb200_set_profiling(lb200_index* ix, int32_t enabled) {
  if (!ix) return set_err(LB200_EINVAL, "null index");
  ix->profiling = enabled != 0;
  return 0;
}
extern "C" int32_t lb200_get_stats(const lb200_index* ix, lb200_stats* out) {
  if (!ix || !out) return set_err(LB200_EINVAL, "null argument");
  std::lock_guard<std::mutex> lk(const_cast<lb200_index*>(ix)->mu);
  *out = ix->stats;
  return 0;
}
extern "C" int32_t lb200_set_scan_variant(lb200_index* ix, int32_t variant) {
  if (!ix || variant < 0 || variant > 3) return set_err(LB200_EINVAL, "variant must be 0..3");
  if ((variant == 1 || variant == 2) && !ix->d_codes)
    return set_err(LB200_EINVAL, "variant %d needs the tile16 layout (create the index with LB200_LAYOUTS=all)", variant);
  if (variant == 3 && !ix->d_codes3) return set_err(LB200_EINVAL, "variant 3 needs num_sub_vectors %% 16 == 0");
  ix->force_variant = variant;
  return 0;
}

#include "api_search.inl"
#include "api_flat.inl"
#include "api_pq4.inl"
#include "api_combined.inl"
#include "api_group.inl"
#include "api_file.inl"
#include "api_file_v3.inl"
#include "api_f16.inl"
#include "api_noalloc.inl"
#include "api_column.inl"
#include "api_build.inl"
#include "api_f16pq.inl"

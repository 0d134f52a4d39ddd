// dropin_host.cpp -- a host that drives an index the way Lance's ANN exec nodes do (bench.py --mode dropin).
//
// Per query (rust/lance/src/io/exec/knn.rs:349-393, 606-634; rust/lance/src/index/vector/ivf/v2.rs:378-382):
//   1. normalise the key iff the index metric is cosine            (knn.rs:366-369)
//   2. index.find_partitions(query)                     -> nprobes partition ids, ONE call   (knn.rs:371)
//   3. for each partition: spawn search_in_partition(part, query, prefilter) with up to
//      get_num_compute_intensive_cpus() calls in flight            (knn.rs:621-632, lance-core/src/utils/tokio.rs:12-31)
//   4. merge the per-partition batches: sort by _distance, fetch k (scanner.rs:2147-2157)
// The same driver runs against two back ends through plain function addresses, so that both are timed the same
// way: the B200 library's C ABI (lb200_find_partitions / lb200_search_partitions with Q = 1, P = 1) and the CPU
// oracle port (lo_find_partitions / lo_search_in_partition) -- the "reference-shaped" CPU baseline of SURVEY 8(d).
// Benchmark / test tooling only: not part of liblance_b200.so.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

extern "C" {
typedef struct {
  int32_t kind;      // 0 = lb200 C ABI, 1 = oracle port
  void* handle;      // lb200_index* | const lo_index*
  void* fn_find;     // lb200_find_partitions | lo_find_partitions
  void* fn_search;   // lb200_search_partitions | lo_search_in_partition
  int32_t metric;    // 0 l2, 1 cosine, 2 dot
  uint32_t dim, nlist;
  const float* centroids;  // oracle only: lo_find_partitions takes the centroid matrix
} dropin_backend;
}

namespace {
typedef int32_t (*lb_find_t)(void*, const float*, uint32_t, uint32_t, uint32_t*, float*);
typedef int32_t (*lb_search_t)(void*, const float*, uint32_t, const uint32_t*, uint32_t, uint32_t, const float*,
                               const float*, const uint64_t* const*, float*, uint64_t*, uint32_t*);
typedef uint32_t (*lo_find_t)(const float*, uint32_t, uint32_t, const float*, uint32_t, int, uint32_t*, float*);
typedef uint32_t (*lo_search_t)(const void*, uint32_t, const float*, uint32_t, int, float, int, float, const uint64_t*,
                                float*, uint64_t*, uint64_t*);

struct Task {
  uint32_t q, j;
};

struct QueryState {
  std::vector<float> key;
  std::vector<uint32_t> parts;
  std::vector<float> dist;      // [P][k]
  std::vector<uint64_t> ids;    // [P][k]
  std::vector<uint32_t> cnt;    // [P]
  std::atomic<uint32_t> left{0};
  std::atomic<int> err{0};
  std::mutex mu;
  std::condition_variable cv;
};

struct Pool {
  const dropin_backend* be;
  uint32_t k;
  std::vector<QueryState>* qs;
  std::deque<Task> tasks;
  std::mutex mu;
  std::condition_variable cv;
  bool stop = false;
  std::vector<std::thread> workers;

  void run_task(const Task& t) {
    QueryState& s = (*qs)[t.q];
    float* d = s.dist.data() + (size_t)t.j * k;
    uint64_t* r = s.ids.data() + (size_t)t.j * k;
    uint32_t c = 0;
    if (be->kind == 0) {
      const uint32_t part = s.parts[t.j];
      int32_t rc = ((lb_search_t)be->fn_search)(be->handle, s.key.data(), 1, &part, 1, k, nullptr, nullptr, nullptr, d, r, &c);
      if (rc) s.err = rc;
    } else {
      c = ((lo_search_t)be->fn_search)(be->handle, s.parts[t.j], s.key.data(), k, 0, 0.f, 0, 0.f, nullptr, d, r, nullptr);
    }
    s.cnt[t.j] = c;
    if (s.left.fetch_sub(1) == 1) {
      std::lock_guard<std::mutex> lk(s.mu);
      s.cv.notify_all();
    }
  }
  void loop() {
    for (;;) {
      Task t;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return stop || !tasks.empty(); });
        if (tasks.empty()) return;
        t = tasks.front();
        tasks.pop_front();
      }
      run_task(t);
    }
  }
};
}  // namespace

// Returns 0, or the first back-end error code.  out_* are [Q][k] / [Q]; lat_ms [Q] = per-query latency
// (find_partitions .. merged result); *wall_s = wall time of the whole run.  `threads` = search_in_partition calls
// in flight (the exec node's buffered(get_num_compute_intensive_cpus())); `clients` = queries in flight (1 = the
// reference's single-query latency shape).
extern "C" int dropin_run(const dropin_backend* be, const float* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                          int threads, int clients, float* out_dist, uint64_t* out_ids, uint32_t* out_cnt,
                          double* lat_ms, double* wall_s) {
  const uint32_t dim = be->dim, P = std::min(nprobes, be->nlist);
  std::vector<QueryState> qs(Q);
  Pool pool;
  pool.be = be;
  pool.k = k;
  pool.qs = &qs;
  threads = std::max(1, threads);
  clients = std::max(1, clients);
  for (int t = 0; t < threads; t++) pool.workers.emplace_back([&pool] { pool.loop(); });
  std::atomic<uint32_t> next{0};
  std::atomic<int> first_err{0};
  auto client = [&] {
    for (;;) {
      const uint32_t q = next.fetch_add(1);
      if (q >= Q) return;
      QueryState& s = qs[q];
      const auto t0 = std::chrono::steady_clock::now();
      s.key.assign(queries + (size_t)q * dim, queries + (size_t)(q + 1) * dim);
      if (be->metric == 1) {  // normalize_arrow (lance-linalg/src/kernels.rs:119-142): x / sqrt(sum x^2), sequential f32 sum
        float ss = 0.0f;
        for (uint32_t i = 0; i < dim; i++) ss += s.key[i] * s.key[i];
        const float nrm = std::sqrt(ss);
        for (uint32_t i = 0; i < dim; i++) s.key[i] = s.key[i] / nrm;
      }
      s.parts.assign(P, 0);
      uint32_t np = P;
      if (be->kind == 0) {
        int32_t rc = ((lb_find_t)be->fn_find)(be->handle, s.key.data(), 1, nprobes, s.parts.data(), nullptr);
        if (rc) { first_err = rc; return; }
      } else {
        np = ((lo_find_t)be->fn_find)(be->centroids, be->nlist, dim, s.key.data(), nprobes, be->metric == 2 ? 2 : 0,
                                      s.parts.data(), nullptr);
      }
      s.dist.assign((size_t)np * k, 0.f);
      s.ids.assign((size_t)np * k, 0);
      s.cnt.assign(np, 0);
      s.left = np;
      {
        std::lock_guard<std::mutex> lk(pool.mu);
        for (uint32_t j = 0; j < np; j++) pool.tasks.push_back(Task{q, j});
      }
      pool.cv.notify_all();
      {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.left.load() == 0; });
      }
      if (s.err) { first_err = s.err.load(); return; }
      // SortExec(_distance).with_fetch(k) over the union of the partition batches; ties by row id (oracle rule)
      std::vector<std::pair<float, uint64_t>> all;
      for (uint32_t j = 0; j < np; j++)
        for (uint32_t i = 0; i < s.cnt[j]; i++) all.emplace_back(s.dist[(size_t)j * k + i], s.ids[(size_t)j * k + i]);
      auto key_of = [](float f) { uint32_t u; memcpy(&u, &f, 4); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); };
      const size_t keep = std::min<size_t>(k, all.size());
      std::partial_sort(all.begin(), all.begin() + keep, all.end(), [&](const auto& a, const auto& b) {
        const uint32_t ka = key_of(a.first), kb = key_of(b.first);  // f32::total_cmp
        return ka != kb ? ka < kb : a.second < b.second;
      });
      for (size_t i = 0; i < keep; i++) {
        out_dist[(size_t)q * k + i] = all[i].first;
        out_ids[(size_t)q * k + i] = all[i].second;
      }
      out_cnt[q] = (uint32_t)keep;
      lat_ms[q] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
  };
  const auto w0 = std::chrono::steady_clock::now();
  std::vector<std::thread> cl;
  for (int c = 0; c < clients; c++) cl.emplace_back(client);
  for (auto& t : cl) t.join();
  *wall_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
  {
    std::lock_guard<std::mutex> lk(pool.mu);
    pool.stop = true;
  }
  pool.cv.notify_all();
  for (auto& t : pool.workers) t.join();
  return first_err.load();
}

/*
 * lance_oracle.c -- CPU restatement of Lance's IVF_PQ / flat-kNN query arithmetic.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (lance_b200/csrc, liblance_b200.so) never links, imports or calls anything here.
 *
 * Parity status: PINNED against the reference's own known-answer tests (see
 * tests/test_oracle_golden.py: l2.rs:316-335, fixture_test.rs:165-231,
 * pq/distance.rs:281-308, pq/storage.rs:741-751, pq.rs:546-592, kmeans.rs:785-810,
 * scanner.rs:2410-2510/2773-2814, knn.rs:667-747, ivf/v2.rs:1192-1302).  The reference
 * and on reference-WRITTEN bytes: all 48 288 PQ codes and partition assignments of the four
 * legacy index files the reference keeps as test data equal lo_pq_quantize /
 * lo_find_partitions on the datasets' raw vectors (tests/test_reader.py).
 * 4-bit PQ (lo_compute_pq_distance_4bit, lo_quantize_distance_table): PARITY UNPINNED -- the
 * reference holds no known-answer vector for it; checked against an independent numpy
 * restatement of the same source lines (tests/test_oracle_pq4.py).
 * The reference itself (Rust, lance v0.23.1) cannot be built in this image (no cargo/rustc);
 * the one piece that compiles from its own source, rust/lance-linalg/src/simd/f16.c, is built
 * by `make ref` into oracle/_ref/libf16_ref.so and pins the f16 restatement at the end of this
 * file within tolerance (tests/test_f16.py).  Tie order of equal distances is unpinned in the
 * reference (arrow-ord 53.3.0 sort_to_indices / slice::sort_unstable / datafusion
 * 44.0.0 TopK are all unstable) and is fixed here as (distance total-order, id) ascending.
 *
 * All arithmetic is IEEE f32, no fast-math, no FMA contraction except where the
 * reference calls multiply_add explicitly (cosine_fast).  Build with
 *   gcc -O3 -march=native -ffp-contract=off -fno-fast-math -shared -fPIC -lpthread
 * The reference's x86_64 build is pinned to target-cpu=haswell (.cargo/config.toml:11-12),
 * i.e. f32x8 = __m256, f32x16 = 2 x __m256 (simd/f32.rs:444-447); the reduce_sum
 * orders below are the AVX2 ones.
 *
 * Paths are relative to /root/reference/rust/.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

/* Minimal dynamic parallel-for on pthreads (no OpenMP runtime needed). */
typedef void (*lo_body_fn)(int64_t i, void *ctx);
typedef struct {
  int64_t n, chunk;
  volatile int64_t next;
  lo_body_fn fn;
  void *ctx;
} lo_pf_t;
static void *lo_pf_worker(void *arg) {
  lo_pf_t *p = (lo_pf_t *)arg;
  for (;;) {
    int64_t s = __atomic_fetch_add(&p->next, p->chunk, __ATOMIC_RELAXED);
    if (s >= p->n) break;
    int64_t e = s + p->chunk < p->n ? s + p->chunk : p->n;
    for (int64_t i = s; i < e; i++) p->fn(i, p->ctx);
  }
  return NULL;
}
int lo_max_threads(void) {
  long n = sysconf(_SC_NPROCESSORS_ONLN);
  return n > 0 ? (int)n : 1;
}
static void lo_parallel_for(int64_t n, int64_t chunk, int threads, lo_body_fn fn, void *ctx) {
  if (threads <= 0) threads = lo_max_threads();
  if (threads > 256) threads = 256;
  lo_pf_t p = {n, chunk > 0 ? chunk : 1, 0, fn, ctx};
  if (threads == 1 || n <= 1) {
    lo_pf_worker(&p);
    return;
  }
  pthread_t th[256];
  int started = 0;
  for (int t = 0; t < threads - 1; t++)
    if (pthread_create(&th[started], NULL, lo_pf_worker, &p) == 0) started++;
  lo_pf_worker(&p);
  for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
}

#define LO_L2 0
#define LO_COSINE 1
#define LO_DOT 2

/* ------------------------------------------------------------------------- */
/* SIMD reduction orders: lance-linalg/src/simd/f32.rs:203-218 (f32x8) and
 * :625-644 (f32x16, non-avx512 branch).                                      */
static inline float reduce_sum8(const float *x) {
  float a0 = x[0] + x[4], a1 = x[1] + x[5], a2 = x[2] + x[6], a3 = x[3] + x[7];
  float b0 = a0 + a2, b1 = a1 + a3;
  return b0 + b1;
}
static inline float reduce_sum16(const float *x) {
  float y[8];
  for (int i = 0; i < 8; i++) y[i] = x[i] + x[i + 8];
  return reduce_sum8(y);
}

/* lance-linalg/src/distance/l2.rs:65-99  l2_scalar::<f32,f32,16>  (L2 for f32, :169-175) */
float lo_l2(const float *x, const float *y, size_t d) {
  size_t nchunk = d / 16, rem0 = nchunk * 16;
  float s = 0.0f;
  for (size_t i = rem0; i < d; i++) {
    float diff = x[i] - y[i];
    s += diff * diff;
  }
  float sums[16];
  for (int i = 0; i < 16; i++) sums[i] = 0.0f;
  for (size_t c = 0; c < nchunk; c++) {
    const float *xc = x + c * 16, *yc = y + c * 16;
    for (int i = 0; i < 16; i++) {
      float diff = xc[i] - yc[i];
      sums[i] += diff * diff;
    }
  }
  float t = 0.0f;
  for (int i = 0; i < 16; i++) t += sums[i];
  return s + t;
}

/* l2.rs:207-216  l2_once::<f32x8,8> / <f32x16,16> */
static inline float l2_once8(const float *x, const float *y) {
  float s[8];
  for (int i = 0; i < 8; i++) {
    float d = x[i] - y[i];
    s[i] = d * d;
  }
  return reduce_sum8(s);
}
static inline float l2_once16(const float *x, const float *y) {
  float s[16];
  for (int i = 0; i < 16; i++) {
    float d = x[i] - y[i];
    s[i] = d * d;
  }
  return reduce_sum16(s);
}

/* l2.rs:177-195 (f32::l2_batch dispatch on dimension) via l2_distance_batch :237 */
void lo_l2_distance_batch(const float *from, const float *to, size_t n, size_t d, float *out) {
  if (d == 8) {
    for (size_t i = 0; i < n; i++) out[i] = l2_once8(from, to + i * 8);
  } else if (d == 16) {
    for (size_t i = 0; i < n; i++) out[i] = l2_once16(from, to + i * 16);
  } else {
    for (size_t i = 0; i < n; i++) out[i] = lo_l2(from, to + i * d, d);
  }
}

/* lance-linalg/src/distance/dot.rs:141-195  Dot for f32 */
float lo_dot(const float *x, const float *y, size_t d) {
  size_t n64 = d / 64, off64 = n64 * 64;
  size_t n8 = (d - off64) / 8, off8 = off64 + n8 * 8;
  float sum = 0.0f;
  for (size_t i = off8; i < d; i++) sum += x[i] * y[i];
  float sum8[8];
  for (int i = 0; i < 8; i++) sum8[i] = 0.0f;
  for (size_t c = 0; c < n8; c++) {
    const float *xc = x + off64 + c * 8, *yc = y + off64 + c * 8;
    for (int i = 0; i < 8; i++) sum8[i] += xc[i] * yc[i];
  }
  float sum16[16];
  for (int i = 0; i < 16; i++) sum16[i] = 0.0f;
  for (size_t c = 0; c < n64; c++) {
    const float *xc = x + c * 64, *yc = y + c * 64;
    for (int i = 0; i < 16; i++) {
      float p1 = xc[i] * yc[i], p2 = xc[16 + i] * yc[16 + i];
      float p3 = xc[32 + i] * yc[32 + i], p4 = xc[48 + i] * yc[48 + i];
      sum16[i] += (p1 + p2) + (p3 + p4);
    }
  }
  return reduce_sum16(sum16) + reduce_sum8(sum8) + sum;
}

/* dot.rs:71-73 dot_distance = 1 - dot ; :216-224 dot_distance_batch */
float lo_dot_distance(const float *x, const float *y, size_t d) { return 1.0f - lo_dot(x, y, d); }
void lo_dot_distance_batch(const float *from, const float *to, size_t n, size_t d, float *out) {
  for (size_t i = 0; i < n; i++) out[i] = 1.0f - lo_dot(from, to + i * d, d);
}

/* lance-linalg/src/distance/norm_l2.rs:102-126  norm_l2_impl::<f32,f32,16> (:88-92) */
float lo_norm_l2(const float *v, size_t d) {
  size_t nchunk = d / 16, rem0 = nchunk * 16;
  float sum = 0.0f;
  for (size_t i = rem0; i < d; i++) sum += v[i] * v[i];
  float sums[16];
  for (int i = 0; i < 16; i++) sums[i] = 0.0f;
  for (size_t c = 0; c < nchunk; c++)
    for (int i = 0; i < 16; i++) sums[i] += v[c * 16 + i] * v[c * 16 + i];
  float t = 0.0f;
  for (int i = 0; i < 16; i++) t += sums[i];
  return sqrtf(sum + t);
}

/* lance-linalg/src/kernels.rs:119-122  normalize: plain sequential sum, NOT norm_l2 */
void lo_normalize(const float *v, size_t d, float *out) {
  float s = 0.0f;
  for (size_t i = 0; i < d; i++) s += v[i] * v[i];
  float n = sqrtf(s);
  for (size_t i = 0; i < d; i++) out[i] = v[i] / n;
}

/* lance-linalg/src/distance/cosine.rs:143-174  cosine_fast for f32 (multiply_add = FMA,
 * simd/f32.rs:317-333,774-799) */
float lo_cosine_fast(const float *x, float x_norm, const float *y, size_t d) {
  size_t unrolled = d / 16 * 16, aligned = d / 8 * 8;
  float yn16[16], xy16[16], yn8[8], xy8[8];
  for (int i = 0; i < 16; i++) yn16[i] = xy16[i] = 0.0f;
  for (int i = 0; i < 8; i++) yn8[i] = xy8[i] = 0.0f;
  for (size_t c = 0; c < unrolled; c += 16)
    for (int i = 0; i < 16; i++) {
      xy16[i] = fmaf(x[c + i], y[c + i], xy16[i]);
      yn16[i] = fmaf(y[c + i], y[c + i], yn16[i]);
    }
  for (size_t c = unrolled; c < aligned; c += 8)
    for (int i = 0; i < 8; i++) {
      xy8[i] = fmaf(x[c + i], y[c + i], xy8[i]);
      yn8[i] = fmaf(y[c + i], y[c + i], yn8[i]);
    }
  float tn = lo_norm_l2(y + aligned, d - aligned);
  float y_norm = reduce_sum16(yn16) + reduce_sum8(yn8) + tn * tn;
  float xy = reduce_sum16(xy16) + reduce_sum8(xy8) + lo_dot(x + aligned, y + aligned, d - aligned);
  return 1.0f - xy / x_norm / sqrtf(y_norm);
}

/* cosine.rs:128-141 cosine_once<S,N> */
static inline float cosine_once(const float *x, float x_norm, const float *y, int n) {
  float y2[16], xy[16];
  for (int i = 0; i < n; i++) {
    y2[i] = y[i] * y[i];
    xy[i] = x[i] * y[i];
  }
  float sxy = n == 8 ? reduce_sum8(xy) : reduce_sum16(xy);
  float sy2 = n == 8 ? reduce_sum8(y2) : reduce_sum16(y2);
  return 1.0f - sxy / x_norm / sqrtf(sy2);
}

/* cosine.rs:201-224 f32::cosine_batch (x_norm = norm_l2(x); dispatch on dimension) */
void lo_cosine_distance_batch(const float *from, const float *to, size_t n, size_t d, float *out) {
  float x_norm = lo_norm_l2(from, d);
  if (d == 8 || d == 16) {
    for (size_t i = 0; i < n; i++) out[i] = cosine_once(from, x_norm, to + i * d, (int)d);
  } else {
    for (size_t i = 0; i < n; i++) out[i] = lo_cosine_fast(from, x_norm, to + i * d, d);
  }
}

/* distance.rs:56-63 DistanceType::arrow_batch_func: one query vs a batch of rows. */
void lo_distance_batch(int metric, const float *from, const float *to, size_t n, size_t d,
                       float *out) {
  if (metric == LO_L2)
    lo_l2_distance_batch(from, to, n, d, out);
  else if (metric == LO_COSINE)
    lo_cosine_distance_batch(from, to, n, d, out);
  else
    lo_dot_distance_batch(from, to, n, d, out);
}

/* ------------------------------------------------------------------------- */
/* Ordering: f32::total_cmp (graph.rs:54-74) on the distance, ties by id.     */
static inline int32_t total_key(float f) {
  int32_t b;
  memcpy(&b, &f, 4);
  return b ^ (int32_t)(((uint32_t)(b >> 31)) >> 1);
}
typedef struct {
  float dist;
  uint64_t id;  /* row id (or row position for flat kNN): the tie-break key */
  uint64_t pos; /* global index position, carried for the refine step */
} cand_t;
static int cand_cmp(const void *a, const void *b) {
  const cand_t *x = (const cand_t *)a, *y = (const cand_t *)b;
  int32_t kx = total_key(x->dist), ky = total_key(y->dist);
  if (kx != ky) return kx < ky ? -1 : 1;
  if (x->id != y->id) return x->id < y->id ? -1 : 1;
  return 0;
}
static inline int cand_less(const cand_t *x, const cand_t *y) { return cand_cmp(x, y) < 0; }

/* Keep the `limit` smallest of a stream in a max-heap; result sorted ascending.
 * Equivalent to "sort everything, take(limit)" under the (dist, id) total order. */
typedef struct {
  cand_t *h;
  size_t n, cap;
} heap_t;
static void heap_push(heap_t *hp, cand_t c) {
  if (hp->cap == 0) return;
  if (hp->n < hp->cap) {
    size_t i = hp->n++;
    hp->h[i] = c;
    while (i > 0) {
      size_t p = (i - 1) / 2;
      if (cand_less(&hp->h[p], &hp->h[i])) {
        cand_t t = hp->h[p];
        hp->h[p] = hp->h[i];
        hp->h[i] = t;
        i = p;
      } else
        break;
    }
  } else if (cand_less(&c, &hp->h[0])) {
    size_t i = 0;
    hp->h[0] = c;
    for (;;) {
      size_t l = 2 * i + 1, r = l + 1, m = i;
      if (l < hp->n && cand_less(&hp->h[m], &hp->h[l])) m = l;
      if (r < hp->n && cand_less(&hp->h[m], &hp->h[r])) m = r;
      if (m == i) break;
      cand_t t = hp->h[m];
      hp->h[m] = hp->h[i];
      hp->h[i] = t;
      i = m;
    }
  }
}

/* ------------------------------------------------------------------------- */
/* lance-linalg/src/kmeans.rs:624-644 kmeans_find_partitions (+ ivf/storage.rs:92-104,
 * lance-index/src/vector/ivf.rs:254-261; Cosine mapped to L2 by the caller,
 * lance/src/index/vector/ivf/v2.rs:403-411).  `query` must already be normalised
 * for a cosine index (v2.rs:370-373).  Returns number of ids written (min(nprobes,nlist)). */
uint32_t lo_find_partitions(const float *centroids, uint32_t nlist, uint32_t d, const float *query,
                            uint32_t nprobes, int metric, uint32_t *out_ids, float *out_dists) {
  float *dists = (float *)malloc(sizeof(float) * (nlist ? nlist : 1));
  if (metric == LO_DOT)
    lo_dot_distance_batch(query, centroids, nlist, d, dists);
  else
    lo_l2_distance_batch(query, centroids, nlist, d, dists);
  uint32_t lim = nprobes < nlist ? nprobes : nlist;
  cand_t *all = (cand_t *)malloc(sizeof(cand_t) * (nlist ? nlist : 1));
  for (uint32_t i = 0; i < nlist; i++) {
    all[i].dist = dists[i];
    all[i].id = i;
  }
  qsort(all, nlist, sizeof(cand_t), cand_cmp);
  for (uint32_t i = 0; i < lim; i++) {
    out_ids[i] = (uint32_t)all[i].id;
    if (out_dists) out_dists[i] = all[i].dist;
  }
  free(all);
  free(dists);
  return lim;
}

/* lance-index/src/vector/pq/distance.rs:17-85 build_distance_table_{l2,dot} (8 bit),
 * codebook slicing pq/utils.rs:54-71.  table is [M][256]. */
void lo_build_distance_table(int metric, const float *codebook, uint32_t M, uint32_t d,
                             const float *query, float *table) {
  uint32_t dsub = d / M;
  for (uint32_t m = 0; m < M; m++) {
    const float *sub = query + (size_t)m * dsub;
    const float *cents = codebook + (size_t)m * 256 * dsub;
    if (metric == LO_DOT)
      lo_dot_distance_batch(sub, cents, 256, dsub, table + (size_t)m * 256);
    else
      lo_l2_distance_batch(sub, cents, 256, dsub, table + (size_t)m * 256);
  }
}

/* The same tables for 4-bit PQ: build_distance_table_{l2,dot}_impl::<4> (pq/distance.rs:24,60): 16 centroids per
 * sub-space, codebook [M][16][dsub], table [M][16]. */
void lo_build_distance_table4(int metric, const float *codebook, uint32_t M, uint32_t d,
                              const float *query, float *table) {
  uint32_t dsub = d / M;
  for (uint32_t m = 0; m < M; m++) {
    const float *sub = query + (size_t)m * dsub;
    const float *cents = codebook + (size_t)m * 16 * dsub;
    if (metric == LO_DOT)
      lo_dot_distance_batch(sub, cents, 16, dsub, table + (size_t)m * 16);
    else
      lo_l2_distance_batch(sub, cents, 16, dsub, table + (size_t)m * 16);
  }
}

/* pq/distance.rs:206-229 quantize_distance_table: qmin = min of all entries (f32::min fold), qmax = max over
 * ADJACENT sub-space pairs (tuple_windows) of (max_a + max_b), entry -> ceil((d-qmin)*255/(qmax-qmin)) as u8
 * (Rust `as u8`: saturating, NaN -> 0). */
void lo_quantize_distance_table(const float *table, uint32_t M, float *qmin_out, float *qmax_out,
                                uint8_t *qt) {
  float qmin = INFINITY, qmax = -INFINITY;
  for (size_t i = 0; i < (size_t)M * 16; i++) qmin = fminf(qmin, table[i]);
  for (uint32_t m = 0; m + 1 < M; m++) {
    float a = -INFINITY, b = -INFINITY;
    for (int j = 0; j < 16; j++) a = fmaxf(a, table[(size_t)m * 16 + j]);
    for (int j = 0; j < 16; j++) b = fmaxf(b, table[(size_t)(m + 1) * 16 + j]);
    qmax = fmaxf(qmax, a + b);
  }
  for (size_t i = 0; i < (size_t)M * 16; i++) {
    float v = ceilf((table[i] - qmin) * 255.0f / (qmax - qmin));
    qt[i] = v != v ? 0 : (v <= 0.0f ? 0 : (v >= 255.0f ? 255 : (uint8_t)v));
  }
  *qmin_out = qmin;
  *qmax_out = qmax;
}

/* pq/distance.rs:138-204 compute_pq_distance_4bit, codes transposed [M/2][n], byte = (code[2j+1] << 4) | code[2j]
 * (pq.rs:166-170).  Rows in whole blocks of 16 go through u8x16: the two looked-up bytes are added with
 * _mm_add_epi8, i.e. WRAPPING u8 (simd/u8.rs:279-287; x86_64 is the reference CPU of this repo, .cargo/config.toml
 * target-cpu=haswell) before widening to f32; the n % 16 tail rows add the two bytes as f32 separately.
 * Then d = d * (qmax - qmin) / 255 + qmin.  n < 16 makes the reference's block loop bound underflow
 * (`num_vectors - 16 + 1` on usize); here such partitions are all tail rows. */
void lo_compute_pq_distance_4bit(const float *table, uint32_t M, const uint8_t *codes, size_t n,
                                 float *out) {
  uint8_t *qt = (uint8_t *)malloc((size_t)M * 16);
  float qmin, qmax;
  lo_quantize_distance_table(table, M, &qmin, &qmax, qt);
  size_t blocks = n >= 16 ? n - n % 16 : 0;
  for (size_t i = 0; i < n; i++) out[i] = 0.0f;
  for (uint32_t j = 0; j < M / 2; j++) {
    const uint8_t *t0 = qt + (size_t)(2 * j) * 16, *t1 = qt + (size_t)(2 * j + 1) * 16;
    const uint8_t *c = codes + (size_t)j * n;
    for (size_t i = 0; i < blocks; i++) out[i] += (float)(uint8_t)(t0[c[i] & 15] + t1[c[i] >> 4]);
    for (size_t i = blocks; i < n; i++) {
      out[i] += (float)t0[c[i] & 15];
      out[i] += (float)t1[c[i] >> 4];
    }
  }
  for (size_t i = 0; i < n; i++) out[i] = out[i] * (qmax - qmin) / 255.0f + qmin;
  free(qt);
}

/* pq/storage.rs:602-617 PQDistCalculator::distance(id) for 4 bit: the UNQUANTISED f32 table, pair sums added in
 * ascending j (prefilter path only). */
float lo_pq_distance_one_4bit(const float *table, uint32_t M, const uint8_t *codes, size_t n, size_t id) {
  float s = 0.0f;
  for (uint32_t j = 0; j < M / 2; j++) {
    uint8_t c = codes[(size_t)j * n + id];
    s += table[(size_t)(2 * j) * 16 + (c & 15)] + table[(size_t)(2 * j + 1) * 16 + (c >> 4)];
  }
  return s;
}

/* pq/distance.rs:102-135 compute_pq_distance (8 bit), codes transposed [M][n]. */
void lo_compute_pq_distance(const float *table, uint32_t M, const uint8_t *codes, size_t n,
                            float *out) {
  for (size_t i = 0; i < n; i++) out[i] = 0.0f;
  for (uint32_t m = 0; m < M; m++) {
    const float *t = table + (size_t)m * 256;
    const uint8_t *c = codes + (size_t)m * n;
    for (size_t i = 0; i < n; i++) out[i] += t[c[i]];
  }
}

/* pq/storage.rs:584-624 PQDistCalculator::distance(id): same sum, one row. */
float lo_pq_distance_one(const float *table, uint32_t M, const uint8_t *codes, size_t n, size_t id) {
  float s = 0.0f;
  for (uint32_t m = 0; m < M; m++) s += table[(size_t)m * 256 + codes[(size_t)m * n + id]];
  return s;
}

/* ------------------------------------------------------------------------- */
typedef struct {
  int32_t metric;
  uint32_t dim, nlist, M;
  const float *centroids;      /* [nlist][dim] */
  const float *codebook;       /* [M][256][dim/M] */
  const uint64_t *part_offsets; /* [nlist+1] */
  const uint8_t *codes;        /* concat over p of [M][n_p] */
  const uint64_t *row_ids;     /* [N] */
  const float *raw_vectors;    /* [N][dim] in row_ids (index) order, or NULL */
  uint32_t num_bits;           /* 0 or 8: 8-bit PQ; 4: codebook [M][16][dim/M], codes [M/2][n_p] packed nibbles */
} lo_index;

/* lance/src/index/vector/ivf/v2.rs:285-301 preprocess_query: the key that reaches the
 * sub-index = key - centroid[part] iff the metric uses residuals (L2|Cosine). */
void lo_preprocess_query(const lo_index *ix, uint32_t part, const float *key, float *out) {
  if (ix->metric == LO_L2 || ix->metric == LO_COSINE) {
    const float *c = ix->centroids + (size_t)part * ix->dim;
    for (uint32_t i = 0; i < ix->dim; i++) out[i] = key[i] - c[i];
  } else {
    memcpy(out, key, sizeof(float) * ix->dim);
  }
}

/* The part of FlatIndex::search (flat/index.rs:76-137) after the distance table exists: distance_all (or the per-row
 * path under a prefilter), [lower, upper) filter, sort, take(k'), local id -> row id.  Shared by the f32 and f16 entry
 * points: from the table on, the element type of the vectors plays no part. */
static uint32_t lo_partition_select(const uint64_t *part_offsets, const uint8_t *all_codes, const uint64_t *row_ids, uint32_t M,
                                    uint32_t num_bits, uint32_t part, const float *table, uint32_t kprime, int has_lower,
                                    float lower, int has_upper, float upper, const uint64_t *bitmask_or_null,
                                    float *out_dist, uint64_t *out_rowid, uint64_t *out_pos) {
  size_t off = part_offsets[part], n = part_offsets[part + 1] - off;
  float *dist = (float *)malloc(sizeof(float) * (n ? n : 1));
  if (num_bits == 4) {
    const uint8_t *codes = all_codes + off * (M / 2);
    if (!bitmask_or_null) {
      lo_compute_pq_distance_4bit(table, M, codes, n, dist);
    } else {
      for (size_t i = 0; i < n; i++)
        if ((bitmask_or_null[i >> 6] >> (i & 63)) & 1) dist[i] = lo_pq_distance_one_4bit(table, M, codes, n, i);
    }
  } else {
    const uint8_t *codes = all_codes + off * M;
    if (!bitmask_or_null) {
      lo_compute_pq_distance(table, M, codes, n, dist);
    } else {
      for (size_t i = 0; i < n; i++)
        if ((bitmask_or_null[i >> 6] >> (i & 63)) & 1) dist[i] = lo_pq_distance_one(table, M, codes, n, i);
    }
  }
  float lb = has_lower ? lower : -3.402823466e+38f, ub = has_upper ? upper : 3.402823466e+38f;
  int bounded = has_lower || has_upper;
  heap_t hp;
  hp.cap = kprime;
  hp.n = 0;
  hp.h = (cand_t *)malloc(sizeof(cand_t) * (kprime ? kprime : 1));
  for (size_t i = 0; i < n; i++) {
    if (bitmask_or_null && !((bitmask_or_null[i >> 6] >> (i & 63)) & 1)) continue;
    if (bounded && !(lb <= dist[i] && dist[i] < ub)) continue;
    /* tie order: reference sorts by distance only (unstable); we pin (dist, row id). */
    cand_t c;
    c.dist = dist[i];
    c.id = row_ids[off + i];
    c.pos = off + i;
    heap_push(&hp, c);
  }
  qsort(hp.h, hp.n, sizeof(cand_t), cand_cmp);
  for (size_t i = 0; i < hp.n; i++) {
    out_dist[i] = hp.h[i].dist;
    out_rowid[i] = hp.h[i].id;
    if (out_pos) out_pos[i] = hp.h[i].pos;
  }
  uint32_t cnt = (uint32_t)hp.n;
  free(hp.h);
  free(dist);
  return cnt;
}

/* lance/src/index/vector/ivf/v2.rs:414-433 search_in_partition:
 *   preprocess_query (:285-301: residual iff L2|Cosine, pq/builder.rs:66-68)
 *   FlatIndex::search (lance-index/src/vector/flat/index.rs:76-137): dist table,
 *   distance_all, [lower,upper) filter, sort, take(k'), local id -> row id.
 * `key` is the (already normalised, for cosine) query.  bitmask_or_null: per-row allow
 * bits of this partition (prefilter path, flat/index.rs:106-122).
 * Returns count (<= kprime); out sorted ascending by (dist, row id). */
uint32_t lo_search_in_partition(const lo_index *ix, uint32_t part, const float *key,
                                uint32_t kprime, int has_lower, float lower, int has_upper,
                                float upper, const uint64_t *bitmask_or_null, float *out_dist,
                                uint64_t *out_rowid, uint64_t *out_pos) {
  uint32_t d = ix->dim, M = ix->M;
  float *q = (float *)malloc(sizeof(float) * d);
  if (ix->metric == LO_L2 || ix->metric == LO_COSINE) {
    const float *c = ix->centroids + (size_t)part * d;
    for (uint32_t i = 0; i < d; i++) q[i] = key[i] - c[i];
  } else {
    memcpy(q, key, sizeof(float) * d);
  }
  float *table = (float *)malloc(sizeof(float) * M * 256);
  if (ix->num_bits == 4)
    lo_build_distance_table4(ix->metric == LO_DOT ? LO_DOT : LO_L2, ix->codebook, M, d, q, table);
  else
    lo_build_distance_table(ix->metric == LO_DOT ? LO_DOT : LO_L2, ix->codebook, M, d, q, table);
  uint32_t cnt = lo_partition_select(ix->part_offsets, ix->codes, ix->row_ids, M, ix->num_bits, part, table, kprime, has_lower,
                                     lower, has_upper, upper, bitmask_or_null, out_dist, out_rowid, out_pos);
  free(table);
  free(q);
  return cnt;
}

/* One query through VectorIndex::search (ivf/v2.rs:368-401) and, when refine_factor>0,
 * the scanner's refine (lance/src/dataset/scanner.rs:1706-1721,2066-2136): exact metric
 * of the INDEX on raw vectors with the ORIGINAL (un-normalised) query, bounds re-applied,
 * top-k.  out arrays sized k*max(refine,1) (ANN stage) -- after refine only k are valid.
 * probes_out (optional) receives the probed partitions in ascending-distance order. */
uint32_t lo_search_one(const lo_index *ix, const float *query, uint32_t k, uint32_t nprobes,
                       uint32_t refine_factor, int has_lower, float lower, int has_upper,
                       float upper, float *out_dist, uint64_t *out_rowid, uint32_t *probes_out) {
  uint32_t d = ix->dim;
  uint32_t kprime = k * (refine_factor ? refine_factor : 1);
  float *key = (float *)malloc(sizeof(float) * d);
  if (ix->metric == LO_COSINE)
    lo_normalize(query, d, key);
  else
    memcpy(key, query, sizeof(float) * d);
  uint32_t P = nprobes < ix->nlist ? nprobes : ix->nlist;
  uint32_t *probes = (uint32_t *)malloc(sizeof(uint32_t) * (P ? P : 1));
  P = lo_find_partitions(ix->centroids, ix->nlist, d, key, nprobes,
                         ix->metric == LO_DOT ? LO_DOT : LO_L2, probes, NULL);
  if (probes_out) memcpy(probes_out, probes, sizeof(uint32_t) * P);
  size_t tot = (size_t)P * kprime;
  float *pd = (float *)malloc(sizeof(float) * (tot ? tot : 1));
  uint64_t *pr = (uint64_t *)malloc(sizeof(uint64_t) * (tot ? tot : 1));
  uint64_t *pp = (uint64_t *)malloc(sizeof(uint64_t) * (tot ? tot : 1));
  cand_t *cps = (cand_t *)malloc(sizeof(cand_t) * (tot ? tot : 1));
  size_t na = 0;
  for (uint32_t i = 0; i < P; i++) {
    uint32_t c = lo_search_in_partition(ix, probes[i], key, kprime, has_lower, lower, has_upper,
                                        upper, NULL, pd + (size_t)i * kprime,
                                        pr + (size_t)i * kprime, pp + (size_t)i * kprime);
    for (uint32_t j = 0; j < c; j++) {
      cps[na].dist = pd[(size_t)i * kprime + j];
      cps[na].id = pr[(size_t)i * kprime + j];
      cps[na].pos = pp[(size_t)i * kprime + j];
      na++;
    }
  }
  /* merge: concat_batches + sort_to_indices(dist, limit=k') (v2.rs:383-400), pinned tie order */
  qsort(cps, na, sizeof(cand_t), cand_cmp);
  uint32_t cnt = (uint32_t)(na < kprime ? na : kprime);
  if (!refine_factor) {
    for (uint32_t i = 0; i < cnt; i++) {
      out_dist[i] = cps[i].dist;
      out_rowid[i] = cps[i].id;
    }
  } else {
    /* refine: TakeExec + flat_knn with metric = index.metric_type() (scanner.rs:1715-1721) */
    cand_t *ref = (cand_t *)malloc(sizeof(cand_t) * (cnt ? cnt : 1));
    uint32_t nr = 0;
    float lb = has_lower ? lower : 0, ub = has_upper ? upper : 0;
    for (uint32_t i = 0; i < cnt; i++) {
      float dist;
      lo_distance_batch(ix->metric, query, ix->raw_vectors + (size_t)cps[i].pos * d, 1, d, &dist);
      if (has_lower && !(dist >= lb)) continue;
      if (has_upper && !(dist < ub)) continue;
      ref[nr].dist = dist;
      ref[nr].id = cps[i].id;
      ref[nr].pos = cps[i].pos;
      nr++;
    }
    qsort(ref, nr, sizeof(cand_t), cand_cmp);
    cnt = nr < k ? nr : k;
    for (uint32_t i = 0; i < cnt; i++) {
      out_dist[i] = ref[i].dist;
      out_rowid[i] = ref[i].id;
    }
    free(ref);
  }
  free(cps);
  free(pp);
  free(pr);
  free(pd);
  free(probes);
  free(key);
  return cnt;
}

/* Batched driver: queries are independent (the reference has no inter-query batching,
 * SURVEY 2a); "throughput-shaped" CPU baseline = queries spread over `threads` threads.
 * out_* are [Q][k'] with k' = k*max(refine,1) stride; out_count [Q]; probes [Q][nprobes] or NULL. */
typedef struct {
  const lo_index *ix;
  const float *queries;
  uint32_t k, nprobes, refine_factor;
  int has_lower, has_upper;
  float lower, upper;
  float *out_dist;
  uint64_t *out_rowid;
  uint32_t *out_count, *probes_out;
} lo_sb_ctx;
static void lo_sb_body(int64_t qi, void *vctx) {
  lo_sb_ctx *c = (lo_sb_ctx *)vctx;
  uint32_t stride = c->k * (c->refine_factor ? c->refine_factor : 1);
  c->out_count[qi] = lo_search_one(
      c->ix, c->queries + (size_t)qi * c->ix->dim, c->k, c->nprobes, c->refine_factor,
      c->has_lower, c->lower, c->has_upper, c->upper, c->out_dist + (size_t)qi * stride,
      c->out_rowid + (size_t)qi * stride,
      c->probes_out ? c->probes_out + (size_t)qi * c->nprobes : NULL);
}
void lo_search_batch(const lo_index *ix, const float *queries, uint32_t Q, uint32_t k,
                     uint32_t nprobes, uint32_t refine_factor, int has_lower, float lower,
                     int has_upper, float upper, int threads, float *out_dist,
                     uint64_t *out_rowid, uint32_t *out_count, uint32_t *probes_out) {
  lo_sb_ctx c = {ix,    queries, k,        nprobes,   refine_factor, has_lower, has_upper,
                 lower, upper,   out_dist, out_rowid, out_count,     probes_out};
  lo_parallel_for(Q, 4, threads, lo_sb_body, &c);
}

/* ------------------------------------------------------------------------- */
/* Flat kNN: KNNVectorDistanceExec (lance/src/io/exec/knn.rs:156-183) -> compute_distance
 * (lance-index/src/vector/flat.rs:28-86) -> SortExec fetch=k -> NOT NULL filter
 * (scanner.rs:2066-2136).  validity_or_null: Arrow validity bitmap (bit set = valid).
 * Null rows sort last and are then dropped, so they can never displace a valid row
 * unless fewer than k valid rows exist.  ids are row positions 0..n-1 (or row_ids[] if given). */
void lo_flat_distance(int metric, const float *query, const float *vectors, size_t n, size_t d,
                      float *out) {
  lo_distance_batch(metric, query, vectors, n, d, out);
}

uint32_t lo_flat_knn_one(int metric, const float *query, const float *vectors,
                         const uint8_t *validity_or_null, const uint64_t *row_ids_or_null,
                         size_t n, size_t d, uint32_t k, int has_lower, float lower, int has_upper,
                         float upper, float *out_dist, uint64_t *out_id) {
  heap_t hp;
  hp.cap = k;
  hp.n = 0;
  hp.h = (cand_t *)malloc(sizeof(cand_t) * (k ? k : 1));
  const size_t B = 8192; /* LANCE_DEFAULT_BATCH_SIZE rows per compute_distance call */
  float *dist = (float *)malloc(sizeof(float) * B);
  for (size_t s = 0; s < n; s += B) {
    size_t m = n - s < B ? n - s : B;
    lo_distance_batch(metric, query, vectors + s * d, m, d, dist);
    for (size_t i = 0; i < m; i++) {
      size_t r = s + i;
      if (validity_or_null && !((validity_or_null[r >> 3] >> (r & 7)) & 1)) continue;
      if (has_lower && !(dist[i] >= lower)) continue;
      if (has_upper && !(dist[i] < upper)) continue;
      cand_t c;
      c.dist = dist[i];
      c.id = row_ids_or_null ? row_ids_or_null[r] : r;
      c.pos = r;
      heap_push(&hp, c);
    }
  }
  qsort(hp.h, hp.n, sizeof(cand_t), cand_cmp);
  for (size_t i = 0; i < hp.n; i++) {
    out_dist[i] = hp.h[i].dist;
    out_id[i] = hp.h[i].id;
  }
  uint32_t cnt = (uint32_t)hp.n;
  free(dist);
  free(hp.h);
  return cnt;
}

typedef struct {
  int metric;
  const float *queries, *vectors;
  const uint8_t *validity;
  const uint64_t *row_ids;
  size_t n, d;
  uint32_t k;
  float *out_dist;
  uint64_t *out_id;
  uint32_t *out_count;
} lo_fk_ctx;
static void lo_fk_body(int64_t qi, void *vctx) {
  lo_fk_ctx *c = (lo_fk_ctx *)vctx;
  c->out_count[qi] = lo_flat_knn_one(c->metric, c->queries + (size_t)qi * c->d, c->vectors,
                                     c->validity, c->row_ids, c->n, c->d, c->k, 0, 0, 0, 0,
                                     c->out_dist + (size_t)qi * c->k, c->out_id + (size_t)qi * c->k);
}
void lo_flat_knn_batch(int metric, const float *queries, uint32_t Q, const float *vectors,
                       const uint8_t *validity_or_null, const uint64_t *row_ids_or_null, size_t n,
                       size_t d, uint32_t k, int threads, float *out_dist, uint64_t *out_id,
                       uint32_t *out_count) {
  lo_fk_ctx c = {metric, queries, vectors, validity_or_null, row_ids_or_null, n, d, k,
                 out_dist, out_id, out_count};
  lo_parallel_for(Q, 1, threads, lo_fk_body, &c);
}

/* PQ quantise (lance-index/src/vector/pq.rs:98-160): per sub-space argmin of the L2
 * (or dot) distance to the 256 centroids; used only by the synthetic index builder's
 * cross-check and golden test #4 (pq.rs:595-632).  out codes row-major [n][M]. */
typedef struct {
  int metric;
  const float *codebook, *vecs;
  uint32_t M, d;
  uint8_t *out_codes;
} lo_pq_ctx;
static void lo_pq_body(int64_t i, void *vctx) {
  lo_pq_ctx *c = (lo_pq_ctx *)vctx;
  uint32_t dsub = c->d / c->M;
  float dist[256];
  for (uint32_t m = 0; m < c->M; m++) {
    const float *sub = c->vecs + (size_t)i * c->d + (size_t)m * dsub;
    const float *cents = c->codebook + (size_t)m * 256 * dsub;
    if (c->metric == LO_DOT)
      lo_dot_distance_batch(sub, cents, 256, dsub, dist);
    else
      lo_l2_distance_batch(sub, cents, 256, dsub, dist);
    /* argmin_value_float, kernels.rs:79-89: strict <, NaN skipped */
    float best = INFINITY;
    int bi = 0;
    for (int j = 0; j < 256; j++)
      if (dist[j] < best) {
        best = dist[j];
        bi = j;
      }
    c->out_codes[(size_t)i * c->M + m] = (uint8_t)bi;
  }
}
void lo_pq_quantize(int metric, const float *codebook, uint32_t M, uint32_t d, const float *vecs,
                    size_t n, uint8_t *out_codes) {
  lo_pq_ctx c = {metric, codebook, vecs, M, d, out_codes};
  lo_parallel_for((int64_t)n, 256, 0, lo_pq_body, &c);
}

/* ProductQuantizer::transform_impl::<4> (lance-index/src/vector/pq.rs:112-190): per sub-space argmin over the 16
 * centroids, two codes per byte as (code[2j+1] << 4) | code[2j].  out_codes row-major [n][M/2]. */
void lo_pq_quantize4(int metric, const float *codebook, uint32_t M, uint32_t d, const float *vecs,
                     size_t n, uint8_t *out_codes) {
  uint32_t dsub = d / M;
  float dist[16];
  for (size_t i = 0; i < n; i++)
    for (uint32_t m = 0; m < M; m++) {
      const float *sub = vecs + i * d + (size_t)m * dsub;
      const float *cents = codebook + (size_t)m * 16 * dsub;
      if (metric == LO_DOT)
        lo_dot_distance_batch(sub, cents, 16, dsub, dist);
      else
        lo_l2_distance_batch(sub, cents, 16, dsub, dist);
      float best = INFINITY;
      int bi = 0;
      for (int j = 0; j < 16; j++)
        if (dist[j] < best) {
          best = dist[j];
          bi = j;
        }
      uint8_t *o = out_codes + i * (M / 2) + m / 2;
      *o = (m & 1) ? (uint8_t)(*o | (bi << 4)) : (uint8_t)bi;
    }
}

/* ------------------------------------------------------------------------- */
/* f16 vectors.  The reference has two CPU behaviours for f16 (l2.rs:136-166, dot.rs:100-137,
 * cosine.rs:88-121, norm_l2.rs:46-76): with the `fp16kernels` feature the C kernels of simd/f16.c (built
 * -ffast-math, so their summation order is the compiler's), otherwise the generic scalar fallbacks restated here,
 * which ARE deterministic: every element is widened to f32 first (`x.as_()`), then
 *   L2      l2_scalar::<f16,f32,16>   (l2.rs:65-99)      == lo_l2 on the widened values
 *   dot     dot_scalar::<f16,f32,32>  (dot.rs:33-61)     remainder first, 32 lane sums, lanes added in order
 *   norm    norm_l2_impl::<f16,f32,32> (norm_l2.rs:102-126)
 *   cosine  cosine_scalar (cosine.rs:233-238): 1 - xy / (x_norm * sqrt(dot(y,y)))
 * oracle/_ref/libf16_ref.so (the reference's own f16.c, compiled by `make ref`) is the cross-check within tolerance. */
typedef _Float16 lo_f16;

float lo_l2_f16(const lo_f16 *x, const lo_f16 *y, size_t d) {
  size_t nchunk = d / 16, rem0 = nchunk * 16;
  float s = 0.0f;
  for (size_t i = rem0; i < d; i++) {
    float diff = (float)x[i] - (float)y[i];
    s += diff * diff;
  }
  float sums[16];
  for (int i = 0; i < 16; i++) sums[i] = 0.0f;
  for (size_t c = 0; c < nchunk; c++)
    for (int i = 0; i < 16; i++) {
      float diff = (float)x[c * 16 + i] - (float)y[c * 16 + i];
      sums[i] += diff * diff;
    }
  float t = 0.0f;
  for (int i = 0; i < 16; i++) t += sums[i];
  return s + t;
}

float lo_dot_f16(const lo_f16 *x, const lo_f16 *y, size_t d) {
  size_t nchunk = d / 32, rem0 = nchunk * 32;
  float s = 0.0f;
  for (size_t i = rem0; i < d; i++) s += (float)x[i] * (float)y[i];
  float sums[32];
  for (int i = 0; i < 32; i++) sums[i] = 0.0f;
  for (size_t c = 0; c < nchunk; c++)
    for (int i = 0; i < 32; i++) sums[i] += (float)x[c * 32 + i] * (float)y[c * 32 + i];
  float t = 0.0f;
  for (int i = 0; i < 32; i++) t += sums[i];
  return s + t;
}

float lo_norm_l2_f16(const lo_f16 *v, size_t d) {
  size_t nchunk = d / 32, rem0 = nchunk * 32;
  float s = 0.0f;
  for (size_t i = rem0; i < d; i++) s += (float)v[i] * (float)v[i];
  float sums[32];
  for (int i = 0; i < 32; i++) sums[i] = 0.0f;
  for (size_t c = 0; c < nchunk; c++)
    for (int i = 0; i < 32; i++) sums[i] += (float)v[c * 32 + i] * (float)v[c * 32 + i];
  float t = 0.0f;
  for (int i = 0; i < 32; i++) t += sums[i];
  return sqrtf(s + t);
}

float lo_cosine_f16(const lo_f16 *x, float x_norm, const lo_f16 *y, size_t d) {
  float y_sq = lo_dot_f16(y, y, d);
  float xy = lo_dot_f16(x, y, d);
  return 1.0f - xy / (x_norm * sqrtf(y_sq));
}

/* {l2,cosine,dot}_distance_arrow_batch on a Float16 FixedSizeList (l2.rs:288-310, cosine.rs:315-340,
 * dot.rs:270-295): one key against n rows. */
void lo_distance_batch_f16(int metric, const lo_f16 *from, const lo_f16 *to, size_t n, size_t d, float *out) {
  float xn = metric == LO_COSINE ? lo_norm_l2_f16(from, d) : 0.0f;
  for (size_t i = 0; i < n; i++) {
    const lo_f16 *y = to + i * d;
    out[i] = metric == LO_L2 ? lo_l2_f16(from, y, d)
           : metric == LO_DOT ? 1.0f - lo_dot_f16(from, y, d)
                              : lo_cosine_f16(from, xn, y, d);
  }
}

/* ------------------------------------------------------------------------------------------------
 * IVF_PQ on Float16 vectors: IVFIndex<FlatIndex, ProductQuantizer> whose centroids, codebook, keys and
 * raw vectors are f16 arrays.  The generic code paths are the f32 ones with T = f16:
 *   find_partitions    kmeans_find_partitions::<f16> (kmeans.rs:581-586,624-644): f16::l2 / f16::dot per centroid
 *   residual           arrow_arith sub on Float16 arrays (v2.rs:294): half's Sub = f16(f32(a) - f32(b))
 *   distance table     build_distance_table_{l2,dot}::<f16> (pq/distance.rs:17-85, storage.rs:500-512): f16::l2_batch is the
 *                      DEFAULT trait impl (per centroid f16::l2 = l2_scalar::<f16,f32,16>, no dimension-8/16 special
 *                      case as f32 has); dot_distance = 1 - dot_scalar::<f16,f32,32>
 *   scan, select       identical to f32 (the table is f32): lo_partition_select
 *   normalize (cosine) kernels.rs:118-121 normalize::<f16>: x.powi(2) per element, sum::<f16>, sqrt, x / norm with the
 *                      half crate's software operators: every operation is done in f32 and rounded to f16, and
 *                      `impl Sum for f16` accumulates in f32 and rounds once (half 2.x binary16.rs).  PARITY UNPINNED for
 *                      this function: the half crate is a dependency that is absent from /root/reference.
 *   refine             flat_knn on the raw f16 rows with the ORIGINAL key: lo_distance_batch_f16
 * f16::l2 / dot / norm_l2 are the deterministic scalar fallbacks above (the fp16kernels C kernels agree within
 * rounding: tests/test_f16.py). */
typedef struct {
  int32_t metric;
  uint32_t dim, nlist, M;
  const lo_f16 *centroids;      /* [nlist][dim] */
  const lo_f16 *codebook;       /* [M][256][dim/M] */
  const uint64_t *part_offsets; /* [nlist+1] */
  const uint8_t *codes;         /* concat over p of [M][n_p] */
  const uint64_t *row_ids;      /* [N] */
  const lo_f16 *raw_vectors;    /* [N][dim] in index order, or NULL */
} lo_index_f16;

void lo_normalize_f16(const lo_f16 *v, size_t d, lo_f16 *out) {
  float ss = 0.0f;
  for (size_t i = 0; i < d; i++) {
    lo_f16 p = (lo_f16)((float)v[i] * (float)v[i]); /* f16::powi(2) */
    ss += (float)p;                                 /* Sum for f16: f32 accumulator */
  }
  lo_f16 s16 = (lo_f16)ss;
  lo_f16 n16 = (lo_f16)sqrtf((float)s16);
  for (size_t i = 0; i < d; i++) out[i] = (lo_f16)((float)v[i] / (float)n16);
}

uint32_t lo_find_partitions_f16(const lo_f16 *centroids, uint32_t nlist, uint32_t d, const lo_f16 *query, uint32_t nprobes,
                                int metric, uint32_t *out_ids, float *out_dists) {
  cand_t *all = (cand_t *)malloc(sizeof(cand_t) * (nlist ? nlist : 1));
  for (uint32_t i = 0; i < nlist; i++) {
    const lo_f16 *c = centroids + (size_t)i * d;
    all[i].dist = metric == LO_DOT ? 1.0f - lo_dot_f16(query, c, d) : lo_l2_f16(query, c, d);
    all[i].id = i;
  }
  uint32_t lim = nprobes < nlist ? nprobes : nlist;
  qsort(all, nlist, sizeof(cand_t), cand_cmp);
  for (uint32_t i = 0; i < lim; i++) {
    out_ids[i] = (uint32_t)all[i].id;
    if (out_dists) out_dists[i] = all[i].dist;
  }
  free(all);
  return lim;
}

void lo_build_distance_table_f16(int metric, const lo_f16 *codebook, uint32_t M, uint32_t d, const lo_f16 *query,
                                 float *table) {
  uint32_t dsub = d / M;
  for (uint32_t m = 0; m < M; m++)
    for (uint32_t c = 0; c < 256; c++) {
      const lo_f16 *y = codebook + ((size_t)m * 256 + c) * dsub;
      table[(size_t)m * 256 + c] = metric == LO_DOT ? 1.0f - lo_dot_f16(query + (size_t)m * dsub, y, dsub)
                                                    : lo_l2_f16(query + (size_t)m * dsub, y, dsub);
    }
}

uint32_t lo_search_in_partition_f16(const lo_index_f16 *ix, uint32_t part, const lo_f16 *key, uint32_t kprime, int has_lower,
                                    float lower, int has_upper, float upper, const uint64_t *bitmask_or_null,
                                    float *out_dist, uint64_t *out_rowid, uint64_t *out_pos) {
  uint32_t d = ix->dim, M = ix->M;
  lo_f16 *q = (lo_f16 *)malloc(sizeof(lo_f16) * d);
  if (ix->metric == LO_L2 || ix->metric == LO_COSINE) {
    const lo_f16 *c = ix->centroids + (size_t)part * d;
    for (uint32_t i = 0; i < d; i++) q[i] = (lo_f16)((float)key[i] - (float)c[i]);
  } else {
    memcpy(q, key, sizeof(lo_f16) * d);
  }
  float *table = (float *)malloc(sizeof(float) * M * 256);
  lo_build_distance_table_f16(ix->metric == LO_DOT ? LO_DOT : LO_L2, ix->codebook, M, d, q, table);
  uint32_t cnt = lo_partition_select(ix->part_offsets, ix->codes, ix->row_ids, M, 8, part, table, kprime, has_lower, lower,
                                     has_upper, upper, bitmask_or_null, out_dist, out_rowid, out_pos);
  free(table);
  free(q);
  return cnt;
}

uint32_t lo_search_one_f16(const lo_index_f16 *ix, const lo_f16 *query, uint32_t k, uint32_t nprobes, uint32_t refine_factor,
                           int has_lower, float lower, int has_upper, float upper, float *out_dist, uint64_t *out_rowid,
                           uint32_t *probes_out) {
  uint32_t d = ix->dim;
  uint32_t kprime = k * (refine_factor ? refine_factor : 1);
  lo_f16 *key = (lo_f16 *)malloc(sizeof(lo_f16) * d);
  if (ix->metric == LO_COSINE)
    lo_normalize_f16(query, d, key);
  else
    memcpy(key, query, sizeof(lo_f16) * d);
  uint32_t P = nprobes < ix->nlist ? nprobes : ix->nlist;
  uint32_t *probes = (uint32_t *)malloc(sizeof(uint32_t) * (P ? P : 1));
  P = lo_find_partitions_f16(ix->centroids, ix->nlist, d, key, nprobes, ix->metric == LO_DOT ? LO_DOT : LO_L2, probes, NULL);
  if (probes_out) memcpy(probes_out, probes, sizeof(uint32_t) * P);
  size_t tot = (size_t)P * kprime;
  float *pd = (float *)malloc(sizeof(float) * (tot ? tot : 1));
  uint64_t *pr = (uint64_t *)malloc(sizeof(uint64_t) * (tot ? tot : 1));
  uint64_t *pp = (uint64_t *)malloc(sizeof(uint64_t) * (tot ? tot : 1));
  cand_t *cps = (cand_t *)malloc(sizeof(cand_t) * (tot ? tot : 1));
  size_t na = 0;
  for (uint32_t i = 0; i < P; i++) {
    uint32_t c = lo_search_in_partition_f16(ix, probes[i], key, kprime, has_lower, lower, has_upper, upper, NULL,
                                            pd + (size_t)i * kprime, pr + (size_t)i * kprime, pp + (size_t)i * kprime);
    for (uint32_t j = 0; j < c; j++) {
      cps[na].dist = pd[(size_t)i * kprime + j];
      cps[na].id = pr[(size_t)i * kprime + j];
      cps[na].pos = pp[(size_t)i * kprime + j];
      na++;
    }
  }
  qsort(cps, na, sizeof(cand_t), cand_cmp);
  uint32_t cnt = (uint32_t)(na < kprime ? na : kprime);
  if (!refine_factor) {
    for (uint32_t i = 0; i < cnt; i++) {
      out_dist[i] = cps[i].dist;
      out_rowid[i] = cps[i].id;
    }
  } else {
    cand_t *ref = (cand_t *)malloc(sizeof(cand_t) * (cnt ? cnt : 1));
    uint32_t nr = 0;
    for (uint32_t i = 0; i < cnt; i++) {
      float dist;
      lo_distance_batch_f16(ix->metric, query, ix->raw_vectors + (size_t)cps[i].pos * d, 1, d, &dist);
      if (has_lower && !(dist >= lower)) continue;
      if (has_upper && !(dist < upper)) continue;
      if (dist != dist) continue; /* IS NOT NULL of the flat_knn plan */
      ref[nr].dist = dist;
      ref[nr].id = cps[i].id;
      ref[nr].pos = cps[i].pos;
      nr++;
    }
    qsort(ref, nr, sizeof(cand_t), cand_cmp);
    cnt = nr < k ? nr : k;
    for (uint32_t i = 0; i < cnt; i++) {
      out_dist[i] = ref[i].dist;
      out_rowid[i] = ref[i].id;
    }
    free(ref);
  }
  free(cps);
  free(pp);
  free(pr);
  free(pd);
  free(probes);
  free(key);
  return cnt;
}

typedef struct {
  const lo_index_f16 *ix;
  const lo_f16 *queries;
  uint32_t k, nprobes, refine_factor;
  int has_lower, has_upper;
  float lower, upper;
  float *out_dist;
  uint64_t *out_rowid;
  uint32_t *out_count, *probes_out;
} lo_sb16_ctx;
static void lo_sb16_body(int64_t qi, void *vctx) {
  lo_sb16_ctx *c = (lo_sb16_ctx *)vctx;
  uint32_t stride = c->k * (c->refine_factor ? c->refine_factor : 1);
  c->out_count[qi] = lo_search_one_f16(c->ix, c->queries + (size_t)qi * c->ix->dim, c->k, c->nprobes, c->refine_factor,
                                       c->has_lower, c->lower, c->has_upper, c->upper, c->out_dist + (size_t)qi * stride,
                                       c->out_rowid + (size_t)qi * stride,
                                       c->probes_out ? c->probes_out + (size_t)qi * c->nprobes : NULL);
}
void lo_search_batch_f16(const lo_index_f16 *ix, const lo_f16 *queries, uint32_t Q, uint32_t k, uint32_t nprobes,
                         uint32_t refine_factor, int has_lower, float lower, int has_upper, float upper, int threads,
                         float *out_dist, uint64_t *out_rowid, uint32_t *out_count, uint32_t *probes_out) {
  lo_sb16_ctx c = {ix, queries, k, nprobes, refine_factor, has_lower, has_upper, lower, upper, out_dist, out_rowid, out_count,
                   probes_out};
  lo_parallel_for(Q, 4, threads, lo_sb16_body, &c);
}

/* ------------------------------------------------------------------------------------------------
 * KMeans training (index build).  Restates lance-linalg/src/kmeans.rs:
 *   KMeansAlgoFloat::compute_membership_and_loss (:208-240): per vector argmin_value_float over
 *     l2_distance_batch / dot_distance_batch against all centroids (None when no distance is < +inf); loss = the
 *     sequential f64 sum of the f32 minima (None counts 0);
 *   KMeansAlgoFloat::to_kmeans (:242-300): per cluster the f32 sum of its members IN DATA ORDER, divided by the count
 *     as f32; empty clusters become NaN and are then re-seeded by split_clusters (:135-148): the FIRST largest cluster
 *     is halved, new = largest * (1 + eps), largest /= (1 + eps), eps = f32::EPSILON;
 *   KMeans::train_kmeans (:440-525): up to max_iters rounds of (membership, update), stop when
 *     |loss - last_loss| / last_loss < tolerance (checked AFTER the update of that round).
 * Unpinned by the reference: the random initialisation (SmallRng::from_entropy) -- the caller supplies the initial
 * centroids; `redos` (default 1) is not restated.  metric: LO_L2 or LO_DOT (cosine callers normalise and use L2,
 * as ivf/builder.rs does). */
typedef struct {
  int metric;
  const float *data, *cents;
  uint32_t dim, k;
  uint32_t *ids;
  float *dists;
} lo_km_ctx;
static void lo_km_body(int64_t i, void *vctx) {
  lo_km_ctx *c = (lo_km_ctx *)vctx;
  float *dist = (float *)malloc(sizeof(float) * c->k);
  const float *v = c->data + (size_t)i * c->dim;
  if (c->metric == LO_DOT)
    lo_dot_distance_batch(v, c->cents, c->k, c->dim, dist);
  else
    lo_l2_distance_batch(v, c->cents, c->k, c->dim, dist);
  float best = INFINITY;
  uint32_t bi = 0xffffffffu;
  for (uint32_t j = 0; j < c->k; j++)
    if (dist[j] < best) {
      best = dist[j];
      bi = j;
    }
  c->ids[i] = bi;
  c->dists[i] = bi == 0xffffffffu ? 0.0f : best;
  free(dist);
}
/* membership (0xffffffff = None) and distance to the chosen centroid, for n vectors */
void lo_compute_membership(int metric, const float *cents, uint32_t k, uint32_t dim, const float *data, size_t n,
                           int threads, uint32_t *ids, float *dists) {
  lo_km_ctx c = {metric, data, cents, dim, k, ids, dists};
  lo_parallel_for((int64_t)n, 64, threads, lo_km_body, &c);
}
/* centroids [k][dim]: in = initial, out = trained.  Returns the number of rounds run; *loss_out = last loss. */
uint32_t lo_kmeans_fit(int metric, const float *data, size_t n, uint32_t dim, uint32_t k, uint32_t max_iters,
                       double tolerance, int threads, float *centroids, double *loss_out) {
  uint32_t *ids = (uint32_t *)malloc(sizeof(uint32_t) * (n ? n : 1));
  float *dists = (float *)malloc(sizeof(float) * (n ? n : 1));
  uint64_t *cnts = (uint64_t *)malloc(sizeof(uint64_t) * k);
  const float one_eps = 1.0f + 1.1920929e-7f;
  double loss = 1.7976931348623157e308, last_loss = 0.0;
  uint32_t it = 0;
  for (it = 1; it <= max_iters; it++) {
    lo_compute_membership(metric, centroids, k, dim, data, n, threads, ids, dists);
    last_loss = 0.0;
    for (size_t i = 0; i < n; i++) last_loss += (double)dists[i];
    for (uint32_t c = 0; c < k; c++) cnts[c] = 0;
    for (size_t j = 0; j < (size_t)k * dim; j++) centroids[j] = 0.0f;
    for (size_t i = 0; i < n; i++) {
      if (ids[i] == 0xffffffffu) continue;
      cnts[ids[i]]++;
      float *c = centroids + (size_t)ids[i] * dim;
      const float *v = data + i * dim;
      for (uint32_t j = 0; j < dim; j++) c[j] += v[j];
    }
    for (uint32_t c = 0; c < k; c++) {
      float *cc = centroids + (size_t)c * dim;
      if (cnts[c] == 0)
        for (uint32_t j = 0; j < dim; j++) cc[j] = NAN;
      else
        for (uint32_t j = 0; j < dim; j++) cc[j] /= (float)cnts[c];
    }
    for (uint32_t i = 0; i < k; i++) { /* split_clusters */
      if (cnts[i] != 0) continue;
      uint32_t largest = 0;
      uint64_t mx = 0; /* argmax: first strictly greater than u64::MIN = 0 ... (all zero -> index stays None; k >= 1 and n >= k make that impossible) */
      for (uint32_t j = 0; j < k; j++)
        if (cnts[j] > mx) {
          mx = cnts[j];
          largest = j;
        }
      cnts[i] = cnts[largest] / 2;
      cnts[largest] /= 2;
      for (uint32_t j = 0; j < dim; j++) {
        centroids[(size_t)i * dim + j] = centroids[(size_t)largest * dim + j] * one_eps;
        centroids[(size_t)largest * dim + j] /= one_eps;
      }
    }
    if (fabs(loss - last_loss) / last_loss < tolerance) break;
    loss = last_loss;
  }
  if (it > max_iters) it = max_iters;
  if (loss_out) *loss_out = last_loss;
  free(ids);
  free(dists);
  free(cnts);
  return it;
}

/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's CPU algorithms on the hot path
 * (batched L2 / inner-product distance evaluation + top-k candidate selection).
 * It is the checker for the CUDA product in hnsw-nsg_b200/csrc; it is never
 * linked into, loaded by, or called from the product library. Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load liboracle.so.
 *
 * Pinning: every function below is checked in tests/test_oracle_pinning.py against
 * (a) the reference's own known-answer vectors (hnswlib/tests/python/bindings_test_spaces.py:11-33,
 *     hnswlib/tests/cpp/data/ (.bin files) -> tests/golden/), and
 * (b) outputs of the UNMODIFIED reference sources compiled into the oracle/_ref/ shared objects
 *     (oracle/Makefile, wrappers oracle/ref_*_capi.cpp) on identical seeded inputs.
 *
 * Flat graph layout shared with the product:
 *   adjacency  u32 [n, gstride], row i = out-neighbours of i in list order, padded with ORC_INVALID.
 *   vectors    f32 [n, stride], stride >= dim (padding, if any, is zero).
 *
 * Where the reference leaves an order unspecified (std::sort / std::priority_queue on equal
 * distances) the oracle uses the total order (distance, id) ascending — documented per function.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_INVALID 0xFFFFFFFFu
#define EXPORT __attribute__((visibility("default")))

/* ------------------------------------------------------------------ distances */
/* distance modes */
enum {
  ORC_L2_SCALAR = 0,     /* hnswlib L2Sqr, space_l2.h:6-20 (what -DNO_MANUAL_VECTORIZATION selects) */
  ORC_IP_SCALAR = 1,     /* hnswlib InnerProductDistance = 1 - dot, space_ip.h:6-19 */
  ORC_L2_HNSW_AVX = 2,   /* hnswlib L2Space dispatch with USE_AVX (no AVX-512), space_l2.h:208-253 */
  ORC_IP_HNSW_AVX = 3,   /* hnswlib InnerProductSpace dispatch with USE_AVX, space_ip.h:342-398 */
  ORC_L2_EFANNA_AVX = 4, /* efanna2e::DistanceL2::compare __AVX__ branch, CNNS/include/nsg/distance.h:25-74 */
  ORC_L2_GPU = 5,        /* summation order of the CUDA kernels (team of lanes, fmaf, xor-butterfly) */
  ORC_IP_GPU = 6,        /* 1 - dot in the CUDA kernels' order */
  ORC_FASTL2_EFANNA_AVX = 7, /* ||x||^2 - 2<q,x>, DistanceFastL2, distance.h:232-322 */
  ORC_L2_SEQ_FMA = 8,    /* order of csrc/bf_knn.cu: sequential over the dimension with fmaf */
  ORC_IP_SEQ_FMA = 9,
  ORC_L2_HYBRID_AVX = 10, /* hnsw_nsg::DistanceL2AVX::compare, hnsw_nsg/include/nsg/distance.h:325-346 */
  ORC_L2_U8 = 11         /* hnswlib L2SqrI / L2SqrI4x over unsigned bytes, space_l2.h:255-292 (see l2_u8 below) */
};

/* space_l2.h:6-20 — sequential, multiply then add (no FMA: built with -ffp-contract=off) */
static float l2_scalar(const float *a, const float *b, size_t qty) {
  float res = 0;
  for (size_t i = 0; i < qty; i++) {
    float t = a[i] - b[i];
    res += t * t;
  }
  return res;
}
/* space_ip.h:6-14 */
static float dot_scalar(const float *a, const float *b, size_t qty) {
  float res = 0;
  for (size_t i = 0; i < qty; i++) res += a[i] * b[i];
  return res;
}

/* space_l2.h:60-91 L2SqrSIMD16ExtAVX: 8 lane accumulators, two 8-blocks per 16, lane sum in index order */
static float l2_hnsw_avx16(const float *a, const float *b, size_t qty) {
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  size_t q16 = (qty >> 4) << 4;
  for (size_t i = 0; i < q16; i += 8)
    for (int l = 0; l < 8; l++) {
      float d = a[i + l] - b[i + l];
      s[l] = s[l] + d * d;
    }
  return s[0] + s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7];
}
/* space_l2.h:165-189 L2SqrSIMD4Ext: 4 SSE lanes */
static float l2_hnsw_sse4(const float *a, const float *b, size_t qty) {
  float s[4] = {0, 0, 0, 0};
  size_t q4 = (qty >> 2) << 2;
  for (size_t i = 0; i < q4; i += 4)
    for (int l = 0; l < 4; l++) {
      float d = a[i + l] - b[i + l];
      s[l] = s[l] + d * d;
    }
  return s[0] + s[1] + s[2] + s[3];
}
/* L2Space::L2Space dispatch, space_l2.h:214-241 (+ residual wrappers :149-160, :191-205) */
static float l2_hnsw_avx(const float *a, const float *b, size_t dim) {
  if (dim % 16 == 0) return l2_hnsw_avx16(a, b, dim);
  if (dim % 4 == 0) return l2_hnsw_sse4(a, b, dim);
  if (dim > 16) {
    size_t q16 = dim >> 4 << 4;
    float res = l2_hnsw_avx16(a, b, q16);
    float tail = l2_scalar(a + q16, b + q16, dim - q16);
    return res + tail;
  }
  if (dim > 4) {
    size_t q4 = dim >> 2 << 2;
    float res = l2_hnsw_sse4(a, b, q4);
    float tail = l2_scalar(a + q4, b + q4, dim - q4);
    return res + tail;
  }
  return l2_scalar(a, b, dim);
}

/* space_ip.h:24-69 InnerProductSIMD4ExtAVX: 8-lane blocks of 16, fold to 4 lanes, 4-wide tail */
static float dot_hnsw_avx4(const float *a, const float *b, size_t qty) {
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  size_t q16 = (qty / 16) * 16, q4 = (qty / 4) * 4;
  size_t i = 0;
  for (; i < q16; i += 8)
    for (int l = 0; l < 8; l++) s[l] = s[l] + a[i + l] * b[i + l];
  float p[4];
  for (int l = 0; l < 4; l++) p[l] = s[l] + s[l + 4];
  for (; i < q4; i += 4)
    for (int l = 0; l < 4; l++) p[l] = p[l] + a[i + l] * b[i + l];
  return p[0] + p[1] + p[2] + p[3];
}
/* space_ip.h InnerProductSIMD16ExtAVX (:210-245): 8 lanes, index-order lane sum */
static float dot_hnsw_avx16(const float *a, const float *b, size_t qty) {
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  size_t q16 = (qty / 16) * 16;
  for (size_t i = 0; i < q16; i += 8)
    for (int l = 0; l < 8; l++) s[l] = s[l] + a[i + l] * b[i + l];
  return s[0] + s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7];
}
/* InnerProductSpace dispatch, space_ip.h:348-385 (+ residual wrappers :313-339) */
static float ip_hnsw_avx(const float *a, const float *b, size_t dim) {
  if (dim % 16 == 0) return 1.0f - dot_hnsw_avx16(a, b, dim);
  if (dim % 4 == 0) return 1.0f - dot_hnsw_avx4(a, b, dim);
  if (dim > 16) {
    size_t q16 = dim >> 4 << 4;
    float res = dot_hnsw_avx16(a, b, q16);
    float tail = dot_scalar(a + q16, b + q16, dim - q16);
    return 1.0f - (res + tail);
  }
  if (dim > 4) {
    size_t q4 = dim >> 2 << 2;
    float res = dot_hnsw_avx4(a, b, q4);
    float tail = dot_scalar(a + q4, b + q4, dim - q4);
    return 1.0f - (res + tail);
  }
  return 1.0f - dot_scalar(a, b, dim);
}

/* efanna2e::DistanceL2::compare, __AVX__ branch (CNNS/include/nsg/distance.h:29-56).
 * D rounds dim up to x8 and the reference then reads past the row end unless the caller padded the
 * rows with zeros (util.h data_align :33-67); the restatement treats elements >= dim as zero. */
static float l2_efanna_avx(const float *a, const float *b, size_t size) {
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  size_t D = (size + 7) & ~(size_t)7, DR = D % 16, DD = D - DR;
  if (DR)
    for (int l = 0; l < 8; l++) {
      size_t i = DD + l;
      float d = (i < size ? a[i] : 0.f) - (i < size ? b[i] : 0.f);
      s[l] = s[l] + d * d;
    }
  for (size_t i = 0; i < DD; i += 8)
    for (int l = 0; l < 8; l++) {
      float d = a[i + l] - b[i + l];
      s[l] = s[l] + d * d;
    }
  return s[0] + s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7];
}
/* efanna2e::DistanceInnerProduct::compare (__AVX__, distance.h:133-171) / DistanceFastL2::norm (:232-258):
 * same lane structure as l2_efanna_avx with a product instead of a squared difference. */
static float dot_efanna_avx(const float *a, const float *b, size_t size) {
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  size_t D = (size + 7) & ~(size_t)7, DR = D % 16, DD = D - DR;
  if (DR)
    for (int l = 0; l < 8; l++) {
      size_t i = DD + l;
      s[l] = s[l] + (i < size ? a[i] : 0.f) * (i < size ? b[i] : 0.f);
    }
  for (size_t i = 0; i < DD; i += 8)
    for (int l = 0; l < 8; l++) s[l] = s[l] + a[i + l] * b[i + l];
  return s[0] + s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7];
}

/* hnsw_nsg::DistanceL2AVX::compare (hnsw_nsg/include/nsg/distance.h:325-346): 8 lanes, mul then add,
 * lanes summed in index order. The reference steps i += 8 while i < size, i.e. it reads past the row
 * end when size % 8 != 0; the restatement treats elements >= size as zero. */
static float l2_hybrid_avx(const float *a, const float *b, size_t size) {
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (size_t i = 0; i < size; i += 8)
    for (int l = 0; l < 8; l++) {
      float d = (i + l < size ? a[i + l] : 0.f) - (i + l < size ? b[i + l] : 0.f);
      s[l] = s[l] + d * d;
    }
  return s[0] + s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7];
}

/* Summation order of the CUDA kernels (hnsw-nsg_b200/csrc/dist.cuh): a row of nch = ceil(dim/4)
 * float4 chunks is split over a team of `team` lanes; lane t owns chunks t, t+team, t+2*team, ...
 * and accumulates x,y,z,w of each chunk in that order with fmaf; lane sums are combined with an
 * xor butterfly (offsets team/2 ... 1). Not a reference function: it exists so that GPU-vs-oracle
 * comparisons are bit-exact on arbitrary float data. */
static int g_gpu_team = 8;
static float gpu_order(const float *a, const float *b, size_t dim, int team, int ip) {
  float lane[32];
  size_t nch = (dim + 3) / 4;
  for (int t = 0; t < team; t++) {
    float acc = 0.f;
    for (size_t c = t; c < nch; c += team)
      for (int j = 0; j < 4; j++) {
        size_t i = c * 4 + j;
        float x = i < dim ? a[i] : 0.f, y = i < dim ? b[i] : 0.f;
        if (ip) acc = fmaf(x, y, acc);
        else { float d = x - y; acc = fmaf(d, d, acc); }
      }
    lane[t] = acc;
  }
  for (int off = team / 2; off >= 1; off >>= 1) {
    float nxt[32];
    for (int t = 0; t < team; t++) nxt[t] = lane[t] + lane[t ^ off];
    memcpy(lane, nxt, sizeof(float) * team);
  }
  return ip ? 1.0f - lane[0] : lane[0];
}

/* csrc/bf_knn.cu accumulates one (query,row) pair sequentially over the dimension with fmaf */
static float seq_fma(const float *a, const float *b, size_t dim, int ip) {
  float acc = 0.f;
  for (size_t i = 0; i < dim; i++) {
    if (ip) acc = fmaf(a[i], b[i], acc);
    else { float d = a[i] - b[i]; acc = fmaf(d, d, acc); }
  }
  return ip ? 1.0f - acc : acc;
}

EXPORT void orc_set_gpu_team(int team) { g_gpu_team = team; }

/* space_l2.h:255-292 (L2SqrI4x unrolls by 4, L2SqrI is the plain loop; int sums, so both give the same value): the
 * rows are dim unsigned bytes behind the float pointers, the distance is an int. It is returned as the float with
 * the same bit pattern: for non-negative ints below 0x7F800000 float order == int order (denormals included; this
 * file is not built with fast-math), +inf stays above every distance, and the HNSW restatements below only ever
 * COMPARE distances - so they serve HierarchicalNSW<int> unchanged. Row strides stay in 4-byte units. */
static float l2_u8(const float *a, const float *b, size_t dim) {
  const unsigned char *x = (const unsigned char *)a, *y = (const unsigned char *)b;
  int res = 0;
  for (size_t i = 0; i < dim; i++) { int d = (int)x[i] - (int)y[i]; res += d * d; }
  float f;
  memcpy(&f, &res, 4);
  return f;
}

EXPORT float orc_dist(int mode, const float *a, const float *b, size_t dim) {
  switch (mode) {
    case ORC_L2_U8: return l2_u8(a, b, dim);
    case ORC_L2_SCALAR: return l2_scalar(a, b, dim);
    case ORC_IP_SCALAR: return 1.0f - dot_scalar(a, b, dim);
    case ORC_L2_HNSW_AVX: return l2_hnsw_avx(a, b, dim);
    case ORC_IP_HNSW_AVX: return ip_hnsw_avx(a, b, dim);
    case ORC_L2_EFANNA_AVX: return l2_efanna_avx(a, b, dim);
    case ORC_L2_GPU: return gpu_order(a, b, dim, g_gpu_team, 0);
    case ORC_IP_GPU: return gpu_order(a, b, dim, g_gpu_team, 1);
    case ORC_L2_SEQ_FMA: return seq_fma(a, b, dim, 0);
    case ORC_IP_SEQ_FMA: return seq_fma(a, b, dim, 1);
    case ORC_L2_HYBRID_AVX: return l2_hybrid_avx(a, b, dim);
    case ORC_FASTL2_EFANNA_AVX: /* distance.h:318-322 with norm = ||b||^2 (:232-258) */
      return -2 * dot_efanna_avx(a, b, dim) + dot_efanna_avx(b, b, dim);
  }
  return NAN;
}

/* ------------------------------------------------------------------ brute-force kNN */
/* hnswlib::BruteforceSearch::searchKnn (bruteforce.h:106-135) and IndexGraph::generate_control_set
 * (efanna_graph/src/index_graph.cpp:138-153): exact scan, keep the k smallest. The reference heap is
 * std::priority_queue<pair<dist,label>> (lexicographic), admits dist <= lastdist and pops the largest
 * pair, so the survivors are the k smallest under (distance, id) and ties go to the smaller id.
 * Output rows ascending by (distance, id) (searchKnnCloserFirst order, hnswlib.h:203-221);
 * if n < k the tail is filled with ORC_INVALID / +inf. */
typedef struct { float d; uint32_t id; } orc_pair;

static int pair_less(orc_pair a, orc_pair b) { return a.d < b.d || (a.d == b.d && a.id < b.id); }
static int pair_cmp(const void *x, const void *y) {
  orc_pair a = *(const orc_pair *)x, b = *(const orc_pair *)y;
  return pair_less(a, b) ? -1 : (pair_less(b, a) ? 1 : 0);
}
/* max-heap on (d,id) */
static void heap_sift_down(orc_pair *h, size_t n, size_t i) {
  for (;;) {
    size_t l = 2 * i + 1, r = l + 1, m = i;
    if (l < n && pair_less(h[m], h[l])) m = l;
    if (r < n && pair_less(h[m], h[r])) m = r;
    if (m == i) return;
    orc_pair t = h[i]; h[i] = h[m]; h[m] = t; i = m;
  }
}
static void heap_sift_up(orc_pair *h, size_t i) {
  while (i > 0) {
    size_t p = (i - 1) / 2;
    if (!pair_less(h[p], h[i])) return;
    orc_pair t = h[i]; h[i] = h[p]; h[p] = t; i = p;
  }
}

EXPORT int orc_bf_knn(const float *base, size_t n, size_t dim, size_t stride, const float *q, size_t nq,
                      size_t qstride, size_t k, int mode, uint32_t *ids, float *dists) {
  if (k == 0) return 0;
#pragma omp parallel
  {
    orc_pair *h = (orc_pair *)malloc(sizeof(orc_pair) * k);
#pragma omp for schedule(dynamic, 4)
    for (long qi = 0; qi < (long)nq; qi++) {
      const float *qv = q + (size_t)qi * qstride;
      size_t m = 0;
      for (size_t i = 0; i < n; i++) {
        orc_pair c = {orc_dist(mode, qv, base + i * stride, dim), (uint32_t)i};
        if (m < k) { h[m] = c; heap_sift_up(h, m); m++; }
        else if (pair_less(c, h[0])) { h[0] = c; heap_sift_down(h, m, 0); }
      }
      qsort(h, m, sizeof(orc_pair), pair_cmp);
      for (size_t j = 0; j < k; j++) {
        ids[(size_t)qi * k + j] = j < m ? h[j].id : ORC_INVALID;
        dists[(size_t)qi * k + j] = j < m ? h[j].d : INFINITY;
      }
    }
    free(h);
  }
  return 0;
}

/* hnswlib::L2SqrI (space_l2.h:281-292; L2SqrI4x :255-279 is the same sum unrolled by 4): int accumulation of
 * (a-b)^2 over unsigned bytes. Integer arithmetic: no rounding, no order dependence. */
static int orc_l2sqr_u8(const uint8_t *a, const uint8_t *b, size_t dim) {
  int res = 0;
  for (size_t i = 0; i < dim; i++) { int d = (int)a[i] - (int)b[i]; res += d * d; }
  return res;
}

/* hnswlib::BruteforceSearch<int>::searchKnn (bruteforce.h:106-135) over L2SpaceI, every query of a batch:
 * the k smallest under (distance, id) (the reference's `dist <= lastdist` admission + pair max-heap eviction
 * keeps the smaller label among equals), ascending. Unfilled slots: ORC_INVALID / INT32_MAX. */
EXPORT int orc_bf_knn_u8(const uint8_t *base, size_t n, size_t dim, size_t stride, const uint8_t *q, size_t nq,
                         size_t qstride, size_t k, uint32_t *ids, int32_t *dists) {
  if (k == 0) return 0;
#pragma omp parallel
  {
    int32_t *hd = (int32_t *)malloc(sizeof(int32_t) * k);
    uint32_t *hi = (uint32_t *)malloc(sizeof(uint32_t) * k);
#pragma omp for schedule(dynamic, 4)
    for (long qi = 0; qi < (long)nq; qi++) {
      const uint8_t *qv = q + (size_t)qi * qstride;
      size_t m = 0;  /* sorted ascending list of the best m <= k, insertion from the back */
      for (size_t i = 0; i < n; i++) {
        int32_t d = orc_l2sqr_u8(qv, base + i * stride, dim);
        if (m == k && !(d < hd[k - 1])) continue;  /* ids arrive ascending: an equal distance never displaces */
        size_t j = m < k ? m++ : k - 1;
        while (j > 0 && hd[j - 1] > d) { hd[j] = hd[j - 1]; hi[j] = hi[j - 1]; j--; }
        hd[j] = d; hi[j] = (uint32_t)i;
      }
      for (size_t j = 0; j < k; j++) {
        ids[(size_t)qi * k + j] = j < m ? hi[j] : ORC_INVALID;
        dists[(size_t)qi * k + j] = j < m ? hd[j] : INT32_MAX;
      }
    }
    free(hd); free(hi);
  }
  return 0;
}

/* ------------------------------------------------------------------ sorted pool (efanna Neighbor) */
/* efanna2e::Neighbor (CNNS/include/nsg/neighbor.h:18-30) */
typedef struct { uint32_t id; float distance; int flag; } orc_nb;

/* efanna2e::InsertIntoPool (neighbor.h:107-135), statement for statement. addr has K+1 slots. */
static int insert_into_pool(orc_nb *addr, unsigned K, orc_nb nn) {
  int left = 0, right = (int)K - 1;
  if (addr[left].distance > nn.distance) {
    memmove(&addr[left + 1], &addr[left], K * sizeof(orc_nb));
    addr[left] = nn;
    return left;
  }
  if (addr[right].distance < nn.distance) {
    addr[K] = nn;
    return (int)K;
  }
  while (left < right - 1) {
    int mid = (left + right) / 2;
    if (addr[mid].distance > nn.distance) right = mid;
    else left = mid;
  }
  while (left > 0) {
    if (addr[left].distance < nn.distance) break;
    if (addr[left].id == nn.id) return (int)K + 1;
    left--;
  }
  if (addr[left].id == nn.id || addr[right].id == nn.id) return (int)K + 1;
  memmove(&addr[right + 1], &addr[right], (K - right) * sizeof(orc_nb));
  addr[right] = nn;
  return right;
}
EXPORT int orc_insert_into_pool(uint32_t *ids, float *dist, int *flags, unsigned K, uint32_t id, float d) {
  orc_nb *p = (orc_nb *)malloc(sizeof(orc_nb) * (K + 1));
  for (unsigned i = 0; i < K; i++) { p[i].id = ids[i]; p[i].distance = dist[i]; p[i].flag = flags[i]; }
  p[K].id = ORC_INVALID; p[K].distance = INFINITY; p[K].flag = 0;
  orc_nb nn = {id, d, 1};
  int r = insert_into_pool(p, K, nn);
  for (unsigned i = 0; i <= K; i++) { ids[i] = p[i].id; dist[i] = p[i].distance; flags[i] = p[i].flag; }
  free(p);
  return r;
}

/* std::sort(Neighbor) compares distance only (neighbor.h:26-28) and is unstable; the oracle orders
 * equal distances by id. */
static int nb_cmp(const void *x, const void *y) {
  const orc_nb *a = (const orc_nb *)x, *b = (const orc_nb *)y;
  if (a->distance < b->distance) return -1;
  if (a->distance > b->distance) return 1;
  return a->id < b->id ? -1 : (a->id > b->id ? 1 : 0);
}

/* random ids for seed fill. rng_mode 0: libc rand() exactly where the reference calls it
 * (index_nsg.cpp:523). rng_mode 1: the product's counter hash (csrc/search.cu seed_hash): stateless,
 * so batched/parallel execution is reproducible. */
static uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
EXPORT uint32_t orc_seed_hash(uint32_t seed, uint32_t query, uint32_t counter) {
  return mix32(mix32(seed ^ (query * 0x9E3779B9U)) + counter * 0x85EBCA6BU);
}

static unsigned row_degree(const uint32_t *row, unsigned gstride) {
  unsigned d = 0;
  while (d < gstride && row[d] != ORC_INVALID) d++;
  return d;
}

/* ------------------------------------------------------------------ NSG beam search */
/* Core of IndexNSG::Search (CNNS/src/nsg/index_nsg.cpp:506-568) and IndexNSG::get_neighbors
 * (:218-285): seeds = first L out-neighbours of ep, random fill up to L, sorted pool of L,
 * expand the first unexpanded entry, InsertIntoPool, k/nk rewind. When `full` is non-NULL every
 * evaluated node is appended to it (get_neighbors' `fullset`, :250,:273) and `flags` is the
 * caller's visited array (Link shares it with sync_prune, :447-448). stats[0] += distance
 * evaluations, stats[1] += expanded nodes. Returns the pool length (L). */
typedef struct { orc_nb *v; size_t n, cap; } nb_vec;
static void nbv_push(nb_vec *v, orc_nb x) {
  if (v->n == v->cap) { v->cap = v->cap ? v->cap * 2 : 1024; v->v = (orc_nb *)realloc(v->v, v->cap * sizeof(orc_nb)); }
  v->v[v->n++] = x;
}

typedef struct {
  const float *data; size_t n, dim, stride;
  const uint32_t *graph; unsigned gstride;
  int mode;
} orc_index;

static unsigned nsg_beam(const orc_index *ix, uint32_t ep, const float *query, unsigned L_in, int rng_mode,
                         uint32_t seed, uint32_t qidx, uint8_t *flags, orc_nb *retset, nb_vec *full,
                         int full_seeds, uint64_t *stats) {
  unsigned L = 0;
  uint32_t *init_ids = (uint32_t *)malloc(sizeof(uint32_t) * (L_in ? L_in : 1));
  const uint32_t *eprow = ix->graph + (size_t)ep * ix->gstride;
  unsigned epdeg = row_degree(eprow, ix->gstride);
  for (unsigned i = 0; i < L_in && i < epdeg; i++) { /* :516-519 / :229-233 */
    init_ids[i] = eprow[i];
    flags[init_ids[i]] = 1;
    L++;
  }
  uint32_t ctr = 0;
  while (L < L_in) { /* :521-527 / :234-240 */
    uint32_t id = rng_mode == 0 ? (uint32_t)rand() % (uint32_t)ix->n
                                : orc_seed_hash(seed, qidx, ctr++) % (uint32_t)ix->n;
    if (flags[id]) continue;
    init_ids[L] = id;
    L++;
    flags[id] = 1;
  }
  for (unsigned i = 0; i < L_in; i++) { /* :530-536 / :243-253 */
    uint32_t id = init_ids[i];
    float d = orc_dist(ix->mode, ix->data + ix->stride * (size_t)id, query, ix->dim);
    retset[i].id = id; retset[i].distance = d; retset[i].flag = 1;
    if (full && full_seeds) nbv_push(full, retset[i]); /* only the flags overload records seeds (:250) */
    if (stats) stats[0]++;
  }
  free(init_ids);
  L = L_in;
  qsort(retset, L, sizeof(orc_nb), nb_cmp); /* :538 */
  int k = 0;
  while (k < (int)L) { /* :539-564 */
    int nk = (int)L;
    if (retset[k].flag) {
      retset[k].flag = 0;
      uint32_t n = retset[k].id;
      const uint32_t *row = ix->graph + (size_t)n * ix->gstride;
      if (stats) stats[1]++;
      for (unsigned m = 0; m < ix->gstride && row[m] != ORC_INVALID; ++m) {
        uint32_t id = row[m];
        if (flags[id]) continue;
        flags[id] = 1;
        float d = orc_dist(ix->mode, query, ix->data + ix->stride * (size_t)id, ix->dim);
        if (stats) stats[0]++;
        orc_nb nn = {id, d, 1};
        if (full) nbv_push(full, nn);
        if (d >= retset[L - 1].distance) continue;
        int r = insert_into_pool(retset, L, nn);
        if (r < nk) nk = r;
      }
    }
    if (nk <= k) k = nk; else ++k;
  }
  return L;
}

/* IndexNSG::Search batched over queries (the reference loops / omp-parallelises outside, e.g.
 * CNNS/tests/cluster_hnsw_nsg_search.cpp:138,207). ids [nq,K]; dists [nq,K] (the reference returns
 * ids only — distances are an extra output, NULL allowed); stats [nq,2] = {n_dist, n_hops} or NULL.
 * rng_mode 0 consumes libc rand() query by query, single-threaded, like the reference under srand(). */
EXPORT int orc_nsg_search(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *graph,
                          unsigned gstride, uint32_t ep, const float *q, size_t nq, size_t qstride, unsigned K,
                          unsigned L, int mode, int rng_mode, uint32_t seed, uint32_t *ids, float *dists,
                          uint64_t *stats) {
  if (L < K || L == 0 || L > n) return -1;
  orc_index ix = {data, n, dim, stride, graph, gstride, mode};
  if (stats) memset(stats, 0, sizeof(uint64_t) * 2 * nq);
#pragma omp parallel if (rng_mode != 0)
  {
    uint8_t *flags = (uint8_t *)malloc(n);
    orc_nb *retset = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
#pragma omp for schedule(dynamic, 8)
    for (long qi = 0; qi < (long)nq; qi++) {
      memset(flags, 0, n);
      nsg_beam(&ix, ep, q + (size_t)qi * qstride, L, rng_mode, seed, (uint32_t)qi, flags, retset, NULL, 0,
               stats ? stats + 2 * qi : NULL);
      for (unsigned i = 0; i < K; i++) { /* :565-567 */
        ids[(size_t)qi * K + i] = retset[i].id;
        if (dists) dists[(size_t)qi * K + i] = retset[i].distance;
      }
    }
    free(flags); free(retset);
  }
  return 0;
}

/* same, with the hash stream index of every query given explicitly (qidx [nq]): what one cluster's share of a
 * routed batch looks like (CNNS/tests/cluster_hnsw_nsg_search.cpp:203-207 runs IndexNSG::Search per (query, cluster);
 * the product hashes with the query's position in the whole batch). */
EXPORT int orc_nsg_search_qidx(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *graph,
                               unsigned gstride, uint32_t ep, const float *q, size_t nq, size_t qstride, unsigned K,
                               unsigned L, int mode, uint32_t seed, const uint32_t *qidx, uint32_t *ids, float *dists) {
  if (L < K || L == 0 || L > n) return -1;
  orc_index ix = {data, n, dim, stride, graph, gstride, mode};
  uint8_t *flags = (uint8_t *)malloc(n);
  orc_nb *retset = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
  for (size_t qi = 0; qi < nq; qi++) {
    memset(flags, 0, n);
    nsg_beam(&ix, ep, q + qi * qstride, L, 1, seed, qidx[qi], flags, retset, NULL, 0, NULL);
    for (unsigned i = 0; i < K; i++) {
      ids[qi * K + i] = retset[i].id;
      if (dists) dists[qi * K + i] = retset[i].distance;
    }
  }
  free(flags); free(retset);
  return 0;
}

/* ------------------------------------------------------------------ HNSW search */
/* Flat HNSW layout (product device layout, built from hnswlib's saveIndex blob by the host shim):
 *   links0      u32 [n, maxM0]   level-0 lists (hnswalg.h get_linklist0 :486), INVALID padded
 *   levels      i32 [n]          element_levels_ (:63)
 *   upper_ptr   u32 [n]          first row of node i in upper_links (valid when levels[i] > 0)
 *   upper_links u32 [rows, maxM] row upper_ptr[i] + (l-1) = list of node i at level l>=1 (:496)
 *   labels      u64 [n]          external labels (getExternalLabel :185-189)
 */
typedef struct {
  const float *data; size_t n, dim, stride;
  const uint32_t *links0; unsigned maxM0;
  const int32_t *levels; const uint32_t *upper_ptr; const uint32_t *upper_links; unsigned maxM;
  uint32_t ep; int maxlevel; int mode;
} orc_hnsw;

/* generic binary heap on (d,id); sign=+1 max-heap (top_candidates), sign=-1 min-heap (candidate_set).
 * The reference uses CompareByFirst (hnswalg.h:68-73), i.e. leaves ties to the heap implementation;
 * the oracle breaks them by id. */
static int hless(orc_pair a, orc_pair b, int sign) { return sign > 0 ? pair_less(a, b) : pair_less(b, a); }
static void hpush(orc_pair *h, size_t *n, orc_pair x, int sign) {
  size_t i = (*n)++;
  h[i] = x;
  while (i > 0) {
    size_t p = (i - 1) / 2;
    if (!hless(h[p], h[i], sign)) break;
    orc_pair t = h[i]; h[i] = h[p]; h[p] = t; i = p;
  }
}
static void hpop(orc_pair *h, size_t *n, int sign) {
  h[0] = h[--(*n)];
  size_t i = 0, m = *n;
  for (;;) {
    size_t l = 2 * i + 1, r = l + 1, b = i;
    if (l < m && hless(h[b], h[l], sign)) b = l;
    if (r < m && hless(h[b], h[r], sign)) b = r;
    if (b == i) break;
    orc_pair t = h[i]; h[i] = h[b]; h[b] = t; i = b;
  }
}

/* upper-layer greedy descent, HierarchicalNSW::searchKnn (hnswalg.h:1275-1303) */
static uint32_t hnsw_descend(const orc_hnsw *ix, const float *q, float *curdist_out, uint64_t *stats) {
  uint32_t cur = ix->ep;
  float curdist = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)cur, ix->dim);
  if (stats) stats[0]++;
  for (int level = ix->maxlevel; level > 0; level--) {
    int changed = 1;
    while (changed) {
      changed = 0;
      const uint32_t *row = ix->upper_links + ((size_t)ix->upper_ptr[cur] + (level - 1)) * ix->maxM;
      if (stats) stats[1]++;
      for (unsigned i = 0; i < ix->maxM && row[i] != ORC_INVALID; i++) {
        uint32_t cand = row[i];
        float d = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)cand, ix->dim);
        if (stats) stats[0]++;
        if (d < curdist) { curdist = d; cur = cand; changed = 1; }
      }
    }
  }
  *curdist_out = curdist;
  return cur;
}

/* searchBaseLayerST<bare_bone=true> (hnswalg.h:309-440), two heaps exactly as the reference.
 * out: up to ef pairs, ascending (d,id). returns count. */
static size_t hnsw_base_twoheap(const orc_hnsw *ix, uint32_t ep, const float *q, size_t ef, uint8_t *visited,
                                orc_pair *top, orc_pair *cand, size_t cand_cap, uint64_t *stats) {
  size_t nt = 0, nc = 0;
  float dist = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)ep, ix->dim);
  if (stats) stats[0]++;
  float lowerBound = dist;
  orc_pair e = {dist, ep};
  hpush(top, &nt, e, +1);
  hpush(cand, &nc, e, -1);
  visited[ep] = 1;
  while (nc) {
    orc_pair cur = cand[0];
    if (cur.d > lowerBound) break; /* :347-349 */
    hpop(cand, &nc, -1);
    const uint32_t *row = ix->links0 + (size_t)cur.id * ix->maxM0;
    if (stats) stats[1]++;
    for (unsigned j = 0; j < ix->maxM0 && row[j] != ORC_INVALID; j++) { /* :377 */
      uint32_t cid = row[j];
      if (visited[cid]) continue;
      visited[cid] = 1;
      float d = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)cid, ix->dim);
      if (stats) stats[0]++;
      if (nt < ef || lowerBound > d) { /* :395 */
        orc_pair c = {d, cid};
        if (nc < cand_cap) hpush(cand, &nc, c, -1);
        hpush(top, &nt, c, +1);
        while (nt > ef) hpop(top, &nt, +1);
        if (nt) lowerBound = top[0].d;
      }
    }
  }
  qsort(top, nt, sizeof(orc_pair), pair_cmp);
  return nt;
}

/* Canonical single-pool form used by the CUDA kernel: one ascending (d,id) pool of <= ef entries,
 * each with an "expanded" flag; expand the best unexpanded entry; stop when none is left. It equals
 * the two-heap form whenever no two pool entries tie on distance at the eviction boundary
 * (an evicted entry has d >= lowerBound and is never expanded unless d == lowerBound). */
static size_t hnsw_base_pool(const orc_hnsw *ix, uint32_t ep, const float *q, size_t ef, uint8_t *visited,
                             orc_pair *pool, uint8_t *expanded, uint64_t *stats) {
  size_t np = 0;
  pool[0].d = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)ep, ix->dim);
  pool[0].id = ep; expanded[0] = 0; np = 1;
  if (stats) stats[0]++;
  visited[ep] = 1;
  for (;;) {
    size_t k = 0;
    while (k < np && expanded[k]) k++;
    if (k == np) break;
    expanded[k] = 1;
    const uint32_t *row = ix->links0 + (size_t)pool[k].id * ix->maxM0;
    if (stats) stats[1]++;
    for (unsigned j = 0; j < ix->maxM0 && row[j] != ORC_INVALID; j++) {
      uint32_t cid = row[j];
      if (visited[cid]) continue;
      visited[cid] = 1;
      orc_pair c = {orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)cid, ix->dim), cid};
      if (stats) stats[0]++;
      if (np < ef || c.d < pool[np - 1].d) {
        size_t pos = np;
        while (pos > 0 && pair_less(c, pool[pos - 1])) pos--;
        size_t last = np < ef ? np : ef - 1; /* index that receives the shifted tail */
        for (size_t t = last; t > pos; t--) { pool[t] = pool[t - 1]; expanded[t] = expanded[t - 1]; }
        pool[pos] = c; expanded[pos] = 0;
        if (np < ef) np++;
      }
    }
  }
  return np;
}

/* HierarchicalNSW::searchKnn batched (hnswalg.h:1270-1324; batch shape of
 * python_bindings/bindings.cpp:612-695): rows ascending by (distance,id), external labels.
 * variant 0 = two-heap form, 1 = single-pool form. Rows with fewer than k results are padded with
 * label ~0 / +inf. stats [nq,2] = {distance evaluations, expanded nodes incl. upper layers}. */
EXPORT int orc_hnsw_search(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *links0,
                           unsigned maxM0, const int32_t *levels, const uint32_t *upper_ptr,
                           const uint32_t *upper_links, unsigned maxM, const uint64_t *labels, uint32_t ep,
                           int maxlevel, const float *q, size_t nq, size_t qstride, size_t k, size_t ef,
                           int mode, int variant, uint64_t *out_labels, float *out_dists, uint64_t *stats) {
  orc_hnsw ix = {data, n, dim, stride, links0, maxM0, levels, upper_ptr, upper_links, maxM, ep, maxlevel, mode};
  size_t efk = ef > k ? ef : k; /* :1309 std::max(ef_, k) */
  if (stats) memset(stats, 0, sizeof(uint64_t) * 2 * nq);
#pragma omp parallel
  {
    uint8_t *visited = (uint8_t *)malloc(n ? n : 1);
    orc_pair *top = (orc_pair *)malloc(sizeof(orc_pair) * (efk + 2));
    orc_pair *cand = (orc_pair *)malloc(sizeof(orc_pair) * (n + 2));
    uint8_t *expd = (uint8_t *)malloc(efk + 2);
#pragma omp for schedule(dynamic, 8)
    for (long qi = 0; qi < (long)nq; qi++) {
      const float *qv = q + (size_t)qi * qstride;
      size_t m = 0;
      if (n) {
        memset(visited, 0, n);
        float cd;
        uint64_t *st = stats ? stats + 2 * qi : NULL;
        uint32_t cur = hnsw_descend(&ix, qv, &cd, st);
        if (st) st[0]--; /* the base layer re-evaluates the entry point; count it once */
        m = variant == 0 ? hnsw_base_twoheap(&ix, cur, qv, efk, visited, top, cand, n + 1, st)
                         : hnsw_base_pool(&ix, cur, qv, efk, visited, top, expd, st);
      }
      for (size_t j = 0; j < k; j++) {
        out_labels[(size_t)qi * k + j] = j < m ? (labels ? labels[top[j].id] : top[j].id) : ~(uint64_t)0;
        out_dists[(size_t)qi * k + j] = j < m ? top[j].d : INFINITY;
      }
    }
    free(visited); free(top); free(cand); free(expd);
  }
  return 0;
}

/* ------------------------------------------------------------------ filtered / deleted-element HNSW search */
/* searchBaseLayerST<bare_bone=false> without a stop condition (hnswalg.h:309-440), the form searchKnn runs when a
 * filter functor is given or elements are marked deleted (:1305-1313). allow[id] != 0 stands for
 * !isMarkedDeleted(id) && (*isIdAllowed)(getExternalLabel(id)) (:325-326, :406-407). Two heaps exactly as the
 * reference: every considered element goes to candidate_set, only allowed ones to top_candidates; the entry point
 * of a disallowed start is queued at FLT_MAX without a distance evaluation (:334-337). */
/* eps / min_cand: the second stop rule of EpsilonSearchStopCondition::should_stop_search (stop_condition.h:247-251)
 * with max_num_candidates_ = ef, for which should_consider_candidate (:255-258) and should_remove_extra (:260-263)
 * are the plain rules below; eps = INFINITY, min_cand = 0 never fires = the search without a stop condition. */
static size_t hnsw_base_twoheap_filtered(const orc_hnsw *ix, uint32_t ep, const float *q, size_t ef,
                                         const uint8_t *allow, uint8_t *visited, orc_pair *top, orc_pair *cand,
                                         uint64_t *stats, float eps, size_t min_cand) {
  size_t nt = 0, nc = 0;
  float lowerBound;
  if (allow[ep]) {
    float dist = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)ep, ix->dim);
    if (stats) stats[0]++;
    lowerBound = dist;
    orc_pair e = {dist, ep};
    hpush(top, &nt, e, +1);
    hpush(cand, &nc, e, -1);
  } else {
    lowerBound = FLT_MAX;
    orc_pair e = {lowerBound, ep};
    hpush(cand, &nc, e, -1);
  }
  visited[ep] = 1;
  while (nc) {
    orc_pair cur = cand[0];
    if (cur.d > lowerBound && nt == ef) break; /* :355, stop_condition.h:243 */
    if (cur.d > eps && nt >= min_cand) break;  /* stop_condition.h:247 */
    hpop(cand, &nc, -1);
    const uint32_t *row = ix->links0 + (size_t)cur.id * ix->maxM0;
    if (stats) stats[1]++;
    for (unsigned j = 0; j < ix->maxM0 && row[j] != ORC_INVALID; j++) {
      uint32_t cid = row[j];
      if (visited[cid]) continue;
      visited[cid] = 1;
      float d = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)cid, ix->dim);
      if (stats) stats[0]++;
      if (nt < ef || lowerBound > d) { /* :395 */
        orc_pair c = {d, cid};
        hpush(cand, &nc, c, -1);
        if (allow[cid]) hpush(top, &nt, c, +1); /* :406-408 */
        while (nt > ef) hpop(top, &nt, +1);
        if (nt) lowerBound = top[0].d;
      }
    }
  }
  qsort(top, nt, sizeof(orc_pair), pair_cmp);
  return nt;
}

/* The CUDA kernel's form of the same search: both heaps as ascending (d,id) arrays, candidate_set bounded by ccap
 * entries. It is the two-heap form verbatim until candidate_set would hold more than ccap entries; then the
 * farthest entry is dropped, and the query is reported inexact (*inexact = 1) as soon as the reference could have
 * expanded a dropped entry: some dropped entry y is not beyond reach (!(y.d > lowerBound && nt == ef)) and the next
 * candidate is not closer than it (or there is none). Without that report the result IS the reference's. */
static void sorted_insert(orc_pair *a, size_t *n, orc_pair x) {
  size_t pos = *n;
  while (pos > 0 && pair_less(x, a[pos - 1])) { a[pos] = a[pos - 1]; pos--; }
  a[pos] = x;
  (*n)++;
}
static size_t hnsw_base_pools_filtered(const orc_hnsw *ix, uint32_t ep, const float *q, size_t ef, const uint8_t *allow,
                                       size_t ccap, uint8_t *visited, orc_pair *top, orc_pair *cand, uint64_t *stats,
                                       int *inexact) {
  size_t nt = 0, nc = 0;
  float lowerBound, dmin = INFINITY;
  *inexact = 0;
  if (allow[ep]) {
    float dist = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)ep, ix->dim);
    if (stats) stats[0]++;
    lowerBound = dist;
    orc_pair e = {dist, ep};
    top[nt++] = e;
    cand[nc++] = e;
  } else {
    lowerBound = FLT_MAX;
    orc_pair e = {lowerBound, ep};
    cand[nc++] = e;
  }
  visited[ep] = 1;
  for (;;) {
    if (dmin < INFINITY && (nc == 0 || dmin <= cand[0].d) && !(dmin > lowerBound && nt == ef)) *inexact = 1;
    if (nc == 0) break;
    orc_pair cur = cand[0];
    if (cur.d > lowerBound && nt == ef) break;
    memmove(cand, cand + 1, sizeof(orc_pair) * --nc);
    const uint32_t *row = ix->links0 + (size_t)cur.id * ix->maxM0;
    if (stats) stats[1]++;
    for (unsigned j = 0; j < ix->maxM0 && row[j] != ORC_INVALID; j++) {
      uint32_t cid = row[j];
      if (visited[cid]) continue;
      visited[cid] = 1;
      float d = orc_dist(ix->mode, q, ix->data + ix->stride * (size_t)cid, ix->dim);
      if (stats) stats[0]++;
      if (nt < ef || lowerBound > d) {
        orc_pair c = {d, cid};
        if (nc == ccap) { /* full: the farthest of (entries, c) goes */
          orc_pair y = c;
          if (pair_less(c, cand[nc - 1])) { y = cand[--nc]; sorted_insert(cand, &nc, c); }
          if (!(nt == ef && y.d > lowerBound) && y.d < dmin) dmin = y.d;
        } else {
          sorted_insert(cand, &nc, c);
        }
        if (allow[cid]) {
          if (nt == ef) {
            if (pair_less(c, top[nt - 1])) { nt--; sorted_insert(top, &nt, c); }
          } else {
            sorted_insert(top, &nt, c);
          }
        }
        if (nt) lowerBound = top[nt - 1].d;
      }
    }
  }
  return nt;
}

/* HierarchicalNSW::searchKnn(query, k, isIdAllowed) batched, for an index with a filter and / or deleted elements.
 * allow [n] by INTERNAL id. variant 0 = two heaps (the reference), 1 = the kernel's bounded arrays (ccap entries of
 * candidate_set; inexact [nq] or NULL). Output as orc_hnsw_search. */
EXPORT int orc_hnsw_search_filtered(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *links0,
                                    unsigned maxM0, const int32_t *levels, const uint32_t *upper_ptr,
                                    const uint32_t *upper_links, unsigned maxM, const uint64_t *labels, uint32_t ep,
                                    int maxlevel, const uint8_t *allow, const float *q, size_t nq, size_t qstride,
                                    size_t k, size_t ef, size_t ccap, int mode, int variant, uint64_t *out_labels,
                                    float *out_dists, uint64_t *stats, uint8_t *inexact) {
  orc_hnsw ix = {data, n, dim, stride, links0, maxM0, levels, upper_ptr, upper_links, maxM, ep, maxlevel, mode};
  size_t efk = ef > k ? ef : k;
  if (stats) memset(stats, 0, sizeof(uint64_t) * 2 * nq);
  if (variant == 1 && ccap == 0) return -1;
#pragma omp parallel
  {
    uint8_t *visited = (uint8_t *)malloc(n ? n : 1);
    orc_pair *top = (orc_pair *)malloc(sizeof(orc_pair) * (efk + 2));
    orc_pair *cand = (orc_pair *)malloc(sizeof(orc_pair) * ((variant == 0 ? n : ccap) + 2));
#pragma omp for schedule(dynamic, 8)
    for (long qi = 0; qi < (long)nq; qi++) {
      const float *qv = q + (size_t)qi * qstride;
      size_t m = 0;
      int bad = 0;
      if (n) {
        memset(visited, 0, n);
        float cd;
        uint64_t *st = stats ? stats + 2 * qi : NULL;
        uint32_t cur = hnsw_descend(&ix, qv, &cd, st);
        if (st && allow[cur]) st[0]--; /* an allowed entry point is evaluated again by the base layer; count it once */
        m = variant == 0 ? hnsw_base_twoheap_filtered(&ix, cur, qv, efk, allow, visited, top, cand, st, INFINITY, 0)
                         : hnsw_base_pools_filtered(&ix, cur, qv, efk, allow, ccap, visited, top, cand, st, &bad);
      }
      if (inexact) inexact[qi] = (uint8_t)bad;
      for (size_t j = 0; j < k; j++) {
        out_labels[(size_t)qi * k + j] = j < m ? (labels ? labels[top[j].id] : top[j].id) : ~(uint64_t)0;
        out_dists[(size_t)qi * k + j] = j < m ? top[j].d : INFINITY;
      }
    }
    free(visited); free(top); free(cand);
  }
  return 0;
}

/* HierarchicalNSW::searchStopConditionClosest(query, EpsilonSearchStopCondition(epsilon, min, max), isIdAllowed)
 * batched (hnswalg.h:1327-1378): greedy descent, searchBaseLayerST<false>(ep, q, 0, filter, &stop) with the stop
 * condition's counters (max plays the part of ef), ascending list, filter_results (stop_condition.h:265-272: drop the
 * tail beyond epsilon, truncate to max). allow [n] by internal id or NULL (everything allowed). The reference puts
 * INTERNAL ids into the result (:1368-1371 copies top_candidates' (dist, tableint) pairs); labels != NULL maps them
 * to labels instead. out_* [nq, max] padded with ~0 / inf, counts [nq]. */
EXPORT int orc_hnsw_search_epsilon(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *links0,
                                   unsigned maxM0, const int32_t *levels, const uint32_t *upper_ptr,
                                   const uint32_t *upper_links, unsigned maxM, const uint64_t *labels, uint32_t ep,
                                   int maxlevel, const uint8_t *allow, const float *q, size_t nq, size_t qstride,
                                   float epsilon, size_t min_cand, size_t max_cand, int mode, uint64_t *out_labels,
                                   float *out_dists, uint32_t *counts, uint64_t *stats) {
  orc_hnsw ix = {data, n, dim, stride, links0, maxM0, levels, upper_ptr, upper_links, maxM, ep, maxlevel, mode};
  if (stats) memset(stats, 0, sizeof(uint64_t) * 2 * nq);
  if (min_cand > max_cand || max_cand == 0) return -1; /* stop_condition.h:227 asserts */
  uint8_t *all = NULL;
  if (!allow) {
    all = (uint8_t *)malloc(n ? n : 1);
    memset(all, 1, n);
    allow = all;
  }
#pragma omp parallel
  {
    uint8_t *visited = (uint8_t *)malloc(n ? n : 1);
    orc_pair *top = (orc_pair *)malloc(sizeof(orc_pair) * (max_cand + 2));
    orc_pair *cand = (orc_pair *)malloc(sizeof(orc_pair) * (n + 2));
#pragma omp for schedule(dynamic, 8)
    for (long qi = 0; qi < (long)nq; qi++) {
      const float *qv = q + (size_t)qi * qstride;
      size_t m = 0;
      if (n) {
        memset(visited, 0, n);
        float cd;
        uint64_t *st = stats ? stats + 2 * qi : NULL;
        uint32_t cur = hnsw_descend(&ix, qv, &cd, st);
        if (st && allow[cur]) st[0]--;
        m = hnsw_base_twoheap_filtered(&ix, cur, qv, max_cand, allow, visited, top, cand, st, epsilon, min_cand);
        while (m && top[m - 1].d > epsilon) m--; /* filter_results */
        if (m > max_cand) m = max_cand;
      }
      counts[qi] = (uint32_t)m;
      for (size_t j = 0; j < max_cand; j++) {
        out_labels[(size_t)qi * max_cand + j] = j < m ? (labels ? labels[top[j].id] : top[j].id) : ~(uint64_t)0;
        out_dists[(size_t)qi * max_cand + j] = j < m ? top[j].d : INFINITY;
      }
    }
    free(visited); free(top); free(cand);
  }
  free(all);
  return 0;
}

/* HierarchicalNSW::searchStopConditionClosest(query, MultiVectorSearchStopCondition(space, num_docs, ef_collection),
 * isIdAllowed) batched (hnswalg.h:1327-1378 -> searchBaseLayerST<false> :309-440; stop_condition.h:146-215). doc [n] =
 * document id of every element (what BaseMultiVectorSpace::get_doc_id reads from the row). The result set is bounded
 * by its number of distinct documents: should_consider_candidate (:191-194), add / remove_point (:167-183),
 * should_remove_extra (:196-199), should_stop_search (:186-189); filter_results drops the farthest results until
 * num_docs documents are left (:201-215). ef_collection = max(ef_collection, num_docs) (:164). Output rows of
 * max_results slots (ids, or labels when labels != NULL), counts [nq]; returns -2 if a row needs more slots. */
EXPORT int orc_hnsw_search_multivector(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *links0,
                                       unsigned maxM0, const int32_t *levels, const uint32_t *upper_ptr,
                                       const uint32_t *upper_links, unsigned maxM, const uint64_t *labels, uint32_t ep,
                                       int maxlevel, const uint8_t *allow, const uint32_t *doc, const float *q, size_t nq,
                                       size_t qstride, size_t num_docs, size_t ef_collection, size_t max_results, int mode,
                                       uint64_t *out_labels, float *out_dists, uint32_t *counts, uint64_t *stats) {
  orc_hnsw ix = {data, n, dim, stride, links0, maxM0, levels, upper_ptr, upper_links, maxM, ep, maxlevel, mode};
  if (stats) memset(stats, 0, sizeof(uint64_t) * 2 * nq);
  if (num_docs == 0 || max_results == 0) return -1;
  if (ef_collection < num_docs) ef_collection = num_docs;
  uint32_t maxdoc = 0;
  for (size_t i = 0; i < n; i++) if (doc[i] > maxdoc) maxdoc = doc[i];
  uint8_t *all = NULL;
  if (!allow) { all = (uint8_t *)malloc(n ? n : 1); memset(all, 1, n); allow = all; }
  int rc = 0;
#pragma omp parallel
  {
    uint8_t *visited = (uint8_t *)malloc(n ? n : 1);
    orc_pair *top = (orc_pair *)malloc(sizeof(orc_pair) * (n + 2));
    orc_pair *cand = (orc_pair *)malloc(sizeof(orc_pair) * (n + 2));
    uint32_t *dc = (uint32_t *)calloc((size_t)maxdoc + 1, 4);
#pragma omp for schedule(dynamic, 8)
    for (long qi = 0; qi < (long)nq; qi++) {
      const float *qv = q + (size_t)qi * qstride;
      size_t nt = 0, nc = 0, ndocs = 0;
      uint64_t *st = stats ? stats + 2 * qi : NULL;
      if (n) {
        memset(visited, 0, n);
        float cdv;
        uint32_t cur = hnsw_descend(&ix, qv, &cdv, st);
        if (st && allow[cur]) st[0]--;
        float lowerBound;
        if (allow[cur]) {
          float dist = orc_dist(mode, qv, data + stride * (size_t)cur, dim);
          if (st) st[0]++;
          lowerBound = dist;
          orc_pair e = {dist, cur};
          hpush(top, &nt, e, +1);
          if (dc[doc[cur]]++ == 0) ndocs++;
          hpush(cand, &nc, e, -1);
        } else {
          lowerBound = FLT_MAX;
          orc_pair e = {lowerBound, cur};
          hpush(cand, &nc, e, -1);
        }
        visited[cur] = 1;
        while (nc) {
          orc_pair c0 = cand[0];
          if (c0.d > lowerBound && ndocs == ef_collection) break;
          hpop(cand, &nc, -1);
          const uint32_t *row = links0 + (size_t)c0.id * maxM0;
          if (st) st[1]++;
          for (unsigned j = 0; j < maxM0 && row[j] != ORC_INVALID; j++) {
            uint32_t cid = row[j];
            if (visited[cid]) continue;
            visited[cid] = 1;
            float d = orc_dist(mode, qv, data + stride * (size_t)cid, dim);
            if (st) st[0]++;
            if (ndocs < ef_collection || lowerBound > d) {
              orc_pair c = {d, cid};
              hpush(cand, &nc, c, -1);
              if (allow[cid]) {
                hpush(top, &nt, c, +1);
                if (dc[doc[cid]]++ == 0) ndocs++;
              }
              while (ndocs > ef_collection) {
                uint32_t id = top[0].id;
                hpop(top, &nt, +1);
                if (--dc[doc[id]] == 0) ndocs--;
              }
              if (nt) lowerBound = top[0].d;
            }
          }
        }
        qsort(top, nt, sizeof(orc_pair), pair_cmp);
        while (ndocs > num_docs) {                  /* filter_results */
          uint32_t id = top[nt - 1].id;
          nt--;
          if (--dc[doc[id]] == 0) ndocs--;
        }
        for (size_t t = 0; t < nt; t++) dc[doc[top[t].id]] = 0;   /* the counters go back to zero for the next query */
      }
      if (nt > max_results) {
#pragma omp atomic write
        rc = -2;
        nt = max_results;
      }
      counts[qi] = (uint32_t)nt;
      for (size_t j = 0; j < max_results; j++) {
        out_labels[(size_t)qi * max_results + j] = j < nt ? (labels ? labels[top[j].id] : top[j].id) : ~(uint64_t)0;
        out_dists[(size_t)qi * max_results + j] = j < nt ? top[j].d : INFINITY;
      }
    }
    free(visited); free(top); free(cand); free(dc);
  }
  free(all);
  return rc;
}

/* ------------------------------------------------------------------ hybrid HNSW_NSG search */
/* IndexNSG::SearchFromEnterpoint (hnsw_nsg/src/index_nsg.cpp:703-782), statement for statement,
 * quirks included:
 *   - the entry point is flagged visited but never evaluated, so it cannot be returned (:711-712);
 *   - the "neighbours of the seeds" pass (:720-728) indexes final_graph_[id] with the running seed
 *     count tmp_l, and a flagged neighbour is skipped by `continue`, which still advances tmp_l and
 *     leaves that init_ids slot at its initial value 0 (std::vector<unsigned> init_ids(L), :707);
 *     such slots are evaluated as node 0 without flagging it (:741-747).
 * rng_mode as in nsg_beam. retset has L+1 slots. */
static void nsg_beam_from_ep(const orc_index *ix, uint32_t ep, const float *query, unsigned L, int rng_mode,
                             uint32_t seed, uint32_t qidx, uint8_t *flags, orc_nb *retset, uint64_t *stats) {
  uint32_t *init_ids = (uint32_t *)calloc(L ? L : 1, sizeof(uint32_t));
  flags[ep] = 1;
  unsigned tmp_l = 0;
  const uint32_t *eprow = ix->graph + (size_t)ep * ix->gstride;
  unsigned epdeg = row_degree(eprow, ix->gstride);
  for (; tmp_l < L && tmp_l < epdeg; tmp_l++) { /* :714-717 */
    init_ids[tmp_l] = eprow[tmp_l];
    flags[init_ids[tmp_l]] = 1;
  }
  unsigned cur = tmp_l;
  for (unsigned i = 0; i < cur && tmp_l < L; i++) { /* :720-728 */
    uint32_t id = init_ids[i];
    const uint32_t *row = ix->graph + (size_t)id * ix->gstride;
    unsigned deg = row_degree(row, ix->gstride);
    for (; tmp_l < L && tmp_l < deg; tmp_l++) {
      if (flags[row[tmp_l]]) continue;
      init_ids[tmp_l] = row[tmp_l];
      flags[init_ids[tmp_l]] = 1;
    }
  }
  uint32_t ctr = 0;
  while (tmp_l < L) { /* :733-739 */
    uint32_t id = rng_mode == 0 ? (uint32_t)rand() % (uint32_t)ix->n
                                : orc_seed_hash(seed, qidx, ctr++) % (uint32_t)ix->n;
    if (flags[id]) continue;
    flags[id] = 1;
    init_ids[tmp_l] = id;
    tmp_l++;
  }
  for (unsigned i = 0; i < L; i++) { /* :741-747 */
    uint32_t id = init_ids[i];
    retset[i].id = id;
    retset[i].distance = orc_dist(ix->mode, ix->data + ix->stride * (size_t)id, query, ix->dim);
    retset[i].flag = 1;
    if (stats) stats[0]++;
  }
  free(init_ids);
  qsort(retset, L, sizeof(orc_nb), nb_cmp); /* :749 */
  int k = 0;
  while (k < (int)L) { /* :750-778 */
    int nk = (int)L;
    if (retset[k].flag) {
      retset[k].flag = 0;
      uint32_t n = retset[k].id;
      const uint32_t *row = ix->graph + (size_t)n * ix->gstride;
      if (stats) stats[1]++;
      for (unsigned m = 0; m < ix->gstride && row[m] != ORC_INVALID; ++m) {
        uint32_t id = row[m];
        if (flags[id]) continue;
        flags[id] = 1;
        float d = orc_dist(ix->mode, query, ix->data + ix->stride * (size_t)id, ix->dim);
        if (stats) stats[0]++;
        if (d >= retset[L - 1].distance) continue;
        orc_nb nn = {id, d, 1};
        int r = insert_into_pool(retset, L, nn);
        if (r < nk) nk = r;
      }
    }
    if (nk <= k) k = nk; else ++k;
  }
}

/* SearchFromEnterpoint for a batch with explicit per-query entry points. ids/dists [nq,K]. */
EXPORT int orc_nsg_search_from(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *graph,
                               unsigned gstride, const uint32_t *entry, const float *q, size_t nq, size_t qstride,
                               unsigned K, unsigned L, int mode, int rng_mode, uint32_t seed, uint32_t *ids,
                               float *dists, uint64_t *stats) {
  if (L < K || L == 0 || L >= n) return -1;
  orc_index ix = {data, n, dim, stride, graph, gstride, mode};
  if (stats) memset(stats, 0, sizeof(uint64_t) * 2 * nq);
#pragma omp parallel if (rng_mode != 0)
  {
    uint8_t *flags = (uint8_t *)malloc(n);
    orc_nb *retset = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
#pragma omp for schedule(dynamic, 8)
    for (long qi = 0; qi < (long)nq; qi++) {
      memset(flags, 0, n);
      nsg_beam_from_ep(&ix, entry[qi], q + (size_t)qi * qstride, L, rng_mode, seed, (uint32_t)qi, flags, retset,
                       stats ? stats + 2 * qi : NULL);
      for (unsigned i = 0; i < K; i++) {
        ids[(size_t)qi * K + i] = retset[i].id;
        if (dists) dists[(size_t)qi * K + i] = retset[i].distance;
      }
    }
    free(flags); free(retset);
  }
  return 0;
}

/* hnsw_nsg::HNSW_NSG::searchKnn (hnsw_nsg/include/index_hnsw_nsg.h:107-151) batched: greedy descent on
 * HNSW levels >= 1 (:113-143, distances in mode_hnsw = the hnswlib space), external label of the node
 * it ends on (:148), then IndexNSG::SearchFromEnterpoint from that label with K = max(k, ef_) (:149)
 * and pool width L = "L_search" (distances in mode_nsg = DistanceL2AVX). ids/dists [nq,K].
 * entry_out (optional) [nq] = the entry labels. stats [nq,2] = {distance evaluations, expanded nodes},
 * upper layers included. */
EXPORT int orc_hybrid_search(const float *data, size_t n, size_t dim, size_t stride, const int32_t *levels,
                             const uint32_t *upper_ptr, const uint32_t *upper_links, unsigned maxM,
                             const uint64_t *labels, uint32_t hnsw_ep, int maxlevel, const uint32_t *graph,
                             unsigned gstride, const float *q, size_t nq, size_t qstride, unsigned K, unsigned L,
                             int mode_hnsw, int mode_nsg, int rng_mode, uint32_t seed, uint32_t *ids, float *dists,
                             uint32_t *entry_out, uint64_t *stats) {
  if (L < K || L == 0 || L >= n) return -1;
  orc_hnsw hx = {data, n, dim, stride, NULL, 0, levels, upper_ptr, upper_links, maxM, hnsw_ep, maxlevel, mode_hnsw};
  orc_index ix = {data, n, dim, stride, graph, gstride, mode_nsg};
  if (stats) memset(stats, 0, sizeof(uint64_t) * 2 * nq);
#pragma omp parallel if (rng_mode != 0)
  {
    uint8_t *flags = (uint8_t *)malloc(n);
    orc_nb *retset = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
#pragma omp for schedule(dynamic, 8)
    for (long qi = 0; qi < (long)nq; qi++) {
      const float *qv = q + (size_t)qi * qstride;
      uint64_t *st = stats ? stats + 2 * qi : NULL;
      float cd;
      uint32_t cur = hnsw_descend(&hx, qv, &cd, st);
      uint32_t entry = labels ? (uint32_t)labels[cur] : cur;
      if (entry_out) entry_out[qi] = entry;
      memset(flags, 0, n);
      nsg_beam_from_ep(&ix, entry, qv, L, rng_mode, seed, (uint32_t)qi, flags, retset, st);
      for (unsigned i = 0; i < K; i++) {
        ids[(size_t)qi * K + i] = retset[i].id;
        if (dists) dists[(size_t)qi * K + i] = retset[i].distance;
      }
    }
    free(flags); free(retset);
  }
  return 0;
}

/* ================================================================== NSG build */
/* IndexNSG::Build (CNNS/src/nsg/index_nsg.cpp:465-504) phase by phase. Graphs are flat u32 [n,gstride]
 * rows padded with ORC_INVALID; the cut graph (SimpleNeighbor[nd*R], :471) is two flat arrays
 * cut_ids/cut_d [n,R] whose unused tail carries ORC_INVALID / -1 (the reference only writes the -1
 * sentinel, :352-354). */

/* IndexNSG::init_graph (:287-303): centroid by sequential fp32 accumulation over rows, divided by
 * nd_; ep0 is the reference's `rand() % nd_` (:299, drawn by the caller); navigating node = nearest
 * pool entry of a beam search for the centroid started at ep0. */
EXPORT uint32_t orc_nsg_init_ep(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *graph,
                                unsigned gstride, uint32_t ep0, unsigned L, int mode, int rng_mode, uint32_t seed,
                                float *center_out) {
  float *center = (float *)calloc(dim, sizeof(float));
  for (size_t i = 0; i < n; i++)
    for (size_t j = 0; j < dim; j++) center[j] += data[i * stride + j];
  for (size_t j = 0; j < dim; j++) center[j] /= (float)(unsigned)n;
  if (center_out) memcpy(center_out, center, dim * sizeof(float));
  orc_index ix = {data, n, dim, stride, graph, gstride, mode};
  uint8_t *flags = (uint8_t *)calloc(n, 1);
  orc_nb *retset = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
  nsg_beam(&ix, ep0, center, L, rng_mode, seed, (uint32_t)n, flags, retset, NULL, 0, NULL);
  uint32_t ep = retset[0].id;
  free(center); free(flags); free(retset);
  return ep;
}

/* the occlusion rule shared by sync_prune (:328-345) and InterInsert (:386-404): scan `pool`
 * (ascending) from index `start`, keep p iff no kept r has p.id == r.id or d(r,p) < d(q,p);
 * stop at `range` kept, at the pool end, or at index maxc. res must hold `range` entries. */
static unsigned occlude_scan(const orc_index *ix, const orc_nb *pool, size_t npool, size_t start, size_t maxc,
                             unsigned range, orc_nb *res, uint64_t *n_eval) {
  unsigned nres = 0;
  if (start >= npool) return 0;
  res[nres++] = pool[start];
  while (nres < range && (++start) < npool && start < maxc) {
    const orc_nb *p = &pool[start];
    int occ = 0;
    for (unsigned t = 0; t < nres; t++) {
      if (p->id == res[t].id) { occ = 1; break; }
      float djk = orc_dist(ix->mode, ix->data + ix->stride * (size_t)res[t].id,
                           ix->data + ix->stride * (size_t)p->id, ix->dim);
      if (n_eval) (*n_eval)++;
      if (djk < p->distance) { occ = 1; break; }
    }
    if (!occ) res[nres++] = *p;
  }
  return nres;
}

/* Link, first loop (:437-457): per node get_neighbors (flags overload, :218-285) over the kNN graph
 * then sync_prune (:305-355). stats [n,3] = {beam distance evaluations, expanded nodes, prune
 * distance evaluations incl. the own-list additions} or NULL.
 * std::sort(pool) (:323) is unstable on equal distances; the oracle orders them by id. */
EXPORT int orc_nsg_link(const float *data, size_t n, size_t dim, size_t stride, const uint32_t *knn,
                        unsigned kstride, uint32_t ep, unsigned L, unsigned R, unsigned C, int mode, int rng_mode,
                        uint32_t seed, uint32_t *cut_ids, float *cut_d, uint64_t *stats) {
  orc_index ix = {data, n, dim, stride, knn, kstride, mode};
  if (stats) memset(stats, 0, sizeof(uint64_t) * 3 * n);
#pragma omp parallel if (rng_mode != 0)
  {
    uint8_t *flags = (uint8_t *)malloc(n);
    orc_nb *retset = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
    orc_nb *res = (orc_nb *)malloc(sizeof(orc_nb) * (R + 1));
    nb_vec full = {0, 0, 0};
#pragma omp for schedule(dynamic, 16)
    for (long q = 0; q < (long)n; q++) {
      uint64_t st[2] = {0, 0}, npr = 0;
      memset(flags, 0, n);
      full.n = 0;
      nsg_beam(&ix, ep, data + stride * (size_t)q, L, rng_mode, seed, (uint32_t)q, flags, retset, &full, 1, st);
      const uint32_t *row = knn + (size_t)q * kstride;
      for (unsigned m = 0; m < kstride && row[m] != ORC_INVALID; m++) { /* :314-321 */
        uint32_t id = row[m];
        if (flags[id]) continue;
        orc_nb nb = {id, orc_dist(mode, data + stride * (size_t)q, data + stride * (size_t)id, dim), 1};
        npr++;
        nbv_push(&full, nb);
      }
      qsort(full.v, full.n, sizeof(orc_nb), nb_cmp);
      size_t start = 0;
      if (full.n && full.v[0].id == (uint32_t)q) start++; /* :326 */
      unsigned nres = occlude_scan(&ix, full.v, full.n, start, C, R, res, &npr);
      for (unsigned t = 0; t < R; t++) {
        cut_ids[(size_t)q * R + t] = t < nres ? res[t].id : ORC_INVALID;
        cut_d[(size_t)q * R + t] = t < nres ? res[t].distance : -1.0f;
      }
      if (stats) { stats[3 * q] = st[0]; stats[3 * q + 1] = st[1]; stats[3 * q + 2] = npr; }
    }
    free(flags); free(retset); free(res); free(full.v);
  }
  return 0;
}

/* IndexNSG::InterInsert over all nodes (:357-424, called serially from Link :459-462 because the
 * `omp for` there is orphaned).
 * variant 0 — the reference, statement for statement, in node order, in place; including its quirk
 *   that a re-pruned row shorter than R keeps its stale tail (no -1 is written after :399-404).
 * variant 1 — the order-independent batch form the CUDA build uses: for every node `des`,
 *   candidates = its own sync_prune row + every n whose sync_prune row lists des (and that des does
 *   not list itself), each with the distance sync_prune stored; if they fit in R the reverse
 *   candidates are appended in ascending n (what the serial form does when no row overflows),
 *   otherwise the candidates are sorted by (distance,id) and pruned with the occlusion rule.
 *   Reads the sync_prune rows only (snapshot), writes new rows; rows are -1/INVALID terminated. */
EXPORT int orc_nsg_interinsert(const float *data, size_t n, size_t dim, size_t stride, unsigned R, int mode,
                               int variant, uint32_t *cut_ids, float *cut_d, uint64_t *n_eval_out) {
  orc_index ix = {data, n, dim, stride, NULL, 0, mode};
  uint64_t n_eval = 0;
  if (variant == 0) {
    orc_nb *tmp = (orc_nb *)malloc(sizeof(orc_nb) * (R + 2));
    orc_nb *res = (orc_nb *)malloc(sizeof(orc_nb) * (R + 2));
    for (size_t nn = 0; nn < n; nn++) {
      uint32_t *src_i = cut_ids + nn * R; float *src_d = cut_d + nn * R;
      for (size_t i = 0; i < R; i++) {
        if (src_d[i] == -1) break;
        orc_nb sn = {(uint32_t)nn, src_d[i], 1};
        size_t des = src_i[i];
        uint32_t *des_i = cut_ids + des * R; float *des_d = cut_d + des * R;
        unsigned nt = 0; int dup = 0;
        for (size_t j = 0; j < R; j++) {
          if (des_d[j] == -1) break;
          if ((uint32_t)nn == des_i[j]) { dup = 1; break; }
          tmp[nt].id = des_i[j]; tmp[nt].distance = des_d[j]; tmp[nt].flag = 1; nt++;
        }
        if (dup) continue;
        tmp[nt++] = sn;
        if (nt > R) {
          qsort(tmp, nt, sizeof(orc_nb), nb_cmp);
          unsigned nres = occlude_scan(&ix, tmp, nt, 0, (size_t)-1, R, res, &n_eval);
          for (unsigned t = 0; t < nres; t++) { des_i[t] = res[t].id; des_d[t] = res[t].distance; }
        } else {
          for (unsigned t = 0; t < R; t++)
            if (des_d[t] == -1) {
              des_i[t] = sn.id; des_d[t] = sn.distance;
              if (t + 1 < R) des_d[t + 1] = -1;
              break;
            }
        }
      }
    }
    free(tmp); free(res);
  } else {
    /* reverse lists (CSR by destination, sources ascending) from the snapshot */
    size_t *off = (size_t *)calloc(n + 1, sizeof(size_t));
    for (size_t i = 0; i < n; i++)
      for (unsigned j = 0; j < R && cut_d[i * R + j] != -1; j++) off[cut_ids[i * R + j] + 1]++;
    for (size_t i = 0; i < n; i++) off[i + 1] += off[i];
    size_t *fill = (size_t *)malloc(sizeof(size_t) * n);
    memcpy(fill, off, sizeof(size_t) * n);
    orc_nb *rev = (orc_nb *)malloc(sizeof(orc_nb) * (off[n] ? off[n] : 1));
    for (size_t i = 0; i < n; i++)
      for (unsigned j = 0; j < R && cut_d[i * R + j] != -1; j++) {
        orc_nb e = {(uint32_t)i, cut_d[i * R + j], 1};
        rev[fill[cut_ids[i * R + j]]++] = e;
      }
    uint32_t *new_i = (uint32_t *)malloc(sizeof(uint32_t) * n * R);
    float *new_d = (float *)malloc(sizeof(float) * n * R);
#pragma omp parallel reduction(+ : n_eval)
    {
      nb_vec cand = {0, 0, 0};
      orc_nb *res = (orc_nb *)malloc(sizeof(orc_nb) * (R + 1));
#pragma omp for schedule(dynamic, 64)
      for (long des = 0; des < (long)n; des++) {
        cand.n = 0;
        unsigned own = 0;
        for (unsigned j = 0; j < R && cut_d[(size_t)des * R + j] != -1; j++, own++) {
          orc_nb e = {cut_ids[(size_t)des * R + j], cut_d[(size_t)des * R + j], 1};
          nbv_push(&cand, e);
        }
        for (size_t e = off[des]; e < off[des + 1]; e++) {
          int dup = 0;
          for (unsigned j = 0; j < own; j++) dup |= cand.v[j].id == rev[e].id;
          if (!dup) nbv_push(&cand, rev[e]);
        }
        unsigned nres;
        const orc_nb *src;
        if (cand.n <= R) { nres = (unsigned)cand.n; src = cand.v; }
        else {
          qsort(cand.v, cand.n, sizeof(orc_nb), nb_cmp);
          uint64_t ne = 0;
          nres = occlude_scan(&ix, cand.v, cand.n, 0, (size_t)-1, R, res, &ne);
          n_eval += ne;
          src = res;
        }
        for (unsigned t = 0; t < R; t++) {
          new_i[(size_t)des * R + t] = t < nres ? src[t].id : ORC_INVALID;
          new_d[(size_t)des * R + t] = t < nres ? src[t].distance : -1.0f;
        }
      }
      free(cand.v); free(res);
    }
    memcpy(cut_ids, new_i, sizeof(uint32_t) * n * R);
    memcpy(cut_d, new_d, sizeof(float) * n * R);
    free(off); free(fill); free(rev); free(new_i); free(new_d);
  }
  if (n_eval_out) *n_eval_out = n_eval;
  return 0;
}

/* compaction of the cut graph into adjacency rows (:473-487), with the reference's length rule
 * (pool_size = index of the last entry before the first -1, plus one). graph [n,gstride], gstride >= R. */
EXPORT int orc_nsg_cut_to_graph(const uint32_t *cut_ids, const float *cut_d, size_t n, unsigned R, uint32_t *graph,
                                unsigned gstride) {
  for (size_t i = 0; i < n; i++) {
    unsigned pool_size = 0;
    for (unsigned j = 0; j < R; j++) {
      if (cut_d[i * R + j] == -1) break;
      pool_size = j;
    }
    pool_size++;
    for (unsigned j = 0; j < gstride; j++) graph[i * gstride + j] = j < pool_size ? cut_ids[i * R + j] : ORC_INVALID;
  }
  return 0;
}

/* IndexNSG::tree_grow (:750-764) with DFS (:684-710) and findroot (:712-749), on a flat graph
 * [n,gstride] (in place). DFS only computes the set reachable from `root`, so the traversal order
 * is immaterial and the oracle marks reachability with an explicit stack. findroot: the first
 * unreached id in index order; get_neighbors (first overload, :150-216 — its fullset holds the
 * expanded nodes' neighbours only, not the seeds) over the current graph; sort; the first reached
 * pool entry becomes the parent and gains the edge parent -> id.
 * Deviation forced by the flat layout: a parent must have a free slot (degree < gstride); the
 * reference's rows are unbounded vectors. With gstride >= the reference's final width both agree.
 * When no pool entry qualifies the reference draws rand() until it hits a reached node (:739-746);
 * rng_mode 1 draws seed_hash(seed, id, 0x10000 + t) instead. Returns the number of added edges;
 * *width_out = max degree (:759-763, never below `width_in` = R, :310). */
static void mark_reachable(const uint32_t *graph, unsigned gstride, uint8_t *flag, uint32_t root, size_t *cnt,
                           uint32_t *stack) {
  size_t sp = 0;
  if (!flag[root]) { flag[root] = 1; (*cnt)++; }
  stack[sp++] = root;
  while (sp) {
    uint32_t v = stack[--sp];
    const uint32_t *row = graph + (size_t)v * gstride;
    for (unsigned i = 0; i < gstride && row[i] != ORC_INVALID; i++)
      if (!flag[row[i]]) { flag[row[i]] = 1; (*cnt)++; stack[sp++] = row[i]; }
  }
}

EXPORT long orc_nsg_tree_grow(const float *data, size_t n, size_t dim, size_t stride, uint32_t *graph,
                              unsigned gstride, uint32_t ep, unsigned L, int mode, int rng_mode, uint32_t seed,
                              unsigned width_in, unsigned *width_out) {
  orc_index ix = {data, n, dim, stride, graph, gstride, mode};
  uint8_t *flag = (uint8_t *)calloc(n, 1), *vis = (uint8_t *)malloc(n);
  uint32_t *stack = (uint32_t *)malloc(sizeof(uint32_t) * n);
  orc_nb *retset = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
  nb_vec full = {0, 0, 0};
  size_t cnt = 0, scan = 0;
  long added = 0;
  uint32_t root = ep;
  while (cnt < n) {
    mark_reachable(graph, gstride, flag, root, &cnt, stack);
    if (cnt >= n) break;
    while (scan < n && flag[scan]) scan++;
    uint32_t id = (uint32_t)scan;
    memset(vis, 0, n);
    full.n = 0;
    nsg_beam(&ix, ep, data + stride * (size_t)id, L, rng_mode, seed, id, vis, retset, &full, 0, NULL);
    qsort(full.v, full.n, sizeof(orc_nb), nb_cmp);
    int found = 0;
    for (size_t i = 0; i < full.n && !found; i++) {
      uint32_t c = full.v[i].id;
      if (flag[c] && row_degree(graph + (size_t)c * gstride, gstride) < gstride) { root = c; found = 1; }
    }
    for (uint32_t t = 0; !found; t++) {
      uint32_t rid = rng_mode == 0 ? (uint32_t)rand() % (uint32_t)n : orc_seed_hash(seed, id, 0x10000u + t) % (uint32_t)n;
      if (flag[rid] && row_degree(graph + (size_t)rid * gstride, gstride) < gstride) { root = rid; found = 1; }
    }
    uint32_t *row = graph + (size_t)root * gstride;
    row[row_degree(row, gstride)] = id;
    added++;
  }
  unsigned w = width_in;
  for (size_t i = 0; i < n; i++) {
    unsigned d = row_degree(graph + i * gstride, gstride);
    if (d > w) w = d;
  }
  if (width_out) *width_out = w;
  free(flag); free(vis); free(stack); free(retset); free(full.v);
  return added;
}

/* ================================================================== shard-result merge */
/* per query: `parts` ascending lists of k (distance,id) -> the kout smallest of the union, ascending
 * distance. What the reference's cluster drivers do on the host with an unordered_map +
 * partial_sort (CNNS/tests/cluster_hnsw_nsg_search.cpp:228-262): ids are global and unique across
 * shards, so no dedup is needed. d/ids [parts,nq,k]. */
typedef struct { float d; uint32_t id; uint32_t ord; } orc_mpair;
static int mpair_cmp(const void *a, const void *b) {
  const orc_mpair *x = (const orc_mpair *)a, *y = (const orc_mpair *)b;
  if (x->d < y->d) return -1;
  if (x->d > y->d) return 1;
  return x->ord < y->ord ? -1 : (x->ord > y->ord ? 1 : 0);
}
/* order: (distance, part, position in the part) - a stable merge of lists that are ascending by distance. With row
 * shards (ids ascending with the part) and (distance,id)-sorted lists this is the (distance,id) order. */
EXPORT int orc_merge_topk(const float *d, const uint32_t *ids, size_t parts, size_t nq, size_t k, size_t kout,
                          float *out_d, uint32_t *out_i) {
  orc_mpair *tmp = (orc_mpair *)malloc(sizeof(orc_mpair) * parts * k);
  for (size_t q = 0; q < nq; q++) {
    size_t m = 0;
    for (size_t s = 0; s < parts; s++)
      for (size_t j = 0; j < k; j++) {
        size_t off = (s * nq + q) * k + j;
        if (ids[off] == ORC_INVALID) continue;
        tmp[m].d = d[off]; tmp[m].id = ids[off]; tmp[m].ord = (uint32_t)(s * k + j); m++;
      }
    qsort(tmp, m, sizeof(orc_mpair), mpair_cmp);
    for (size_t j = 0; j < kout; j++) {
      out_i[q * kout + j] = j < m ? tmp[j].id : ORC_INVALID;
      if (out_d) out_d[q * kout + j] = j < m ? tmp[j].d : INFINITY;
    }
  }
  free(tmp);
  return 0;
}

/* ================================================================== NN-descent (efanna_graph IndexGraph) */
/* std::mt19937 (the generator behind GenRandom's callers: index_graph.cpp:124,181; index_random.h:19).
 * Standard MT19937: seeding x[i] = 1812433253 * (x[i-1] ^ (x[i-1] >> 30)) + i, tempering 11/7/15/18. */
typedef struct { uint32_t mt[624]; int idx; } orc_mt;
static void mt_seed(orc_mt *m, uint32_t s) {
  m->mt[0] = s;
  for (int i = 1; i < 624; i++) m->mt[i] = 1812433253U * (m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) + (uint32_t)i;
  m->idx = 624;
}
static uint32_t mt_next(void *st) {
  orc_mt *m = (orc_mt *)st;
  if (m->idx >= 624) {
    for (int i = 0; i < 624; i++) {
      uint32_t y = (m->mt[i] & 0x80000000U) | (m->mt[(i + 1) % 624] & 0x7fffffffU);
      m->mt[i] = m->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1U) ? 0x9908b0dfU : 0U);
    }
    m->idx = 0;
  }
  uint32_t y = m->mt[m->idx++];
  y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680U; y ^= (y << 15) & 0xefc60000U; y ^= y >> 18;
  return y;
}
EXPORT void orc_mt19937(uint32_t seed, uint32_t *out, size_t count) {
  orc_mt m; mt_seed(&m, seed);
  for (size_t i = 0; i < count; i++) out[i] = mt_next(&m);
}
/* the product's stateless stream: draw c of stream (seed, node) = orc_seed_hash(seed, node, c) */
typedef struct { uint32_t seed, node, ctr; } orc_hs;
static uint32_t hs_next(void *st) {
  orc_hs *s = (orc_hs *)st;
  return orc_seed_hash(s->seed, s->node, s->ctr++);
}
static int u32_cmp(const void *a, const void *b) {
  uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
  return x < y ? -1 : (x > y ? 1 : 0);
}
/* efanna2e::GenRandom (efanna_graph/include/efanna2e/util.h:17-31): `size` distinct ids in [0,N) */
static void gen_random(uint32_t (*next)(void *), void *st, uint32_t *addr, unsigned size, unsigned N) {
  for (unsigned i = 0; i < size; ++i) addr[i] = next(st) % (N - size);
  qsort(addr, size, sizeof(uint32_t), u32_cmp);
  for (unsigned i = 1; i < size; ++i)
    if (addr[i] <= addr[i - 1]) addr[i] = addr[i - 1] + 1;
  unsigned off = next(st) % N;
  for (unsigned i = 0; i < size; ++i) addr[i] = (addr[i] + off) % N;
}
EXPORT void orc_gen_random(int rng, uint32_t seed, uint32_t node, uint32_t *addr, unsigned size, unsigned N) {
  if (rng == 0) { orc_mt m; mt_seed(&m, seed); gen_random(mt_next, &m, addr, size, N); }
  else { orc_hs s = {seed, node, 0}; gen_random(hs_next, &s, addr, size, N); }
}

/* libstdc++ heap algorithms over Neighbor with operator< = distance only (neighbor.h:26-28), restated
 * (bits/stl_heap.h __push_heap / __adjust_heap / make_heap / pop_heap): the element order inside the
 * heap decides what `pool.back()` is in IndexGraph::update (index_graph.cpp:74,88). */
static void std_push_heap_(orc_nb *f, long hole, long top, orc_nb v) {
  long parent = (hole - 1) / 2;
  while (hole > top && f[parent].distance < v.distance) {
    f[hole] = f[parent]; hole = parent; parent = (hole - 1) / 2;
  }
  f[hole] = v;
}
static void std_adjust_heap(orc_nb *f, long hole, long len, orc_nb v) {
  long top = hole, child = hole;
  while (child < (len - 1) / 2) {
    child = 2 * (child + 1);
    if (f[child].distance < f[child - 1].distance) child--;
    f[hole] = f[child]; hole = child;
  }
  if ((len & 1) == 0 && child == (len - 2) / 2) {
    child = 2 * (child + 1);
    f[hole] = f[child - 1]; hole = child - 1;
  }
  std_push_heap_(f, hole, top, v);
}
static void std_make_heap(orc_nb *f, long len) {
  if (len < 2) return;
  for (long parent = (len - 2) / 2;; parent--) {
    orc_nb v = f[parent];
    std_adjust_heap(f, parent, len, v);
    if (parent == 0) return;
  }
}
static void std_push_heap(orc_nb *f, long len) { orc_nb v = f[len - 1]; std_push_heap_(f, len - 1, 0, v); }
static void std_pop_heap(orc_nb *f, long len) {
  if (len < 2) return;
  orc_nb v = f[len - 1];
  f[len - 1] = f[0];
  std_adjust_heap(f, 0, len - 1, v);
}

/* efanna2e::nhood (efanna_graph/include/efanna2e/neighbor.h:32-88) with flat storage */
typedef struct {
  orc_nb *pool; unsigned psize;           /* capacity L */
  unsigned M;
  uint32_t *nn_new, *nn_old, *rnn_new, *rnn_old;
  unsigned n_new, n_old, n_rnew, n_rold;
} orc_hood;

/* nhood::insert (neighbor.h:57-72), statement for statement (the mutex is dropped: single thread) */
static void hood_insert(orc_hood *h, unsigned L, uint32_t id, float dist, uint64_t *n_ins) {
  if (dist > h->pool[0].distance) return;
  for (unsigned i = 0; i < h->psize; i++)
    if (id == h->pool[i].id) return;
  orc_nb nn = {id, dist, 1};
  if (h->psize < L) {
    h->pool[h->psize++] = nn;
    std_push_heap(h->pool, h->psize);
  } else {
    std_pop_heap(h->pool, h->psize);
    h->pool[h->psize - 1] = nn;
    std_push_heap(h->pool, h->psize);
  }
  if (n_ins) (*n_ins)++;
}

/* IndexGraph::Build (index_graph.cpp:265-294) = InitializeGraph (:175-201) + NNDescent (:122-136: iter x
 * {join :22-33, update :34-120}) + top-K extraction, with the IndexRandom initializer
 * (index_random.cpp:24-27), executed as ONE thread (OMP_NUM_THREADS=1 is the only reproducible
 * schedule of the reference) and consuming libc rand() exactly where the reference does
 * (:181 mt19937 seed, :124 mt19937 seed, :81/:92 reservoir slot). IndexRandom's own generator is a
 * default-constructed std::mt19937 (seed 5489, index_random.h:19).
 * std::sort on equal distances is unspecified: ordered by id here (pinning data has no ties).
 * knn [n,K]; info (optional) [4] = {distance evaluations, pool insertions, 0, 0}. */
EXPORT int orc_nndescent_ref(const float *data, size_t n, size_t dim, size_t stride, unsigned K, unsigned L,
                             unsigned iter, unsigned S, unsigned R, int mode, uint32_t *knn, uint64_t *info) {
  if (n <= 2 * S || n <= 100 || S + 1 > L || K > L || R == 0) return -1;
  unsigned cap_new = (2 * S > S + R ? 2 * S : S + R), cap_old = L + R;
  orc_hood *g = (orc_hood *)calloc(n, sizeof(orc_hood));
  uint64_t n_dist = 0, n_ins = 0;
  orc_mt rng; mt_seed(&rng, (uint32_t)rand());                 /* :181 */
  for (size_t i = 0; i < n; i++) {                               /* :182-184, nhood ctor neighbor.h:41-47 */
    orc_hood *h = &g[i];
    h->pool = (orc_nb *)malloc(sizeof(orc_nb) * L);
    h->nn_new = (uint32_t *)malloc(sizeof(uint32_t) * cap_new);
    h->nn_old = (uint32_t *)malloc(sizeof(uint32_t) * cap_old);
    h->rnn_new = (uint32_t *)malloc(sizeof(uint32_t) * R);
    h->rnn_old = (uint32_t *)malloc(sizeof(uint32_t) * R);
    h->M = S;
    h->n_new = 2 * S;
    gen_random(mt_next, &rng, h->nn_new, 2 * S, (unsigned)n);
  }
  orc_mt init_rng; mt_seed(&init_rng, 5489U);
  uint32_t *tmp = (uint32_t *)malloc(sizeof(uint32_t) * (S + 1));
  for (size_t i = 0; i < n; i++) {                               /* :185-200 */
    gen_random(mt_next, &init_rng, tmp, S + 1, (unsigned)n);     /* IndexRandom::Search */
    orc_hood *h = &g[i];
    for (unsigned j = 0; j < S; j++) {
      uint32_t id = tmp[j];
      if (id == i) continue;
      float d = orc_dist(mode, data + i * stride, data + (size_t)id * stride, dim);
      n_dist++;
      orc_nb nn = {id, d, 1};
      h->pool[h->psize++] = nn;
    }
    std_make_heap(h->pool, h->psize);
  }
  free(tmp);
  { /* NNDescent :124-128: seeds a second generator and draws the control points (not used by the result) */
    orc_mt crng; mt_seed(&crng, (uint32_t)rand());
    uint32_t ctrl[100];
    gen_random(mt_next, &crng, ctrl, 100, (unsigned)n);
  }
  for (unsigned it = 0; it < iter; it++) {
    /* ---- join (:22-33, nhood::join neighbor.h:74-87) */
    for (size_t nd = 0; nd < n; nd++) {
      orc_hood *h = &g[nd];
      for (unsigned a = 0; a < h->n_new; a++) {
        uint32_t i = h->nn_new[a];
        for (unsigned b = 0; b < h->n_new + h->n_old; b++) {
          uint32_t j = b < h->n_new ? h->nn_new[b] : h->nn_old[b - h->n_new];
          if (b < h->n_new && !(i < j)) continue;
          if (i == j) continue;
          float d = orc_dist(mode, data + (size_t)i * stride, data + (size_t)j * stride, dim);
          n_dist++;
          hood_insert(&g[i], L, j, d, &n_ins);
          hood_insert(&g[j], L, i, d, &n_ins);
        }
      }
    }
    /* ---- update (:34-120) */
    for (size_t i = 0; i < n; i++) g[i].n_new = g[i].n_old = 0;  /* :38-47 */
    for (size_t nd = 0; nd < n; nd++) {                          /* :49-64 */
      orc_hood *nn = &g[nd];
      qsort(nn->pool, nn->psize, sizeof(orc_nb), nb_cmp);
      if (nn->psize > L) nn->psize = L;
      unsigned maxl = nn->M + S < nn->psize ? nn->M + S : nn->psize;
      unsigned c = 0, l = 0;
      while (l < maxl && c < S) {
        if (nn->pool[l].flag) ++c;
        ++l;
      }
      nn->M = l;
    }
    for (size_t nd = 0; nd < n; nd++) {                          /* :66-98 */
      orc_hood *h = &g[nd];
      for (unsigned l = 0; l < h->M; ++l) {
        orc_nb *nn = &h->pool[l];
        orc_hood *o = &g[nn->id];
        if (nn->flag) {
          h->nn_new[h->n_new++] = nn->id;
          if (nn->distance > o->pool[o->psize - 1].distance) {
            if (o->n_rnew < R) o->rnn_new[o->n_rnew++] = (uint32_t)nd;
            else o->rnn_new[(unsigned)rand() % R] = (uint32_t)nd;
          }
          nn->flag = 0;
        } else {
          h->nn_old[h->n_old++] = nn->id;
          if (nn->distance > o->pool[o->psize - 1].distance) {
            if (o->n_rold < R) o->rnn_old[o->n_rold++] = (uint32_t)nd;
            else o->rnn_old[(unsigned)rand() % R] = (uint32_t)nd;
          }
        }
      }
      std_make_heap(h->pool, h->psize);
    }
    for (size_t i = 0; i < n; i++) {                             /* :100-119 (the shuffles never trigger: size <= R) */
      orc_hood *h = &g[i];
      memcpy(h->nn_new + h->n_new, h->rnn_new, sizeof(uint32_t) * h->n_rnew); h->n_new += h->n_rnew;
      memcpy(h->nn_old + h->n_old, h->rnn_old, sizeof(uint32_t) * h->n_rold); h->n_old += h->n_rold;
      if (h->n_old > R * 2) h->n_old = R * 2;
      h->n_rnew = h->n_rold = 0;
    }
  }
  int rc = 0;
  for (size_t i = 0; i < n; i++) {                               /* :275-291 */
    orc_hood *h = &g[i];
    qsort(h->pool, h->psize, sizeof(orc_nb), nb_cmp);
    for (unsigned j = 0; j < K; j++) {
      if (j < h->psize) knn[i * K + j] = h->pool[j].id;
      else { knn[i * K + j] = ORC_INVALID; rc = 1; }             /* the reference reads past the pool here */
    }
  }
  for (size_t i = 0; i < n; i++) { free(g[i].pool); free(g[i].nn_new); free(g[i].nn_old); free(g[i].rnn_new); free(g[i].rnn_old); }
  free(g);
  if (info) { info[0] = n_dist; info[1] = n_ins; info[2] = info[3] = 0; }
  return rc;
}

/* ---- NN-descent, batch form: what the CUDA path (csrc/nndescent.cu) computes, restated serially.
 * Same phases, sizes and sampling rules as IndexGraph::Build above, made independent of execution order:
 *  - a pool is the sorted list of the <= L smallest keys (distance bits << 32 | id << 1 | flag) it has been
 *    offered, deduplicated by id, where an offer is admitted only if distance <= the pool's largest distance
 *    at the start of the join (nhood::insert's `dist > pool.front().distance` test, neighbor.h:59, evaluated
 *    against the join's starting state: the heap top never grows during a join). A set function of the offers.
 *  - the reverse lists keep, per node, the R offers with the smallest (hash(seed_it, node, source), source)
 *    instead of "first R, then overwrite slot rand() % R" (index_graph.cpp:78-82,89-94);
 *  - `pool.back()` (:76,:88): in the reference's one-thread schedule node nd reads the pool of node o while it
 *    is still sorted when o > nd (back() = largest distance) and after std::make_heap (:97) when o < nd
 *    (back() = the heap's last leaf). With distinct distances the sorted rank of that leaf depends only on the
 *    pool size p: leaf_rank[p], computed by heapifying 0..p-1. The batch form uses exactly that rule, which
 *    depends on ids and not on execution order (measured: comparing against the largest distance everywhere
 *    loses 0.04 recall on uniform 32-d data at matched parameters);
 *  - GenRandom draws from the stateless stream orc_seed_hash(seed, node, draw) instead of a shared mt19937.
 * knn [n,K] (ORC_INVALID past the pool), knn_d [n,K] or NULL, info [4] = {distance evaluations, offers that
 * passed the static test, 0, 0}. */
typedef struct {
  uint64_t *pool; unsigned psize, M;
  uint32_t *fwd_new, *fwd_old; unsigned n_fnew, n_fold;
  uint64_t *rnn_new, *rnn_old; unsigned n_rnew, n_rold;
} orc_bhood;
static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
/* keep the `cap` smallest keys, sorted ascending; dedup on (low word >> 1) when dedup != 0. returns 1 if inserted */
static int topl_insert(uint64_t *list, unsigned *size, unsigned cap, uint64_t key, int dedup) {
  unsigned sz = *size, pos = 0;
  uint32_t id = (uint32_t)key >> 1;
  for (unsigned i = 0; i < sz; i++) {
    if (dedup && ((uint32_t)list[i] >> 1) == id) return 0;
    if (list[i] < key) pos++;
  }
  if (pos >= cap) return 0;
  unsigned last = sz < cap ? sz : cap - 1;
  for (unsigned i = last; i > pos; i--) list[i] = list[i - 1];
  list[pos] = key;
  if (sz < cap) *size = sz + 1;
  return 1;
}
#define NND_SEED_POOL 0x68E31DA4U
/* init_graph [n, init_k] (or NULL): IndexGraph::RefineGraph (index_graph.cpp:235-262) = InitializeGraph_Refine
 * (:203-232) + NNDescent: the pools start from the rows of an initializer's graph (ids sorted, duplicates and the
 * node itself dropped, every entry flagged new) instead of S random ids; nn_new is the nhood ctor's random draw
 * either way. Rows may be padded with ORC_INVALID; init_k <= L. */
EXPORT int orc_nndescent_refine(const float *data, size_t n, size_t dim, size_t stride, unsigned K, unsigned L,
                                unsigned iter, unsigned S, unsigned R, uint32_t seed, int mode, const uint32_t *init_graph,
                                unsigned init_k, uint32_t *knn, float *knn_d, uint64_t *info);
EXPORT int orc_nndescent_batch(const float *data, size_t n, size_t dim, size_t stride, unsigned K, unsigned L,
                               unsigned iter, unsigned S, unsigned R, uint32_t seed, int mode, uint32_t *knn,
                               float *knn_d, uint64_t *info) {
  return orc_nndescent_refine(data, n, dim, stride, K, L, iter, S, R, seed, mode, NULL, 0, knn, knn_d, info);
}
EXPORT int orc_nndescent_refine(const float *data, size_t n, size_t dim, size_t stride, unsigned K, unsigned L,
                                unsigned iter, unsigned S, unsigned R, uint32_t seed, int mode, const uint32_t *init_graph,
                                unsigned init_k, uint32_t *knn, float *knn_d, uint64_t *info) {
  if (n <= 2 * S + 1 || S + 1 > L || K > L || R == 0 || S == 0 || n >= (1u << 31)) return -1;
  if (init_graph && (init_k == 0 || init_k > L)) return -1;
  orc_bhood *g = (orc_bhood *)calloc(n, sizeof(orc_bhood));
  uint32_t *oldmax = (uint32_t *)malloc(sizeof(uint32_t) * n);
  unsigned cap_new = (2 * S > S + R ? 2 * S : S + R), cap_old = 2 * R;
  uint32_t *lst = (uint32_t *)malloc(sizeof(uint32_t) * (cap_new + L + R));
  uint32_t *tmp = (uint32_t *)malloc(sizeof(uint32_t) * (S + 1));
  uint64_t n_dist = 0, n_offer = 0;
  /* leaf_rank[p]: sorted rank of the element std::make_heap leaves at position p-1 of an ascending array */
  unsigned *leaf_rank = (unsigned *)calloc(L + 1, sizeof(unsigned));
  uint32_t *leafd = (uint32_t *)malloc(sizeof(uint32_t) * n);
  {
    orc_nb *hp = (orc_nb *)malloc(sizeof(orc_nb) * (L + 1));
    for (unsigned p = 1; p <= L; p++) {
      for (unsigned i = 0; i < p; i++) { hp[i].id = i; hp[i].distance = (float)i; hp[i].flag = 0; }
      std_make_heap(hp, p);
      leaf_rank[p] = hp[p - 1].id;
    }
    free(hp);
  }
  for (size_t i = 0; i < n; i++) {
    orc_bhood *h = &g[i];
    h->pool = (uint64_t *)malloc(sizeof(uint64_t) * L);
    h->fwd_new = (uint32_t *)malloc(sizeof(uint32_t) * 2 * S);
    h->fwd_old = (uint32_t *)malloc(sizeof(uint32_t) * L);
    h->rnn_new = (uint64_t *)malloc(sizeof(uint64_t) * R);
    h->rnn_old = (uint64_t *)malloc(sizeof(uint64_t) * R);
    h->M = S;
    orc_hs s1 = {seed, (uint32_t)i, 0};
    gen_random(hs_next, &s1, h->fwd_new, 2 * S, (unsigned)n);    /* nhood ctor, neighbor.h:41-47 */
    h->n_fnew = 2 * S;
    if (init_graph) {                                             /* InitializeGraph_Refine, :215-226 */
      for (unsigned j = 0; j < init_k; j++) {
        uint32_t id = init_graph[i * init_k + j];
        if (id == ORC_INVALID || id == i || id >= n) continue;
        float d = orc_dist(mode, data + i * stride, data + (size_t)id * stride, dim);
        n_dist++;
        topl_insert(h->pool, &h->psize, L, ((uint64_t)f2u(d) << 32) | (id << 1) | 1u, 1);
      }
    }
    if (!init_graph || h->psize == 0) {
      orc_hs s2 = {seed ^ NND_SEED_POOL, (uint32_t)i, 0};
      gen_random(hs_next, &s2, tmp, S + 1, (unsigned)n);         /* IndexRandom::Search, :189 */
      for (unsigned j = 0; j < S; j++) {
        if (tmp[j] == i) continue;
        float d = orc_dist(mode, data + i * stride, data + (size_t)tmp[j] * stride, dim);
        n_dist++;
        topl_insert(h->pool, &h->psize, L, ((uint64_t)f2u(d) << 32) | (tmp[j] << 1) | 1u, 1);
      }
    }
  }
  for (unsigned it = 0; it < iter; it++) {
    for (size_t i = 0; i < n; i++) oldmax[i] = (uint32_t)(g[i].pool[g[i].psize - 1] >> 32);
    /* ---- join */
    for (size_t nd = 0; nd < n; nd++) {
      orc_bhood *h = &g[nd];
      unsigned nn = 0, no = 0;
      for (unsigned a = 0; a < h->n_fnew; a++) lst[nn++] = h->fwd_new[a];
      for (unsigned a = 0; a < h->n_rnew; a++) lst[nn++] = (uint32_t)h->rnn_new[a];
      for (unsigned a = 0; a < h->n_fold && no < cap_old; a++) lst[nn + no++] = h->fwd_old[a];
      for (unsigned a = 0; a < h->n_rold && no < cap_old; a++) lst[nn + no++] = (uint32_t)h->rnn_old[a];
      for (unsigned a = 0; a < nn; a++)
        for (unsigned b = 0; b < nn + no; b++) {
          if (b < nn && b <= a) continue;                        /* each unordered new x new pair once */
          uint32_t i = lst[a], j = lst[b];
          if (i == j) continue;
          float d = orc_dist(mode, data + (size_t)i * stride, data + (size_t)j * stride, dim);
          n_dist++;
          uint32_t db = f2u(d);
          if (db <= oldmax[i]) { n_offer++; topl_insert(g[i].pool, &g[i].psize, L, ((uint64_t)db << 32) | (j << 1) | 1u, 1); }
          if (db <= oldmax[j]) { n_offer++; topl_insert(g[j].pool, &g[j].psize, L, ((uint64_t)db << 32) | (i << 1) | 1u, 1); }
        }
    }
    /* ---- update */
    for (size_t nd = 0; nd < n; nd++) {
      orc_bhood *h = &g[nd];
      unsigned maxl = h->M + S < h->psize ? h->M + S : h->psize, c = 0, l = 0;
      while (l < maxl && c < S) { if (h->pool[l] & 1u) ++c; ++l; }
      h->M = l;
      oldmax[nd] = (uint32_t)(h->pool[h->psize - 1] >> 32);      /* pool.back() of a sorted pool */
      leafd[nd] = (uint32_t)(h->pool[leaf_rank[h->psize]] >> 32); /* pool.back() of a heapified pool */
      h->n_fnew = h->n_fold = h->n_rnew = h->n_rold = 0;
    }
    uint32_t seed_new = seed + (2 * it + 1) * 0x9E3779B1U, seed_old = seed + (2 * it + 2) * 0x9E3779B1U;
    for (size_t nd = 0; nd < n; nd++) {
      orc_bhood *h = &g[nd];
      for (unsigned l = 0; l < h->M; l++) {
        uint64_t e = h->pool[l];
        uint32_t id = (uint32_t)e >> 1, db = (uint32_t)(e >> 32);
        orc_bhood *o = &g[id];
        if (e & 1u) {
          h->fwd_new[h->n_fnew++] = id;
          if (db > (id < nd ? leafd[id] : oldmax[id]))
            topl_insert(o->rnn_new, &o->n_rnew, R, ((uint64_t)orc_seed_hash(seed_new, id, (uint32_t)nd) << 32) | nd, 0);
          h->pool[l] = e & ~(uint64_t)1;
        } else {
          h->fwd_old[h->n_fold++] = id;
          if (db > (id < nd ? leafd[id] : oldmax[id]))
            topl_insert(o->rnn_old, &o->n_rold, R, ((uint64_t)orc_seed_hash(seed_old, id, (uint32_t)nd) << 32) | nd, 0);
        }
      }
    }
  }
  int rc = 0;
  for (size_t i = 0; i < n; i++)
    for (unsigned j = 0; j < K; j++) {
      if (j < g[i].psize) {
        knn[i * K + j] = (uint32_t)g[i].pool[j] >> 1;
        if (knn_d) knn_d[i * K + j] = u2f((uint32_t)(g[i].pool[j] >> 32));
      } else {
        knn[i * K + j] = ORC_INVALID; rc = 1;
        if (knn_d) knn_d[i * K + j] = INFINITY;
      }
    }
  for (size_t i = 0; i < n; i++) { free(g[i].pool); free(g[i].fwd_new); free(g[i].fwd_old); free(g[i].rnn_new); free(g[i].rnn_old); }
  free(g); free(oldmax); free(lst); free(tmp); free(leaf_rank); free(leafd);
  if (info) { info[0] = n_dist; info[1] = n_offer; info[2] = info[3] = 0; }
  return rc;
}

/* ================================================================== HNSW construction (hnswlib addPoint) */
/* std::default_random_engine (libstdc++: minstd_rand0, x <- 16807 x mod 2^31-1) driving
 * std::uniform_real_distribution<double>(0,1) = std::generate_canonical<double,53>: two draws,
 * (x1-1) + (x2-1)*R over R^2 with R = 2147483646 (bits/random.tcc), then
 * HierarchicalNSW::getRandomLevel (hnswalg.h:206-210): (int)(-log(u) * mult), mult = 1/log(M) (:114). */
EXPORT void orc_hnsw_levels(size_t n, unsigned M, uint32_t seed, int32_t *levels) {
  uint64_t x = seed % 2147483647ULL;
  if (x == 0) x = 1;
  const long double r = 2147483646.0L;
  const double mult = 1.0 / log(1.0 * M);
  for (size_t i = 0; i < n; i++) {
    double sum = 0.0, tmp = 1.0;
    for (int k = 0; k < 2; k++) {
      x = (x * 16807ULL) % 2147483647ULL;
      sum += (double)(x - 1) * tmp;
      tmp = (double)((long double)tmp * r);
    }
    double u = sum / tmp;
    if (u >= 1.0) u = nextafter(1.0, 0.0);
    levels[i] = (int)(-log(u) * mult);
  }
}

typedef struct {
  const float *data; size_t n, dim, stride; int mode;
  unsigned M, maxM, maxM0, efc;
  uint32_t *links0, *cnt0;            /* [n,maxM0], [n] */
  const int32_t *levels; uint32_t *upper_ptr, *upper_links, *cntu;   /* [n], [n], [rows,maxM], [rows] */
  uint32_t ep; int maxlevel;
  uint32_t *visited, tag;             /* epoch array like VisitedList (visited_list_pool.h:10-31) */
  uint64_t n_dist;
} hb_t;
static uint32_t *hb_row(hb_t *h, uint32_t node, int level, uint32_t **cnt, unsigned *cap) {
  if (level == 0) { *cnt = &h->cnt0[node]; *cap = h->maxM0; return h->links0 + (size_t)node * h->maxM0; }
  size_t r = (size_t)h->upper_ptr[node] + (level - 1);
  *cnt = &h->cntu[r]; *cap = h->maxM;
  return h->upper_links + r * h->maxM;
}
static float hb_dist(hb_t *h, const float *a, uint32_t id) {
  h->n_dist++;
  return orc_dist(h->mode, a, h->data + h->stride * (size_t)id, h->dim);
}
/* searchBaseLayer (hnswalg.h:225-305; no deleted elements). variant 0: the reference's two heaps;
 * variant 1: the CUDA kernel's single ascending (d,id) pool with expanded flags (see hnsw_base_pool).
 * out: <= efc pairs ascending (d,id); returns the count. */
static size_t hb_search(hb_t *h, uint32_t ep, const float *q, int level, int variant, orc_pair *top, orc_pair *cand,
                        uint8_t *expanded) {
  const size_t ef = h->efc;
  uint32_t tag = ++h->tag;
  size_t nt = 0, nc = 0;
  uint32_t *cnt; unsigned cap;
  float dist = hb_dist(h, q, ep);
  orc_pair e = {dist, ep};
  h->visited[ep] = tag;
  if (variant == 0) {
    float lowerBound = dist;
    hpush(top, &nt, e, +1);
    hpush(cand, &nc, e, -1);
    while (nc) {
      orc_pair cur = cand[0];
      if (cur.d > lowerBound && nt == ef) break;                  /* :245-247 */
      hpop(cand, &nc, -1);
      const uint32_t *row = hb_row(h, cur.id, level, &cnt, &cap);
      for (unsigned j = 0; j < *cnt; j++) {
        uint32_t cid = row[j];
        if (h->visited[cid] == tag) continue;
        h->visited[cid] = tag;
        float d = hb_dist(h, q, cid);
        if (nt < ef || lowerBound > d) {                          /* :284 */
          orc_pair c = {d, cid};
          hpush(cand, &nc, c, -1);
          hpush(top, &nt, c, +1);
          if (nt > ef) hpop(top, &nt, +1);
          lowerBound = top[0].d;
        }
      }
    }
    qsort(top, nt, sizeof(orc_pair), pair_cmp);
    return nt;
  }
  top[0] = e; expanded[0] = 0; nt = 1;
  for (;;) {
    size_t k = 0;
    while (k < nt && expanded[k]) k++;
    if (k == nt) break;
    expanded[k] = 1;
    const uint32_t *row = hb_row(h, top[k].id, level, &cnt, &cap);
    for (unsigned j = 0; j < *cnt; j++) {
      uint32_t cid = row[j];
      if (h->visited[cid] == tag) continue;
      h->visited[cid] = tag;
      orc_pair c = {hb_dist(h, q, cid), cid};
      if (nt < ef || c.d < top[nt - 1].d) {
        size_t pos = nt;
        while (pos > 0 && pair_less(c, top[pos - 1])) pos--;
        size_t last = nt < ef ? nt : ef - 1;
        for (size_t t = last; t > pos; t--) { top[t] = top[t - 1]; expanded[t] = expanded[t - 1]; }
        top[pos] = c; expanded[pos] = 0;
        if (nt < ef) nt++;
      }
    }
  }
  return nt;
}
/* getNeighborsByHeuristic2 (hnswalg.h:443-483): fewer than M candidates -> all kept; otherwise scan in
 * ascending distance (ties by id here; the reference's pair heap pops the larger id first) and keep c iff
 * no kept r has d(r,c) < d(q,c); stop at M. In/out: c[0..n) -> the kept ones, ascending. returns count. */
static size_t hb_heuristic(hb_t *h, orc_pair *c, size_t n, size_t M) {
  qsort(c, n, sizeof(orc_pair), pair_cmp);
  if (n < M) return n;
  size_t kept = 0;
  for (size_t i = 0; i < n && kept < M; i++) {
    int good = 1;
    for (size_t r = 0; r < kept; r++) {
      float cur = hb_dist(h, h->data + h->stride * (size_t)c[r].id, c[i].id);
      if (cur < c[i].d) { good = 0; break; }
    }
    if (good) c[kept++] = c[i];
  }
  return kept;
}
static void hb_write_row(uint32_t *row, uint32_t *cnt, const orc_pair *sel, size_t k) {
  for (size_t i = 0; i < k; i++) row[i] = sel[k - 1 - i].id;      /* popped from a max-heap: farthest first (:514-517,:611-615) */
  *cnt = (uint32_t)k;
}
/* greedy descent of addPoint (hnswalg.h:1214-1239) from level `from` down to `to`+1 */
static uint32_t hb_descend(hb_t *h, const float *q, uint32_t cur, int from, int to) {
  float curdist = hb_dist(h, q, cur);
  uint32_t *cnt; unsigned cap;
  for (int level = from; level > to; level--) {
    int changed = 1;
    while (changed) {
      changed = 0;
      const uint32_t *row = hb_row(h, cur, level, &cnt, &cap);
      unsigned size = *cnt;
      for (unsigned i = 0; i < size; i++) {
        uint32_t cand = row[i];
        float d = hb_dist(h, q, cand);
        if (d < curdist) { curdist = d; cur = cand; changed = 1; }
      }
    }
  }
  return cur;
}

/* Builds the HNSW graph of rows 0..n-1 inserted in index order (hnswlib addPoint loop, e.g.
 * hnswlib/tests/cpp/sift_1m.cpp:321-336 run with one thread). Flat outputs as parse_hnsw / b200ann_index_set_hnsw:
 * links0 [n,maxM0], levels [n] (input: from orc_hnsw_levels), upper_ptr [n], upper_links [rows,maxM] with
 * rows = sum(levels), all padded with ORC_INVALID. ep_out/maxlevel_out.
 * variant 0: the reference statement for statement, one point at a time (pinned to the compiled reference).
 * variant 1: the CUDA path's batch form (csrc/hnsw_build.cu): points are inserted in batches of
 *   min(bmax, max(1, inserted/ratio)), ratio at most 8 while fewer than min(2048, n/16) rows are inserted (a point whose level
 *   exceeds the current top level is a batch of its own);
 *   every point of a batch searches the graph as it was BEFORE the batch (single-pool beam) and selects its M
 *   links with the heuristic; then, per (neighbour, level), all new reverse links are merged at once: appended in
 *   ascending id if they fit in Mmax, else the heuristic runs over existing + new (the reference does the same
 *   one reverse link at a time, :554-627). Order independent inside a batch.
 * info [4] = {distance evaluations, batches, 0, 0}. */
typedef struct { uint32_t row_node; int level; uint32_t src; float d; } hb_rev;
static int rev_cmp(const void *x, const void *y) {
  const hb_rev *a = (const hb_rev *)x, *b = (const hb_rev *)y;
  if (a->row_node != b->row_node) return a->row_node < b->row_node ? -1 : 1;
  if (a->level != b->level) return a->level < b->level ? -1 : 1;
  return a->src < b->src ? -1 : (a->src > b->src ? 1 : 0);
}
EXPORT size_t orc_hnsw_batch_end(size_t s, size_t n, const int32_t *levels, int maxlevel, unsigned ratio, unsigned bmax) {
  if (s == 0 || levels[s] > maxlevel) return s + 1;
  /* the first min(2048, n/16) rows: batches of up to 1/8 (hnsw_build.cu hnsw_run_batches) */
  if (s < (n / 16 < 2048 ? n / 16 : 2048) && ratio > 8) ratio = 8;
  size_t b = s / ratio;
  if (b < 1) b = 1;
  if (b > bmax) b = bmax;
  size_t e = s + b < n ? s + b : n;
  for (size_t i = s + 1; i < e; i++)
    if (levels[i] > maxlevel) return i;
  return e;
}
EXPORT int orc_hnsw_build(const float *data, size_t n, size_t dim, size_t stride, unsigned M, unsigned efc, int mode,
                          int variant, unsigned ratio, unsigned bmax, const int32_t *levels, uint32_t *links0,
                          uint32_t *upper_ptr, uint32_t *upper_links, uint32_t *ep_out, int *maxlevel_out, uint64_t *info) {
  hb_t h; memset(&h, 0, sizeof h);
  h.data = data; h.n = n; h.dim = dim; h.stride = stride; h.mode = mode;
  h.M = M; h.maxM = M; h.maxM0 = 2 * M; h.efc = efc > M ? efc : M;   /* hnswalg.h:104-107 */
  h.links0 = links0; h.levels = levels; h.upper_ptr = upper_ptr; h.upper_links = upper_links;
  size_t rows = 0;
  for (size_t i = 0; i < n; i++) { upper_ptr[i] = (uint32_t)rows; rows += (size_t)levels[i]; }
  h.cnt0 = (uint32_t *)calloc(n, 4); h.cntu = (uint32_t *)calloc(rows ? rows : 1, 4);
  h.visited = (uint32_t *)calloc(n, 4);
  size_t capc = h.efc + h.maxM0 + 8;
  orc_pair *top = (orc_pair *)malloc(sizeof(orc_pair) * (capc + 1)), *cand = (orc_pair *)malloc(sizeof(orc_pair) * (n + 1));
  uint8_t *expanded = (uint8_t *)malloc(capc + 1);
  orc_pair *tmp = (orc_pair *)malloc(sizeof(orc_pair) * (h.maxM0 + bmax + 8));
  hb_rev *rev = NULL; size_t nrev = 0, caprev = 0;
  uint64_t batches = 0;
  h.ep = ORC_INVALID; h.maxlevel = -1;
  size_t s = 0;
  while (s < n) {
    size_t e = variant == 0 ? s + 1 : orc_hnsw_batch_end(s, n, levels, h.maxlevel, ratio, bmax);
    batches++;
    nrev = 0;
    for (size_t p = s; p < e; p++) {
      const float *q = data + stride * p;
      int curlevel = levels[p];
      if (h.ep == ORC_INVALID) continue;                           /* first element: nothing to link (:1255-1258) */
      uint32_t cur = h.ep;
      if (curlevel < h.maxlevel) cur = hb_descend(&h, q, cur, h.maxlevel, curlevel);
      for (int level = curlevel < h.maxlevel ? curlevel : h.maxlevel; level >= 0; level--) {
        size_t nt = hb_search(&h, cur, q, level, variant, top, cand, expanded);
        size_t k = hb_heuristic(&h, top, nt, h.M);                 /* mutuallyConnectNewElement :513 */
        uint32_t *cnt; unsigned cap;
        uint32_t *row = hb_row(&h, (uint32_t)p, level, &cnt, &cap);
        hb_write_row(row, cnt, top, k);
        cur = top[0].id;                                           /* next_closest_entry_point = closest selected (:520) */
        for (size_t i = 0; i < k; i++) {
          if (variant == 0) {                                      /* :554-627, one reverse link at a time */
            uint32_t sn = top[k - 1 - i].id;
            uint32_t *ocnt; unsigned ocap;
            uint32_t *orow = hb_row(&h, sn, level, &ocnt, &ocap);
            if (*ocnt < ocap) { orow[(*ocnt)++] = (uint32_t)p; continue; }
            size_t m = 0;
            tmp[m].d = hb_dist(&h, q, sn); tmp[m].id = (uint32_t)p; m++;
            for (unsigned j = 0; j < *ocnt; j++) {
              tmp[m].d = hb_dist(&h, data + stride * (size_t)orow[j], sn); tmp[m].id = orow[j]; m++;
            }
            size_t kk = hb_heuristic(&h, tmp, m, ocap);
            hb_write_row(orow, ocnt, tmp, kk);
          } else {
            if (nrev == caprev) { caprev = caprev ? caprev * 2 : 4096; rev = (hb_rev *)realloc(rev, caprev * sizeof(hb_rev)); }
            rev[nrev].row_node = top[i].id; rev[nrev].level = level; rev[nrev].src = (uint32_t)p; rev[nrev].d = top[i].d;
            nrev++;
          }
        }
      }
    }
    if (variant != 0 && nrev) {
      qsort(rev, nrev, sizeof(hb_rev), rev_cmp);
      for (size_t a = 0; a < nrev;) {
        size_t b = a;
        while (b < nrev && rev[b].row_node == rev[a].row_node && rev[b].level == rev[a].level) b++;
        uint32_t sn = rev[a].row_node;
        uint32_t *ocnt; unsigned ocap;
        uint32_t *orow = hb_row(&h, sn, rev[a].level, &ocnt, &ocap);
        if (*ocnt + (b - a) <= ocap) {
          for (size_t t = a; t < b; t++) orow[(*ocnt)++] = rev[t].src;
        } else {
          size_t m = 0;
          for (unsigned j = 0; j < *ocnt; j++) {
            tmp[m].d = hb_dist(&h, data + stride * (size_t)sn, orow[j]); tmp[m].id = orow[j]; m++;
          }
          for (size_t t = a; t < b; t++) { tmp[m].d = rev[t].d; tmp[m].id = rev[t].src; m++; }
          size_t kk = hb_heuristic(&h, tmp, m, ocap);
          hb_write_row(orow, ocnt, tmp, kk);
        }
        a = b;
      }
    }
    for (size_t p = s; p < e; p++)
      if (h.ep == ORC_INVALID || levels[p] > h.maxlevel) { h.ep = (uint32_t)p; h.maxlevel = levels[p]; }  /* :1256-1265 */
    s = e;
  }
  for (size_t i = 0; i < n; i++)
    for (unsigned j = h.cnt0[i]; j < h.maxM0; j++) links0[i * h.maxM0 + j] = ORC_INVALID;
  for (size_t r = 0; r < rows; r++)
    for (unsigned j = h.cntu[r]; j < h.maxM; j++) upper_links[r * h.maxM + j] = ORC_INVALID;
  *ep_out = h.ep; *maxlevel_out = h.maxlevel;
  if (info) { info[0] = h.n_dist; info[1] = batches; info[2] = info[3] = 0; }
  free(h.cnt0); free(h.cntu); free(h.visited); free(top); free(cand); free(expanded); free(tmp); free(rev);
  return 0;
}

/* ------------------------------------------------------------------ KD-tree initializer */
/* efanna2e::IndexKDtree::Build (CNNS/efanna_graph/src/index_kdtree.cpp:279-435): nTrees randomized KD-trees (meanSplit
 * :17-67 over the first SAMPLE_NUM+1 = 101 points of a node, selectDivision :110-133 = a random one of the RAND_DIM = 5
 * highest-variance dimensions, planeSplit :69-108, leaves of <= TNS = 10 points, DFSbuild :135-160), then below the
 * merge level ml every internal node is "conquered" (mergeSubGraphs :221-275): each point of its smaller child walks
 * the other child's subtree to a leaf (SearchToLeaf :207-218) and the distances to that leaf's points are offered to
 * both ends' K-lists. The reference is not reproducible (std::random_shuffle, mt19937 seeded with the OpenMP thread
 * number, lock order), so this is the batch form the CUDA path computes (csrc/kdtree.cu), equal to it bit for bit:
 *   shuffle    tree t lists the points in the order  id_j = (a_t * j + b_t) mod n  (a_t odd-ish, coprime to n)
 *   split      as meanSplit, the random pick = orc_seed_hash(seed ^ KD_SEED, t, start * 31 + count) % num; the
 *              partition is the stable three-way one (< cutval | == cutval | > cutval), the split index the
 *              reference's rule (:54-62)
 *   lists      a K-list is the set of the K smallest (distance, id) it was offered (what a bounded heap holds,
 *              in any order); distances in the sequential-fma order (ORC_L2_SEQ_FMA)
 *   output     rows ascending (distance, id); a row with fewer than K entries is filled with distinct ids
 *              orc_seed_hash(seed ^ KD_SEED, i, 1000 + draw) % n not yet present (:413-424 fills with rng() % N), their
 *              distance +inf.
 * info [4] = {distance evaluations, tree nodes, leaves, deepest level}. */
#define KD_SEED 0x4B445452U
#define KD_TNS 10
#define KD_SAMPLE 101
#define KD_RAND_DIM 5
typedef struct { uint32_t start, end, left; int cutdim; float cutval; int depth; } kd_node;
static uint32_t kd_gcd(uint32_t a, uint32_t b) { while (b) { uint32_t t = a % b; a = b; b = t; } return a; }
/* multiplier of tree t's permutation: the first value >= hash % n that is coprime to n (1 for n == 1) */
static uint32_t kd_mult(uint32_t seed, uint32_t t, uint32_t n) {
  uint32_t a = orc_seed_hash(seed ^ KD_SEED, t, 0x7001u) % n;
  if (a == 0) a = 1;
  while (kd_gcd(a, n) != 1) a = a + 1 == n ? 1 : a + 1;
  return a;
}
/* meanSplit over ids[0..count): returns the split index, sets cutdim / cutval; ids are partitioned (stable 3-way) */
static uint32_t kd_split(const float *data, size_t dim, size_t stride, uint32_t *ids, uint32_t count, uint32_t pick_hash,
                         float *mean, float *var, uint32_t *tmp, int *cutdim_out, float *cutval_out) {
  uint32_t cnt = count < KD_SAMPLE ? count : KD_SAMPLE;
  for (size_t k = 0; k < dim; k++) { mean[k] = 0.f; var[k] = 0.f; }
  for (uint32_t j = 0; j < cnt; j++) {
    const float *v = data + (size_t)ids[j] * stride;
    for (size_t k = 0; k < dim; k++) mean[k] += v[k];
  }
  float div = 1.0f / (float)cnt;
  for (size_t k = 0; k < dim; k++) mean[k] *= div;
  for (uint32_t j = 0; j < cnt; j++) {
    const float *v = data + (size_t)ids[j] * stride;
    for (size_t k = 0; k < dim; k++) { float d = v[k] - mean[k]; var[k] += d * d; }
  }
  size_t topind[KD_RAND_DIM];
  int num = 0;
  for (size_t i = 0; i < dim; i++) {               /* selectDivision :110-133 */
    if (num < KD_RAND_DIM || var[i] > var[topind[num - 1]]) {
      if (num < KD_RAND_DIM) topind[num++] = i; else topind[num - 1] = i;
      int j = num - 1;
      while (j > 0 && var[topind[j]] > var[topind[j - 1]]) { size_t t = topind[j]; topind[j] = topind[j - 1]; topind[j - 1] = t; --j; }
    }
  }
  int cutdim = (int)topind[pick_hash % (uint32_t)num];
  float cutval = mean[cutdim];
  uint32_t lim1 = 0, lim2 = 0, o = 0;
  for (uint32_t j = 0; j < count; j++) { float x = data[(size_t)ids[j] * stride + cutdim]; lim1 += x < cutval; lim2 += x <= cutval; }
  for (int cls = 0; cls < 3; cls++)
    for (uint32_t j = 0; j < count; j++) {
      float x = data[(size_t)ids[j] * stride + cutdim];
      int c = x < cutval ? 0 : (x == cutval ? 1 : 2);
      if (c == cls) tmp[o++] = ids[j];
    }
  memcpy(ids, tmp, sizeof(uint32_t) * count);
  uint32_t index;
  if (lim1 > count / 2) index = lim1; else if (lim2 < count / 2) index = lim2; else index = count / 2;
  if (lim1 == count || lim2 == 0) index = count / 2;
  *cutdim_out = cutdim; *cutval_out = cutval;
  return index;
}
EXPORT int orc_kdtree_graph(const float *data, size_t n, size_t dim, size_t stride, unsigned K, unsigned nTrees, unsigned ml,
                            uint32_t seed, uint32_t *knn, float *knn_d, uint64_t *info) {
  if (n < 2 || K == 0 || K >= n || nTrees == 0 || n >= (1u << 31)) return -1;
  uint64_t *lists = (uint64_t *)malloc(sizeof(uint64_t) * n * K);
  unsigned *lsize = (unsigned *)calloc(n, sizeof(unsigned));
  uint32_t *ids = (uint32_t *)malloc(4 * n), *tmp = (uint32_t *)malloc(4 * n);
  float *mean = (float *)malloc(4 * dim), *var = (float *)malloc(4 * dim);
  size_t capn = 2 * n + 8;
  kd_node *nodes = (kd_node *)malloc(sizeof(kd_node) * capn);
  uint64_t n_dist = 0, n_nodes = 0, n_leaves = 0, deepest = 0;
  for (unsigned t = 0; t < nTrees; t++) {
    const uint32_t a = kd_mult(seed, t, (uint32_t)n), b = orc_seed_hash(seed ^ KD_SEED, t, 0x7002u) % (uint32_t)n;
    for (size_t j = 0; j < n; j++) ids[j] = (uint32_t)(((uint64_t)a * j + b) % n);
    /* level-synchronous build: nodes[lo, hi) is the current level */
    size_t nn = 0, lo = 0, hi = 1;
    nodes[nn].start = 0; nodes[nn].end = (uint32_t)n; nodes[nn].left = ORC_INVALID; nodes[nn].depth = 0; nodes[nn].cutdim = -1; nn++;
    while (lo < hi) {
      for (size_t x = lo; x < hi; x++) {
        kd_node *nd = &nodes[x];
        uint32_t count = nd->end - nd->start;
        if ((uint64_t)nd->depth > deepest) deepest = (uint64_t)nd->depth;
        if (count <= KD_TNS) { n_leaves++; continue; }
        uint32_t pick = orc_seed_hash(seed ^ KD_SEED, t, nd->start * 31u + count);
        uint32_t idx = kd_split(data, dim, stride, ids + nd->start, count, pick, mean, var, tmp, &nd->cutdim, &nd->cutval);
        nd->left = (uint32_t)nn;
        nodes[nn].start = nd->start; nodes[nn].end = nd->start + idx; nodes[nn].left = ORC_INVALID; nodes[nn].depth = nd->depth + 1; nodes[nn].cutdim = -1; nn++;
        nodes[nn].start = nd->start + idx; nodes[nn].end = nd->end; nodes[nn].left = ORC_INVALID; nodes[nn].depth = nd->depth + 1; nodes[nn].cutdim = -1; nn++;
      }
      lo = hi; hi = nn;
    }
    n_nodes += nn;
    /* conquer: every point walks its own path; below the merge level the points of the smaller child search the other child */
    for (size_t pos = 0; pos < n; pos++) {
      const uint32_t p = ids[pos];
      const float *pv = data + (size_t)p * stride;
      uint32_t x = 0;
      while (nodes[x].left != ORC_INVALID) {
        const uint32_t L = nodes[x].left, R = L + 1;
        const int inL = pos < nodes[L].end;
        const uint32_t numL = nodes[L].end - nodes[L].start, numR = nodes[R].end - nodes[R].start;
        const int searchers_left = numL < numR;          /* :229-237 */
        if ((unsigned)nodes[x].depth >= ml && inL == searchers_left) {
          uint32_t y = inL ? R : L;
          while (nodes[y].left != ORC_INVALID) y = pv[nodes[y].cutdim] < nodes[y].cutval ? nodes[y].left : nodes[y].left + 1;
          for (uint32_t j = nodes[y].start; j < nodes[y].end; j++) {
            const uint32_t q = ids[j];
            float d = seq_fma(data + (size_t)q * stride, pv, dim, 0);
            n_dist++;
            uint64_t db = (uint64_t)f2u(d) << 32;
            topl_insert(lists + (size_t)q * K, &lsize[q], K, db | ((uint64_t)p << 1), 1);
            topl_insert(lists + (size_t)p * K, &lsize[p], K, db | ((uint64_t)q << 1), 1);
          }
        }
        x = inL ? L : R;
      }
    }
  }
  for (size_t i = 0; i < n; i++) {
    unsigned sz = lsize[i], ctr = 0;
    uint64_t *l = lists + i * K;
    for (unsigned j = 0; j < sz; j++) { knn[i * K + j] = (uint32_t)l[j] >> 1; if (knn_d) knn_d[i * K + j] = u2f((uint32_t)(l[j] >> 32)); }
    while (sz < K) {                                       /* :413-424 */
      uint32_t id = orc_seed_hash(seed ^ KD_SEED, (uint32_t)i, 1000u + ctr++) % (uint32_t)n;
      int dup = 0;
      for (unsigned j = 0; j < sz; j++) dup |= knn[i * K + j] == id;
      if (dup) continue;
      knn[i * K + sz] = id; if (knn_d) knn_d[i * K + sz] = INFINITY;
      sz++;
    }
  }
  if (info) { info[0] = n_dist; info[1] = n_nodes; info[2] = n_leaves; info[3] = deepest; }
  free(lists); free(lsize); free(ids); free(tmp); free(mean); free(var); free(nodes);
  return 0;
}

/* ------------------------------------------------------------------ HNSW updatePoint */
/* HierarchicalNSW::updatePoint(dataPoint, internalId, updateNeighborProbability = 1.0) + repairConnectionsForUpdate
 * (hnswalg.h:993-1140; what addPoint does for a label that already exists, :1160-1174), for nb elements ids[0..nb)
 * whose new vectors are newvec [nb, stride]. data [n, stride] holds the current vectors and is updated in place;
 * links0 / upper_links [in/out] are the flat graph (padded with ORC_INVALID); levels, upper_ptr, ep, maxlevel do not
 * change. No element is marked deleted.
 * variant 0: the reference, one element at a time:
 *   A. per layer <= level(p): sCand = {p} + one-hop + two-hop neighbours (:1006-1027); every one-hop neighbour gets
 *      a NEW list: the min(ef_construction, |sCand \ {neigh}|) closest of sCand to it (:1029-1052) through the
 *      heuristic with Mmax (:1055), written farthest first (:1057-1068).
 *   B. repairConnectionsForUpdate: greedy descent to level(p)+1, then per level searchBaseLayer from the running entry
 *      point, p itself filtered out (:1112-1118), mutuallyConnectNewElement(isUpdate = true): heuristic with M, p's row
 *      rewritten, p appended to each selected neighbour unless already there (:548-560), heuristic on overflow.
 * variant 1: the CUDA path's batch form (csrc/hnsw_build.cu hnsw_update_launch): all new vectors are stored first;
 *   phase A for all elements on the graph as it was before the call (a row claimed by several updated elements is
 *   computed for the LAST of them in `ids` order), rows staged and then written; phase B for all elements on the
 *   graph phase A left (single-pool beam), own rows staged and then written, reverse links grouped per (neighbour,
 *   level): sources already present skipped, appended in ascending id if they fit, else heuristic over existing + new.
 * info [2] = {distance evaluations, rows rewritten by phase A}. */
static size_t hb_row_get(hb_t *h, uint32_t node, int level, uint32_t **row) {
  uint32_t *cnt; unsigned cap;
  *row = hb_row(h, node, level, &cnt, &cap);
  return *cnt;
}
/* phase A for one (p, layer, neigh): the new list of neigh -> sel[0..return) ascending (d,id) */
static size_t hb_update_neighbor(hb_t *h, uint32_t p, int layer, uint32_t neigh, orc_pair *buf, size_t bufcap) {
  uint32_t *r1, *r2;
  size_t n1 = hb_row_get(h, p, layer, &r1), m = 0;
  const float *nv = h->data + h->stride * (size_t)neigh;
  unsigned cap = layer == 0 ? h->maxM0 : h->maxM;
  (void)bufcap;
  /* sCand = {p} + one-hop + two-hop, as a multiset; duplicates carry identical (d,id) and are dropped after the sort */
  if (p != neigh) { buf[m].d = hb_dist(h, nv, p); buf[m].id = p; m++; }
  for (size_t a = 0; a < n1; a++) {
    uint32_t w = r1[a];
    if (w != neigh) { buf[m].d = hb_dist(h, nv, w); buf[m].id = w; m++; }
    size_t n2 = hb_row_get(h, w, layer, &r2);
    for (size_t b = 0; b < n2; b++)
      if (r2[b] != neigh) { buf[m].d = hb_dist(h, nv, r2[b]); buf[m].id = r2[b]; m++; }
  }
  qsort(buf, m, sizeof(orc_pair), pair_cmp);
  size_t u = 0;
  for (size_t i = 0; i < m; i++)
    if (u == 0 || buf[u - 1].id != buf[i].id || buf[u - 1].d != buf[i].d) buf[u++] = buf[i];
  if (u > h->efc) u = h->efc;                                       /* elementsToKeep closest (:1034-1050) */
  return hb_heuristic(h, buf, u, cap);
}
EXPORT int orc_hnsw_update(float *data, size_t n, size_t dim, size_t stride, unsigned M, unsigned efc, int mode, int variant,
                           const int32_t *levels, uint32_t *links0, uint32_t *upper_ptr, uint32_t *upper_links, size_t rows,
                           uint32_t ep, int maxlevel, const uint32_t *ids, const float *newvec, size_t nb, uint64_t *info) {
  hb_t h; memset(&h, 0, sizeof h);
  h.data = data; h.n = n; h.dim = dim; h.stride = stride; h.mode = mode;
  h.M = M; h.maxM = M; h.maxM0 = 2 * M; h.efc = efc > M ? efc : M;
  h.links0 = links0; h.levels = levels; h.upper_ptr = upper_ptr; h.upper_links = upper_links;
  h.ep = ep; h.maxlevel = maxlevel;
  h.cnt0 = (uint32_t *)calloc(n, 4); h.cntu = (uint32_t *)calloc(rows ? rows : 1, 4);
  h.visited = (uint32_t *)calloc(n, 4);
  for (size_t i = 0; i < n; i++) { unsigned c = 0; while (c < h.maxM0 && links0[i * h.maxM0 + c] != ORC_INVALID) c++; h.cnt0[i] = c; }
  for (size_t r = 0; r < rows; r++) { unsigned c = 0; while (c < h.maxM && upper_links[r * h.maxM + c] != ORC_INVALID) c++; h.cntu[r] = c; }
  const size_t bufcap = 1 + (size_t)h.maxM0 * (h.maxM0 + 1) + 8;
  orc_pair *buf = (orc_pair *)malloc(sizeof(orc_pair) * bufcap);
  size_t capc = h.efc + h.maxM0 + 8;
  orc_pair *top = (orc_pair *)malloc(sizeof(orc_pair) * (capc + 1)), *cand = (orc_pair *)malloc(sizeof(orc_pair) * (n + 1));
  uint8_t *expanded = (uint8_t *)malloc(capc + 1);
  orc_pair *tmp = (orc_pair *)malloc(sizeof(orc_pair) * (h.maxM0 + nb + 8));
  uint64_t rewritten = 0;
  const size_t row_bytes = stride * sizeof(float);
  if (variant == 0) {
    for (size_t i = 0; i < nb; i++) {
      const uint32_t p = ids[i];
      memcpy(data + stride * (size_t)p, newvec + stride * i, row_bytes);
      if (ep == p && n == 1) continue;
      const float *q = data + stride * (size_t)p;
      const int el = levels[p];
      for (int layer = 0; layer <= el; layer++) {                   /* phase A */
        uint32_t *r1;
        size_t n1 = hb_row_get(&h, p, layer, &r1);
        if (n1 == 0) continue;
        /* sCand is fixed before any list of this layer is rewritten: compute all new lists, then write them */
        uint32_t *neighs = (uint32_t *)malloc(4 * n1);
        orc_pair *lists = (orc_pair *)malloc(sizeof(orc_pair) * n1 * h.maxM0);
        size_t *lens = (size_t *)malloc(sizeof(size_t) * n1);
        memcpy(neighs, r1, 4 * n1);
        for (size_t a = 0; a < n1; a++) {
          lens[a] = hb_update_neighbor(&h, p, layer, neighs[a], buf, bufcap);
          memcpy(lists + a * h.maxM0, buf, sizeof(orc_pair) * lens[a]);
        }
        for (size_t a = 0; a < n1; a++) {
          uint32_t *cnt; unsigned cap;
          uint32_t *row = hb_row(&h, neighs[a], layer, &cnt, &cap);
          hb_write_row(row, cnt, lists + a * h.maxM0, lens[a]);
          rewritten++;
        }
        free(neighs); free(lists); free(lens);
      }
      uint32_t cur = ep;                                            /* phase B */
      if (el < maxlevel) cur = hb_descend(&h, q, cur, maxlevel, el);
      for (int level = el; level >= 0; level--) {
        size_t nt = hb_search(&h, cur, q, level, 0, top, cand, expanded), f = 0;
        for (size_t t = 0; t < nt; t++) if (top[t].id != p) top[f++] = top[t];
        if (f == 0) continue;
        size_t k = hb_heuristic(&h, top, f, h.M);
        uint32_t *cnt; unsigned cap;
        uint32_t *row = hb_row(&h, p, level, &cnt, &cap);
        hb_write_row(row, cnt, top, k);
        cur = top[0].id;
        for (size_t t = 0; t < k; t++) {
          uint32_t sn = top[k - 1 - t].id;
          uint32_t *ocnt; unsigned ocap;
          uint32_t *orow = hb_row(&h, sn, level, &ocnt, &ocap);
          int present = 0;
          for (unsigned j = 0; j < *ocnt; j++) present |= orow[j] == p;
          if (present) continue;                                    /* :548-560 */
          if (*ocnt < ocap) { orow[(*ocnt)++] = p; continue; }
          size_t m = 0;
          tmp[m].d = hb_dist(&h, q, sn); tmp[m].id = p; m++;
          for (unsigned j = 0; j < *ocnt; j++) { tmp[m].d = hb_dist(&h, data + stride * (size_t)orow[j], sn); tmp[m].id = orow[j]; m++; }
          size_t kk = hb_heuristic(&h, tmp, m, ocap);
          hb_write_row(orow, ocnt, tmp, kk);
        }
      }
    }
  } else {
    for (size_t i = 0; i < nb; i++) memcpy(data + stride * (size_t)ids[i], newvec + stride * i, row_bytes);
    /* phase A on the snapshot: claim = last batch position per (row); staged rows */
    const size_t nrow = n + rows;
    uint32_t *claim = (uint32_t *)calloc(nrow, 4);
    for (size_t i = 0; i < nb; i++) {
      const uint32_t p = ids[i];
      for (int layer = 0; layer <= levels[p]; layer++) {
        uint32_t *r1; size_t n1 = hb_row_get(&h, p, layer, &r1);
        for (size_t a = 0; a < n1; a++) {
          size_t ridx = layer == 0 ? r1[a] : n + upper_ptr[r1[a]] + (layer - 1);
          claim[ridx] = (uint32_t)i + 1;
        }
      }
    }
    typedef struct { uint32_t node; int layer; size_t len; } stg_t;
    stg_t *stg = NULL; orc_pair *stgl = NULL; size_t ns = 0, caps = 0;
    for (size_t i = 0; i < nb; i++) {
      const uint32_t p = ids[i];
      for (int layer = 0; layer <= levels[p]; layer++) {
        uint32_t *r1; size_t n1 = hb_row_get(&h, p, layer, &r1);
        for (size_t a = 0; a < n1; a++) {
          size_t ridx = layer == 0 ? r1[a] : n + upper_ptr[r1[a]] + (layer - 1);
          if (claim[ridx] != (uint32_t)i + 1) continue;
          if (ns == caps) { caps = caps ? caps * 2 : 1024; stg = (stg_t *)realloc(stg, caps * sizeof(stg_t)); stgl = (orc_pair *)realloc(stgl, caps * h.maxM0 * sizeof(orc_pair)); }
          stg[ns].node = r1[a]; stg[ns].layer = layer;
          stg[ns].len = hb_update_neighbor(&h, p, layer, r1[a], buf, bufcap);
          memcpy(stgl + ns * h.maxM0, buf, sizeof(orc_pair) * stg[ns].len);
          ns++;
          claim[ridx] = 0;   /* a row listed twice by the same element is computed once */
        }
      }
    }
    for (size_t t = 0; t < ns; t++) {
      uint32_t *cnt; unsigned cap;
      uint32_t *row = hb_row(&h, stg[t].node, stg[t].layer, &cnt, &cap);
      hb_write_row(row, cnt, stgl + t * h.maxM0, stg[t].len);
    }
    rewritten = ns;
    free(claim); free(stg); free(stgl);
    /* phase B on the graph phase A left: own rows staged, reverse links grouped */
    const unsigned nlev = (unsigned)maxlevel + 1;
    orc_pair *own = (orc_pair *)malloc(sizeof(orc_pair) * nb * nlev * (h.M ? h.M : 1));
    size_t *ownk = (size_t *)calloc(nb * nlev, sizeof(size_t));
    uint8_t *ownset = (uint8_t *)calloc(nb * nlev, 1);
    hb_rev *rev = NULL; size_t nrev = 0, caprev = 0;
    for (size_t i = 0; i < nb; i++) {
      const uint32_t p = ids[i];
      const float *q = data + stride * (size_t)p;
      if (ep == p && n == 1) continue;
      const int el = levels[p];
      uint32_t cur = ep;
      if (el < maxlevel) cur = hb_descend(&h, q, cur, maxlevel, el);
      for (int level = el; level >= 0; level--) {
        size_t nt = hb_search(&h, cur, q, level, 1, top, cand, expanded), f = 0;
        for (size_t t = 0; t < nt; t++) if (top[t].id != p) top[f++] = top[t];
        if (f == 0) continue;
        size_t k = hb_heuristic(&h, top, f, h.M);
        memcpy(own + (i * nlev + level) * h.M, top, sizeof(orc_pair) * k);
        ownk[i * nlev + level] = k; ownset[i * nlev + level] = 1;
        cur = top[0].id;
        for (size_t t = 0; t < k; t++) {
          if (nrev == caprev) { caprev = caprev ? caprev * 2 : 4096; rev = (hb_rev *)realloc(rev, caprev * sizeof(hb_rev)); }
          rev[nrev].row_node = top[t].id; rev[nrev].level = level; rev[nrev].src = p; rev[nrev].d = top[t].d;
          nrev++;
        }
      }
    }
    for (size_t i = 0; i < nb; i++)
      for (unsigned level = 0; level < nlev; level++)
        if (ownset[i * nlev + level]) {
          uint32_t *cnt; unsigned cap;
          uint32_t *row = hb_row(&h, ids[i], (int)level, &cnt, &cap);
          hb_write_row(row, cnt, own + (i * nlev + level) * h.M, ownk[i * nlev + level]);
        }
    if (nrev) {
      qsort(rev, nrev, sizeof(hb_rev), rev_cmp);
      for (size_t a = 0; a < nrev;) {
        size_t b = a;
        while (b < nrev && rev[b].row_node == rev[a].row_node && rev[b].level == rev[a].level) b++;
        uint32_t sn = rev[a].row_node;
        uint32_t *ocnt; unsigned ocap;
        uint32_t *orow = hb_row(&h, sn, rev[a].level, &ocnt, &ocap);
        size_t m = 0, knew = 0;
        for (size_t t = a; t < b; t++) {
          int present = 0;
          for (unsigned j = 0; j < *ocnt; j++) present |= orow[j] == rev[t].src;
          for (size_t t2 = a; t2 < t; t2++) present |= rev[t2].src == rev[t].src;   /* the same element twice in `ids` */
          if (!present) { tmp[h.maxM0 + knew].d = rev[t].d; tmp[h.maxM0 + knew].id = rev[t].src; knew++; }
        }
        if (*ocnt + knew <= ocap) {
          for (size_t t = 0; t < knew; t++) orow[(*ocnt)++] = tmp[h.maxM0 + t].id;
        } else {
          for (unsigned j = 0; j < *ocnt; j++) { tmp[m].d = hb_dist(&h, data + stride * (size_t)sn, orow[j]); tmp[m].id = orow[j]; m++; }
          for (size_t t = 0; t < knew; t++) tmp[m++] = tmp[h.maxM0 + t];
          size_t kk = hb_heuristic(&h, tmp, m, ocap);
          hb_write_row(orow, ocnt, tmp, kk);
        }
        a = b;
      }
    }
    free(own); free(ownk); free(ownset); free(rev);
  }
  for (size_t i = 0; i < n; i++)
    for (unsigned j = h.cnt0[i]; j < h.maxM0; j++) links0[i * h.maxM0 + j] = ORC_INVALID;
  for (size_t r = 0; r < rows; r++)
    for (unsigned j = h.cntu[r]; j < h.maxM; j++) upper_links[r * h.maxM + j] = ORC_INVALID;
  if (info) { info[0] = h.n_dist; info[1] = rewritten; }
  free(h.cnt0); free(h.cntu); free(h.visited); free(buf); free(top); free(cand); free(expanded); free(tmp);
  return 0;
}

// TEST INFRASTRUCTURE ONLY — never linked into the product library.
// Thin extern "C" wrapper over the UNMODIFIED reference hnswlib headers
// (/root/reference/hnswlib/hnswlib/*.h), compiled into oracle/_ref/libref_hnsw.so
// by oracle/Makefile. Used by tests/ to pin oracle/oracle.c and as the
// "reference" CPU baseline in bench.py --impl reference.
#include "hnswlib/hnswlib.h"
#include <omp.h>
#include <cstring>
#include <vector>

using namespace hnswlib;

struct RefHnsw {
  SpaceInterface<float> *space;
  HierarchicalNSW<float> *index;
  size_t dim;
};

static SpaceInterface<float> *make_space(int metric, size_t dim) {
  if (metric == 1) return new InnerProductSpace(dim);
  return new L2Space(dim);
}

extern "C" {

float ref_dist(int metric, const float *a, const float *b, size_t dim) {
  SpaceInterface<float> *s = make_space(metric, dim);
  float d = s->get_dist_func()(a, b, s->get_dist_func_param());
  delete s;
  return d;
}

// BruteforceSearch::searchKnn (bruteforce.h:106-135), one query per call like the reference.
// out rows are ascending distance (searchKnnCloserFirst order, hnswlib.h:203-221).
int ref_bf_knn(int metric, const float *base, size_t n, size_t dim, const float *q, size_t nq,
               size_t k, unsigned long long *labels, float *dists, int threads) {
  SpaceInterface<float> *s = make_space(metric, dim);
  BruteforceSearch<float> bf(s, n);
  for (size_t i = 0; i < n; i++) bf.addPoint(base + i * dim, i);
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 8)
  for (long i = 0; i < (long)nq; i++) {
    auto r = bf.searchKnn(q + i * dim, k);
    size_t m = r.size();
    for (size_t j = m; j < k; j++) { labels[i * k + j] = (unsigned long long)-1; dists[i * k + j] = -1.f; }
    while (!r.empty()) {
      --m;
      labels[i * k + m] = r.top().second;
      dists[i * k + m] = r.top().first;
      r.pop();
    }
  }
  delete s;
  return 0;
}

// BruteforceSearch<int> over L2SpaceI (space_l2.h:294-323): u8 rows, int distances.
int ref_bf_knn_u8(const unsigned char *base, size_t n, size_t dim, const unsigned char *q, size_t nq, size_t k,
                  unsigned long long *labels, int *dists, int threads) {
  L2SpaceI space(dim);
  BruteforceSearch<int> bf(&space, n);
  for (size_t i = 0; i < n; i++) bf.addPoint(base + i * dim, i);
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 8)
  for (long i = 0; i < (long)nq; i++) {
    auto r = bf.searchKnn(q + i * dim, k);
    size_t m = r.size();
    for (size_t j = m; j < k; j++) { labels[i * k + j] = (unsigned long long)-1; dists[i * k + j] = -1; }
    while (!r.empty()) {
      --m;
      labels[i * k + m] = r.top().second;
      dists[i * k + m] = r.top().first;
      r.pop();
    }
  }
  return 0;
}

void *ref_hnsw_new(int metric, size_t dim, size_t max_elements, size_t M, size_t efc, size_t seed) {
  RefHnsw *h = new RefHnsw;
  h->dim = dim;
  h->space = make_space(metric, dim);
  h->index = new HierarchicalNSW<float>(h->space, max_elements, M, efc, seed);
  return h;
}

void *ref_hnsw_load(int metric, size_t dim, const char *path) {
  RefHnsw *h = new RefHnsw;
  h->dim = dim;
  h->space = make_space(metric, dim);
  try {
    h->index = new HierarchicalNSW<float>(h->space, std::string(path));
  } catch (std::exception &e) {
    delete h->space; delete h; return nullptr;
  }
  return h;
}

void ref_hnsw_free(void *p) {
  RefHnsw *h = (RefHnsw *)p;
  delete h->index; delete h->space; delete h;
}

// omp-parallel addPoint loop, the way hnswlib/tests/cpp/sift_1m.cpp:321-336 drives it
// (first point inserted alone, rest in parallel). threads=1 gives a reproducible graph.
int ref_hnsw_add(void *p, const float *data, size_t n, size_t label0, int threads) {
  RefHnsw *h = (RefHnsw *)p;
  if (threads <= 0) threads = omp_get_max_threads();
  size_t start = 0;
  if (h->index->cur_element_count == 0 && n > 0) { h->index->addPoint(data, label0); start = 1; }
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64)
  for (long i = (long)start; i < (long)n; i++) h->index->addPoint(data + i * h->dim, label0 + i);
  return 0;
}

int ref_hnsw_search(void *p, const float *q, size_t nq, size_t k, size_t ef,
                    unsigned long long *labels, float *dists, int threads,
                    long long *hops, long long *ndist) {
  RefHnsw *h = (RefHnsw *)p;
  h->index->setEf(ef);
  h->index->metric_hops = 0;
  h->index->metric_distance_computations = 0;
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
  for (long i = 0; i < (long)nq; i++) {
    auto r = h->index->searchKnn(q + i * h->dim, k);
    size_t m = r.size();
    for (size_t j = m; j < k; j++) { labels[i * k + j] = (unsigned long long)-1; dists[i * k + j] = -1.f; }
    while (!r.empty()) {
      --m;
      labels[i * k + m] = r.top().second;
      dists[i * k + m] = r.top().first;
      r.pop();
    }
  }
  if (hops) *hops = h->index->metric_hops;
  if (ndist) *ndist = h->index->metric_distance_computations;
  return 0;
}

// searchKnn(query, k, isIdAllowed) (hnswalg.h:1270-1324 -> searchBaseLayerST<false>, :309-440): allow [max label + 1]
// by LABEL (NULL = no functor, only the deleted marks filter)
struct BitmapFilter : public BaseFilterFunctor {
  const unsigned char *allow;
  explicit BitmapFilter(const unsigned char *a) : allow(a) {}
  bool operator()(labeltype id) override { return allow[id] != 0; }
};
int ref_hnsw_search_filtered(void *p, const float *q, size_t nq, size_t k, size_t ef, const unsigned char *allow,
                             unsigned long long *labels, float *dists, int threads) {
  RefHnsw *h = (RefHnsw *)p;
  h->index->setEf(ef);
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
  for (long i = 0; i < (long)nq; i++) {
    BitmapFilter f(allow);
    auto r = h->index->searchKnn(q + i * h->dim, k, allow ? &f : nullptr);
    size_t m = r.size();
    for (size_t j = m; j < k; j++) { labels[i * k + j] = (unsigned long long)-1; dists[i * k + j] = -1.f; }
    while (!r.empty()) {
      --m;
      labels[i * k + m] = r.top().second;
      dists[i * k + m] = r.top().first;
      r.pop();
    }
  }
  return 0;
}
// searchStopConditionClosest(query, EpsilonSearchStopCondition(eps, min, max), isIdAllowed) (hnswalg.h:1327-1378,
// stop_condition.h:219-275), the UNMODIFIED reference. ids [nq, max] = what the reference returns in .second (internal
// ids, :1368-1371), padded with ~0 / -1; counts [nq]. allow by LABEL or NULL.
int ref_hnsw_search_epsilon(void *p, const float *q, size_t nq, float eps, size_t min_cand, size_t max_cand,
                            const unsigned char *allow, unsigned long long *ids, float *dists, unsigned *counts, int threads) {
  RefHnsw *h = (RefHnsw *)p;
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
  for (long i = 0; i < (long)nq; i++) {
    BitmapFilter f(allow);
    hnswlib::EpsilonSearchStopCondition<float> stop(eps, min_cand, max_cand);
    auto r = h->index->searchStopConditionClosest(q + i * h->dim, stop, allow ? &f : nullptr);
    size_t m = r.size();
    counts[i] = (unsigned)m;
    for (size_t j = 0; j < max_cand; j++) {
      ids[i * max_cand + j] = j < m ? (unsigned long long)r[j].second : (unsigned long long)-1;
      dists[i * max_cand + j] = j < m ? r[j].first : -1.f;
    }
  }
  return 0;
}
// markDelete / unmarkDelete (hnswalg.h:853-911)
int ref_hnsw_mark_deleted(void *p, size_t label, int on) {
  RefHnsw *h = (RefHnsw *)p;
  try {
    if (on) h->index->markDelete(label); else h->index->unmarkDelete(label);
  } catch (std::exception &e) { return -1; }
  return 0;
}

int ref_hnsw_save(void *p, const char *path) {
  ((RefHnsw *)p)->index->saveIndex(path);
  return 0;
}

// introspection so tests can compare graphs edge by edge
long long ref_hnsw_info(void *p, int what) {
  HierarchicalNSW<float> *ix = ((RefHnsw *)p)->index;
  switch (what) {
    case 0: return (long long)ix->cur_element_count;
    case 1: return ix->maxlevel_;
    case 2: return ix->enterpoint_node_;
    case 3: return (long long)ix->maxM_;
    case 4: return (long long)ix->maxM0_;
    case 5: return (long long)ix->M_;
    case 6: return (long long)ix->ef_construction_;
  }
  return -1;
}
int ref_hnsw_level(void *p, size_t id) { return ((RefHnsw *)p)->index->element_levels_[id]; }
int ref_hnsw_links(void *p, size_t id, int level, unsigned *out) {
  HierarchicalNSW<float> *ix = ((RefHnsw *)p)->index;
  unsigned *ll = ix->get_linklist_at_level((tableint)id, level);
  int cnt = ix->getListCount(ll);
  memcpy(out, ll + 1, cnt * sizeof(unsigned));
  return cnt;
}
}

// ---- multi-vector spaces + MultiVectorSearchStopCondition (stop_condition.h:9-215): rows = [dim floats][u32 doc id]
struct RefHnswMV {
  BaseMultiVectorSpace<unsigned> *space;
  HierarchicalNSW<float> *index;
  size_t dim;
};
extern "C" {
void *ref_hnswmv_new(int metric, size_t dim, size_t max_elements, size_t M, size_t efc, size_t seed) {
  RefHnswMV *h = new RefHnswMV;
  h->dim = dim;
  if (metric == 1) h->space = new MultiVectorInnerProductSpace<unsigned>(dim);
  else h->space = new MultiVectorL2Space<unsigned>(dim);
  h->index = new HierarchicalNSW<float>(h->space, max_elements, M, efc, seed);
  return h;
}
void ref_hnswmv_free(void *p) {
  RefHnswMV *h = (RefHnswMV *)p;
  delete h->index; delete h->space; delete h;
}
// data [n, dim] floats + doc [n]: the rows the reference stores are assembled here (vector, then the document id)
int ref_hnswmv_add(void *p, const float *data, const unsigned *doc, size_t n, size_t label0) {
  RefHnswMV *h = (RefHnswMV *)p;
  std::vector<char> row(h->space->get_data_size());
  for (size_t i = 0; i < n; i++) {
    memcpy(row.data(), data + i * h->dim, h->dim * sizeof(float));
    h->space->set_doc_id(row.data(), doc[i]);
    h->index->addPoint(row.data(), label0 + i);
  }
  return 0;
}
int ref_hnswmv_search(void *p, const float *q, size_t nq, size_t num_docs, size_t ef_collection, size_t max_results,
                      const unsigned char *allow, unsigned long long *ids, float *dists, unsigned *counts, int threads) {
  RefHnswMV *h = (RefHnswMV *)p;
  if (threads <= 0) threads = omp_get_max_threads();
  int rc = 0;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
  for (long i = 0; i < (long)nq; i++) {
    BitmapFilter f(allow);
    MultiVectorSearchStopCondition<unsigned, float> stop(*h->space, num_docs, ef_collection);
    auto r = h->index->searchStopConditionClosest(q + i * h->dim, stop, allow ? &f : nullptr);
    size_t m = r.size();
    if (m > max_results) { rc = -2; m = max_results; }
    counts[i] = (unsigned)m;
    for (size_t j = 0; j < max_results; j++) {
      ids[i * max_results + j] = j < m ? (unsigned long long)r[j].second : (unsigned long long)-1;
      dists[i * max_results + j] = j < m ? r[j].first : -1.f;
    }
  }
  return rc;
}
int ref_hnswmv_save(void *p, const char *path) {
  try { ((RefHnswMV *)p)->index->saveIndex(path); } catch (std::exception &e) { return -1; }
  return 0;
}
}

// ---- HierarchicalNSW<int> over L2SpaceI (space_l2.h:255-323): unsigned-byte rows, int distances. Same drivers as
// the float wrappers above, on the UNMODIFIED reference templates instantiated with dist_t = int.
struct RefHnswI {
  L2SpaceI *space;
  HierarchicalNSW<int> *index;
  size_t dim;
};
extern "C" {
void *ref_hnswi_new(size_t dim, size_t max_elements, size_t M, size_t efc, size_t seed) {
  RefHnswI *h = new RefHnswI;
  h->dim = dim;
  h->space = new L2SpaceI(dim);
  h->index = new HierarchicalNSW<int>(h->space, max_elements, M, efc, seed);
  return h;
}
void ref_hnswi_free(void *p) {
  RefHnswI *h = (RefHnswI *)p;
  delete h->index; delete h->space; delete h;
}
int ref_hnswi_add(void *p, const unsigned char *data, size_t n, size_t label0, int threads) {
  RefHnswI *h = (RefHnswI *)p;
  if (threads <= 0) threads = omp_get_max_threads();
  size_t start = 0;
  if (h->index->cur_element_count == 0 && n > 0) { h->index->addPoint(data, label0); start = 1; }
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64)
  for (long i = (long)start; i < (long)n; i++) h->index->addPoint(data + i * h->dim, label0 + i);
  return 0;
}
// allow by LABEL or NULL (then plain searchKnn, the bare-bone path)
int ref_hnswi_search(void *p, const unsigned char *q, size_t nq, size_t k, size_t ef, const unsigned char *allow,
                     unsigned long long *labels, int *dists, int threads) {
  RefHnswI *h = (RefHnswI *)p;
  h->index->setEf(ef);
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
  for (long i = 0; i < (long)nq; i++) {
    BitmapFilter f(allow);
    auto r = h->index->searchKnn(q + i * h->dim, k, allow ? &f : nullptr);
    size_t m = r.size();
    for (size_t j = m; j < k; j++) { labels[i * k + j] = (unsigned long long)-1; dists[i * k + j] = 0x7FFFFFFF; }
    while (!r.empty()) {
      --m;
      labels[i * k + m] = r.top().second;
      dists[i * k + m] = r.top().first;
      r.pop();
    }
  }
  return 0;
}
int ref_hnswi_search_epsilon(void *p, const unsigned char *q, size_t nq, float eps, size_t min_cand, size_t max_cand,
                             unsigned long long *ids, int *dists, unsigned *counts, int threads) {
  RefHnswI *h = (RefHnswI *)p;
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
  for (long i = 0; i < (long)nq; i++) {
    EpsilonSearchStopCondition<int> stop(eps, min_cand, max_cand);
    auto r = h->index->searchStopConditionClosest(q + i * h->dim, stop, nullptr);
    size_t m = r.size();
    counts[i] = (unsigned)m;
    for (size_t j = 0; j < max_cand; j++) {
      ids[i * max_cand + j] = j < m ? (unsigned long long)r[j].second : (unsigned long long)-1;
      dists[i * max_cand + j] = j < m ? r[j].first : 0x7FFFFFFF;
    }
  }
  return 0;
}
int ref_hnswi_save(void *p, const char *path) {
  try { ((RefHnswI *)p)->index->saveIndex(path); } catch (std::exception &e) { return -1; }
  return 0;
}
void *ref_hnswi_load(size_t dim, const char *path) {
  RefHnswI *h = new RefHnswI;
  h->dim = dim;
  h->space = new L2SpaceI(dim);
  try {
    h->index = new HierarchicalNSW<int>(h->space, std::string(path));
  } catch (std::exception &e) {
    delete h->space; delete h; return nullptr;
  }
  return h;
}
}

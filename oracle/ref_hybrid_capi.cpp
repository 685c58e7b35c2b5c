// TEST INFRASTRUCTURE ONLY — never linked into the product library.
// extern "C" wrapper over the UNMODIFIED reference hybrid index
// (/root/reference/hnsw_nsg/include/index_hnsw_nsg.h, hnsw_nsg/src/{index,index_nsg}.cpp),
// compiled into oracle/_ref/libref_hybrid*.so by oracle/Makefile.
#include "index_hnsw_nsg.h"
#include <omp.h>
#include <cstdlib>
#include <string>
#include <vector>

using namespace hnsw_nsg;

struct RefHybrid {
  L2Space *space;
  HNSW_NSG<float> *index;
  size_t dim;
};

extern "C" {

// DistanceL2AVX::compare (hnsw_nsg/include/nsg/distance.h:325-346)
float ref_hybrid_dist(const float *a, const float *b, unsigned dim) {
  DistanceL2AVX d;
  return d.compare(a, b, dim);
}

// the build sequence of hnsw_nsg/tests/test_hnsw_nsg_search.cpp:318-351: addPoint for every row
// (label = row id), Build_NSG from a kNN-graph file, saveIndex(prefix) -> prefix_hnsw.bin, prefix_nsg.bin
int ref_hybrid_build(float *data, size_t n, size_t dim, const char *knn_path, unsigned L, unsigned R, unsigned C,
                     unsigned M, unsigned efc, const char *prefix, unsigned seed, int threads) {
  if (threads <= 0) threads = omp_get_max_threads();
  omp_set_num_threads(threads);
  srand(seed);
  Parameters p;
  p.Set<unsigned>("L", L);
  p.Set<unsigned>("R", R);
  p.Set<unsigned>("C", C);
  p.Set<std::string>("nn_graph_path", knn_path);
  L2Space space(dim);
  HNSW_NSG<float> ix(&space, (unsigned)dim, n, data, nullptr, p, M, efc, 20);
  if (n) ix.addPoint(data, 0, false);
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64)
  for (long i = 1; i < (long)n; i++) ix.addPoint(data + (size_t)i * dim, (labeltype)i, false);
  ix.Build_NSG(data, p);
  ix.saveIndex(prefix);
  return 0;
}

void *ref_hybrid_load(float *data, size_t n, size_t dim, const char *prefix, unsigned L_search) {
  RefHybrid *h = new RefHybrid;
  h->dim = dim;
  h->space = new L2Space(dim);
  Parameters p;
  p.Set<unsigned>("L_search", L_search);
  p.Set<unsigned>("P_search", L_search);
  try {
    h->index = new HNSW_NSG<float>(h->space, (unsigned)dim, n, data, nullptr, p, std::string(prefix) + "_hnsw.bin",
                                   std::string(prefix) + "_nsg.bin");
  } catch (std::exception &e) {
    delete h->space; delete h; return nullptr;
  }
  // IndexNSG::Load leaves data_ null and SearchFromEnterpoint reads the seeds through data_
  // (hnsw_nsg/src/index_nsg.cpp:744); IndexNSG::Search binds data_ = x (:446), so one throw-away
  // Search call makes the loaded index usable without touching the reference.
  unsigned tmp = 0;
  Parameters one;
  one.Set<unsigned>("L_search", 1);
  h->index->nsg_.Search(data, data, 1, one, &tmp);
  return h;
}

void ref_hybrid_free(void *p) {
  RefHybrid *h = (RefHybrid *)p;
  delete h->index; delete h->space; delete h;
}

// HNSW_NSG::searchKnn (index_hnsw_nsg.h:107-151), one query per call. Kout = max(k, ef) ids per
// query. The reference's searchKnn appends to a member vector (`res`, :99) so it is not thread-safe:
// single-threaded here; the vector is cleared between batches to bound memory.
int ref_hybrid_search(void *p, const float *q, size_t nq, size_t k, size_t ef, unsigned *ids, unsigned seed) {
  RefHybrid *h = (RefHybrid *)p;
  h->index->setEf(ef);
  size_t kout = k > ef ? k : ef;
  srand(seed);
  for (size_t i = 0; i < nq; i++) {
    std::vector<unsigned> r = h->index->searchKnn(q + i * h->dim, k, nullptr);
    for (size_t j = 0; j < kout; j++) ids[i * kout + j] = j < r.size() ? r[j] : 0xFFFFFFFFu;
    h->index->res.clear();
  }
  return 0;
}

// IndexNSG::SearchFromEnterpoint (hnsw_nsg/src/index_nsg.cpp:703-782) with explicit entry points
int ref_hybrid_search_from(void *p, float *data, const float *q, size_t nq, size_t K, unsigned L_search,
                           const unsigned *entry, unsigned *ids, unsigned seed) {
  RefHybrid *h = (RefHybrid *)p;
  Parameters par;
  par.Set<unsigned>("L_search", L_search);
  srand(seed);
  for (size_t i = 0; i < nq; i++)
    h->index->nsg_.SearchFromEnterpoint(q + i * h->dim, data, K, par, ids + i * K, entry[i]);
  return 0;
}
}

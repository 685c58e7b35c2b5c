// TEST INFRASTRUCTURE ONLY — never linked into the product library.
// extern "C" wrapper over the UNMODIFIED reference NSG sources
// (/root/reference/CNNS/src/nsg/{index_nsg,index}.cpp, CNNS/include/nsg/*.h),
// compiled into oracle/_ref/libref_nsg.so by oracle/Makefile.
#include "index_nsg.h"
#include <omp.h>
#include <cstdlib>
#include <vector>

using namespace efanna2e;

extern "C" {

float ref_nsg_dist_l2(const float *a, const float *b, unsigned dim) {
  DistanceL2 d;
  return d.compare(a, b, dim);
}

// IndexNSG::Build (index_nsg.cpp:465-504) + Save (:37-49).
int ref_nsg_build(const float *data, size_t n, size_t dim, const char *knn_path, unsigned L, unsigned R,
                  unsigned C, const char *out_path, unsigned seed, int threads) {
  if (threads > 0) omp_set_num_threads(threads);
  srand(seed);
  IndexNSG index(dim, n, L2, nullptr);
  Parameters p;
  p.Set<unsigned>("L", L);
  p.Set<unsigned>("R", R);
  p.Set<unsigned>("C", C);
  p.Set<std::string>("nn_graph_path", knn_path);
  index.Build(n, data, p);
  index.Save(out_path);
  return 0;
}

void *ref_nsg_load(size_t n, size_t dim, const char *path) {
  IndexNSG *ix = new IndexNSG(dim, n, L2, nullptr);
  ix->Load(path);
  return ix;
}
void ref_nsg_free(void *p) { delete (IndexNSG *)p; }

// IndexNSG::Search (index_nsg.cpp:506-568), queries in a loop.
// threads=1 + srand(seed) => rand()-filled seeds are reproducible.
int ref_nsg_search(void *p, const float *data, const float *q, size_t nq, size_t dim, size_t K,
                   unsigned L_search, unsigned *ids, unsigned seed, int threads) {
  IndexNSG *ix = (IndexNSG *)p;
  Parameters par;
  par.Set<unsigned>("L_search", L_search);
  par.Set<unsigned>("P_search", L_search);
  srand(seed);
  if (threads <= 0) threads = omp_get_max_threads();
  if (threads == 1) {
    for (size_t i = 0; i < nq; i++) ix->Search(q + i * dim, data, K, par, ids + i * K);
  } else {
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
    for (long i = 0; i < (long)nq; i++) ix->Search(q + i * dim, data, K, par, ids + i * K);
  }
  return 0;
}

// OptimizeGraph (:659-682) consumes final_graph_; call once, then ref_nsg_search_opt.
int ref_nsg_optimize(void *p, float *data) {
  ((IndexNSG *)p)->OptimizeGraph(data);
  return 0;
}
int ref_nsg_search_opt(void *p, const float *q, size_t nq, size_t dim, size_t K, unsigned L_search,
                       unsigned *ids, unsigned seed, int threads) {
  IndexNSG *ix = (IndexNSG *)p;
  Parameters par;
  par.Set<unsigned>("L_search", L_search);
  par.Set<unsigned>("P_search", L_search);
  srand(seed);
  if (threads <= 0) threads = omp_get_max_threads();
  if (threads == 1) {
    for (size_t i = 0; i < nq; i++) ix->SearchWithOptGraph(q + i * dim, K, par, ids + i * K);
  } else {
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
    for (long i = 0; i < (long)nq; i++) ix->SearchWithOptGraph(q + i * dim, K, par, ids + i * K);
  }
  return 0;
}
}

// TEST INFRASTRUCTURE ONLY — never linked into the product library.
// extern "C" wrapper over the UNMODIFIED reference NN-descent sources
// (/root/reference/CNNS/efanna_graph/src/{index_graph,index_random,index}.cpp),
// compiled into oracle/_ref/libref_efanna.so by oracle/Makefile.
#include <efanna2e/index_graph.h>
#include <efanna2e/index_random.h>
#include <omp.h>
#include <cstdlib>

using namespace efanna2e;

extern "C" {
// IndexGraph::Build (index_graph.cpp:265-294) with the IndexRandom initializer, then Save (:348-357).
int ref_nndescent_build(const float *data, size_t n, size_t dim, unsigned K, unsigned L, unsigned iter,
                        unsigned S, unsigned R, const char *out_path, unsigned seed, int threads) {
  if (threads > 0) omp_set_num_threads(threads);
  srand(seed);
  IndexRandom init(dim, n);
  IndexGraph index(dim, n, L2, (Index *)(&init));
  Parameters p;
  p.Set<unsigned>("K", K);
  p.Set<unsigned>("L", L);
  p.Set<unsigned>("iter", iter);
  p.Set<unsigned>("S", S);
  p.Set<unsigned>("R", R);
  index.Build(n, data, p);
  index.Save(out_path);
  return 0;
}
}

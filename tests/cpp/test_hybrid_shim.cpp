// Drives the hybrid shim the way hnsw_nsg/tests/test_hnsw_nsg_search.cpp:300-400 drives the reference:
// HNSW_NSG build ctor, addPoint loop, kNN graph file, Build_NSG, setEf, searchKnn per query, saveIndex(prefix),
// then the load ctor on the saved pair of files. Exit code 0 iff recall@10 > 0.9 against brute force, the batched
// and single-query forms agree, and the reloaded index answers identically.
#include <b200ann/hnsw_nsg.hpp>
#include <cstdio>
#include <random>
#include <set>

#define CHECK(c) do { if (!(c)) { printf("CHECK failed: %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

int main(int argc, char **argv) {
  std::string dir = argc > 1 ? argv[1] : "/tmp";
  const size_t n = 6000, dim = 32, nq = 200, k = 10;
  std::mt19937 rng(47);
  std::uniform_real_distribution<float> u(0.f, 1.f);
  std::vector<float> base(n * dim), q(nq * dim);
  for (auto &v : base) v = u(rng);
  for (auto &v : q) v = u(rng);
  std::string kpath = dir + "/hy.graph", prefix = dir + "/hyshim";
  {  // kNN graph for the NSG part (efanna IndexGraph, NN-descent on the GPU)
    efanna2e::IndexRandom init(dim, n);
    efanna2e::IndexGraph kg(dim, n, efanna2e::L2, (efanna2e::Index *)(&init));
    efanna2e::Parameters bp;
    bp.Set<unsigned>("K", 40); bp.Set<unsigned>("L", 60); bp.Set<unsigned>("iter", 8); bp.Set<unsigned>("S", 10); bp.Set<unsigned>("R", 60);
    kg.Build(n, base.data(), bp);
    kg.Save(kpath.c_str());
  }
  hnsw_nsg::L2Space space(dim);
  efanna2e::Parameters params;
  params.Set<unsigned>("L_search", 60);
  hnsw_nsg::HNSW_NSG<float> hybrid(&space, (unsigned)dim, n, base.data(), q.data(), params, 16, 200, 20);
  for (size_t i = 0; i < n; i++) hybrid.addPoint(base.data() + i * dim, i, false);
  efanna2e::Parameters np;
  np.Set<unsigned>("L", 40); np.Set<unsigned>("R", 30); np.Set<unsigned>("C", 300);
  np.Set<std::string>("nn_graph_path", kpath);
  hybrid.Build_NSG(base.data(), np);
  hybrid.setEf(10);
  std::vector<unsigned> batch(nq * k);
  hybrid.searchKnnBatch(q.data(), nq, k, batch.data());
  size_t hit = 0;
  for (size_t i = 0; i < nq; i++) {
    std::vector<std::pair<float, unsigned>> d(n);
    for (size_t j = 0; j < n; j++) {
      float s = 0;
      for (size_t t = 0; t < dim; t++) { float x = q[i * dim + t] - base[j * dim + t]; s += x * x; }
      d[j] = {s, (unsigned)j};
    }
    std::partial_sort(d.begin(), d.begin() + k, d.end());
    std::set<unsigned> truth;
    for (size_t j = 0; j < k; j++) truth.insert(d[j].second);
    for (size_t j = 0; j < k; j++) hit += truth.count(batch[i * k + j]);
  }
  double recall = (double)hit / (nq * k);
  printf("hybrid shim: recall@10 %.4f (L_search 60)\n", recall);
  CHECK(recall > 0.9);
  hybrid.setEf(25);  // K = max(k, ef) ids per query (index_hnsw_nsg.h:147)
  auto r = hybrid.searchKnn(q.data(), k, nullptr);
  CHECK(r.size() == 25);
  hybrid.setEf(10);
  hybrid.saveIndex(prefix);
  hnsw_nsg::HNSW_NSG<float> loaded(&space, (unsigned)dim, n, base.data(), q.data(), params, prefix + "_hnsw.bin", prefix + "_nsg.bin");
  loaded.setEf(10);
  std::vector<unsigned> again(nq * k);
  loaded.searchKnnBatch(q.data(), nq, k, again.data());
  CHECK(again == batch);
  printf("hybrid shim ok\n");
  return 0;
}

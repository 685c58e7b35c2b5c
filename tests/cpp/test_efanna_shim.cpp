// Drives the efanna2e shim the way the reference's drivers do (CNNS/tests/nndescent_nsg.cpp:35-128,
// CNNS/efanna_graph/tests/test_nndescent.cpp): IndexRandom + IndexGraph::Build/Save, IndexNSG::Build/
// Save/Load/Search. Prints recall@10 against brute force; exit code 0 iff recall > 0.9 and the
// reloaded index answers identically.
#include <b200ann/efanna2e.hpp>
#include <algorithm>
#include <cstdlib>
#include <random>
#include <set>

int main(int argc, char **argv) {
  const char *dir = argc > 1 ? argv[1] : "/tmp";
  const size_t n = 5000, dim = 32, nq = 100, K = 10;
  std::mt19937 rng(47);
  std::uniform_real_distribution<float> u(0.f, 1.f);
  std::vector<float> base(n * dim), q(nq * dim);
  for (auto &v : base) v = u(rng);
  for (auto &v : q) v = u(rng);
  std::string kpath = std::string(dir) + "/shim.graph", gpath = std::string(dir) + "/shim.nsg";

  efanna2e::IndexRandom init(dim, n);
  efanna2e::IndexGraph kg(dim, n, efanna2e::L2, (efanna2e::Index *)(&init));
  efanna2e::Parameters bp;
  bp.Set<unsigned>("K", 40); bp.Set<unsigned>("L", 60); bp.Set<unsigned>("iter", 8); bp.Set<unsigned>("S", 10); bp.Set<unsigned>("R", 100);
  kg.Build(n, base.data(), bp);
  kg.Save(kpath.c_str());

  efanna2e::IndexNSG index(dim, n, efanna2e::L2, nullptr);
  efanna2e::Parameters p;
  p.Set<unsigned>("L", 40); p.Set<unsigned>("R", 30); p.Set<unsigned>("C", 300);
  p.Set<std::string>("nn_graph_path", kpath);
  index.Build(n, base.data(), p);
  index.Save(gpath.c_str());

  efanna2e::Parameters sp;
  sp.Set<unsigned>("L_search", 60); sp.Set<unsigned>("P_search", 60);
  std::vector<unsigned> res(nq * K), res2(nq * K), res3(nq * K);
  for (size_t i = 0; i < nq; i++) index.Search(q.data() + i * dim, base.data(), K, sp, res.data() + i * K);
  index.SearchBatch(q.data(), nq, K, sp, res2.data());

  efanna2e::IndexNSG loaded(dim, n, efanna2e::L2, nullptr);
  loaded.Load(gpath.c_str());
  // the pseudo-random pool seeds are a counter hash of (seed, position in the batch, draw) - the reference draws
  // them with rand() (index_nsg.cpp:523) - so a query's result is a function of its batch position: compare
  // singles with singles and batches with batches, across Save/Load
  std::vector<unsigned> res4(nq * K);
  for (size_t i = 0; i < nq; i++) loaded.Search(q.data() + i * dim, base.data(), K, sp, res4.data() + i * K);
  loaded.SearchBatch(q.data(), nq, K, sp, res3.data());

  // brute force truth
  size_t hit = 0;
  for (size_t i = 0; i < nq; i++) {
    std::vector<std::pair<float, unsigned>> d(n);
    for (size_t j = 0; j < n; j++) {
      float s = 0;
      for (size_t t = 0; t < dim; t++) { float x = q[i * dim + t] - base[j * dim + t]; s += x * x; }
      d[j] = {s, (unsigned)j};
    }
    std::partial_sort(d.begin(), d.begin() + K, d.end());
    std::set<unsigned> truth;
    for (size_t t = 0; t < K; t++) truth.insert(d[t].second);
    for (size_t t = 0; t < K; t++) hit += truth.count(res[i * K + t]);
  }
  double recall = (double)hit / (nq * K);
  bool same = res == res4 && res2 == res3;
  // KD-tree initializer + RefineGraph (efanna_graph/tests/test_kdtree_graph.cpp; index_graph.cpp:235-262): the refined graph
  // has the requested shape, lists no node in its own row, and is closer to the exact graph than the same NN-descent
  // effort from a random start
  {
    efanna2e::IndexKDtree kd(dim, n, efanna2e::L2, nullptr);
    efanna2e::Parameters kp;
    kp.Set<unsigned>("K", 40); kp.Set<unsigned>("nTrees", 8); kp.Set<unsigned>("mLevel", 4);
    kd.Build(n, base.data(), kp);
    kd.Save((std::string(dir) + "/shim_kd.graph").c_str());
    efanna2e::IndexGraph warm(dim, n, efanna2e::L2, (efanna2e::Index *)(&kd)), cold(dim, n, efanna2e::L2, (efanna2e::Index *)(&init)),
        exact(dim, n, efanna2e::L2, (efanna2e::Index *)(&init));
    efanna2e::Parameters rp;
    rp.Set<unsigned>("K", 20); rp.Set<unsigned>("L", 40); rp.Set<unsigned>("iter", 2); rp.Set<unsigned>("S", 8); rp.Set<unsigned>("R", 30);
    warm.RefineGraph(base.data(), rp);
    cold.Build(n, base.data(), rp);
    efanna2e::Parameters ep;
    ep.Set<unsigned>("K", 20); ep.Set<unsigned>("exact", 1);
    exact.Build(n, base.data(), ep);
    auto overlap = [&](const efanna2e::IndexGraph &g) {
      size_t h = 0;
      for (size_t i = 0; i < 500; i++) {
        std::set<unsigned> t(exact.Graph().begin() + i * 20, exact.Graph().begin() + (i + 1) * 20);
        for (size_t j = 0; j < 20; j++) { h += t.count(g.Graph()[i * 20 + j]); if (g.Graph()[i * 20 + j] == i) return -1.0; }
      }
      return (double)h / (500 * 20);
    };
    double rw = overlap(warm), rc = overlap(cold);
    printf("kd-tree init + RefineGraph: graph recall %.4f (random start, same effort: %.4f)\n", rw, rc);
    if (!(warm.K() == 20 && warm.Graph().size() == n * 20 && rw > rc && rc >= 0)) return 1;
  }
  bool threw = false;
  try { efanna2e::Parameters e; e.Get<unsigned>("missing"); } catch (std::invalid_argument &) { threw = true; }
  printf("recall@10 %.4f batch==single==reloaded %d width %u ep %u missing-key-throws %d\n", recall, (int)same, index.GetWidth(),
         index.GetEntryPoint(), (int)threw);
  return (recall > 0.9 && same && threw) ? 0 : 1;
}

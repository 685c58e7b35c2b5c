// Label order vs arrival order in the hybrid shim: the reference's driver adds points under
// `#pragma omp parallel for` (hnsw_nsg/tests/test_hnsw_nsg_search.cpp:331-333), so internal ids are not labels, while
// the NSG's ids are rows of `data` = labels (hnsw_nsg/include/index_hnsw_nsg.h:145-150).
//   mode "build <dir>":  addPoint in a shuffled, parallel order; every stored row must find ITSELF (its data-row id,
//                        not its arrival number) and recall@10 must exceed 0.9.
//   mode "load <prefix> <base.f32> <n> <dim> <queries.f32> <nq> <L_search> <k> <out.u32>": load a pair of files written
//                        by the UNMODIFIED reference (built with several threads) and write the ids found.
#include <b200ann/hnsw_nsg.hpp>
#include <cstdio>
#include <numeric>
#include <random>
#include <set>

#define CHECK(c) do { if (!(c)) { printf("CHECK failed: %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

static std::vector<float> read_f32(const char *path, size_t count) {
  std::vector<float> v(count);
  FILE *f = fopen(path, "rb");
  if (!f || fread(v.data(), 4, count, f) != count) { printf("cannot read %s\n", path); exit(2); }
  fclose(f);
  return v;
}

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  std::string mode = argc >= 3 ? argv[1] : "build";
  if (mode == "load") {
    std::string prefix = argv[2];
    size_t n = atol(argv[4]), dim = atol(argv[5]), nq = atol(argv[7]);
    unsigned L = (unsigned)atoi(argv[8]);
    size_t k = atol(argv[9]);
    std::vector<float> base = read_f32(argv[3], n * dim), q = read_f32(argv[6], nq * dim);
    hnsw_nsg::L2Space space(dim);
    efanna2e::Parameters params;
    params.Set<unsigned>("L_search", L);
    params.Set<unsigned>("seed", 1);
    hnsw_nsg::HNSW_NSG<float> ix(&space, (unsigned)dim, n, base.data(), q.data(), params, prefix + "_hnsw.bin", prefix + "_nsg.bin");
    ix.setEf(k);
    std::vector<unsigned> ids(nq * k);
    ix.searchKnnBatch(q.data(), nq, k, ids.data());
    FILE *f = fopen(argv[10], "wb");
    fwrite(ids.data(), 4, ids.size(), f);
    fclose(f);
    printf("hybrid order load ok\n");
    return 0;
  }
  std::string dir = argc >= 3 ? argv[2] : argv[1];
  const size_t n = 6000, dim = 32, nq = 200, k = 10;
  std::mt19937 rng(48);
  std::uniform_real_distribution<float> u(0.f, 1.f);
  std::vector<float> base(n * dim), q(nq * dim);
  for (auto &v : base) v = u(rng);
  for (auto &v : q) v = u(rng);
  std::string kpath = dir + "/hyo.graph";
  {
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
  std::vector<size_t> order(n);
  std::iota(order.begin(), order.end(), 0);
  std::shuffle(order.begin(), order.end(), rng);
#pragma omp parallel for schedule(dynamic, 16)
  for (long i = 0; i < (long)n; i++) hybrid.addPoint(base.data() + order[i] * dim, order[i], false);
  efanna2e::Parameters np;
  np.Set<unsigned>("L", 40); np.Set<unsigned>("R", 30); np.Set<unsigned>("C", 300);
  np.Set<std::string>("nn_graph_path", kpath);
  hybrid.Build_NSG(base.data(), np);
  hybrid.setEf(10);
  // every stored row finds itself under its data-row id
  std::vector<unsigned> self(500 * k);
  hybrid.searchKnnBatch(base.data(), 500, k, self.data());
  size_t own = 0;
  for (size_t i = 0; i < 500; i++) own += self[i * k] == i;
  printf("hybrid order: %zu / 500 rows find themselves first\n", own);
  // not all 500: SearchFromEnterpoint marks its entry point visited without ever evaluating it
  // (hnsw_nsg/src/index_nsg.cpp:711-712), so a row the HNSW descent lands on exactly cannot be returned - the
  // reference's own behaviour. Arrival numbers instead of row ids would give ~0.
  CHECK(own >= 400);
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
  printf("hybrid order: recall@10 %.4f\n", recall);
  CHECK(recall > 0.9);
  // labels that are not row numbers are refused (the reference would read `data` out of bounds)
  hnsw_nsg::HNSW_NSG<float> bad(&space, (unsigned)dim, 100, base.data(), q.data(), params, 16, 200, 20);
  for (size_t i = 0; i < 100; i++) bad.addPoint(base.data() + i * dim, i + 7, false);
  bool threw = false;
  try { bad.Build_NSG(base.data(), np); } catch (const std::runtime_error &) { threw = true; }
  CHECK(threw);
  printf("hybrid order ok\n");
  return 0;
}

// Mirrors hnswlib/tests/cpp/searchKnnCloserFirst_test.cpp:18-68 (ordering equivalence searchKnn vs
// searchKnnCloserFirst, Bruteforce as truth) and searchKnnWithFilter_test.cpp:33-88 (filter honoured)
// against the b200ann hnswlib shim; then loads an index file written by the REFERENCE hnswlib
// (argv[1], dim argv[2]) and checks recall of the GPU search against the shim's brute force and that
// saveIndex round-trips. Exit code 0 iff every assertion holds.
#include <b200ann/hnswlib.hpp>
#include <algorithm>
#include <cassert>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <set>

#define CHECK(c) do { if (!(c)) { printf("CHECK failed: %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

struct PickDivisibleIds : hnswlib::BaseFilterFunctor {
  unsigned divisor;
  explicit PickDivisibleIds(unsigned d) : divisor(d) {}
  bool operator()(hnswlib::labeltype id) override { return id % divisor == 0; }
};

int main(int argc, char **argv) {
  {  // ---- searchKnnCloserFirst_test: d=4 n=100 nq=10 k=10, seed 47
    int d = 4; size_t n = 100, nq = 10, k = 10;
    std::vector<float> data(n * d), query(nq * d);
    std::mt19937 rng; rng.seed(47);
    std::uniform_real_distribution<> distrib;
    for (auto &v : data) v = (float)distrib(rng);
    for (auto &v : query) v = (float)distrib(rng);
    hnswlib::L2Space space(d);
    hnswlib::AlgorithmInterface<float> *alg_brute = new hnswlib::BruteforceSearch<float>(&space, 2 * n);
    for (size_t i = 0; i < n; ++i) alg_brute->addPoint(data.data() + d * i, i);
    for (size_t j = 0; j < nq; ++j) {
      const void *p = query.data() + j * d;
      auto gd = alg_brute->searchKnn(p, k);
      auto res = alg_brute->searchKnnCloserFirst(p, k);
      CHECK(gd.size() == res.size() && gd.size() == k);
      size_t t = gd.size();
      while (!gd.empty()) { CHECK(gd.top() == res[--t]); gd.pop(); }
      for (size_t i = 1; i < res.size(); i++) CHECK(res[i - 1].first <= res[i].first);
      // truth by the Space's own distance function
      std::vector<std::pair<float, size_t>> all;
      size_t dd = d;
      for (size_t i = 0; i < n; i++) all.push_back({space.get_dist_func()(p, data.data() + d * i, &dd), i});
      std::sort(all.begin(), all.end());
      for (size_t i = 0; i < k; i++) CHECK(all[i].second == res[i].second);
    }
    // ---- searchKnnWithFilter_test: only ids divisible by 3
    PickDivisibleIds pick(3);
    for (size_t j = 0; j < nq; ++j) {
      auto res = alg_brute->searchKnnCloserFirst(query.data() + j * d, k, &pick);
      CHECK(res.size() == k);
      for (auto &r : res) CHECK(r.second % 3 == 0);
    }
    bool threw = false;
    try { for (size_t i = 0; i < 3 * n; i++) alg_brute->addPoint(data.data(), 1000 + i); } catch (std::runtime_error &) { threw = true; }
    CHECK(threw);  // capacity error like bruteforce.h:73-75
    delete alg_brute;
  }
  if (argc > 2) {  // ---- reference-built index file
    size_t dim = (size_t)atoi(argv[2]);
    hnswlib::L2Space space(dim);
    hnswlib::HierarchicalNSW<float> index(&space, std::string(argv[1]));
    size_t n = index.cur_element_count, nq = 100, k = 10;
    hnswlib::BruteforceSearch<float> bf(&space, n);
    for (size_t i = 0; i < n; i++) bf.addPoint(index.getDataByInternalId((hnswlib::tableint)i), index.getExternalLabel((hnswlib::tableint)i));
    std::mt19937 rng(3);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    std::vector<float> q(nq * dim);
    for (auto &v : q) v = u(rng);
    index.setEf(100);
    size_t hit = 0;
    std::vector<hnswlib::labeltype> lab(nq * k); std::vector<float> dist(nq * k);
    index.searchKnnBatch(q.data(), nq, k, lab.data(), dist.data());
    for (size_t j = 0; j < nq; j++) {
      auto truth = bf.searchKnnCloserFirst(q.data() + j * dim, k);
      auto got = index.searchKnnCloserFirst(q.data() + j * dim, k);
      CHECK(got.size() == k);
      std::set<hnswlib::labeltype> ts;
      for (auto &t : truth) ts.insert(t.second);
      for (size_t i = 0; i < k; i++) { hit += ts.count(got[i].second); CHECK(got[i].second == lab[j * k + i]); }
    }
    double recall = (double)hit / (nq * k);
    printf("reference-built index: n=%zu recall@10 %.4f\n", n, recall);
    CHECK(recall > 0.95);
    std::string out = std::string(argv[1]) + ".resaved";
    index.saveIndex(out);
    hnswlib::HierarchicalNSW<float> again(&space, out);
    std::vector<hnswlib::labeltype> lab2(nq * k); std::vector<float> dist2(nq * k);
    again.setEf(100);
    again.searchKnnBatch(q.data(), nq, k, lab2.data(), dist2.data());
    CHECK(lab == lab2 && dist == dist2);
    bool threw = false;
    try { hnswlib::HierarchicalNSW<float> bad(&space, std::string("/nonexistent/file.bin")); } catch (std::runtime_error &) { threw = true; }
    CHECK(threw);
    // ---- searchKnnWithFilter_test.cpp:33-88 on the HNSW index: only labels divisible by 3 come back, PickNothing
    // gives an empty result; then markDelete (hnswalg.h:853-911): deleted labels never come back, unknown / repeated
    // labels throw like the reference, the marks survive saveIndex -> load, unmarkDelete restores the answers
    PickDivisibleIds pick3(3);
    struct PickNothing : hnswlib::BaseFilterFunctor { bool operator()(hnswlib::labeltype) override { return false; } } none;
    for (size_t j = 0; j < 20; j++) {
      auto res = index.searchKnnCloserFirst(q.data() + j * dim, k, &pick3);
      CHECK(res.size() == k);
      for (auto &r : res) CHECK(r.second % 3 == 0);
      for (size_t i = 1; i < res.size(); i++) CHECK(res[i - 1].first <= res[i].first);
      CHECK(index.searchKnn(q.data() + j * dim, k, &none).empty());
    }
    std::set<hnswlib::labeltype> gone;
    for (size_t j = 0; j < nq; j++) for (size_t i = 0; i < 3; i++) gone.insert(lab[j * k + i]);
    for (auto l : gone) index.markDelete(l);
    CHECK(index.getDeletedCount() == gone.size());
    threw = false; try { index.markDelete(*gone.begin()); } catch (std::runtime_error &) { threw = true; } CHECK(threw);
    threw = false; try { index.markDelete(987654321); } catch (std::runtime_error &) { threw = true; } CHECK(threw);
    std::vector<hnswlib::labeltype> labd(nq * k); std::vector<float> distd(nq * k);
    index.searchKnnBatch(q.data(), nq, k, labd.data(), distd.data(), &pick3);   // deleted marks AND a filter
    for (size_t i = 0; i < nq * k; i++) CHECK(labd[i] % 3 == 0 && !gone.count(labd[i]));
    index.searchKnnBatch(q.data(), nq, k, labd.data(), distd.data());           // deleted marks only
    for (size_t i = 0; i < nq * k; i++) CHECK(!gone.count(labd[i]));
    for (size_t j = 0; j < nq; j++) {  // what was 4th is first now
      CHECK(labd[j * k] == lab[j * k + 3] || gone.count(lab[j * k + 3]));
    }
    std::string outd = std::string(argv[1]) + ".deleted";
    index.saveIndex(outd);
    hnswlib::HierarchicalNSW<float> marked(&space, outd);
    CHECK(marked.getDeletedCount() == gone.size());
    marked.setEf(100);
    std::vector<hnswlib::labeltype> labm(nq * k); std::vector<float> distm(nq * k);
    marked.searchKnnBatch(q.data(), nq, k, labm.data(), distm.data());
    CHECK(labm == labd && distm == distd);
    {  // queries + answers for the python driver, which asks the unmodified reference the same question
      FILE *f = fopen((std::string(argv[1]) + ".deleted.q").c_str(), "wb");
      fwrite(q.data(), sizeof(float), q.size(), f); fclose(f);
      std::vector<unsigned long long> l64(labd.begin(), labd.end());
      f = fopen((std::string(argv[1]) + ".deleted.labels").c_str(), "wb");
      fwrite(l64.data(), 8, l64.size(), f); fclose(f);
    }
    for (auto l : gone) index.unmarkDelete(l);
    threw = false; try { index.unmarkDelete(*gone.begin()); } catch (std::runtime_error &) { threw = true; } CHECK(threw);
    index.searchKnnBatch(q.data(), nq, k, labd.data(), distd.data());
    CHECK(labd == lab && distd == dist);
  }
  {  // ---- construction through the reference API (hnswlib/examples/cpp/example_search.cpp): addPoint loop from
     // several threads, searchKnn of every stored point finds itself, capacity error, saveIndex -> load ctor
    int dim = 16; size_t n = 3000;
    std::mt19937 rng(47);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    std::vector<float> data(n * dim);
    for (auto &v : data) v = u(rng);
    hnswlib::L2Space space(dim);
    hnswlib::HierarchicalNSW<float> *alg = new hnswlib::HierarchicalNSW<float>(&space, n, 16, 200);
#pragma omp parallel for
    for (long i = 0; i < (long)n; i++) alg->addPoint(data.data() + (size_t)i * dim, (hnswlib::labeltype)(i + 10));
    bool threw = false;
    try { alg->addPoint(data.data(), 999999); } catch (std::runtime_error &) { threw = true; }
    CHECK(threw);  // hnswalg.h:1178
    threw = false;
    try { alg->resizeIndex(n - 1); } catch (std::runtime_error &) { threw = true; }
    CHECK(threw);  // :642
    alg->resizeIndex(n + 1);
    alg->addPoint(data.data(), 999999);  // fits now; same vector as label 10
    CHECK(alg->cur_element_count == n + 1);
    uint64_t info[8];
    alg->build(info);
    CHECK(alg->cur_element_count == n + 1 && alg->maxlevel_ >= 1 && alg->M_ == 16 && alg->maxM0_ == 32);
    float correct = 0;
    for (size_t i = 0; i < n; i++) {
      auto r = alg->searchKnn(data.data() + i * dim, 1);
      CHECK(r.size() == 1);
      if (r.top().second == alg->getExternalLabel((hnswlib::tableint)i) || r.top().first == 0.f) correct++;
    }
    printf("GPU-built index: n=%zu self-recall %.4f batches %llu\n", n, correct / n, (unsigned long long)info[3]);
    CHECK(correct / n > 0.99);  // example_search.cpp prints this recall
    if (argc > 1) {
      std::string out = std::string(argv[1]) + ".gpubuilt";
      alg->saveIndex(out);
      hnswlib::HierarchicalNSW<float> again(&space, out);
      CHECK(again.cur_element_count == n + 1 && again.maxlevel_ == alg->maxlevel_ && again.enterpoint_node_ == alg->enterpoint_node_);
      std::vector<hnswlib::labeltype> a(10 * 5), b(10 * 5); std::vector<float> da(50), db(50);
      alg->setEf(50); again.setEf(50);
      alg->searchKnnBatch(data.data(), 10, 5, a.data(), da.data());
      again.searchKnnBatch(data.data(), 10, 5, b.data(), db.data());
      CHECK(a == b && da == db);
    }
    delete alg;
  }
  printf("hnswlib shim ok\n");
  return 0;
}

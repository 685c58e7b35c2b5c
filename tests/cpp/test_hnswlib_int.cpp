// HierarchicalNSW<int> over L2SpaceI through include/b200ann/hnswlib.hpp (hnswlib/hnswlib/space_l2.h:255-323: rows of
// unsigned bytes, int squared-L2 distances; the same hnswalg.h template instantiated with dist_t = int).
//   test_hnswlib_int                      build on the GPU, recall against BruteforceSearch<int>, save -> load round trip,
//                                         filter functor, markDelete, epsilon search
//   test_hnswlib_int <ref.bin> <dim> <expect.bin>
//                                         load an index file the UNMODIFIED reference wrote; searchKnn must return the
//                                         labels / int distances in expect.bin (nq, k, ef, queries, labels, dists)
#include <b200ann/hnswlib.hpp>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

static int against_reference(const char *path, size_t dim, const char *expect) {
  hnswlib::L2SpaceI space(dim);
  hnswlib::HierarchicalNSW<int> ix(&space, std::string(path));
  FILE *f = fopen(expect, "rb");
  REQUIRE(f);
  unsigned long long hdr[3];
  REQUIRE(fread(hdr, 8, 3, f) == 3);
  const size_t nq = hdr[0], k = hdr[1], ef = hdr[2];
  std::vector<unsigned char> q(nq * dim);
  std::vector<unsigned long long> lab(nq * k);
  std::vector<int> d(nq * k);
  REQUIRE(fread(q.data(), 1, q.size(), f) == q.size());
  REQUIRE(fread(lab.data(), 8, lab.size(), f) == lab.size());
  REQUIRE(fread(d.data(), 4, d.size(), f) == d.size());
  fclose(f);
  ix.setEf(ef);
  std::vector<hnswlib::labeltype> gl(nq * k);
  std::vector<int> gd(nq * k);
  ix.searchKnnBatch(q.data(), nq, k, gl.data(), gd.data());
  size_t label_diff = 0;
  for (size_t i = 0; i < nq * k; i++) {
    REQUIRE(gd[i] == d[i]);                     // int distances: exact
    label_diff += gl[i] != lab[i];              // labels: equal except among equal distances
  }
  REQUIRE(label_diff * 200 <= nq * k);
  for (size_t i = 0; i < nq; i += 7) {          // one query at a time gives the same as the batch
    auto r = ix.searchKnnCloserFirst(&q[i * dim], k);
    REQUIRE(r.size() == k);
    for (size_t j = 0; j < k; j++) REQUIRE(r[j].first == gd[i * k + j] && r[j].second == gl[i * k + j]);
  }
  printf("hnswlib<int> vs reference file ok (%zu label ties)\n", label_diff);
  return 0;
}

int main(int argc, char **argv) {
  try {
    if (argc == 4) return against_reference(argv[1], (size_t)atol(argv[2]), argv[3]);
    const size_t n = 8000, dim = 128, nq = 200, k = 10;
    // clustered bytes (64 centres + noise): iid bytes in 128 dimensions have no neighbourhood structure to recall
    std::mt19937 rng(11);
    std::gamma_distribution<float> g(1.2f, 24.f);
    std::uniform_int_distribution<int> noise(-24, 24);
    std::vector<unsigned char> centres(64 * dim), base(n * dim), q(nq * dim);
    for (auto &v : centres) v = (unsigned char)std::min(255.f, g(rng));
    auto draw = [&](unsigned char *row, size_t c) {
      for (size_t j = 0; j < dim; j++) row[j] = (unsigned char)std::max(0, std::min(255, (int)centres[c * dim + j] + noise(rng)));
    };
    for (size_t i = 0; i < n; i++) draw(&base[i * dim], i % 64);
    for (size_t i = 0; i < nq; i++) draw(&q[i * dim], (i * 7) % 64);
    hnswlib::L2SpaceI space(dim);
    hnswlib::HierarchicalNSW<int> hnsw(&space, n, 16, 100);
    hnswlib::BruteforceSearch<int> brute(&space, n);
    for (size_t i = 0; i < n; i++) { hnsw.addPoint(&base[i * dim], 1000 + i); brute.addPoint(&base[i * dim], 1000 + i); }
    hnsw.setEf(100);
    size_t hit = 0;
    for (size_t i = 0; i < nq; i++) {
      auto a = hnsw.searchKnnCloserFirst(&q[i * dim], k);
      auto b = brute.searchKnnCloserFirst(&q[i * dim], k);
      REQUIRE(a.size() == k && b.size() == k);
      for (auto &x : a) for (auto &y : b) hit += x.second == y.second;
      for (size_t j = 1; j < k; j++) REQUIRE(a[j - 1].first <= a[j].first);
    }
    REQUIRE((double)hit / (nq * k) > 0.9);
    // every element is its own nearest neighbour at distance 0 (a sample)
    for (size_t i = 0; i < n; i += 41) {
      auto r = hnsw.searchKnnCloserFirst(&base[i * dim], 1);
      REQUIRE(r.size() == 1 && r[0].first == 0);
    }
    // save -> load: identical answers
    std::string path = std::string(argc > 1 ? argv[1] : "/tmp") + "/hnsw_int.bin";
    hnsw.saveIndex(path);
    hnswlib::HierarchicalNSW<int> back(&space, path);
    back.setEf(100);
    REQUIRE(back.cur_element_count == n);
    for (size_t i = 0; i < nq; i += 5) {
      auto a = hnsw.searchKnnCloserFirst(&q[i * dim], k), b = back.searchKnnCloserFirst(&q[i * dim], k);
      REQUIRE(a == b);
    }
    std::vector<unsigned char> v = back.getDataByLabel<unsigned char>(1000 + 17);
    REQUIRE(v.size() == dim && std::equal(v.begin(), v.end(), &base[17 * dim]));
    // filter functor and markDelete
    struct Odd : hnswlib::BaseFilterFunctor { bool operator()(hnswlib::labeltype l) override { return l % 2 == 1; } } odd;
    auto r = hnsw.searchKnnCloserFirst(&q[0], k, &odd);
    REQUIRE(r.size() == k);
    for (auto &x : r) REQUIRE(x.second % 2 == 1);
    auto first = hnsw.searchKnnCloserFirst(&q[0], 1);
    hnsw.markDelete(first[0].second);
    auto second = hnsw.searchKnnCloserFirst(&q[0], 1);
    REQUIRE(second[0].second != first[0].second && second[0].first >= first[0].first);
    hnsw.unmarkDelete(first[0].second);
    // range search: int distances, everything within epsilon
    hnswlib::EpsilonSearchStopCondition<int> stop((float)(first[0].first * 2), 10, 200);
    auto rs = hnsw.searchStopConditionClosest(&q[0], stop);
    REQUIRE(!rs.empty() && rs[0].first == first[0].first);
    for (auto &x : rs) REQUIRE(x.first <= first[0].first * 2);
    // appended points are linked into the live graph
    std::vector<unsigned char> extra(50 * dim);
    for (size_t i = 0; i < 50; i++) draw(&extra[i * dim], i % 64);
    hnswlib::HierarchicalNSW<int> grow(&space, n + 50, 16, 100);
    for (size_t i = 0; i < n; i++) grow.addPoint(&base[i * dim], i);
    grow.setEf(100);
    grow.searchKnnCloserFirst(&q[0], 1);   // builds
    for (size_t i = 0; i < 50; i++) grow.addPoint(&extra[i * dim], n + i);
    size_t self = 0;
    for (size_t i = 0; i < 50; i++) {
      auto e = grow.searchKnnCloserFirst(&extra[i * dim], 1);
      self += e[0].second == n + i && e[0].first == 0;
    }
    REQUIRE(self >= 48);
    printf("hnswlib<int> shim ok\n");
    return 0;
  } catch (std::exception &e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 1;
  }
}

// Range search through include/b200ann/hnswlib.hpp: HierarchicalNSW::searchStopConditionClosest with an
// EpsilonSearchStopCondition, driven the way the reference's own test drives it
// (hnswlib/tests/cpp/epsilon_search_test.cpp: 10000 x 16 uniform points, epsilon^2 = 1, min 2000 / max 10000 candidates;
// every result inside the ball, recall of the ball > 0.95 against brute force; then every element finds itself at
// distance 0 with a tiny epsilon). The batched form must return what the one-query calls return.
#include <b200ann/hnswlib.hpp>
#include <cstdio>
#include <random>
#include <unordered_set>
#include <vector>

#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

int main() {
  try {
    const size_t n = 10000, dim = 16, nq = 100;
    const float eps2 = 1.0f;
    const size_t max_cand = n, min_cand = 2000;
    std::mt19937 rng(47);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    std::vector<float> data(n * dim), q(nq * dim);
    for (auto &v : data) v = u(rng);
    for (auto &v : q) v = u(rng);
    hnswlib::L2Space space(dim);
    hnswlib::HierarchicalNSW<float> hnsw(&space, n, 16, 200);
    hnswlib::BruteforceSearch<float> brute(&space, n);
    for (size_t i = 0; i < n; i++) { hnsw.addPoint(&data[i * dim], i); brute.addPoint(&data[i * dim], i); }

    std::vector<hnswlib::labeltype> bl(nq * max_cand);
    std::vector<float> bd(nq * max_cand);
    std::vector<unsigned> bc(nq);
    hnsw.searchEpsilonBatch(q.data(), nq, eps2, min_cand, max_cand, bl.data(), bd.data(), bc.data(), nullptr, true);
    size_t in_ball = 0;
    for (size_t i = 0; i < nq; i++) {
      hnswlib::EpsilonSearchStopCondition<float> stop(eps2, min_cand, max_cand);
      auto res = hnsw.searchStopConditionClosest(&q[i * dim], stop);
      REQUIRE(res.size() == bc[i]);
      std::unordered_set<hnswlib::labeltype> got;
      for (size_t j = 0; j < res.size(); j++) {
        REQUIRE(res[j].first >= 0 && res[j].first <= eps2);
        REQUIRE(j == 0 || res[j - 1].first <= res[j].first);
        REQUIRE(res[j].second == bl[i * max_cand + j] && res[j].first == bd[i * max_cand + j]);
        got.insert(res[j].second);
      }
      auto all = brute.searchKnn(&q[i * dim], n);
      size_t truth = 0, hit = 0;
      while (!all.empty()) {
        if (all.top().first < eps2) { truth++; hit += got.count(all.top().second); }
        all.pop();
      }
      in_ball += truth;
      if (truth == 0) { REQUIRE(hit == 0); continue; }
      REQUIRE((float)hit / truth > 0.95f);
    }
    REQUIRE(in_ball > 0);
    printf("ball recall ok (%zu points in %zu balls)\n", in_ball, nq);

    // every element finds itself: batched over all n, plus a sample of one-query calls
    const float eps_small = 0.0001f;
    const size_t min_small = 500, max_small = 600;
    std::vector<hnswlib::labeltype> sl(n * max_small);
    std::vector<float> sd(n * max_small);
    std::vector<unsigned> sc(n);
    hnsw.searchEpsilonBatch(data.data(), n, eps_small, min_small, max_small, sl.data(), sd.data(), sc.data());
    for (size_t i = 0; i < n; i++) REQUIRE(sc[i] >= 1 && sd[i * max_small] == 0.f && sl[i * max_small] == i);
    for (size_t i = 0; i < n; i += 97) {
      hnswlib::EpsilonSearchStopCondition<float> stop(eps_small, min_small, max_cand);
      auto res = hnsw.searchStopConditionClosest(hnsw.getDataByInternalId((hnswlib::tableint)i), stop);
      REQUIRE(!res.empty() && res[0].first == 0.f);
    }
    // a filter is honoured; a foreign stop condition is refused
    struct Even : hnswlib::BaseFilterFunctor { bool operator()(hnswlib::labeltype l) override { return l % 2 == 0; } } even;
    hnswlib::EpsilonSearchStopCondition<float> stop(eps2, 200, 1000);
    auto res = hnsw.searchStopConditionClosest(&q[0], stop, &even);
    REQUIRE(!res.empty());
    for (auto &pr : res) REQUIRE(pr.second % 2 == 0 && pr.first <= eps2);
    bool threw = false;
    try { hnswlib::EpsilonSearchStopCondition<float> bad(1.f, 10, 5); } catch (std::runtime_error &) { threw = true; }
    REQUIRE(threw);
    printf("epsilon search ok\n");
    return 0;
  } catch (std::exception &e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 1;
  }
}

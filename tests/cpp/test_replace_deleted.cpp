// allow_replace_deleted / addPoint(..., replace_deleted = true) through include/b200ann/hnswlib.hpp, following the
// reference's contract (hnswlib/hnswlib/hnswalg.h:954-992, :1160-1167) and its test tests/python/bindings_test_replace.py:
// deleted slots are reused, the index does not grow, replaced labels disappear, new labels are found.
#include <b200ann/hnswlib.hpp>
#include <algorithm>
#include <cstdio>
#include <random>
#include <vector>

#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

template <typename F>
static bool throws_with(F f, const char *what) {
  try { f(); } catch (const std::runtime_error &e) { return std::string(e.what()).find(what) != std::string::npos; }
  return false;
}

int main() {
  try {
    const size_t n = 2000, dim = 16, ndel = 100;
    std::mt19937 rng(3);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    std::vector<float> base(n * dim), extra(2 * ndel * dim);
    for (auto &v : base) v = u(rng);
    for (auto &v : extra) v = u(rng);
    hnswlib::L2Space space(dim);
    {  // disabled in the constructor -> the reference's exception
      hnswlib::HierarchicalNSW<float> off(&space, n, 16, 100);
      REQUIRE(throws_with([&] { off.addPoint(&base[0], 0, true); }, "Replacement of deleted elements is disabled in constructor"));
    }
    hnswlib::HierarchicalNSW<float> ix(&space, n, 16, 100, 100, true);
    for (size_t i = 0; i < n; i++) ix.addPoint(&base[i * dim], i);
    ix.setEf(100);
    for (size_t i = 0; i < ndel; i++) ix.markDelete(i * 3);
    REQUIRE(ix.getDeletedCount() == ndel && ix.cur_element_count == n);
    // an update of a deleted label is refused while replacement is enabled (:1163-1167)
    REQUIRE(throws_with([&] { ix.addPoint(&base[0], 0); }, "Can't use addPoint to update deleted elements"));
    // the index is full: a plain add throws, a replacing add reuses the deleted slots
    REQUIRE(throws_with([&] { ix.addPoint(&extra[0], 100000); }, "exceeds the specified limit"));
    for (size_t i = 0; i < ndel; i++) ix.addPoint(&extra[i * dim], 100000 + i, true);
    REQUIRE(ix.getDeletedCount() == 0 && ix.cur_element_count == n);
    // replaced labels are gone, new labels are their own nearest neighbour at distance 0
    REQUIRE(throws_with([&] { ix.markDelete(0); }, "Label not found"));
    REQUIRE(throws_with([&] { ix.getDataByLabel<float>(0); }, "Label not found"));
    {
      std::vector<float> v = ix.getDataByLabel<float>(100000 + 7);
      REQUIRE(v.size() == dim && std::equal(v.begin(), v.end(), &extra[7 * dim]));
    }
    size_t miss_new = 0, miss_old = 0;  // graph search is approximate: allow a stray miss, never a replaced label
    for (size_t i = 0; i < ndel; i++) {
      auto r = ix.searchKnnCloserFirst(&extra[i * dim], 1);
      REQUIRE(r.size() == 1);
      if (!(r[0].second == 100000 + i && r[0].first == 0.f)) miss_new++;
    }
    for (size_t i = 0; i < n; i++) {  // surviving originals still answer for themselves, replaced ones never appear
      auto r = ix.searchKnnCloserFirst(&base[i * dim], 1);
      REQUIRE(r.size() == 1);
      bool replaced = i % 3 == 0 && i / 3 < ndel;
      if (replaced) REQUIRE(r[0].second != i);
      else if (r[0].second != i) miss_old++;
      REQUIRE(r[0].second >= 100000 || !(r[0].second % 3 == 0 && r[0].second / 3 < ndel));
    }
    REQUIRE(miss_new <= 1 && miss_old <= 4);
    // no vacant slot left: a replacing add is a plain add (:976-978) and hits the capacity limit
    REQUIRE(throws_with([&] { ix.addPoint(&extra[ndel * dim], 200000, true); }, "exceeds the specified limit"));
    ix.markDelete(5);
    ix.addPoint(&extra[ndel * dim], 200000, true);
    auto r = ix.searchKnnCloserFirst(&extra[ndel * dim], 1);
    REQUIRE(r[0].second == 200000);
    printf("replace_deleted ok\n");
    return 0;
  } catch (const std::exception &e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 3;
  }
}

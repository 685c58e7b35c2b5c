// Drives hnswlib::BruteforceSearch<int> over hnswlib::L2SpaceI through include/b200ann/hnswlib.hpp the way a user of
// the reference drives the reference (bruteforce.h, space_l2.h:294-323), and checks every answer against a scalar
// restatement of L2SqrI in this file: labels and int distances must be identical, ties to the earlier insertion.
// usage: test_l2spacei_shim <tmpdir>
#include <b200ann/hnswlib.hpp>
#include <algorithm>
#include <cstdio>
#include <random>
#include <vector>

static int l2i(const unsigned char *a, const unsigned char *b, size_t d) {
  int s = 0;
  for (size_t i = 0; i < d; i++) { int t = (int)a[i] - (int)b[i]; s += t * t; }
  return s;
}

int main(int argc, char **argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s tmpdir\n", argv[0]); return 2; }
  try {
    const size_t n = 3000, dim = 50, nq = 40, k = 10;
    std::mt19937 rng(7);
    std::vector<unsigned char> base(n * dim), q(nq * dim);
    for (auto &v : base) v = (unsigned char)(rng() % 7);  // few values: many exact ties
    for (auto &v : q) v = (unsigned char)(rng() % 7);
    hnswlib::L2SpaceI space(dim);
    hnswlib::BruteforceSearch<int> bf(&space, n);
    for (size_t i = 0; i < n; i++) bf.addPoint(&base[i * dim], 1000 + i);
    auto check = [&](hnswlib::BruteforceSearch<int> &ix, size_t live) {
      for (size_t qi = 0; qi < nq; qi++) {
        std::vector<std::pair<int, size_t>> truth;
        for (size_t i = 0; i < live; i++) truth.emplace_back(l2i(&q[qi * dim], &base[i * dim], dim), 1000 + i);
        std::sort(truth.begin(), truth.end());
        auto r = ix.searchKnnCloserFirst(&q[qi * dim], k);
        if (r.size() != k) return 1;
        for (size_t j = 0; j < k; j++)
          if (r[j].first != truth[j].first || r[j].second != truth[j].second) {
            fprintf(stderr, "query %zu rank %zu: got (%d,%zu) want (%d,%zu)\n", qi, j, r[j].first, (size_t)r[j].second,
                    truth[j].first, truth[j].second);
            return 1;
          }
      }
      return 0;
    };
    if (check(bf, n)) return 1;
    // file round trip in the reference's format
    std::string path = std::string(argv[1]) + "/bf_u8.bin";
    bf.saveIndex(path);
    hnswlib::BruteforceSearch<int> bf2(&space, 1);
    bf2.loadIndex(path, &space);
    if (bf2.cur_element_count != n || check(bf2, n)) return 1;
    // capacity error like bruteforce.h:73-75
    bool threw = false;
    try { bf.addPoint(&base[0], 999999); } catch (const std::runtime_error &) { threw = true; }
    if (!threw) return 1;
    // removePoint of the last element keeps the others in place
    bf.removePoint(1000 + n - 1);
    if (bf.cur_element_count != n - 1 || check(bf, n - 1)) return 1;
    printf("l2spacei shim ok\n");
    return 0;
  } catch (const std::exception &e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 3;
  }
}

// The flow of the reference's updates test (hnswlib/tests/cpp/updates_test.cpp:177-260, "update" mode) through
// include/b200ann/hnswlib.hpp: an index is built from a dummy batch, then EVERY label is added again with the vector of
// the final batch - addPoint on an existing label = updatePoint (hnswalg.h:1160-1174) - and an ef sweep must reach
// recall > 0.99 against the golden ground truth of the final batch (:149-172). Here the updates run in place on the
// GPU (b200ann_hnsw_update_batch) when the next search flushes them.
//   test_updates <dir>    dir holds config.txt, batch_dummy_00.bin, batch_final.bin, queries.bin, gt.bin
#include <b200ann/hnswlib.hpp>
#include <chrono>
#include <cstdio>
#include <fstream>
#include <unordered_set>
#include <vector>

#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

template <typename T>
static std::vector<T> load(const std::string &path, size_t count) {
  std::vector<T> v(count);
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("cannot open " + path);
  f.read((char *)v.data(), (std::streamsize)(count * sizeof(T)));
  if (!f) throw std::runtime_error("short file " + path);
  return v;
}

int main(int argc, char **argv) {
  try {
    REQUIRE(argc == 2);
    { hnswlib::L2Space s4(4); hnswlib::HierarchicalNSW<float> probe(&s4, 1, 4, 10); }   // no GPU: fails here, loudly
    const std::string dir = std::string(argv[1]) + "/";
    int N, mult, NQ, d, K;
    {
      std::ifstream c(dir + "config.txt");
      REQUIRE((bool)(c >> N >> mult >> NQ >> d >> K));
    }
    std::vector<float> dummy = load<float>(dir + "batch_dummy_00.bin", (size_t)N * d);
    std::vector<float> fin = load<float>(dir + "batch_final.bin", (size_t)N * d);
    std::vector<float> q = load<float>(dir + "queries.bin", (size_t)NQ * d);
    std::vector<int> gt = load<int>(dir + "gt.bin", (size_t)NQ * K);
    hnswlib::L2Space space(d);
    hnswlib::HierarchicalNSW<float> alg(&space, N + 1, 16, 200);
    for (int i = 0; i < N; i++) alg.addPoint(&dummy[(size_t)i * d], (size_t)i);
    alg.setEf(10);
    REQUIRE(alg.searchKnn(&q[0], K).size() == (size_t)K);   // the graph of the dummy batch exists on the GPU now
    uint64_t info[8];
    auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < N; i++) alg.addPoint(&fin[(size_t)i * d], (size_t)i);   // every element moves
    alg.build(info);                                                            // flushes the updates in place
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    REQUIRE(alg.cur_element_count == (size_t)N);
    printf("updated %d elements in place in %.3f s (%llu chunks, %llu neighbour lists rewritten)\n", N, sec,
           (unsigned long long)info[3], (unsigned long long)info[4]);
    std::vector<size_t> efs = {1};
    for (int i = K; i < 30; i++) efs.push_back(i);
    for (int i = 30; i < 400; i += 10) efs.push_back(i);
    bool passed = false;
    std::vector<hnswlib::labeltype> lab((size_t)NQ * K);
    std::vector<float> dist((size_t)NQ * K);
    for (size_t ef : efs) {
      alg.setEf(ef);
      alg.searchKnnBatch(q.data(), NQ, K, lab.data(), dist.data());
      size_t correct = 0;
      for (int i = 0; i < NQ; i++) {
        std::unordered_set<int> truth(gt.begin() + (size_t)i * K, gt.begin() + (size_t)(i + 1) * K);
        for (int j = 0; j < K; j++) correct += truth.count((int)lab[(size_t)i * K + j]);
      }
      const double recall = (double)correct / ((double)NQ * K);
      if (recall > 0.99) { printf("ef %zu: recall %.4f\n", ef, recall); passed = true; break; }
    }
    REQUIRE(passed);
    // the vectors the index holds are the final batch
    std::vector<float> v = alg.getDataByLabel<float>(12345);
    REQUIRE(std::equal(v.begin(), v.end(), &fin[(size_t)12345 * d]));
    printf("updates ok\n");
    return 0;
  } catch (std::exception &e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 1;
  }
}

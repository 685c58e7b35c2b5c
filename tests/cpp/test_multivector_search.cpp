// The flow of the reference's hnswlib/tests/cpp/multivector_search_test.cpp through include/b200ann/hnswlib.hpp: 1000
// elements of 16 dimensions, each tagged with one of 50 document ids (MultiVectorL2Space<unsigned>), searched with a
// MultiVectorSearchStopCondition(space, num_docs = 10, ef_collection = 15): every answer covers exactly num_docs
// documents (:86), the overall recall against brute force is > 0.95 (:104) and every element finds itself (:120).
#include <b200ann/hnswlib.hpp>
#include <cstdio>
#include <random>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)
typedef unsigned int docidtype;

int main() {
  try {
    const int dim = 16, n = 1000, M = 16, efc = 200, nq = 100, num_docs = 10, ef_collection = 15;
    hnswlib::MultiVectorL2Space<docidtype> space(dim);
    hnswlib::HierarchicalNSW<float> hnsw(&space, n, M, efc);
    hnswlib::BruteforceSearch<float> brute(&space, n);
    std::mt19937 rng(47);
    std::uniform_real_distribution<float> real(0.f, 1.f);
    std::uniform_int_distribution<docidtype> docid(0, 49);
    const size_t dps = space.get_data_size();
    REQUIRE(dps == dim * sizeof(float) + sizeof(docidtype));
    std::vector<char> data(dps * n);
    std::unordered_map<hnswlib::labeltype, docidtype> doc_of;
    for (int i = 0; i < n; i++) {
      char *row = data.data() + i * dps;
      for (int j = 0; j < dim; j++) { float v = real(rng); memcpy(row + j * sizeof(float), &v, sizeof v); }
      space.set_doc_id(row, docid(rng));
      REQUIRE(space.get_doc_id(row) <= 49);
      hnsw.addPoint(row, i);
      brute.addPoint(row, i);
      doc_of[i] = space.get_doc_id(row);
    }
    double correct = 0, total = 0;
    for (int i = 0; i < nq; i++) {
      std::vector<float> q(dim);
      for (auto &v : q) v = real(rng);
      hnswlib::MultiVectorSearchStopCondition<docidtype, float> stop(space, num_docs, ef_collection);
      auto res = hnsw.searchStopConditionClosest(q.data(), stop);
      std::unordered_set<docidtype> docs;
      std::unordered_set<hnswlib::labeltype> labels;
      for (size_t j = 0; j < res.size(); j++) {
        REQUIRE(j == 0 || res[j - 1].first <= res[j].first);
        labels.insert(res[j].second);
        docs.insert(doc_of[res[j].second]);
      }
      REQUIRE(docs.size() == (size_t)num_docs);
      auto gt = brute.searchKnnCloserFirst(q.data(), n);
      std::unordered_set<docidtype> gt_docs;
      for (size_t j = 0; j < gt.size() && gt_docs.size() < (size_t)num_docs; j++) {
        correct += labels.count(gt[j].second);
        gt_docs.insert(doc_of[gt[j].second]);
        total += 1;
      }
    }
    printf("random elements search recall : %.4f\n", correct / total);
    REQUIRE(correct / total > 0.95);
    int self = 0;
    for (int i = 0; i < n; i++) {
      hnswlib::MultiVectorSearchStopCondition<docidtype, float> stop(space, num_docs, ef_collection);
      auto res = hnsw.searchStopConditionClosest(data.data() + i * dps, stop);
      self += !res.empty() && res[0].second == (hnswlib::labeltype)i;
    }
    printf("same elements search recall : %.4f\n", (double)self / n);
    REQUIRE((double)self / n > 0.99);
    // the document id travels with the row: file round trip, getDataByLabel
    hnsw.saveIndex("/tmp/b200ann_mv_test.bin");
    hnswlib::HierarchicalNSW<float> back(&space, std::string("/tmp/b200ann_mv_test.bin"));
    std::vector<float> row = back.getDataByLabel<float>(7);
    REQUIRE(row.size() == (size_t)dim + 1 && space.get_doc_id(row.data()) == doc_of[7]);
    hnswlib::MultiVectorSearchStopCondition<docidtype, float> stop(space, 3, 3);
    auto a = hnsw.searchStopConditionClosest(data.data(), stop), b = back.searchStopConditionClosest(data.data(), stop);
    REQUIRE(a == b && !a.empty());
    printf("multivector search ok\n");
    return 0;
  } catch (std::exception &e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 1;
  }
}

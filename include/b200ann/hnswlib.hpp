// Host-side C++ mirror of the reference's hnswlib API for the hot path, over the b200ann C ABI.
// Same names, signatures, argument meaning and error behaviour as
//   hnswlib/hnswlib/hnswlib.h:125-221    labeltype, BaseFilterFunctor, DISTFUNC, SpaceInterface, AlgorithmInterface
//   hnswlib/hnswlib/space_l2.h:208-253   L2Space            hnswlib/hnswlib/space_ip.h:342-398  InnerProductSpace
//   hnswlib/hnswlib/bruteforce.h:10-172  BruteforceSearch   hnswlib/hnswlib/hnswalg.h:18-1411   HierarchicalNSW
// Header-only; link with -lb200ann. No CPU fallback on the search / build path: the distance functions a
// Space hands out exist only because SpaceInterface promises a callable (the index identifies the
// metric by the Space's type; a device kernel cannot call a host function pointer, so user-defined
// spaces are rejected with std::runtime_error).
//
// Added batched entry points in the [nq,k] row shape of python_bindings/bindings.cpp:612-695:
//   searchKnnBatch(queries, nq, k, labels[nq*k], dists[nq*k])   (ascending distance; throws if a row
//   has fewer than k results, bindings.cpp:646-648)
#ifndef B200ANN_HNSWLIB_HPP
#define B200ANN_HNSWLIB_HPP
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <mutex>
#include <queue>
#include <stdexcept>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>
#include "../b200ann.h"

namespace hnswlib {

typedef size_t labeltype;
typedef unsigned int tableint;
typedef unsigned int linklistsizeint;

class BaseFilterFunctor {
 public:
  virtual bool operator()(hnswlib::labeltype) { return true; }
  virtual ~BaseFilterFunctor() {}
};

// hnswlib.h:134-150: the stop-condition interface. The walk runs on the GPU, which knows ONE condition, the epsilon
// range search below; the virtual hooks are kept so that reference code compiles and a host-side caller can still
// drive a condition by hand.
template <typename dist_t>
class BaseSearchStopCondition {
 public:
  virtual void add_point_to_result(labeltype label, const void *datapoint, dist_t dist) = 0;
  virtual void remove_point_from_result(labeltype label, const void *datapoint, dist_t dist) = 0;
  virtual bool should_stop_search(dist_t candidate_dist, dist_t lowerBound) = 0;
  virtual bool should_consider_candidate(dist_t candidate_dist, dist_t lowerBound) = 0;
  virtual bool should_remove_extra() = 0;
  virtual void filter_results(std::vector<std::pair<dist_t, labeltype>> &candidates) = 0;
  virtual ~BaseSearchStopCondition() {}
};

// stop_condition.h:218-275, same constructor and hooks; epsilon() / min_num_candidates() / max_num_candidates() are
// what HierarchicalNSW::searchStopConditionClosest hands to b200ann_hnsw_search_epsilon.
template <typename dist_t>
class EpsilonSearchStopCondition : public BaseSearchStopCondition<dist_t> {
  float epsilon_;
  size_t min_num_candidates_, max_num_candidates_, curr_num_items_;

 public:
  EpsilonSearchStopCondition(float epsilon, size_t min_num_candidates, size_t max_num_candidates)
      : epsilon_(epsilon), min_num_candidates_(min_num_candidates), max_num_candidates_(max_num_candidates), curr_num_items_(0) {
    if (min_num_candidates > max_num_candidates)  // an assert in the reference (:227)
      throw std::runtime_error("EpsilonSearchStopCondition: min_num_candidates > max_num_candidates");
  }
  float epsilon() const { return epsilon_; }
  size_t min_num_candidates() const { return min_num_candidates_; }
  size_t max_num_candidates() const { return max_num_candidates_; }
  void add_point_to_result(labeltype, const void *, dist_t) override { curr_num_items_ += 1; }
  void remove_point_from_result(labeltype, const void *, dist_t) override { curr_num_items_ -= 1; }
  bool should_stop_search(dist_t candidate_dist, dist_t lowerBound) override {
    if (candidate_dist > lowerBound && curr_num_items_ == max_num_candidates_) return true;
    return candidate_dist > epsilon_ && curr_num_items_ >= min_num_candidates_;
  }
  bool should_consider_candidate(dist_t candidate_dist, dist_t lowerBound) override {
    return curr_num_items_ < max_num_candidates_ || lowerBound > candidate_dist;
  }
  bool should_remove_extra() override { return curr_num_items_ > max_num_candidates_; }
  void filter_results(std::vector<std::pair<dist_t, labeltype>> &candidates) override {
    while (!candidates.empty() && candidates.back().first > epsilon_) candidates.pop_back();
    while (candidates.size() > max_num_candidates_) candidates.pop_back();
  }
};

template <typename MTYPE>
using DISTFUNC = MTYPE (*)(const void *, const void *, const void *);

template <typename MTYPE>
class SpaceInterface {
 public:
  virtual size_t get_data_size() = 0;
  virtual DISTFUNC<MTYPE> get_dist_func() = 0;
  virtual void *get_dist_func_param() = 0;
  virtual ~SpaceInterface() {}
};

namespace detail {
inline float l2_host(const void *a, const void *b, const void *dimp) {
  const float *x = (const float *)a, *y = (const float *)b;
  size_t d = *(const size_t *)dimp;
  float s = 0;
  for (size_t i = 0; i < d; i++) { float t = x[i] - y[i]; s += t * t; }
  return s;
}
inline float ip_host(const void *a, const void *b, const void *dimp) {
  const float *x = (const float *)a, *y = (const float *)b;
  size_t d = *(const size_t *)dimp;
  float s = 0;
  for (size_t i = 0; i < d; i++) s += x[i] * y[i];
  return 1.0f - s;
}
inline void check(b200ann_handle_t h, int rc, const char *what) {
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + (h ? b200ann_last_error(h) : "no CUDA device (b200ann has no CPU fallback)"));
}
struct Gpu {
  b200ann_handle_t h = nullptr;
  explicit Gpu(int device = 0) { check(nullptr, b200ann_create(device, &h), "b200ann_create"); }
  ~Gpu() { b200ann_destroy(h); }
  Gpu(const Gpu &) = delete;
  Gpu &operator=(const Gpu &) = delete;
};
}  // namespace detail

class L2Space : public SpaceInterface<float> {
  size_t data_size_, dim_;
 public:
  explicit L2Space(size_t dim) : data_size_(dim * sizeof(float)), dim_(dim) {}
  size_t get_data_size() override { return data_size_; }
  DISTFUNC<float> get_dist_func() override { return detail::l2_host; }
  void *get_dist_func_param() override { return &dim_; }
};

class InnerProductSpace : public SpaceInterface<float> {
  size_t data_size_, dim_;
 public:
  explicit InnerProductSpace(size_t dim) : data_size_(dim * sizeof(float)), dim_(dim) {}
  size_t get_data_size() override { return data_size_; }
  DISTFUNC<float> get_dist_func() override { return detail::ip_host; }
  void *get_dist_func_param() override { return &dim_; }
};

namespace detail {
inline int l2i_host(const void *a, const void *b, const void *dimp) {
  const unsigned char *x = (const unsigned char *)a, *y = (const unsigned char *)b;
  size_t d = *(const size_t *)dimp;
  int s = 0;
  for (size_t i = 0; i < d; i++) { int t = (int)x[i] - (int)y[i]; s += t * t; }
  return s;
}
}  // namespace detail

// hnswlib::L2SpaceI (space_l2.h:294-323): unsigned-byte vectors, int squared L2. Served on the GPU by
// BruteforceSearch<int> (b200ann_bf_knn_u8); HierarchicalNSW<int> is not built.
class L2SpaceI : public SpaceInterface<int> {
  size_t data_size_, dim_;
 public:
  explicit L2SpaceI(size_t dim) : data_size_(dim * sizeof(unsigned char)), dim_(dim) {}
  size_t get_data_size() override { return data_size_; }
  DISTFUNC<int> get_dist_func() override { return detail::l2i_host; }
  void *get_dist_func_param() override { return &dim_; }
};

// stop_condition.h:9-122: spaces whose elements carry a document id behind the vector (several vectors per document).
// get_data_size() = dim * 4 + sizeof(DOCIDTYPE); the distance reads the vector only. (The reference's
// MultiVectorInnerProductSpace never sets its dim_, :85-122; this one does.)
template <typename DOCIDTYPE>
class BaseMultiVectorSpace : public SpaceInterface<float> {
 public:
  virtual DOCIDTYPE get_doc_id(const void *datapoint) = 0;
  virtual void set_doc_id(void *datapoint, DOCIDTYPE doc_id) = 0;
  virtual int b200ann_metric() const = 0;
};
namespace detail {
inline float mv_unused_dist(const void *, const void *, const void *) {
  throw std::runtime_error("b200ann: distances are evaluated on the GPU; the host DISTFUNC of this space is not callable");
}
}  // namespace detail
template <typename DOCIDTYPE, int METRIC>
class MultiVectorSpaceT : public BaseMultiVectorSpace<DOCIDTYPE> {
  size_t data_size_, vector_size_, dim_;

 public:
  explicit MultiVectorSpaceT(size_t dim) : data_size_(dim * sizeof(float) + sizeof(DOCIDTYPE)), vector_size_(dim * sizeof(float)), dim_(dim) {
    static_assert(sizeof(DOCIDTYPE) == 4, "b200ann: document ids are 32-bit (rows stay 4-byte aligned)");
  }
  size_t get_data_size() override { return data_size_; }
  DISTFUNC<float> get_dist_func() override { return detail::mv_unused_dist; }
  void *get_dist_func_param() override { return &dim_; }
  DOCIDTYPE get_doc_id(const void *datapoint) override { DOCIDTYPE v; std::memcpy(&v, (const char *)datapoint + vector_size_, sizeof v); return v; }
  void set_doc_id(void *datapoint, DOCIDTYPE doc_id) override { std::memcpy((char *)datapoint + vector_size_, &doc_id, sizeof doc_id); }
  int b200ann_metric() const override { return METRIC; }
};
template <typename DOCIDTYPE> using MultiVectorL2Space = MultiVectorSpaceT<DOCIDTYPE, B200ANN_L2>;
template <typename DOCIDTYPE> using MultiVectorInnerProductSpace = MultiVectorSpaceT<DOCIDTYPE, B200ANN_IP>;

// stop_condition.h:146-215, same constructor; the walk itself runs on the GPU (b200ann_hnsw_search_multivector), which
// reads num_docs_to_search() / ef_collection() and the document ids of the index's rows. The virtual hooks count
// documents like the reference's for a host-side caller.
template <typename DOCIDTYPE, typename dist_t>
class MultiVectorSearchStopCondition : public BaseSearchStopCondition<dist_t> {
  size_t curr_num_docs_, num_docs_to_search_, ef_collection_;
  std::unordered_map<DOCIDTYPE, size_t> doc_counter_;
  std::priority_queue<std::pair<dist_t, DOCIDTYPE>> search_results_;
  BaseMultiVectorSpace<DOCIDTYPE> &space_;

 public:
  MultiVectorSearchStopCondition(BaseMultiVectorSpace<DOCIDTYPE> &space, size_t num_docs_to_search, size_t ef_collection = 10)
      : curr_num_docs_(0), num_docs_to_search_(num_docs_to_search), ef_collection_(std::max(ef_collection, num_docs_to_search)), space_(space) {}
  size_t num_docs_to_search() const { return num_docs_to_search_; }
  size_t ef_collection() const { return ef_collection_; }
  void add_point_to_result(labeltype, const void *datapoint, dist_t dist) override {
    DOCIDTYPE doc_id = space_.get_doc_id(datapoint);
    if (doc_counter_[doc_id] == 0) curr_num_docs_ += 1;
    search_results_.emplace(dist, doc_id);
    doc_counter_[doc_id] += 1;
  }
  void remove_point_from_result(labeltype, const void *datapoint, dist_t) override {
    DOCIDTYPE doc_id = space_.get_doc_id(datapoint);
    doc_counter_[doc_id] -= 1;
    if (doc_counter_[doc_id] == 0) curr_num_docs_ -= 1;
    search_results_.pop();
  }
  bool should_stop_search(dist_t candidate_dist, dist_t lowerBound) override { return candidate_dist > lowerBound && curr_num_docs_ == ef_collection_; }
  bool should_consider_candidate(dist_t candidate_dist, dist_t lowerBound) override { return curr_num_docs_ < ef_collection_ || lowerBound > candidate_dist; }
  bool should_remove_extra() override { return curr_num_docs_ > ef_collection_; }
  void filter_results(std::vector<std::pair<dist_t, labeltype>> &candidates) override {
    while (curr_num_docs_ > num_docs_to_search_) {
      DOCIDTYPE doc_id = search_results_.top().second;
      doc_counter_[doc_id] -= 1;
      if (doc_counter_[doc_id] == 0) curr_num_docs_ -= 1;
      search_results_.pop();
      candidates.pop_back();
    }
  }
};

namespace detail {
// metric + dim of a Space; anything but the two known spaces is rejected
inline void space_info(SpaceInterface<float> *s, int &metric, size_t &dim) {
  if (dynamic_cast<L2Space *>(s)) metric = B200ANN_L2;
  else if (dynamic_cast<InnerProductSpace *>(s)) metric = B200ANN_IP;
  else if (auto *mv = dynamic_cast<BaseMultiVectorSpace<unsigned int> *>(s)) metric = mv->b200ann_metric();
  else throw std::runtime_error("b200ann: only L2Space, InnerProductSpace and the multi-vector spaces run on the GPU (a device kernel cannot call a host DISTFUNC)");
  dim = *(size_t *)s->get_dist_func_param();
}
// HierarchicalNSW<int>: L2SpaceI (unsigned-byte rows, int squared L2) -> a B200ANN_L2_U8 index
inline void space_info(SpaceInterface<int> *s, int &metric, size_t &dim) {
  if (!dynamic_cast<L2SpaceI *>(s)) throw std::runtime_error("b200ann: L2SpaceI is the only SpaceInterface<int> that runs on the GPU");
  metric = B200ANN_L2_U8;
  dim = *(size_t *)s->get_dist_func_param();
}
}  // namespace detail

template <typename dist_t>
class AlgorithmInterface {
 public:
  virtual void addPoint(const void *datapoint, labeltype label, bool replace_deleted = false) = 0;
  virtual std::priority_queue<std::pair<dist_t, labeltype>> searchKnn(const void *, size_t, BaseFilterFunctor *isIdAllowed = nullptr) const = 0;
  // closer first (hnswlib.h:203-221)
  virtual std::vector<std::pair<dist_t, labeltype>> searchKnnCloserFirst(const void *query_data, size_t k,
                                                                        BaseFilterFunctor *isIdAllowed = nullptr) const {
    auto ret = searchKnn(query_data, k, isIdAllowed);
    std::vector<std::pair<dist_t, labeltype>> result(ret.size());
    size_t sz = ret.size();
    while (!ret.empty()) { result[--sz] = ret.top(); ret.pop(); }
    return result;
  }
  virtual void saveIndex(const std::string &location) = 0;
  virtual ~AlgorithmInterface() {}
};

// hnswlib::BruteforceSearch (bruteforce.h:10-172): exact scan with top-k, the reference's own
// ground-truth oracle. Points are staged on the host by addPoint (the reference copies them too,
// bruteforce.h:79-81) and uploaded once before the next search.
template <typename dist_t>
class BruteforceSearch : public AlgorithmInterface<dist_t> {
 public:
  size_t maxelements_, cur_element_count, dim_;
  int metric_;

  BruteforceSearch(SpaceInterface<dist_t> *s, size_t maxElements, int device = 0) : maxelements_(maxElements), cur_element_count(0), gpu_(device) {
    detail::space_info(s, metric_, dim_);
    data_.resize(maxElements * dim_);
    labels_.resize(maxElements);
  }
  ~BruteforceSearch() { if (ix_) b200ann_index_destroy(ix_); }

  void addPoint(const void *datapoint, labeltype label, bool = false) override {
    std::unique_lock<std::mutex> lock(mu_);
    size_t idx;
    auto it = pos_.find(label);
    if (it != pos_.end()) idx = it->second;
    else {
      if (cur_element_count >= maxelements_) throw std::runtime_error("The number of elements exceeds the specified limit\n");  // bruteforce.h:73-75
      idx = cur_element_count++;
      pos_[label] = idx;
    }
    labels_[idx] = label;
    std::memcpy(&data_[idx * dim_], datapoint, dim_ * sizeof(float));
    dirty_ = true;
  }
  void removePoint(labeltype label) {  // bruteforce.h:86-103: the last element takes the hole
    std::unique_lock<std::mutex> lock(mu_);
    auto it = pos_.find(label);
    if (it == pos_.end()) return;
    size_t idx = it->second, last = cur_element_count - 1;
    pos_.erase(it);
    if (idx != last) {
      std::memcpy(&data_[idx * dim_], &data_[last * dim_], dim_ * sizeof(float));
      labels_[idx] = labels_[last];
      pos_[labels_[idx]] = idx;
    }
    cur_element_count--;
    dirty_ = true;
  }

  // [nq,k] rows ascending by (distance, insertion position)
  void searchKnnBatch(const float *queries, size_t nq, size_t k, labeltype *labels, dist_t *dists) const {
    std::unique_lock<std::mutex> lock(mu_);
    if (k > cur_element_count) throw std::runtime_error("BruteforceSearch: k > number of elements");  // assert at bruteforce.h:108
    upload();
    std::vector<uint32_t> ids(nq * k);
    detail::check(gpu_.h, b200ann_bf_knn(gpu_.h, ix_, queries, nq, dim_, k, ids.data(), dists), "BruteforceSearch::searchKnn");
    for (size_t i = 0; i < nq * k; i++) labels[i] = labels_[ids[i]];
  }
  std::priority_queue<std::pair<dist_t, labeltype>> searchKnn(const void *query_data, size_t k,
                                                              BaseFilterFunctor *isIdAllowed = nullptr) const override {
    std::priority_queue<std::pair<dist_t, labeltype>> res;
    if (cur_element_count == 0) return res;
    size_t kk = k < cur_element_count ? k : cur_element_count;
    for (;;) {
      std::vector<labeltype> lab(kk);
      std::vector<dist_t> d(kk);
      searchKnnBatch((const float *)query_data, 1, kk, lab.data(), d.data());
      res = std::priority_queue<std::pair<dist_t, labeltype>>();
      for (size_t i = 0; i < kk && res.size() < k; i++)
        if (!isIdAllowed || (*isIdAllowed)(lab[i])) res.emplace(d[i], lab[i]);
      if (res.size() >= k || kk == cur_element_count || !isIdAllowed) break;
      kk = kk * 4 < cur_element_count ? kk * 4 : cur_element_count;  // filtered: widen the exact scan
    }
    return res;
  }

  // bruteforce.h:138-171: [maxelements][size_per_element][cur_element_count] then rows [vector][label]
  void saveIndex(const std::string &location) override {
    std::ofstream out(location, std::ios::binary);
    size_t spe = dim_ * sizeof(float) + sizeof(labeltype);
    out.write((char *)&maxelements_, sizeof(size_t));
    out.write((char *)&spe, sizeof(size_t));
    out.write((char *)&cur_element_count, sizeof(size_t));
    for (size_t i = 0; i < maxelements_; i++) {
      out.write((const char *)&data_[i * dim_], dim_ * sizeof(float));
      out.write((const char *)&labels_[i], sizeof(labeltype));
    }
  }
  void loadIndex(const std::string &location, SpaceInterface<dist_t> *s) {
    std::ifstream in(location, std::ios::binary);
    if (!in) throw std::runtime_error("Cannot open file");
    detail::space_info(s, metric_, dim_);
    size_t spe;
    in.read((char *)&maxelements_, sizeof(size_t));
    in.read((char *)&spe, sizeof(size_t));
    in.read((char *)&cur_element_count, sizeof(size_t));
    data_.resize(maxelements_ * dim_);
    labels_.resize(maxelements_);
    pos_.clear();
    for (size_t i = 0; i < maxelements_; i++) {
      in.read((char *)&data_[i * dim_], dim_ * sizeof(float));
      in.read((char *)&labels_[i], sizeof(labeltype));
      if (i < cur_element_count) pos_[labels_[i]] = i;
    }
    dirty_ = true;
  }

 private:
  void upload() const {
    if (!dirty_ && ix_) return;
    if (ix_) { b200ann_index_destroy(ix_); ix_ = nullptr; }
    detail::check(gpu_.h, b200ann_index_create(gpu_.h, data_.data(), cur_element_count, dim_, dim_, metric_, &ix_), "b200ann_index_create");
    dirty_ = false;
  }
  detail::Gpu gpu_;
  std::vector<float> data_;
  std::vector<labeltype> labels_;
  std::unordered_map<labeltype, size_t> pos_;
  mutable b200ann_index_t ix_ = nullptr;
  mutable bool dirty_ = true;
  mutable std::mutex mu_;
};

// BruteforceSearch<int> over L2SpaceI: rows of dim unsigned bytes, int distances; same contract and file format
// (bruteforce.h:138-171 with size_per_element = dim + sizeof(labeltype)) as the float class above. Distances and
// labels are bit-identical to the reference's (integer arithmetic).
template <>
class BruteforceSearch<int> : public AlgorithmInterface<int> {
 public:
  size_t maxelements_, cur_element_count, dim_;

  BruteforceSearch(SpaceInterface<int> *s, size_t maxElements, int device = 0) : maxelements_(maxElements), cur_element_count(0), gpu_(device) {
    set_space(s);
    data_.resize(maxElements * dim_);
    labels_.resize(maxElements);
  }
  ~BruteforceSearch() { if (dev_) b200ann_free_device(dev_); }

  void addPoint(const void *datapoint, labeltype label, bool = false) override {
    std::unique_lock<std::mutex> lock(mu_);
    size_t idx;
    auto it = pos_.find(label);
    if (it != pos_.end()) idx = it->second;
    else {
      if (cur_element_count >= maxelements_) throw std::runtime_error("The number of elements exceeds the specified limit\n");  // bruteforce.h:73-75
      idx = cur_element_count++;
      pos_[label] = idx;
    }
    labels_[idx] = label;
    std::memcpy(&data_[idx * dim_], datapoint, dim_);
    dirty_ = true;
  }
  void removePoint(labeltype label) {  // bruteforce.h:86-103
    std::unique_lock<std::mutex> lock(mu_);
    auto it = pos_.find(label);
    if (it == pos_.end()) return;
    size_t idx = it->second, last = cur_element_count - 1;
    pos_.erase(it);
    if (idx != last) {
      std::memcpy(&data_[idx * dim_], &data_[last * dim_], dim_);
      labels_[idx] = labels_[last];
      pos_[labels_[idx]] = idx;
    }
    cur_element_count--;
    dirty_ = true;
  }

  // [nq,k] rows ascending by (distance, insertion position)
  void searchKnnBatch(const unsigned char *queries, size_t nq, size_t k, labeltype *labels, int *dists) const {
    std::unique_lock<std::mutex> lock(mu_);
    if (k > cur_element_count) throw std::runtime_error("BruteforceSearch: k > number of elements");  // assert at bruteforce.h:108
    upload();
    std::vector<uint32_t> ids(nq * k);
    detail::check(gpu_.h, b200ann_bf_knn_u8(gpu_.h, dev_, cur_element_count, dim_, dim_, queries, nq, dim_, k, ids.data(), dists),
                  "BruteforceSearch<int>::searchKnn");
    for (size_t i = 0; i < nq * k; i++) labels[i] = labels_[ids[i]];
  }
  std::priority_queue<std::pair<int, labeltype>> searchKnn(const void *query_data, size_t k,
                                                           BaseFilterFunctor *isIdAllowed = nullptr) const override {
    std::priority_queue<std::pair<int, labeltype>> res;
    if (cur_element_count == 0) return res;
    size_t kk = k < cur_element_count ? k : cur_element_count;
    for (;;) {
      std::vector<labeltype> lab(kk);
      std::vector<int> d(kk);
      searchKnnBatch((const unsigned char *)query_data, 1, kk, lab.data(), d.data());
      res = std::priority_queue<std::pair<int, labeltype>>();
      for (size_t i = 0; i < kk && res.size() < k; i++)
        if (!isIdAllowed || (*isIdAllowed)(lab[i])) res.emplace(d[i], lab[i]);
      if (res.size() >= k || kk == cur_element_count || !isIdAllowed) break;
      kk = kk * 4 < cur_element_count ? kk * 4 : cur_element_count;  // filtered: widen the exact scan
    }
    return res;
  }

  void saveIndex(const std::string &location) override {
    std::ofstream out(location, std::ios::binary);
    size_t spe = dim_ + sizeof(labeltype);
    out.write((char *)&maxelements_, sizeof(size_t));
    out.write((char *)&spe, sizeof(size_t));
    out.write((char *)&cur_element_count, sizeof(size_t));
    for (size_t i = 0; i < maxelements_; i++) {
      out.write((const char *)&data_[i * dim_], dim_);
      out.write((const char *)&labels_[i], sizeof(labeltype));
    }
  }
  void loadIndex(const std::string &location, SpaceInterface<int> *s) {
    std::ifstream in(location, std::ios::binary);
    if (!in) throw std::runtime_error("Cannot open file");
    set_space(s);
    size_t spe;
    in.read((char *)&maxelements_, sizeof(size_t));
    in.read((char *)&spe, sizeof(size_t));
    in.read((char *)&cur_element_count, sizeof(size_t));
    data_.resize(maxelements_ * dim_);
    labels_.resize(maxelements_);
    pos_.clear();
    for (size_t i = 0; i < maxelements_; i++) {
      in.read((char *)&data_[i * dim_], dim_);
      in.read((char *)&labels_[i], sizeof(labeltype));
      if (i < cur_element_count) pos_[labels_[i]] = i;
    }
    dirty_ = true;
  }

 private:
  void set_space(SpaceInterface<int> *s) {
    if (!dynamic_cast<L2SpaceI *>(s)) throw std::runtime_error("b200ann: BruteforceSearch<int> needs L2SpaceI (a device kernel cannot call a host DISTFUNC)");
    dim_ = *(size_t *)s->get_dist_func_param();
  }
  void upload() const {  // the rows stay resident on the device between searches
    if (!dirty_ && dev_) return;
    if (dev_) { b200ann_free_device(dev_); dev_ = nullptr; }
    detail::check(gpu_.h, b200ann_upload_u8(gpu_.h, data_.data(), cur_element_count * dim_, &dev_), "b200ann_upload_u8");
    dirty_ = false;
  }
  detail::Gpu gpu_;
  std::vector<unsigned char> data_;
  std::vector<labeltype> labels_;
  std::unordered_map<labeltype, size_t> pos_;
  mutable uint8_t *dev_ = nullptr;
  mutable bool dirty_ = true;
  mutable std::mutex mu_;
};

// hnswlib::HierarchicalNSW (hnswalg.h:18-1411): build with the construction ctor + addPoint, or load an index
// written by HierarchicalNSW::saveIndex (:685-713); search it on the GPU; write it back in the same format.
// addPoint copies the vector (like :1204) into a host staging area; the graph is built on the GPU
// (b200ann_hnsw_build: same level sequence as the reference, batched insertion) by build(), or lazily by the
// first searchKnn / saveIndex after an addPoint. Points added to a built index are linked into the live device graph
// in place on the next use (b200ann_hnsw_insert_batch: what addPoint does to a live hnswlib index, :1153-1267).
// addPoint with an existing label replaces that element's vector (the reference's updatePoint, :1158-1175, re-links
// in place; here an update or a replaced deleted slot re-links everything on the next use).
template <typename dist_t>
class HierarchicalNSW : public AlgorithmInterface<dist_t> {
 public:
  // same public fields the reference exposes and the hybrid index pokes (index_hnsw_nsg.h:110-148)
  size_t max_elements_ = 0, cur_element_count = 0, M_ = 0, maxM_ = 0, maxM0_ = 0, ef_construction_ = 0, ef_ = 10;
  double mult_ = 0;
  int maxlevel_ = -1;
  tableint enterpoint_node_ = 0;
  std::vector<int> element_levels_;   // valid after saveIndex / loadIndex (the graph lives on the device)
  // hnswalg.h:65-66: distance evaluations and expanded nodes of every search so far
  mutable std::atomic<long> metric_distance_computations{0};
  mutable std::atomic<long> metric_hops{0};

  // construction ctor (hnswalg.h:89-144)
  HierarchicalNSW(SpaceInterface<dist_t> *s, size_t max_elements, size_t M = 16, size_t ef_construction = 200,
                  size_t random_seed = 100, bool allow_replace_deleted = false, int device = 0)
      : gpu_(device), allow_replace_deleted_(allow_replace_deleted) {
    detail::space_info(s, metric_, dim_); set_geometry(s->get_data_size());
    max_elements_ = max_elements;
    M_ = M <= 10000 ? M : 10000;                       // :98-104
    maxM_ = M_; maxM0_ = M_ * 2;
    ef_construction_ = ef_construction > M_ ? ef_construction : M_;   // :107
    mult_ = 1 / std::log(1.0 * (double)M_);            // :114
    seed_ = random_seed;
  }
  // load ctor (hnswalg.h:78-86)
  HierarchicalNSW(SpaceInterface<dist_t> *s, const std::string &location, bool = false, size_t max_elements = 0,
                  bool allow_replace_deleted = false, int device = 0)
      : gpu_(device), allow_replace_deleted_(allow_replace_deleted) {
    detail::space_info(s, metric_, dim_); set_geometry(s->get_data_size());
    loadIndex(location, s, max_elements);
  }
  ~HierarchicalNSW() { if (ix_) b200ann_index_destroy(ix_); }

  void setEf(size_t ef) { ef_ = ef; }  // hnswalg.h:173

  // hnswalg.h:716-822. Throws std::runtime_error("Cannot open file") like :720.
  void loadIndex(const std::string &location, SpaceInterface<dist_t> *s, size_t max_elements_i = 0) {
    std::ifstream in(location, std::ios::binary);
    if (!in.is_open()) throw std::runtime_error("Cannot open file");
    detail::space_info(s, metric_, dim_); set_geometry(s->get_data_size());
    size_t offsetLevel0, size_data_per_element, label_offset, offsetData;
    rd(in, offsetLevel0); rd(in, max_elements_); rd(in, cur_element_count); rd(in, size_data_per_element);
    rd(in, label_offset); rd(in, offsetData); rd(in, maxlevel_); rd(in, enterpoint_node_); rd(in, maxM_); rd(in, maxM0_);
    rd(in, M_); rd(in, mult_); rd(in, ef_construction_);
    if (!in) throw std::runtime_error("Index seems to be corrupted or unsupported");
    if (max_elements_i > max_elements_) max_elements_ = max_elements_i;
    if (label_offset < offsetData || label_offset - offsetData != data_size_) throw std::runtime_error("index file dimension does not match the space");
    const size_t n = cur_element_count;
    // the layout the reference writes (hnswalg.h:116-123) and the size check of its loader (:740-760)
    if (maxM0_ == 0 || maxM_ == 0 || maxM0_ > (1u << 20) || maxM_ > (1u << 20) || n > max_elements_ || offsetLevel0 != 0 ||
        offsetData != maxM0_ * 4 + 4 || size_data_per_element != label_offset + sizeof(labeltype) || (n > 0 && enterpoint_node_ >= n))
      throw std::runtime_error("Index seems to be corrupted or unsupported");
    {
      const std::streampos here = in.tellg();
      in.seekg(0, std::ios::end);
      const std::streampos total = in.tellg();
      in.seekg(here);
      if ((unsigned long long)(total - here) < (unsigned long long)n * size_data_per_element + (unsigned long long)n * 4)
        throw std::runtime_error("Index seems to be corrupted or unsupported");
    }
    std::vector<char> blob(n * size_data_per_element);
    in.read(blob.data(), (std::streamsize)blob.size());
    vec_.resize(n * rw_); labels_.resize(n); links0_.assign(n * maxM0_, B200ANN_INVALID);
    deleted_.assign(n, 0); num_deleted_ = 0; deleted_set_.clear();
    label_lookup_.clear();
    for (size_t i = 0; i < n; i++) {
      const char *row = blob.data() + i * size_data_per_element;
      unsigned short cnt = *(const unsigned short *)(row + offsetLevel0);
      if (cnt > maxM0_) throw std::runtime_error("Index seems to be corrupted or unsupported");
      if (*(const unsigned char *)(row + offsetLevel0 + 2) & DELETE_MARK) {  // :811-815
        deleted_[i] = 1; num_deleted_++;
        if (allow_replace_deleted_) deleted_set_.insert((tableint)i);
      }
      std::memcpy(&links0_[i * maxM0_], row + offsetLevel0 + 4, (size_t)cnt * 4);
      std::memcpy(&vec_[i * rw_], row + offsetData, data_size_);
      labeltype lab;
      std::memcpy(&lab, row + label_offset, sizeof(labeltype));
      labels_[i] = lab;
      label_lookup_[lab] = i;
    }
    const size_t slpe = maxM_ * 4 + 4;  // size_links_per_element_ (:113)
    element_levels_.assign(n, 0); upper_ptr_.assign(n, 0); upper_links_.clear();
    for (size_t i = 0; i < n; i++) {
      unsigned int sz;
      rd(in, sz);
      if (!in) throw std::runtime_error("Index seems to be corrupted or unsupported");
      if (!sz) continue;
      if (sz % slpe != 0 || sz / slpe > 64) throw std::runtime_error("Index seems to be corrupted or unsupported");
      int lv = (int)(sz / slpe);
      element_levels_[i] = lv;
      upper_ptr_[i] = (uint32_t)(upper_links_.size() / maxM_);
      std::vector<char> ll(sz);
      in.read(ll.data(), sz);
      if (!in) throw std::runtime_error("Index seems to be corrupted or unsupported");
      for (int l = 0; l < lv; l++) {
        const char *p = ll.data() + (size_t)l * slpe;
        unsigned short cnt = *(const unsigned short *)p;
        if (cnt > maxM_) throw std::runtime_error("Index seems to be corrupted or unsupported");
        size_t base = upper_links_.size();
        upper_links_.resize(base + maxM_, B200ANN_INVALID);
        std::memcpy(&upper_links_[base], p + 4, (size_t)cnt * 4);
      }
    }
    needs_rebuild_ = false;
    built_count_ = n;
    host_graph_valid_ = true;
    upload();
  }

  // hnswalg.h:685-713, byte-compatible with the reference's loader
  void saveIndex(const std::string &location) override {
    { std::unique_lock<std::mutex> lock(mu_); build_locked(nullptr); sync_host_graph_locked(); }
    std::ofstream out(location, std::ios::binary);
    const size_t n = cur_element_count;
    size_t size_links_level0 = maxM0_ * 4 + 4, data_size = data_size_;
    size_t size_data_per_element = size_links_level0 + data_size + sizeof(labeltype);
    size_t offsetLevel0 = 0, offsetData = size_links_level0, label_offset = size_links_level0 + data_size;
    wr(out, offsetLevel0); wr(out, max_elements_); wr(out, cur_element_count); wr(out, size_data_per_element);
    wr(out, label_offset); wr(out, offsetData); wr(out, maxlevel_); wr(out, enterpoint_node_); wr(out, maxM_); wr(out, maxM0_);
    wr(out, M_); wr(out, mult_); wr(out, ef_construction_);
    std::vector<char> row(size_data_per_element);
    for (size_t i = 0; i < n; i++) {
      std::memset(row.data(), 0, row.size());
      unsigned short cnt = 0;
      while (cnt < maxM0_ && links0_[i * maxM0_ + cnt] != B200ANN_INVALID) cnt++;
      std::memcpy(row.data(), &cnt, 2);
      if (deleted_[i]) row[2] = (char)DELETE_MARK;  // the mark lives in the upper half of the list-size word (:864-878)
      std::memcpy(row.data() + 4, &links0_[i * maxM0_], (size_t)cnt * 4);
      std::memcpy(row.data() + offsetData, &vec_[i * rw_], data_size);
      labeltype lab = labels_[i];
      std::memcpy(row.data() + label_offset, &lab, sizeof(labeltype));
      out.write(row.data(), (std::streamsize)row.size());
    }
    const size_t slpe = maxM_ * 4 + 4;
    for (size_t i = 0; i < n; i++) {
      unsigned int sz = element_levels_[i] > 0 ? (unsigned int)(slpe * element_levels_[i]) : 0;
      wr(out, sz);
      for (int l = 0; l < element_levels_[i]; l++) {
        std::vector<char> ll(slpe, 0);
        const uint32_t *src = &upper_links_[((size_t)upper_ptr_[i] + l) * maxM_];
        unsigned short cnt = 0;
        while (cnt < maxM_ && src[cnt] != B200ANN_INVALID) cnt++;
        std::memcpy(ll.data(), &cnt, 2);
        std::memcpy(ll.data() + 4, src, (size_t)cnt * 4);
        out.write(ll.data(), (std::streamsize)slpe);
      }
    }
  }

  // hnswalg.h:954-992 / :1153-1267. Thread-safe (the drivers call it from omp / ParallelFor loops); internal ids are
  // handed out in arrival order like cur_element_count++ (:1181).
  // replace_deleted (:954-992, needs allow_replace_deleted in the ctor): the new point takes the slot (internal id,
  // vector, label) of an element marked deleted instead of growing the index; with no deleted element it is a plain
  // add. The reference takes *deleted_elements.begin() of an unordered_set (an unspecified one); here the smallest
  // deleted internal id. Like the reference, the slot is re-linked in place (updatePoint -> b200ann_hnsw_update_batch on next use).
  void addPoint(const void *datapoint, labeltype label, bool replace_deleted = false) override {
    if (!allow_replace_deleted_ && replace_deleted) throw std::runtime_error("Replacement of deleted elements is disabled in constructor");  // :955-957
    std::unique_lock<std::mutex> lock(mu_);
    if (replace_deleted && !deleted_set_.empty()) {
      tableint id = *deleted_set_.begin();
      deleted_set_.erase(deleted_set_.begin());
      label_lookup_.erase(labels_[id]);      // :982-987
      labels_[id] = label;
      label_lookup_[label] = id;
      deleted_[id] = 0; num_deleted_--;      // unmarkDeletedInternal :989
      std::memcpy(&vec_[id * rw_], datapoint, data_size_);
      if (id < built_count_) { pending_upd_.insert(id); pending_lab_.insert(id); }  // :989-991 updatePoint on the slot
      return;
    }
    auto it = label_lookup_.find(label);
    if (it != label_lookup_.end()) {
      if (allow_replace_deleted_ && deleted_[it->second])  // :1163-1167
        throw std::runtime_error("Can't use addPoint to update deleted elements if replacement of deleted elements is enabled.");
      if (deleted_[it->second]) { deleted_[it->second] = 0; num_deleted_--; }  // :1166-1168: an update un-deletes
      std::memcpy(&vec_[it->second * rw_], datapoint, data_size_);
      if (it->second < built_count_) pending_upd_.insert((tableint)it->second);  // a linked element moved: updatePoint (:1172)
      return;
    }
    if (cur_element_count >= max_elements_) throw std::runtime_error("The number of elements exceeds the specified limit");  // :1178
    label_lookup_[label] = cur_element_count;
    vec_.resize(vec_.size() + rw_, 0.f);
    std::memcpy(&vec_[vec_.size() - rw_], datapoint, data_size_);
    labels_.push_back(label);
    deleted_.push_back(0);
    cur_element_count++;   // linked into the device graph by the next build_locked()
  }
  // hnswalg.h:853-911 markDelete / :893-922 unmarkDelete: the element stays in the graph and is walked through, but is
  // never returned (searchBaseLayerST<false>); same exceptions as the reference
  void markDelete(labeltype label) {
    std::unique_lock<std::mutex> lock(mu_);
    auto it = label_lookup_.find(label);
    if (it == label_lookup_.end()) throw std::runtime_error("Label not found");
    if (deleted_[it->second]) throw std::runtime_error("The requested to delete element is already deleted");
    deleted_[it->second] = 1; num_deleted_++;
    if (allow_replace_deleted_) deleted_set_.insert((tableint)it->second);  // :879-882
  }
  void unmarkDelete(labeltype label) {
    std::unique_lock<std::mutex> lock(mu_);
    auto it = label_lookup_.find(label);
    if (it == label_lookup_.end()) throw std::runtime_error("Label not found");
    if (!deleted_[it->second]) throw std::runtime_error("The requested to undelete element is not deleted");
    deleted_[it->second] = 0; num_deleted_--;
    if (allow_replace_deleted_) deleted_set_.erase((tableint)it->second);  // :921-924
  }
  // hnswalg.h:640-675: only the capacity changes here (the staging vectors grow on demand)
  void resizeIndex(size_t new_max_elements) {
    std::unique_lock<std::mutex> lock(mu_);
    if (new_max_elements < cur_element_count) throw std::runtime_error("Cannot resize, max element is less than the current number of elements");
    max_elements_ = new_max_elements;
  }
  bool isMarkedDeleted(tableint internal_id) const { return deleted_[internal_id] != 0; }
  // hnswalg.h:825-847: the stored vector of a label; unknown and deleted labels throw alike
  template <typename data_t>
  std::vector<data_t> getDataByLabel(labeltype label) const {
    std::unique_lock<std::mutex> lock(mu_);
    auto it = label_lookup_.find(label);
    if (it == label_lookup_.end() || deleted_[it->second]) throw std::runtime_error("Label not found");
    const data_t *p = (const data_t *)&vec_[it->second * rw_];
    return std::vector<data_t>(p, p + data_size_ / sizeof(data_t));
  }
  size_t getDeletedCount() const { return num_deleted_; }
  size_t getCurrentElementCount() const { return cur_element_count; }
  size_t getMaxElements() const { return max_elements_; }
  // builds the graph of everything added so far on the GPU (no-op when up to date). info8: b200ann_hnsw_build's.
  void build(uint64_t *info8 = nullptr) {
    std::unique_lock<std::mutex> lock(mu_);
    build_locked(info8);
  }

  // [nq,k] rows, ascending distance. Without a filter and without deleted elements: b200ann_hnsw_search (throws when
  // some row has fewer than k results, bindings.cpp:646-648). Otherwise searchBaseLayerST<false> on the GPU
  // (b200ann_hnsw_search_filtered): the functor is evaluated ONCE per batch over all labels into a bitmap; short
  // rows are padded with ~0 / +inf.
  void searchKnnBatch(const void *queries, size_t nq, size_t k, labeltype *labels, dist_t *dists,
                      BaseFilterFunctor *isIdAllowed = nullptr) const {
    std::vector<uint64_t> lab(nq * k);
    std::unique_lock<std::mutex> lock(mu_);
    build_locked(nullptr);
    if (!ix_) throw std::runtime_error("HierarchicalNSW::searchKnnBatch: the index is empty");
    std::vector<uint32_t> st(nq * 2, 0u);
    if (!num_deleted_ && !isIdAllowed) {  // :1306 bare_bone_search
      detail::check(gpu_.h, b200ann_hnsw_search(gpu_.h, ix_, (const float *)queries, nq, dim_, k, ef_, lab.data(), (float *)dists, st.data()), "HierarchicalNSW::searchKnn");
    } else {
      std::vector<uint32_t> allow = allow_bitmap(isIdAllowed);
      detail::check(gpu_.h, b200ann_hnsw_search_filtered(gpu_.h, ix_, (const float *)queries, nq, dim_, k, ef_, allow.data(), lab.data(), (float *)dists,
                                                         st.data(), nullptr), "HierarchicalNSW::searchKnn (filtered)");
    }
    count_metrics(st);
    for (size_t i = 0; i < nq * k; i++) labels[i] = (labeltype)lab[i];
  }
  // hnswalg.h:1270-1324: max-heap of (distance,label), at most k entries; no exception when fewer than k exist
  std::priority_queue<std::pair<dist_t, labeltype>> searchKnn(const void *query_data, size_t k,
                                                              BaseFilterFunctor *isIdAllowed = nullptr) const override {
    std::priority_queue<std::pair<dist_t, labeltype>> res;
    if (cur_element_count == 0) return res;
    std::vector<uint64_t> lab(k);
    std::vector<dist_t> d(k);
    std::vector<uint32_t> st(2, 0u);
    {
      std::unique_lock<std::mutex> lock(mu_);
      build_locked(nullptr);
      if (!num_deleted_ && !isIdAllowed) {
        size_t kk = k < cur_element_count ? k : cur_element_count;
        int rc = b200ann_hnsw_search(gpu_.h, ix_, (const float *)query_data, 1, dim_, kk, ef_, lab.data(), (float *)d.data(), st.data());
        if (rc != 0 && rc != B200ANN_ERR_FEWER) detail::check(gpu_.h, rc, "HierarchicalNSW::searchKnn");
        for (size_t i = kk; i < k; i++) lab[i] = ~(uint64_t)0;
      } else {
        std::vector<uint32_t> allow = allow_bitmap(isIdAllowed);
        detail::check(gpu_.h, b200ann_hnsw_search_filtered(gpu_.h, ix_, (const float *)query_data, 1, dim_, k, ef_, allow.data(),
                                                           lab.data(), (float *)d.data(), st.data(), nullptr), "HierarchicalNSW::searchKnn (filtered)");
      }
    }
    count_metrics(st);
    for (size_t i = 0; i < k; i++)
      if (lab[i] != ~(uint64_t)0) res.emplace(d[i], (labeltype)lab[i]);
    return res;
  }

  // hnswalg.h:1327-1378 with an EpsilonSearchStopCondition: (distance, id) pairs, closest first, every distance <=
  // epsilon, at most max_num_candidates of them. Like the reference, .second is the INTERNAL id (it copies
  // top_candidates' pairs without getExternalLabel, :1368-1371); searchEpsilonBatch(..., internal_ids = false) gives
  // labels. Any other condition (the multi-vector one) has no GPU form and throws.
  std::vector<std::pair<dist_t, labeltype>> searchStopConditionClosest(const void *query_data,
                                                                       BaseSearchStopCondition<dist_t> &stop_condition,
                                                                       BaseFilterFunctor *isIdAllowed = nullptr) const {
    std::vector<std::pair<dist_t, labeltype>> result;
    if (cur_element_count == 0) return result;
    if (auto *mv = dynamic_cast<MultiVectorSearchStopCondition<unsigned int, dist_t> *>(&stop_condition))
      return search_multivector(query_data, mv->num_docs_to_search(), mv->ef_collection(), isIdAllowed);
    auto *eps = dynamic_cast<EpsilonSearchStopCondition<dist_t> *>(&stop_condition);
    if (!eps) throw std::runtime_error("searchStopConditionClosest: EpsilonSearchStopCondition and MultiVectorSearchStopCondition<unsigned int> run on the GPU");
    const size_t mx = eps->max_num_candidates();
    if (mx == 0) return result;
    std::vector<labeltype> ids(mx);
    std::vector<dist_t> d(mx);
    unsigned cnt = 0;
    searchEpsilonBatch(query_data, 1, eps->epsilon(), eps->min_num_candidates(), mx, ids.data(), d.data(), &cnt,
                       isIdAllowed, true);
    result.resize(cnt);
    for (unsigned i = 0; i < cnt; i++) result[i] = std::make_pair(d[i], ids[i]);
    return result;
  }
  // MultiVectorSearchStopCondition (stop_condition.h:146-215) for one query: the document ids are read from the rows
  // (behind the vector, where set_doc_id put them); the result list grows until it fits (a document may own many rows)
  std::vector<std::pair<dist_t, labeltype>> search_multivector(const void *query_data, size_t num_docs, size_t ef_collection,
                                                               BaseFilterFunctor *isIdAllowed) const {
    std::vector<std::pair<dist_t, labeltype>> result;
    std::unique_lock<std::mutex> lock(mu_);
    build_locked(nullptr);
    if (!ix_) return result;
    if (data_size_ != dim_ * sizeof(float) + 4) throw std::runtime_error("MultiVectorSearchStopCondition needs an index over a multi-vector space");
    const size_t n = cur_element_count;
    std::vector<uint32_t> doc(n), allow;
    for (size_t i = 0; i < n; i++) std::memcpy(&doc[i], (const char *)&vec_[i * rw_] + dim_ * sizeof(float), 4);
    const bool filtered = num_deleted_ || isIdAllowed;
    if (filtered) allow = allow_bitmap(isIdAllowed);
    size_t mx = std::min<size_t>(n, std::max<size_t>(256, 8 * ef_collection));
    for (;;) {
      std::vector<uint64_t> lab(mx);
      std::vector<dist_t> d(mx);
      uint32_t cnt = 0, st[2] = {0, 0};
      int rc = b200ann_hnsw_search_multivector(gpu_.h, ix_, (const float *)query_data, 1, dim_, doc.data(), num_docs, ef_collection, mx,
                                               filtered ? allow.data() : nullptr, 1, lab.data(), (float *)d.data(), &cnt, st);
      if (rc == B200ANN_ERR_CAPACITY && mx < n) { mx = std::min(n, mx * 4); continue; }
      detail::check(gpu_.h, rc, "HierarchicalNSW::searchStopConditionClosest (multi-vector)");
      count_metrics(std::vector<uint32_t>(st, st + 2));
      result.resize(cnt);
      for (uint32_t i = 0; i < cnt; i++) result[i] = std::make_pair(d[i], (labeltype)lab[i]);
      return result;
    }
  }
  // batched form: ids / dists [nq, max_num_candidates] ascending, counts[q] valid entries per row, the rest ~0 / +inf
  void searchEpsilonBatch(const void *queries, size_t nq, float epsilon, size_t min_num_candidates, size_t max_num_candidates,
                          labeltype *ids, dist_t *dists, unsigned *counts, BaseFilterFunctor *isIdAllowed = nullptr,
                          bool internal_ids = false) const {
    std::vector<uint64_t> lab(nq * max_num_candidates);
    std::vector<uint32_t> st(nq * 2, 0u), cnt(nq, 0u);
    std::unique_lock<std::mutex> lock(mu_);
    build_locked(nullptr);
    if (!ix_) throw std::runtime_error("HierarchicalNSW::searchStopConditionClosest: the index is empty");
    std::vector<uint32_t> allow;
    const bool filtered = num_deleted_ || isIdAllowed;
    if (filtered) allow = allow_bitmap(isIdAllowed);
    detail::check(gpu_.h, b200ann_hnsw_search_epsilon(gpu_.h, ix_, (const float *)queries, nq, dim_, epsilon, min_num_candidates, max_num_candidates,
                                                      filtered ? allow.data() : nullptr, internal_ids ? 1 : 0, lab.data(), (float *)dists,
                                                      cnt.data(), st.data()),
                  "HierarchicalNSW::searchStopConditionClosest");
    count_metrics(st);
    for (size_t i = 0; i < lab.size(); i++) ids[i] = (labeltype)lab[i];
    for (size_t i = 0; i < nq; i++) counts[i] = cnt[i];
  }

  // b200ann extension for the hybrid index: renumber the elements so that internal id == label. The reference's
  // hybrid enters its NSG at getExternalLabel(node) and the NSG's ids are rows of the caller's `data`
  // (index_hnsw_nsg.h:145-150), whatever order addPoint ran in - its driver adds under `omp parallel for`
  // (hnsw_nsg/tests/test_hnsw_nsg_search.cpp:331-333), so internal id != label there. One device index holds both
  // graphs here, so its rows must be in label order. Throws unless the labels are a permutation of 0..n-1 (the
  // reference would read `data` out of bounds). Returns false when nothing had to move.
  bool relabel_to_labels() {
    std::unique_lock<std::mutex> lock(mu_);
    build_locked(nullptr);
    const size_t n = cur_element_count;
    bool identity = true;
    for (size_t i = 0; i < n && identity; i++) identity = labels_[i] == i;
    if (identity) return false;
    sync_host_graph_locked();
    identity = true;
    std::vector<unsigned char> seen(n, 0);
    for (size_t i = 0; i < n; i++) {
      const size_t l = labels_[i];
      if (l >= n || seen[l]) throw std::runtime_error("HNSW_NSG: labels must be the row numbers 0..n-1 of the NSG's base set");
      seen[l] = 1;
      identity &= l == i;
    }
    if (identity) return false;
    auto map = [&](uint32_t v) { return v == B200ANN_INVALID ? v : (uint32_t)labels_[v]; };
    std::vector<float> vec(vec_.size());
    std::vector<uint32_t> l0(links0_.size()), up(n);
    std::vector<int> lv(n);
    std::vector<unsigned char> del(n);
    for (size_t i = 0; i < n; i++) {
      const size_t t = labels_[i];
      std::memcpy(&vec[t * rw_], &vec_[i * rw_], rw_ * sizeof(float));
      for (size_t j = 0; j < maxM0_; j++) l0[t * maxM0_ + j] = map(links0_[i * maxM0_ + j]);
      lv[t] = element_levels_[i]; up[t] = upper_ptr_[i]; del[t] = deleted_[i];
    }
    for (auto &v : upper_links_) v = map(v);
    std::set<tableint> ds;
    for (tableint d : deleted_set_) ds.insert((tableint)labels_[d]);
    enterpoint_node_ = (tableint)labels_[enterpoint_node_];
    vec_.swap(vec); links0_.swap(l0); upper_ptr_.swap(up); element_levels_.swap(lv); deleted_.swap(del); deleted_set_.swap(ds);
    for (size_t i = 0; i < n; i++) { labels_[i] = i; label_lookup_[i] = i; }
    upload();
    return true;
  }

  labeltype getExternalLabel(tableint internal_id) const { return labels_[internal_id]; }
  const float *getDataByInternalId(tableint internal_id) const { return &vec_[(size_t)internal_id * rw_]; }  // byte rows for <int>
  b200ann_handle_t handle() const { return gpu_.h; }
  b200ann_index_t device_index() const { return ix_; }
  size_t generation() const { return gen_; }  // bumped whenever the device index is re-created

 private:
  // bit i set <=> !isMarkedDeleted(i) && (*isIdAllowed)(getExternalLabel(i)) (hnswalg.h:325-326, :406-407)
  std::vector<uint32_t> allow_bitmap(BaseFilterFunctor *f) const {
    const size_t n = cur_element_count;
    std::vector<uint32_t> bits((n + 31) / 32, 0u);
    for (size_t i = 0; i < n; i++)
      if (!deleted_[i] && (!f || (*f)(labels_[i]))) bits[i >> 5] |= 1u << (i & 31);
    return bits;
  }
  static const unsigned char DELETE_MARK = 0x01;  // hnswalg.h:21
  std::vector<unsigned char> deleted_;
  size_t num_deleted_ = 0;
  bool allow_replace_deleted_ = false;
  std::set<tableint> deleted_set_;  // internal ids marked deleted (kept only with allow_replace_deleted_, like :71)
  template <typename T> static void rd(std::istream &in, T &v) { in.read((char *)&v, sizeof(T)); }
  template <typename T> static void wr(std::ostream &out, const T &v) { out.write((const char *)&v, sizeof(T)); }
  void upload() {
    gen_++;
    if (ix_) { b200ann_index_destroy(ix_); ix_ = nullptr; }
    const size_t n = cur_element_count;
    if (n == 0) return;
    detail::check(gpu_.h, create_index(n, &ix_), "b200ann_index_create");
    std::vector<int32_t> lv(element_levels_.begin(), element_levels_.end());
    std::vector<uint64_t> lab(labels_.begin(), labels_.end());
    detail::check(gpu_.h, b200ann_index_set_hnsw(ix_, links0_.data(), maxM0_, lv.data(), upper_ptr_.data(),
                                                 upper_links_.empty() ? nullptr : upper_links_.data(), upper_links_.size() / (maxM_ ? maxM_ : 1),
                                                 maxM_, lab.data(), enterpoint_node_, maxlevel_), "b200ann_index_set_hnsw");
  }
  void count_metrics(const std::vector<uint32_t> &st) const {
    long nd = 0, nh = 0;
    for (size_t i = 0; i + 1 < st.size(); i += 2) { nd += st[i]; nh += st[i + 1]; }
    metric_distance_computations += nd;
    metric_hops += nh;
  }
  // Brings the device graph up to date with the staged points; caller holds mu_. Points appended since the last call
  // are linked into the live graph in place (b200ann_hnsw_insert_batch); an updated vector or a replaced deleted
  // slot re-links everything (b200ann_hnsw_build).
  void build_locked(uint64_t *info8) const {
    HierarchicalNSW *self = const_cast<HierarchicalNSW *>(this);
    const size_t n = cur_element_count;
    if (n == 0) return;
    if (ix_ && !needs_rebuild_ && built_count_ == n && pending_upd_.empty()) return;
    if (ix_ && !needs_rebuild_ && built_count_ > 0) {
      if (built_count_ < n) {  // appended points: linked into the live graph
        std::vector<uint64_t> lab(labels_.begin() + built_count_, labels_.end());
        detail::check(gpu_.h, b200ann_hnsw_insert_batch(gpu_.h, ix_, &vec_[built_count_ * rw_], row_stride(), lab.data(), n - built_count_,
                                                        ef_construction_, info8), "HierarchicalNSW::addPoint (b200ann_hnsw_insert_batch)");
      }
      if (!pending_upd_.empty()) {  // moved points (addPoint on an existing label, a replaced deleted slot): updatePoint in place
        std::vector<uint32_t> ids(pending_upd_.begin(), pending_upd_.end());
        std::vector<float> rows(ids.size() * rw_);
        for (size_t i = 0; i < ids.size(); i++) std::memcpy(&rows[i * rw_], &vec_[(size_t)ids[i] * rw_], rw_ * sizeof(float));
        detail::check(gpu_.h, b200ann_hnsw_update_batch(gpu_.h, ix_, ids.data(), rows.data(), row_stride(), ids.size(), info8),
                      "HierarchicalNSW::addPoint (b200ann_hnsw_update_batch)");
        if (!pending_lab_.empty()) {
          std::vector<uint32_t> lid(pending_lab_.begin(), pending_lab_.end());
          std::vector<uint64_t> lab(lid.size());
          for (size_t i = 0; i < lid.size(); i++) lab[i] = labels_[lid[i]];
          detail::check(gpu_.h, b200ann_index_set_labels(ix_, lid.data(), lab.data(), lid.size()), "b200ann_index_set_labels");
        }
      }
    } else {
      if (self->ix_) { b200ann_index_destroy(self->ix_); self->ix_ = nullptr; }
      detail::check(gpu_.h, create_index(n, &self->ix_), "b200ann_index_create");
      std::vector<uint64_t> lab(labels_.begin(), labels_.end());
      detail::check(gpu_.h, b200ann_hnsw_build(gpu_.h, ix_, M_, ef_construction_, (uint32_t)seed_, 0, 0, lab.data(), info8),
                    "HierarchicalNSW::addPoint (b200ann_hnsw_build)");
    }
    uint64_t shape[5];
    detail::check(gpu_.h, b200ann_index_hnsw_shape(ix_, shape), "b200ann_index_hnsw_shape");
    self->maxlevel_ = (int)shape[3];
    self->enterpoint_node_ = (tableint)shape[4];
    self->gen_++;
    self->built_count_ = n;
    self->needs_rebuild_ = false;
    self->host_graph_valid_ = false;
    self->pending_upd_.clear();
    self->pending_lab_.clear();
  }
  // host copy of the graph (saveIndex, relabel_to_labels): fetched from the device only when needed; caller holds mu_
  void sync_host_graph_locked() const {
    if (host_graph_valid_ || !ix_) return;
    HierarchicalNSW *self = const_cast<HierarchicalNSW *>(this);
    const size_t n = cur_element_count;
    uint64_t shape[5];
    detail::check(gpu_.h, b200ann_index_hnsw_shape(ix_, shape), "b200ann_index_hnsw_shape");
    self->links0_.assign(n * maxM0_, B200ANN_INVALID);
    self->upper_ptr_.assign(n, 0);
    self->upper_links_.assign((size_t)shape[2] * maxM_, B200ANN_INVALID);
    std::vector<int32_t> lv(n);
    detail::check(gpu_.h, b200ann_index_get_hnsw(ix_, self->links0_.data(), lv.data(), self->upper_ptr_.data(),
                                                 shape[2] ? self->upper_links_.data() : nullptr, nullptr), "b200ann_index_get_hnsw");
    self->element_levels_.assign(lv.begin(), lv.end());
    self->host_graph_valid_ = true;
  }
  detail::Gpu gpu_;
  int metric_ = 0;
  size_t dim_ = 0;
  size_t data_size_ = 0;  // bytes per element: dim floats, or dim unsigned bytes for HierarchicalNSW<int> / L2SpaceI
  size_t rw_ = 0;         // floats of vec_ one row occupies (byte rows are padded to 4 bytes)
  // space_data_size = SpaceInterface::get_data_size(): dim floats, dim bytes, or (multi-vector spaces) dim floats + a
  // 32-bit document id that travels with the row (file format, getDataByLabel) and is skipped by the distance
  void set_geometry(size_t space_data_size) {
    const bool u8 = metric_ == B200ANN_L2_U8;
    data_size_ = u8 ? dim_ : space_data_size;
    if (!u8 && (data_size_ < dim_ * sizeof(float) || data_size_ % 4 != 0)) throw std::runtime_error("b200ann: unsupported element size of the space");
    rw_ = (data_size_ + 3) / 4;
  }
  size_t row_stride() const { return metric_ == B200ANN_L2_U8 ? rw_ * 4 : rw_; }  // elements between rows of vec_
  int create_index(size_t n, b200ann_index_t *out) const {
    if (metric_ == B200ANN_L2_U8) return b200ann_index_create_u8(gpu_.h, (const uint8_t *)vec_.data(), n, dim_, rw_ * 4, out);
    return b200ann_index_create(gpu_.h, vec_.data(), n, dim_, rw_, metric_, out);
  }
  size_t seed_ = 100;
  size_t gen_ = 0;
  std::vector<float> vec_;
  std::vector<labeltype> labels_;
  std::unordered_map<labeltype, size_t> label_lookup_;
  std::vector<uint32_t> links0_, upper_ptr_, upper_links_;
  b200ann_index_t ix_ = nullptr;
  mutable size_t built_count_ = 0;        // elements linked into the device graph
  mutable bool needs_rebuild_ = false;    // the device graph must be built from scratch on next use
  mutable std::set<tableint> pending_upd_;  // linked elements whose vector changed: b200ann_hnsw_update_batch on next use
  mutable std::set<tableint> pending_lab_;  // ... whose label changed as well (replaced deleted slots)
  mutable bool host_graph_valid_ = false; // links0_ / upper_* / element_levels_ mirror the device graph
  mutable std::mutex mu_;
};

}  // namespace hnswlib
#endif

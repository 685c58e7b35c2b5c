// C++ host shim mirroring the reference's hybrid index over the b200ann C ABI.
// Same class name, signatures, argument meaning and error behaviour as
//   hnsw_nsg/include/index_hnsw_nsg.h:7-160   hnsw_nsg::HNSW_NSG<dist_t>
// (HNSW levels >= 1 choose an entry point, the NSG graph replaces HNSW layer 0). Header-only; link with
// -lb200ann. Everything runs on the GPU: addPoint -> b200ann_hnsw_build (batched, on first use), Build_NSG ->
// b200ann_nsg_build, searchKnn -> b200ann_hybrid_search (ONE kernel: greedy descent + SearchFromEnterpoint).
//
// Differences a maintainer should know (results are unchanged):
//   - the reference's `res` member grows by one vector per query forever (index_hnsw_nsg.h:100); not kept.
//   - `data_` may be null in the reference's load ctor path if the caller passes null; here the vectors are the
//     ones the HNSW part owns (addPoint copies them / the index file holds them), `data_` is informational.
//   - searchKnnBatch(queries, nq, k, ids) is the batched form the GPU wants; searchKnn is a batch of one.
#ifndef B200ANN_HNSW_NSG_HPP
#define B200ANN_HNSW_NSG_HPP
#include <algorithm>
#include <string>
#include <vector>
#include "efanna2e.hpp"
#include "hnswlib.hpp"

namespace hnsw_nsg {

using efanna2e::Parameters;
using hnswlib::BaseFilterFunctor;
using hnswlib::InnerProductSpace;
using hnswlib::L2Space;
using hnswlib::labeltype;
using hnswlib::SpaceInterface;
using hnswlib::tableint;
template <typename dist_t>
using HierarchicalNSW = hnswlib::HierarchicalNSW<dist_t>;

template <typename dist_t>
class HNSW_NSG {
 public:
  HierarchicalNSW<dist_t> hnsw_;
  float *data_;
  float *query_data_;
  size_t dim_;
  Parameters parameters_;   // "L_search" (+ optional "seed") used by every search
  unsigned point_num_;

  // build ctor (index_hnsw_nsg.h:36-46)
  HNSW_NSG(SpaceInterface<dist_t> *space, unsigned int dim, size_t max_elements, float *data, float *query_data,
           const Parameters &parameters, unsigned hnsw_M = 16, size_t hnsw_ef_construction = 200, unsigned nsg_width = 20,
           int device = 0)
      : hnsw_(space, max_elements, hnsw_M, hnsw_ef_construction, 100, false, device),
        data_(data), query_data_(query_data), dim_(dim), parameters_(parameters), point_num_((unsigned)max_elements) {
    (void)nsg_width;
  }
  // load ctor (:48-61): "<prefix>_hnsw.bin" written by HierarchicalNSW::saveIndex, "<prefix>_nsg.bin" by IndexNSG::Save
  HNSW_NSG(SpaceInterface<dist_t> *space, unsigned int dim, size_t max_elements, float *data, float *query_data,
           const Parameters &parameters, std::string hnsw_index_path, std::string nsg_index_path, int device = 0)
      : hnsw_(space, hnsw_index_path, false, max_elements, false, device),
        data_(data), query_data_(query_data), dim_(dim), parameters_(parameters), point_num_((unsigned)max_elements) {
    load_nsg(nsg_index_path.c_str());
  }
  ~HNSW_NSG() {}

  void setEf(size_t ef) { hnsw_.ef_ = ef; }  // :68-70

  // :79-82: points go to the HNSW part only; labels must be the row numbers of the NSG's base set
  void addPoint(const void *data_point, labeltype label, bool replace_deleted = false) {
    hnsw_.addPoint(data_point, label, replace_deleted);
  }

  // :72-74 -> IndexNSG::Build (hnsw_nsg/src/index_nsg.cpp): "L", "R", "C", "nn_graph_path" (efanna .graph file).
  // The NSG is built over the vectors the HNSW part holds, renumbered so that internal id == label == row of `data`
  // (the reference's driver adds points under omp parallel for, so arrival order is not row order).
  void Build_NSG(const float *data, Parameters &parameters) {
    (void)data;  // == the rows addPoint copied, in label order after the renumbering below
    hnsw_.build();
    hnsw_.relabel_to_labels();  // NSG ids are rows of `data` = labels (index_hnsw_nsg.h:145-150): internal id := label
    checked_gen_ = hnsw_.generation();
    std::string path = parameters.template Get<std::string>("nn_graph_path");
    unsigned L = parameters.template Get<unsigned>("L"), R = parameters.template Get<unsigned>("R"),
             C = parameters.template Get<unsigned>("C");
    size_t kn; unsigned k;
    std::vector<uint32_t> knn = efanna2e::load_nn_graph_file(path.c_str(), kn, k);
    if (kn != hnsw_.cur_element_count) throw std::runtime_error("nn_graph_path holds a different number of nodes than the index");
    uint64_t info[8];
    hnswlib::detail::check(hnsw_.handle(), b200ann_nsg_build(hnsw_.handle(), hnsw_.device_index(), knn.data(), k, L, R, C,
                                                             parameters.template Get<unsigned>("seed", 0u), B200ANN_INVALID, info),
                           "HNSW_NSG::Build_NSG");
    ep_ = (unsigned)info[0]; width_ = (unsigned)info[1];
    fetch_graph();
    installed_gen_ = hnsw_.generation();
  }

  // :107-151: K = max(k, ef_) ids (NSG-local ids == labels); filter is ignored by the reference too
  std::vector<unsigned> searchKnn(const void *query, size_t k, BaseFilterFunctor *filter = nullptr) {
    (void)filter;
    if (hnsw_.cur_element_count == 0) return {};
    size_t K = std::max(k, hnsw_.ef_);
    std::vector<unsigned> ids(K);
    searchKnnBatch((const float *)query, 1, k, ids.data());
    return ids;
  }
  // ids [nq, max(k, ef_)]
  void searchKnnBatch(const float *queries, size_t nq, size_t k, unsigned *ids) {
    size_t K = std::max(k, hnsw_.ef_);
    install();
    hnswlib::detail::check(hnsw_.handle(),
                           b200ann_hybrid_search(hnsw_.handle(), hnsw_.device_index(), queries, nq, dim_, K,
                                                 parameters_.template Get<unsigned>("L_search"),
                                                 parameters_.template Get<unsigned>("seed", 0u), ids, nullptr, nullptr, nullptr),
                           "HNSW_NSG::searchKnn");
  }

  // :153-159
  void saveIndex(const std::string &location) {
    hnsw_.saveIndex(location + "_hnsw.bin");
    if (graph_.empty()) throw std::runtime_error("HNSW_NSG::saveIndex: no NSG graph (Build_NSG or the load ctor first)");
    std::ofstream out((location + "_nsg.bin").c_str(), std::ios::binary | std::ios::out);  // IndexNSG::Save, index_nsg.cpp:37-49
    if (!out) throw std::runtime_error("cannot open " + location + "_nsg.bin");
    out.write((const char *)&width_, 4);
    out.write((const char *)&ep_, 4);
    const size_t n = graph_.size() / gs_;
    for (size_t i = 0; i < n; i++) {
      unsigned k = 0;
      while (k < gs_ && graph_[i * gs_ + k] != B200ANN_INVALID) k++;
      out.write((const char *)&k, 4);
      out.write((const char *)&graph_[i * gs_], (std::streamsize)k * 4);
    }
  }

  unsigned nsg_entry_point() const { return ep_; }
  unsigned nsg_width() const { return width_; }

 private:
  void load_nsg(const char *filename) {  // IndexNSG::Load, index_nsg.cpp:51-68
    std::ifstream in(filename, std::ios::binary);
    if (!in) throw std::runtime_error(std::string("cannot open ") + filename);
    in.read((char *)&width_, 4);
    in.read((char *)&ep_, 4);
    std::vector<std::vector<uint32_t>> rows;
    unsigned k, maxk = 1;
    while (in.read((char *)&k, 4)) {
      std::vector<uint32_t> r(k);
      in.read((char *)r.data(), (std::streamsize)k * 4);
      maxk = std::max(maxk, k);
      rows.push_back(std::move(r));
    }
    gs_ = ((size_t)maxk + 7) / 8 * 8;
    graph_.assign(rows.size() * gs_, B200ANN_INVALID);
    for (size_t i = 0; i < rows.size(); i++) std::copy(rows[i].begin(), rows[i].end(), graph_.begin() + i * gs_);
    installed_gen_ = (size_t)-1;
  }
  void fetch_graph() {
    uint32_t ep, w;
    b200ann_index_t ix = hnsw_.device_index();
    hnswlib::detail::check(hnsw_.handle(), b200ann_index_get_graph(ix, nullptr, 0, &ep, &w), "b200ann_index_get_graph");
    gs_ = w ? w : 1;
    graph_.resize(hnsw_.cur_element_count * gs_);
    hnswlib::detail::check(hnsw_.handle(), b200ann_index_get_graph(ix, graph_.data(), gs_, &ep, &w), "b200ann_index_get_graph");
  }
  // both graphs live in ONE device index; re-install the NSG when the HNSW part rebuilt its index
  void install() {
    hnsw_.build();
    if (hnsw_.generation() != checked_gen_) {  // e.g. a reference-written file whose points were added in parallel
      hnsw_.relabel_to_labels();
      checked_gen_ = hnsw_.generation();
    }
    b200ann_index_t ix = hnsw_.device_index();
    if (!ix) throw std::runtime_error("HNSW_NSG: empty index");
    if (hnsw_.generation() == installed_gen_) return;
    if (graph_.empty()) throw std::runtime_error("HNSW_NSG::searchKnn: no NSG graph (Build_NSG or the load ctor first)");
    if (graph_.size() / gs_ != hnsw_.cur_element_count) throw std::runtime_error("HNSW_NSG: the NSG and HNSW parts hold different numbers of points");
    hnswlib::detail::check(hnsw_.handle(), b200ann_index_set_graph(ix, graph_.data(), gs_, ep_), "b200ann_index_set_graph");
    installed_gen_ = hnsw_.generation();
  }
  std::vector<uint32_t> graph_;  // host copy of the NSG adjacency [n, gs_]
  size_t gs_ = 1;
  unsigned ep_ = 0, width_ = 0;
  size_t installed_gen_ = (size_t)-1;  // generation of the HNSW part's device index that holds the NSG graph
  size_t checked_gen_ = (size_t)-1;    // generation whose internal ids were checked / renumbered to equal the labels
};

}  // namespace hnsw_nsg
#endif

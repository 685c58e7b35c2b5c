// Host-side C++ mirror of the reference's efanna2e API for the hot path, over the b200ann C ABI.
// Same class names, signatures, argument meaning and error behaviour as
//   CNNS/include/nsg/parameters.h:15-57   efanna2e::Parameters
//   CNNS/include/nsg/index.h:19-52        efanna2e::Index
//   CNNS/include/nsg/index_nsg.h:17-85    efanna2e::IndexNSG   (Build/Search/Save/Load/OptimizeGraph/SearchWithOptGraph)
//   CNNS/efanna_graph/include/efanna2e/index_graph.h:23-72  efanna2e::IndexGraph (Build/Save/Load)
// so a driver written against the reference (e.g. CNNS/tests/nndescent_nsg.cpp) compiles against
// this header unchanged and runs on the GPU. Header-only; link with -lb200ann. No CPU fallback:
// constructing an index without a CUDA device throws std::runtime_error.
//
// Added (batched) entry points, modelled on hnswlib's python batch shape (bindings.cpp:612-695):
//   IndexNSG::SearchBatch(queries, nq, K, params, indices[nq*K])
#ifndef B200ANN_EFANNA2E_HPP
#define B200ANN_EFANNA2E_HPP
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <unordered_map>
#include <vector>
#include "../b200ann.h"

namespace efanna2e {

enum Metric { L2 = 0, INNER_PRODUCT = 1, FAST_L2 = 2, PQ = 3 };  // distance.h:11-16

// string -> string map with typed accessors (parameters.h:15-57); Get throws
// std::invalid_argument on a missing key and std::runtime_error on a bad conversion
class Parameters {
 public:
  template <typename T>
  void Set(const std::string &name, const T &value) {
    std::ostringstream os;
    os << value;
    kv_[name] = os.str();
  }
  template <typename T>
  T Get(const std::string &name) const {
    auto it = kv_.find(name);
    if (it == kv_.end()) throw std::invalid_argument("Invalid parameter name.");
    std::istringstream is(it->second);
    T v;
    if (!(is >> v) || !is.eof())
      throw std::runtime_error("Failed to convert value '" + it->second + "' to type: " + typeid(T).name());
    return v;
  }
  template <typename T>
  T Get(const std::string &name, const T &dflt) const {
    return kv_.count(name) ? Get<T>(name) : dflt;
  }
  bool Has(const std::string &name) const { return kv_.count(name) != 0; }

 private:
  std::unordered_map<std::string, std::string> kv_;
};

namespace detail {
inline void check(b200ann_handle_t h, int rc, const char *what) {
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + (h ? b200ann_last_error(h) : "no CUDA device (b200ann has no CPU fallback)"));
}
// one handle per index object (handles are per GPU and per host thread; calls are serialised per handle)
struct Gpu {
  b200ann_handle_t h = nullptr;
  explicit Gpu(int device = 0) { check(nullptr, b200ann_create(device, &h), "b200ann_create"); }
  ~Gpu() { b200ann_destroy(h); }
  Gpu(const Gpu &) = delete;
  Gpu &operator=(const Gpu &) = delete;
};
}  // namespace detail

class Index {
 public:
  explicit Index(const size_t dimension, const size_t n, Metric metric) : dimension_(dimension), nd_(n), has_built(false), metric_(metric) {}
  virtual ~Index() {}
  virtual void Build(size_t n, const float *data, const Parameters &parameters) = 0;
  virtual void Search(const float *query, const float *x, size_t k, const Parameters &parameters, unsigned *indices) = 0;
  virtual void Save(const char *filename) = 0;
  virtual void Load(const char *filename) = 0;
  bool HasBuilt() const { return has_built; }
  size_t GetDimension() const { return dimension_; }
  size_t GetSizeOfDataset() const { return nd_; }
  const float *GetDataset() const { return data_; }

 protected:
  const size_t dimension_;
  const float *data_ = nullptr;
  size_t nd_;
  bool has_built;
  Metric metric_;
};

// efanna kNN-graph file: per node [u32 K][u32 x K] (index_graph.cpp:348-377, read by
// IndexNSG::Load_nn_graph, index_nsg.cpp:128-148)
inline std::vector<uint32_t> load_nn_graph_file(const char *filename, size_t &n, unsigned &k) {
  std::ifstream in(filename, std::ios::binary);
  if (!in) throw std::runtime_error(std::string("cannot open kNN graph file ") + filename);
  in.read((char *)&k, 4);
  in.seekg(0, std::ios::end);
  size_t fsize = (size_t)in.tellg();
  n = fsize / ((size_t)k + 1) / 4;
  std::vector<uint32_t> raw((size_t)n * (k + 1));
  in.seekg(0, std::ios::beg);
  in.read((char *)raw.data(), raw.size() * 4);
  std::vector<uint32_t> flat((size_t)n * k);
  for (size_t i = 0; i < n; i++)
    for (unsigned j = 0; j < k; j++) flat[i * k + j] = raw[i * (k + 1) + 1 + j];
  return flat;
}

// efanna2e::IndexNSG (CNNS/include/nsg/index_nsg.h:17-85, CNNS/src/nsg/index_nsg.cpp)
class IndexNSG : public Index {
 public:
  explicit IndexNSG(const size_t dimension, const size_t n, Metric m, Index *initializer, int device = 0)
      : Index(dimension, n, m), initializer_(initializer), gpu_(device) {}
  virtual ~IndexNSG() {
    if (ix_) b200ann_index_destroy(ix_);
    delete initializer_;  // the reference owns its initializer (index_nsg.cpp:22-35)
  }

  // IndexNSG::Build (:465-504). Parameters: "L", "R", "C", "nn_graph_path"; `data` is borrowed and
  // copied to HBM once. The kNN graph file is what IndexGraph::Save writes.
  void Build(size_t n, const float *data, const Parameters &parameters) override {
    std::string path = parameters.Get<std::string>("nn_graph_path");
    unsigned L = parameters.Get<unsigned>("L"), R = parameters.Get<unsigned>("R"), C = parameters.Get<unsigned>("C");
    size_t kn; unsigned k;
    std::vector<uint32_t> knn = load_nn_graph_file(path.c_str(), kn, k);
    if (kn != n) throw std::runtime_error("nn_graph_path holds a different number of nodes than n");
    BuildFromKnn(n, data, knn.data(), k, L, R, C, parameters.Get<unsigned>("seed", 0u));
  }
  // same, kNN graph given in memory: knn [n,K] host or device memory
  void BuildFromKnn(size_t n, const float *data, const uint32_t *knn, unsigned K, unsigned L, unsigned R, unsigned C,
                    unsigned seed = 0) {
    nd_ = n;
    bind(data);
    uint64_t info[8];
    detail::check(gpu_.h, b200ann_nsg_build(gpu_.h, ix_, knn, K, L, R, C, seed, B200ANN_INVALID, info), "IndexNSG::Build");
    ep_ = (unsigned)info[0];
    width = (unsigned)info[1];
    host_graph_.clear();
    has_built = true;
  }

  // IndexNSG::Search (:506-568): "L_search"; K ids into indices; x is the base set (borrowed)
  void Search(const float *query, const float *x, size_t K, const Parameters &parameters, unsigned *indices) override {
    SearchBatchOn(x, query, 1, K, parameters, indices);
  }
  // batch of nq queries [nq,dim] -> indices [nq,K]
  void SearchBatch(const float *queries, size_t nq, size_t K, const Parameters &parameters, unsigned *indices) {
    SearchBatchOn(data_, queries, nq, K, parameters, indices);
  }
  // OptimizeGraph (:659-682) interleaves vectors and lists for the CPU cache; on the GPU the index
  // is already resident, so this only binds `data`. SearchWithOptGraph (:570-657) == Search.
  void OptimizeGraph(float *data) { bind(data); }
  void SearchWithOptGraph(const float *query, size_t K, const Parameters &parameters, unsigned *indices) {
    SearchBatchOn(data_, query, 1, K, parameters, indices);
  }

  // .nsg file (:37-68): [u32 width][u32 ep] then per node [u32 k][u32 x k]
  void Save(const char *filename) override {
    std::vector<uint32_t> g;
    size_t gs = fetch_graph(g);
    std::ofstream out(filename, std::ios::binary | std::ios::out);
    if (!out) throw std::runtime_error(std::string("cannot open ") + filename);
    out.write((const char *)&width, 4);
    out.write((const char *)&ep_, 4);
    for (size_t i = 0; i < nd_; i++) {
      unsigned k = 0;
      while (k < gs && g[i * gs + k] != B200ANN_INVALID) k++;
      out.write((const char *)&k, 4);
      out.write((const char *)&g[i * gs], (std::streamsize)k * 4);
    }
  }
  void Load(const char *filename) override {
    std::ifstream in(filename, std::ios::binary);
    if (!in) throw std::runtime_error(std::string("cannot open ") + filename);
    in.read((char *)&width, 4);
    in.read((char *)&ep_, 4);
    std::vector<std::vector<uint32_t>> rows;
    unsigned k, maxk = 1;
    while (in.read((char *)&k, 4)) {
      std::vector<uint32_t> r(k);
      in.read((char *)r.data(), (std::streamsize)k * 4);
      if (k > maxk) maxk = k;
      rows.push_back(std::move(r));
    }
    nd_ = rows.size();
    host_gs_ = ((size_t)maxk + 7) / 8 * 8;
    host_graph_.assign(nd_ * host_gs_, B200ANN_INVALID);
    for (size_t i = 0; i < nd_; i++)
      for (size_t j = 0; j < rows[i].size(); j++) host_graph_[i * host_gs_ + j] = rows[i][j];
    graph_dirty_ = true;
    has_built = true;
  }

  unsigned GetWidth() const { return width; }
  unsigned GetEntryPoint() const { return ep_; }
  b200ann_handle_t handle() const { return gpu_.h; }
  b200ann_index_t device_index() const { return ix_; }

 protected:
  Index *initializer_ = nullptr;

 private:
  // (re)create the device index when the base pointer changes (Search rebinds data_ = x, :509)
  void bind(const float *data) {
    if (!data) throw std::runtime_error("IndexNSG: no base vectors");
    if (data == data_ && ix_) return;
    std::vector<uint32_t> keep;
    size_t gs = 0;
    if (ix_ && !graph_dirty_ && has_built) { gs = fetch_graph(keep); host_graph_.swap(keep); host_gs_ = gs; graph_dirty_ = true; }
    if (ix_) { b200ann_index_destroy(ix_); ix_ = nullptr; }
    detail::check(gpu_.h, b200ann_index_create(gpu_.h, data, nd_, dimension_, dimension_,
                                                metric_ == INNER_PRODUCT ? B200ANN_IP : B200ANN_L2, &ix_), "b200ann_index_create");
    data_ = data;
  }
  void sync_graph() {
    if (!graph_dirty_) return;
    detail::check(gpu_.h, b200ann_index_set_graph(ix_, host_graph_.data(), host_gs_, ep_), "b200ann_index_set_graph");
    graph_dirty_ = false;
    host_graph_.clear(); host_graph_.shrink_to_fit();
  }
  size_t fetch_graph(std::vector<uint32_t> &g) {
    if (graph_dirty_ || !ix_) { g = host_graph_; return host_gs_; }
    uint32_t ep, w;
    detail::check(gpu_.h, b200ann_index_get_graph(ix_, nullptr, 0, &ep, &w), "b200ann_index_get_graph");
    size_t gs = w ? w : 1;
    g.resize(nd_ * gs);
    detail::check(gpu_.h, b200ann_index_get_graph(ix_, g.data(), gs, &ep, &w), "b200ann_index_get_graph");
    return gs;
  }
  void SearchBatchOn(const float *x, const float *queries, size_t nq, size_t K, const Parameters &parameters, unsigned *indices) {
    if (!has_built) throw std::runtime_error("IndexNSG::Search: index has no graph (Build or Load first)");
    unsigned L = parameters.Get<unsigned>("L_search");
    bind(x);
    sync_graph();
    detail::check(gpu_.h, b200ann_nsg_search(gpu_.h, ix_, queries, nq, dimension_, K, L, parameters.Get<unsigned>("seed", 0u),
                                             indices, nullptr, nullptr), "IndexNSG::Search");
  }

  detail::Gpu gpu_;
  b200ann_index_t ix_ = nullptr;
  std::vector<uint32_t> host_graph_;  // graph waiting for its vectors (after Load) or for a re-bind
  size_t host_gs_ = 0;
  bool graph_dirty_ = false;
  unsigned width = 0;
  unsigned ep_ = 0;
};

// efanna2e::IndexKDtree (efanna_graph/include/efanna2e/index_kdtree.h, src/index_kdtree.cpp:279-435; driver
// efanna_graph/tests/test_kdtree_graph.cpp): the KD-tree initializer. Build(n, data, {"K", "nTrees", "mLevel"}) computes
// an initial kNN graph on the GPU (b200ann_kdtree_graph); Save writes it like the reference ([K][K ids] per row, :436-445).
// Hand it to IndexGraph as the initializer and call RefineGraph. "seed" (default 0) replaces the reference's generators.
class IndexKDtree : public Index {
 public:
  explicit IndexKDtree(const size_t dimension, const size_t n, Metric m, Index *initializer, int device = 0)
      : Index(dimension, n, m), gpu_(device) { (void)initializer; }
  void Build(size_t n, const float *data, const Parameters &parameters) override {
    K_ = parameters.Get<unsigned>("K");
    nd_ = n; data_ = data;
    b200ann_index_t ix = nullptr;
    detail::check(gpu_.h, b200ann_index_create(gpu_.h, data, n, dimension_, dimension_, B200ANN_L2, &ix), "b200ann_index_create");
    knn_.resize(n * (size_t)K_);
    int rc = b200ann_kdtree_graph(gpu_.h, ix, K_, parameters.Get<unsigned>("nTrees"), parameters.Get<unsigned>("mLevel"),
                                  parameters.Get<unsigned>("seed", 0u), knn_.data(), nullptr, info_);
    b200ann_index_destroy(ix);
    detail::check(gpu_.h, rc, "IndexKDtree::Build");
    has_built = true;
  }
  void Search(const float *, const float *, size_t, const Parameters &, unsigned *) override {}  // empty in the reference too (:450-456)
  void Save(const char *filename) override {
    std::ofstream out(filename, std::ios::binary | std::ios::out);
    if (!out) throw std::runtime_error(std::string("cannot open ") + filename);
    for (size_t i = 0; i < nd_; i++) {
      out.write((const char *)&K_, 4);
      out.write((const char *)&knn_[i * K_], (std::streamsize)K_ * 4);
    }
  }
  void Load(const char *) override {}  // empty in the reference (:447-448)
  const std::vector<uint32_t> &Graph() const { return knn_; }  // [n,K] row-major
  unsigned K() const { return K_; }
  const uint64_t *BuildInfo() const { return info_; }

 private:
  detail::Gpu gpu_;
  std::vector<uint32_t> knn_;
  unsigned K_ = 0;
  uint64_t info_[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

// efanna2e::IndexGraph (efanna_graph/include/efanna2e/index_graph.h:23-72): Build produces the kNN
// graph NSG consumes. Like the reference (index_graph.cpp:265-294) Build runs NN-descent with "K","L",
// "iter","S","R" (b200ann_nndescent: InitializeGraph from random ids, iter x {join, update}, top K).
// Two additions: "seed" (default 0) replaces the process-global rand(), and "exact"=1 (or a Build
// without "iter") computes the EXACT K-NN graph with the tensor-core brute-force kernel instead
// (b200ann_knn_graph) - what NN-descent converges towards, and faster than it below a few million rows.
class IndexGraph : public Index {
 public:
  explicit IndexGraph(const size_t dimension, const size_t n, Metric m, Index *initializer, int device = 0)
      : Index(dimension, n, m), initializer_(initializer), gpu_(device) {}
  virtual ~IndexGraph() {}  // does not own the initializer (index_graph.cpp:18-20)

  void Build(size_t n, const float *data, const Parameters &parameters) override {
    K_ = parameters.Get<unsigned>("K");
    nd_ = n; data_ = data;
    b200ann_index_t ix = nullptr;
    detail::check(gpu_.h, b200ann_index_create(gpu_.h, data, n, dimension_, dimension_,
                                                metric_ == INNER_PRODUCT ? B200ANN_IP : B200ANN_L2, &ix), "b200ann_index_create");
    knn_.resize(n * (size_t)K_);
    int rc;
    bool exact = parameters.Get<unsigned>("exact", 0u) != 0 || !parameters.Has("iter");
    if (exact) {
      rc = b200ann_knn_graph(gpu_.h, ix, K_, knn_.data());
    } else {
      rc = b200ann_nndescent(gpu_.h, ix, K_, parameters.Get<unsigned>("L"), parameters.Get<unsigned>("iter"),
                             parameters.Get<unsigned>("S"), parameters.Get<unsigned>("R"),
                             parameters.Get<unsigned>("seed", 0u), knn_.data(), nullptr, info_);
    }
    b200ann_index_destroy(ix);
    detail::check(gpu_.h, rc, "IndexGraph::Build");
    has_built = true;
  }
  // index_graph.cpp:235-262: NN-descent started from the initializer's graph (an IndexKDtree that has been built, or
  // this object's own graph from a previous Build / Load) instead of from random ids -> b200ann_nndescent_refine
  void RefineGraph(const float *data, const Parameters &parameters) {
    const std::vector<uint32_t> *g = &knn_;
    unsigned gk = K_;
    if (auto *kd = dynamic_cast<IndexKDtree *>(initializer_)) {
      if (!kd->HasBuilt()) throw std::runtime_error("IndexGraph::RefineGraph: the initializer has not been built");
      g = &kd->Graph(); gk = kd->K();
    }
    if (g->empty() || g->size() != nd_ * (size_t)gk) {
      if (g == &knn_) throw std::runtime_error("IndexGraph::RefineGraph: no initial graph (an IndexKDtree initializer, or Build / Load first)");
      nd_ = g->size() / gk;
    }
    data_ = data;
    const unsigned K = parameters.Get<unsigned>("K");
    b200ann_index_t ix = nullptr;
    detail::check(gpu_.h, b200ann_index_create(gpu_.h, data, nd_, dimension_, dimension_, B200ANN_L2, &ix), "b200ann_index_create");
    std::vector<uint32_t> out(nd_ * (size_t)K);
    int rc = b200ann_nndescent_refine(gpu_.h, ix, K, parameters.Get<unsigned>("L"), parameters.Get<unsigned>("iter"),
                                      parameters.Get<unsigned>("S"), parameters.Get<unsigned>("R"), parameters.Get<unsigned>("seed", 0u),
                                      g->data(), gk, out.data(), nullptr, info_);
    b200ann_index_destroy(ix);
    detail::check(gpu_.h, rc, "IndexGraph::RefineGraph");
    knn_.swap(out); K_ = K;
    has_built = true;
  }
  // Search over a bare kNN graph is not on the drivers' path (index_graph.cpp:296-346 is only used by
  // GraphAdd); kept for interface completeness
  void Search(const float *, const float *, size_t, const Parameters &, unsigned *) override {
    throw std::runtime_error("IndexGraph::Search is not provided; build an IndexNSG over the graph");
  }
  void Save(const char *filename) override {  // index_graph.cpp:348-357
    std::ofstream out(filename, std::ios::binary | std::ios::out);
    if (!out) throw std::runtime_error(std::string("cannot open ") + filename);
    for (size_t i = 0; i < nd_; i++) {
      out.write((const char *)&K_, 4);
      out.write((const char *)&knn_[i * K_], (std::streamsize)K_ * 4);
    }
  }
  void Load(const char *filename) override {  // index_graph.cpp:359-377
    size_t n;
    knn_ = load_nn_graph_file(filename, n, K_);
    nd_ = n;
    has_built = true;
  }
  const std::vector<uint32_t> &Graph() const { return knn_; }  // [n,K] row-major
  unsigned K() const { return K_; }
  const uint64_t *BuildInfo() const { return info_; }            // b200ann_nndescent's info (0 after an exact build)

 protected:
  Index *initializer_;

 private:
  detail::Gpu gpu_;
  std::vector<uint32_t> knn_;
  unsigned K_ = 0;
  uint64_t info_[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

// efanna2e::IndexRandom (efanna_graph/src/index_random.cpp): the NN-descent initializer object the
// drivers construct and hand to IndexGraph; nothing to do on the GPU path.
class IndexRandom : public Index {
 public:
  IndexRandom(const size_t dimension, const size_t n) : Index(dimension, n, L2) { has_built = true; }
  void Build(size_t, const float *, const Parameters &) override {}
  void Search(const float *, const float *, size_t, const Parameters &, unsigned *) override {}
  void Save(const char *) override {}
  void Load(const char *) override {}
};

}  // namespace efanna2e
#endif

// Shared host-side plumbing for the b200ann C ABI: handle / index structs, error capture,
// host-or-device pointer staging. No torch types; plain CUDA runtime.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <random>
#include <string>
#include <vector>
#include "../../include/b200ann.h"

#define INVALID_ID 0xFFFFFFFFu

struct b200ann_handle {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  uint64_t launches = 0;
  float last_ms = 0.f;
  int visited_bits = 10;
  bool visited_bits_auto = true;  // search v2 sizes the table itself until b200ann_set_visited_bits is called (0 = auto again)
  int link_visited_bits = 0;    // 0 = auto. NSG Link is occupancy-bound by shared memory: 1M x 128 K=200 L=40 C=500 takes
                                // 1.45 / 1.30 / 1.23 / 1.28 s with 2^11 / 2^10 / 2^9 / 2^8 slots (more warps per SM against more
                                // re-evaluations); at dim 960 2^10 wins. auto = 2^9 for dim <= 128, else 2^10; HNSW build 2^11
  int sm_count = 148;
  int knn_mode = 0;          // 0 auto, 1 fp32 SIMT (bf_knn.cu), 2 tensor cores (knn_tc.cu)
  int knn_slack = -1;        // shortlist slack of the tensor path; < 0 = default
  uint64_t knn_stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t *work = nullptr;  // device work-queue counter of the persistent search kernel (search.cu)
  int persistent = 1;        // 1: search warps pull queries from the queue; 0: one query per warp, one launch slot each
  const uint32_t *query_order = nullptr;  // b200ann_set_query_order: device permutation of the next search's work items
  size_t query_order_n = 0;
  int search_prefetch = 1;   // beam.cuh prefetch_next_row (B200ANN_SEARCH_PREFETCH=0 turns it off for A/B runs)
  int visited_compact = 0;   // 1: 16-bit visited-set entries when n allows - twice the slots in the same shared memory, but ~3x the
                             // instructions per probe; measured slower than 32-bit ids with a few % more re-evaluations (B200ANN_VISITED_COMPACT)
  int search_version = 0;    // 0: auto (v2 for L <= 64, else v1); 2: beam2.cuh (TMA row ring, register pool); 1: round-1 kernel (B200ANN_SEARCH_VERSION)
  int search_mode = 0;       // v2: 0 = auto split between one query per warp / per block, 1 = per warp only, 2 = per block only (B200ANN_SEARCH_MODE)
  int search_tail = -1;      // v2 experiments: queries run one per block at the end of a batch (-1 = the launcher's rule; B200ANN_SEARCH_TAIL)
  int ring_stages = 0;       // v2: stages of the per-warp row ring (0 = default 3; B200ANN_RING_STAGES)
  uint32_t id_base = 0;      // b200ann_set_id_base: added to every result id of the NSG-family searches (row shards)
  int search_ctas_per_sm = 0;  // cap on resident blocks per SM of the persistent search kernel (0 = occupancy limit)
  bool async_mode = false;   // b200ann_set_async: device-to-device search / merge calls do not synchronize the stream
};

struct b200ann_index {
  b200ann_handle *h = nullptr;
  size_t n = 0, dim = 0, stride = 0;  // stride in floats, multiple of 4
  int metric = 0;
  float *vec = nullptr;  // [n, stride]
  // flat NSG / kNN graph
  uint32_t *graph = nullptr;
  size_t gstride = 0;
  uint32_t ep = 0, width = 0;
  // flat HNSW
  uint32_t *links0 = nullptr;
  size_t maxM0 = 0;
  int32_t *levels = nullptr;
  uint32_t *upper_ptr = nullptr, *upper_links = nullptr;
  size_t upper_rows = 0, maxM = 0;
  uint64_t *labels = nullptr;
  uint32_t hnsw_ep = 0;
  int maxlevel = -1;
  bool has_hnsw = false;
  // in-place growth (b200ann_hnsw_insert_batch): capacities in elements (0 = exactly what the size needs), the owner of
  // every upper row, the level generator that continues across appends (hnswalg.h:122 level_generator_), M / efC
  size_t cap_vec = 0, cap_links0 = 0, cap_levels = 0, cap_uptr = 0, cap_upper = 0, cap_owner = 0, cap_labels = 0;
  uint32_t *upper_owner = nullptr;
  std::default_random_engine level_rng{100};
  size_t hnsw_M = 0, hnsw_efc = 0;
  // tensor-core kNN cache (knn_tc.cu): bf16 hi/lo copies [n, tc_dpad], fp32 row norms
  void *tc_hi = nullptr, *tc_lo = nullptr;
  void *tc_ext = nullptr;   // [n, 16] bf16: -||row||^2 / 2 as three bf16 terms in columns 0-2 (knn_tc.cu: the norm rides the contraction)
  float *tc_norms = nullptr;
  size_t tc_dpad = 0;
  bool tc_inexact = false;
  float tc_ny_max = 0.f;
};

// B200ANN_L2_U8 indexes (unsigned-byte rows, int squared L2 = hnswlib::L2SpaceI): rows are padded to 16 bytes,
// `stride` stays in 4-byte units so that `vec + id * stride` addresses a row for both element types
inline bool ix_u8(const b200ann_index *ix) { return ix->metric == B200ANN_L2_U8; }
inline int ix_nch(const b200ann_index *ix) { return (int)(ix_u8(ix) ? (ix->dim + 15) / 16 : (ix->dim + 3) / 4); }  // 16-byte chunks per row
inline size_t ix_row_bytes(const b200ann_index *ix) { return ix_u8(ix) ? ix->dim : ix->dim * 4; }               // payload bytes per row

inline int set_err(b200ann_handle *h, int code, const std::string &msg) {
  if (h) h->err = msg;
  return code;
}

// entry points that only exist for float rows
#define FLOAT_ONLY(h, ix, name)                                                                                        \
  do {                                                                                                                 \
    if ((ix) && (ix)->metric == B200ANN_L2_U8)                                                                         \
      return set_err(h, B200ANN_ERR_ARG, name ": the index holds unsigned-byte rows (b200ann_index_create_u8); only the HNSW build / search entry points take it"); \
  } while (0)

#define CU_CHECK(h, call)                                                                         \
  do {                                                                                            \
    cudaError_t e__ = (call);                                                                     \
    if (e__ != cudaSuccess) {                                                                     \
      char b__[512];                                                                              \
      snprintf(b__, sizeof b__, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
               __LINE__);                                                                         \
      return set_err(h, B200ANN_ERR_CUDA, b__);                                                   \
    }                                                                                             \
  } while (0)

inline bool is_device_ptr(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// RAII device scratch buffer (freed with the stream ordered allocator)
template <typename T>
struct DevBuf {
  T *p = nullptr;
  cudaStream_t s = nullptr;
  cudaError_t alloc(size_t count, cudaStream_t st) {
    s = st;
    return cudaMallocAsync((void **)&p, (count ? count : 1) * sizeof(T), st);
  }
  ~DevBuf() {
    if (p) cudaFreeAsync(p, s);
  }
};

// Input staging: returns a device pointer for `src` (copying host memory to a scratch buffer).
template <typename T>
struct InArg {
  DevBuf<T> buf;
  const T *d = nullptr;
  cudaError_t stage(const T *src, size_t count, cudaStream_t st) {
    if (!src) { d = nullptr; return cudaSuccess; }
    if (is_device_ptr(src)) { d = src; return cudaSuccess; }
    cudaError_t e = buf.alloc(count, st);
    if (e != cudaSuccess) return e;
    d = buf.p;
    return cudaMemcpyAsync(buf.p, src, count * sizeof(T), cudaMemcpyHostToDevice, st);
  }
};

// Output staging: kernels write to .d; finish() copies back when the destination is host memory.
template <typename T>
struct OutArg {
  DevBuf<T> buf;
  T *d = nullptr, *dst = nullptr;
  size_t count = 0;
  bool host = false;
  cudaError_t stage(T *out, size_t cnt, cudaStream_t st) {
    dst = out; count = cnt;
    if (!out) { d = nullptr; return cudaSuccess; }
    if (is_device_ptr(out)) { d = out; return cudaSuccess; }
    host = true;
    cudaError_t e = buf.alloc(cnt, st);
    d = buf.p;
    return e;
  }
  cudaError_t finish(cudaStream_t st) {
    if (host && dst) return cudaMemcpyAsync(dst, d, count * sizeof(T), cudaMemcpyDeviceToHost, st);
    return cudaSuccess;
  }
};

// kernel launchers implemented in the .cu files (all asynchronous on h->stream)
int launch_nsg_search(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                      size_t K, size_t L, uint32_t seed, uint32_t *ids, float *dists, uint32_t *stats,
                      uint32_t qidx_base = 0);
int launch_nsg_search_from(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                           size_t K, size_t L, uint32_t seed, const uint32_t *entry, uint32_t *entry_out, uint32_t *ids,
                           float *dists, uint32_t *stats);
int launch_nsg_search_clustered(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                                size_t K, size_t L, uint32_t seed, const uint32_t *ep, const uint32_t *base,
                                const uint32_t *size, uint32_t *ids, float *dists, uint32_t *stats);
int launch_hnsw_search(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                       size_t k, size_t ef, uint64_t *labels, float *dists, uint32_t *stats);
int launch_gather_probe(b200ann_handle *h, const b200ann_index *ix, uint64_t rows, int inflight, int wpb, int bpsm,
                        uint32_t seed, float *sink);
int launch_hnsw_search_filtered(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                                size_t k, size_t ef, const uint32_t *allow_bits, size_t ccap, bool big,
                                const uint8_t *only, uint64_t *labels, float *dists, uint32_t *stats, uint8_t *inexact,
                                size_t *ccap_used, float eps = INFINITY, size_t min_cand = 0, uint32_t *counts = nullptr,
                                bool internal_ids = false, const uint32_t *doc = nullptr, size_t efcoll = 0, size_t docs_keep = 0);
int launch_merge_topk(b200ann_handle *h, const float *d, const uint32_t *ids, size_t parts, size_t nq, size_t k, size_t kout,
                      float *out_d, uint32_t *out_i);
int launch_bf_knn(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                  size_t k, bool exclude_self, uint32_t *ids, float *dists, const uint32_t *self_ids = nullptr,
                  uint32_t self_base = 0);
// u8 rows, integer squared L2 (bf_knn_u8.cu); device pointers only
int launch_bf_knn_u8(b200ann_handle *h, const uint8_t *base, size_t n, size_t dim, size_t bstride, const uint8_t *q,
                     size_t nq, size_t qstride, size_t k, uint32_t *ids, int32_t *dists);
// tensor-core path (knn_tc.cu); same contract. knn_dispatch picks the path from h->knn_mode.
int launch_knn_tc(b200ann_handle *h, b200ann_index *ix, const float *q, size_t nq, size_t qstride, size_t k,
                  bool exclude_self, uint32_t *ids, float *dists, uint32_t self_base = 0);
bool knn_tc_supported(const b200ann_index *ix, size_t nq, size_t k);
void knn_tc_free(b200ann_index *ix);

// HNSW construction (hnsw_build.cu): builds and installs the graph in the index
int hnsw_build_launch(b200ann_handle *h, b200ann_index *ix, size_t M, size_t efc, uint32_t seed, size_t ratio, size_t bmax,
                      const uint64_t *labels_host_or_dev, uint64_t *info8);
int hnsw_update_launch(b200ann_handle *h, b200ann_index *ix, const uint32_t *ids, size_t nb, const void *data, size_t dstride_bytes,
                       uint64_t *info8);
int hnsw_insert_launch(b200ann_handle *h, b200ann_index *ix, const void *data, size_t dstride_bytes, size_t nb,
                       const uint64_t *labels_host_or_dev, size_t ratio, size_t bmax, uint64_t *info8);

// NN-descent (nndescent.cu); device pointers only
int nndescent_launch(b200ann_handle *h, const b200ann_index *ix, size_t K, size_t L, size_t iter, size_t S, size_t R,
                     uint32_t seed, uint32_t *knn, float *knn_d, uint64_t *info8, const uint32_t *init_graph = nullptr,
                     size_t init_k = 0);

int kdtree_launch(b200ann_handle *h, const b200ann_index *ix, size_t K, size_t nTrees, size_t ml, uint32_t seed, uint32_t *knn,
                  float *knn_d, uint64_t *info8);

// NSG build phases (nsg_build.cu); device pointers only
int nsg_centroid_launch(b200ann_handle *h, const b200ann_index *ix, float *center);
int nsg_link_launch(b200ann_handle *h, const b200ann_index *ix, const uint32_t *knn, size_t kstride, uint32_t ep,
                    size_t L, size_t R, size_t C, uint32_t seed, uint32_t *cut_ids, float *cut_d, uint32_t *stats);
int nsg_interinsert_launch(b200ann_handle *h, const b200ann_index *ix, size_t R, uint32_t *cut_ids, float *cut_d,
                           unsigned long long *n_eval_dev);
int nsg_finish(b200ann_handle *h, b200ann_index *ix, const uint32_t *cut_ids, const float *cut_d, size_t R, uint32_t ep,
               size_t L, uint32_t seed, size_t slack, uint32_t *added_out);

// extern "C" boundary of libb200ann.so (declared in include/b200ann.h). Host-side only: argument
// checks, host<->device staging, stream/event bookkeeping. All arithmetic is in the kernels.
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <cstring>
#include <algorithm>
#include <vector>
#include "common.cuh"

namespace {

struct CallTimer {  // device time of the kernels of one API call
  b200ann_handle *h;
  explicit CallTimer(b200ann_handle *hh) : h(hh) { cudaEventRecord(h->ev0, h->stream); }
  void stop() { cudaEventRecord(h->ev1, h->stream); }
  cudaError_t sync() {
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e == cudaSuccess) cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return e;
  }
};

template <typename T>
cudaError_t upload(T **dst, const T *src, size_t count, cudaStream_t st) {
  cudaError_t e = cudaMalloc((void **)dst, (count ? count : 1) * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpyAsync(*dst, src, count * sizeof(T), cudaMemcpyDefault, st);
}

__global__ void pad_rows_kernel(const float *src, size_t sstride, float *dst, size_t dstride, size_t n, size_t dim) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t total = n * dstride;
  for (; i < total; i += (size_t)gridDim.x * blockDim.x) {
    size_t r = i / dstride, c = i % dstride;
    dst[i] = c < dim ? src[r * sstride + c] : 0.f;
  }
}

// u8 indexes: distances are int32 bit patterns in the float-typed output; an unfilled slot holds the kernels' +inf
// pattern and becomes INT32_MAX like b200ann_bf_knn_u8's
__global__ void u8_pad_kernel(float *d, size_t count) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (; i < count; i += (size_t)gridDim.x * blockDim.x)
    if (__float_as_uint(d[i]) == 0x7F800000u) d[i] = __uint_as_float(0x7FFFFFFFu);
}

// labels of single elements (a replaced deleted slot gets a new label, hnswalg.h:982-987)
__global__ void set_labels_kernel(uint64_t *labels, const uint32_t *ids, const uint64_t *lab, size_t nb) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < nb) labels[ids[i]] = lab[i];
}
__global__ void iota_labels_kernel(uint64_t *p, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = i;
}

__global__ void regraph_kernel(const uint32_t *src, size_t sstride, uint32_t *dst, size_t dstride, size_t n,
                               uint32_t *width) {
  size_t r = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (r >= n) return;
  uint32_t deg = 0;
  bool ended = false;
  for (size_t c = 0; c < dstride; c++) {
    uint32_t v = (c < sstride && !ended) ? src[r * sstride + c] : INVALID_ID;
    if (v == INVALID_ID) ended = true; else deg++;
    dst[r * dstride + c] = v;
  }
  if (width) atomicMax(width, deg);
}

}  // namespace

// Query rows of a search call. Float index: [nq, qstride] floats, staged like any input. u8 index: q points at
// unsigned-byte rows qstride BYTES apart; the kernels address rows in 4-byte units, so rows that are on the host,
// unaligned or at an odd stride are copied to a device buffer with a 4-byte aligned pitch.
struct QueryArg {
  InArg<float> f;
  DevBuf<uint8_t> b;
  const float *d = nullptr;
  size_t qstride = 0;  // in 4-byte units
  cudaError_t stage(const b200ann_index *ix, const float *q, size_t nq, size_t qs, cudaStream_t st) {
    if (!ix_u8(ix)) {
      cudaError_t e = f.stage(q, nq * qs, st);
      d = f.d; qstride = qs;
      return e;
    }
    if (is_device_ptr(q) && qs % 4 == 0 && ((uintptr_t)q & 3) == 0) { d = q; qstride = qs / 4; return cudaSuccess; }
    const size_t pitch = (ix->dim + 3) / 4 * 4;
    cudaError_t e = b.alloc(nq * pitch, st);
    if (e != cudaSuccess) return e;
    d = (const float *)b.p; qstride = pitch / 4;
    return cudaMemcpy2DAsync(b.p, pitch, q, qs, ix->dim, nq, cudaMemcpyDefault, st);
  }
};
static int u8_fix_pad(b200ann_handle *h, const b200ann_index *ix, float *dists_dev, size_t count) {
  if (!ix_u8(ix) || !dists_dev || !count) return 0;
  u8_pad_kernel<<<(unsigned)((count + 255) / 256 < 4096 ? (count + 255) / 256 : 4096), 256, 0, h->stream>>>(dists_dev, count);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

extern "C" {

int b200ann_create(int device, b200ann_handle_t *out) {
  if (!out) return B200ANN_ERR_ARG;
  *out = nullptr;
  int cnt = 0;
  if (cudaGetDeviceCount(&cnt) != cudaSuccess || cnt <= 0 || device < 0 || device >= cnt) {
    cudaGetLastError();
    return B200ANN_ERR_CUDA;  // no CPU fallback: the product needs a CUDA device
  }
  b200ann_handle *h = new b200ann_handle;
  h->device = device;
  if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess) {
    delete h;
    return B200ANN_ERR_CUDA;
  }
  cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (const char *e = getenv("B200ANN_PERSISTENT")) h->persistent = atoi(e);  // tuning knob: 0 = one launch slot per query
  if (const char *e = getenv("B200ANN_SEARCH_VERSION")) h->search_version = atoi(e);
  if (const char *e = getenv("B200ANN_RING_STAGES")) h->ring_stages = atoi(e);
  if (const char *e = getenv("B200ANN_SEARCH_MODE")) h->search_mode = atoi(e);
  if (const char *e = getenv("B200ANN_SEARCH_TAIL")) h->search_tail = atoi(e);
  if (const char *e = getenv("B200ANN_SEARCH_PREFETCH")) h->search_prefetch = atoi(e);
  if (const char *e = getenv("B200ANN_VISITED_COMPACT")) h->visited_compact = atoi(e);
  if (const char *e = getenv("B200ANN_VISITED_BITS")) { h->visited_bits = atoi(e); h->visited_bits_auto = false; }
  if (const char *e = getenv("B200ANN_SEARCH_CTAS_PER_SM")) h->search_ctas_per_sm = atoi(e);
  {  // keep the stream-ordered scratch pool mapped between calls (default: released at every synchronize)
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      uint64_t keep = ~(uint64_t)0;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    cudaGetLastError();
  }
  *out = h;
  return 0;
}

int b200ann_destroy(b200ann_handle_t h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  if (h->work) cudaFree(h->work);
  cudaEventDestroy(h->ev0);
  cudaEventDestroy(h->ev1);
  cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

const char *b200ann_last_error(b200ann_handle_t h) { return h ? h->err.c_str() : "null handle"; }
void *b200ann_stream(b200ann_handle_t h) { return h ? (void *)h->stream : nullptr; }
uint64_t b200ann_launch_count(b200ann_handle_t h) { return h ? h->launches : 0; }
float b200ann_last_kernel_ms(b200ann_handle_t h) { return h ? h->last_ms : 0.f; }

int b200ann_set_visited_bits(b200ann_handle_t h, int bits) {
  if (!h) return B200ANN_ERR_ARG;
  if (bits == 0) { h->visited_bits = 10; h->visited_bits_auto = true; return 0; }
  if (bits < 5 || bits > 14) return set_err(h, B200ANN_ERR_ARG, "visited_bits must be 0 (auto) or in [5,14]");
  h->visited_bits = bits; h->visited_bits_auto = false;
  return 0;
}

int b200ann_set_search_options(b200ann_handle_t h, int prefetch, int visited_compact, int ctas_per_sm) {
  if (!h) return B200ANN_ERR_ARG;
  if (prefetch >= 0) h->search_prefetch = prefetch != 0;
  if (visited_compact >= 0) h->visited_compact = visited_compact != 0;
  if (ctas_per_sm >= 0) h->search_ctas_per_sm = ctas_per_sm;
  return 0;
}

int b200ann_set_search_kernel(b200ann_handle_t h, int version, int ring_stages, int mode) {
  if (!h) return B200ANN_ERR_ARG;
  if (mode >= 0) {
    if (mode > 2) return set_err(h, B200ANN_ERR_ARG, "set_search_kernel: mode must be 0 (auto), 1 (query per warp) or 2 (query per block)");
    h->search_mode = mode;
  }
  if (version >= 0) {
    if (version > 2) return set_err(h, B200ANN_ERR_ARG, "set_search_kernel: version must be 0 (auto), 1 or 2");
    h->search_version = version;
  }
  if (ring_stages >= 0) {
    if (ring_stages > 8) return set_err(h, B200ANN_ERR_ARG, "set_search_kernel: ring_stages must be in [0,8]");
    h->ring_stages = ring_stages;
  }
  return 0;
}

int b200ann_set_id_base(b200ann_handle_t h, uint32_t id_base) {
  if (!h) return B200ANN_ERR_ARG;
  h->id_base = id_base;
  return 0;
}

int b200ann_set_query_order(b200ann_handle_t h, const uint32_t *order, size_t n) {
  if (!h) return B200ANN_ERR_ARG;
  if (order && !is_device_ptr(order)) return set_err(h, B200ANN_ERR_ARG, "set_query_order: order must be device memory");
  h->query_order = order; h->query_order_n = order ? n : 0;
  return 0;
}

int b200ann_probe_gather(b200ann_handle_t h, b200ann_index_t ix, uint64_t rows, int inflight, int warps_per_block,
                         int blocks_per_sm, uint32_t seed, float *ms) {
  FLOAT_ONLY(h, ix, "probe_gather");
  if (!h || !ix || !ms) return B200ANN_ERR_ARG;
  CU_CHECK(h, cudaSetDevice(h->device));
  DevBuf<float> sink;
  CU_CHECK(h, sink.alloc(1, h->stream));
  CallTimer t(h);
  int rc = launch_gather_probe(h, ix, rows, inflight, warps_per_block, blocks_per_sm, seed, sink.p);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, t.sync());
  *ms = h->last_ms;
  return 0;
}

int b200ann_index_create(b200ann_handle_t h, const float *data, size_t n, size_t dim, size_t stride, int metric,
                         b200ann_index_t *out) {
  if (!h || !out) return B200ANN_ERR_ARG;
  *out = nullptr;
  if (!data || n == 0 || dim == 0 || stride < dim) return set_err(h, B200ANN_ERR_ARG, "index_create: bad data/n/dim/stride");
  if (n >= 0x7fffffffu) return set_err(h, B200ANN_ERR_ARG, "index_create: n must be < 2^31");
  if (metric != B200ANN_L2 && metric != B200ANN_IP) return set_err(h, B200ANN_ERR_ARG, "index_create: unknown metric");
  CU_CHECK(h, cudaSetDevice(h->device));
  b200ann_index *ix = new b200ann_index;
  ix->h = h; ix->n = n; ix->dim = dim; ix->stride = (dim + 3) / 4 * 4; ix->metric = metric;
  cudaError_t e = cudaMalloc((void **)&ix->vec, n * ix->stride * sizeof(float));
  if (e == cudaSuccess) {
    if (is_device_ptr(data)) {
      pad_rows_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(data, stride, ix->vec, ix->stride, n, dim);
      h->launches++;
      e = cudaGetLastError();
    } else {
      if (ix->stride != dim) e = cudaMemsetAsync(ix->vec, 0, n * ix->stride * sizeof(float), h->stream);
      if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(ix->vec, ix->stride * sizeof(float), data, stride * sizeof(float), dim * sizeof(float), n,
                              cudaMemcpyHostToDevice, h->stream);
    }
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) {
    if (ix->vec) cudaFree(ix->vec);
    delete ix;
    return set_err(h, B200ANN_ERR_CUDA, std::string("index_create: ") + cudaGetErrorString(e));
  }
  *out = ix;
  return 0;
}

int b200ann_index_create_u8(b200ann_handle_t h, const uint8_t *data, size_t n, size_t dim, size_t stride, b200ann_index_t *out) {
  if (!h || !out) return B200ANN_ERR_ARG;
  *out = nullptr;
  if (!data || n == 0 || dim == 0 || stride < dim) return set_err(h, B200ANN_ERR_ARG, "index_create_u8: bad data/n/dim/stride");
  if (n >= 0x7fffffffu) return set_err(h, B200ANN_ERR_ARG, "index_create_u8: n must be < 2^31");
  if (dim > 8192) return set_err(h, B200ANN_ERR_ARG, "index_create_u8: dim > 8192 is not supported");
  CU_CHECK(h, cudaSetDevice(h->device));
  b200ann_index *ix = new b200ann_index;
  const size_t rowbytes = (dim + 15) / 16 * 16;
  ix->h = h; ix->n = n; ix->dim = dim; ix->stride = rowbytes / 4; ix->metric = B200ANN_L2_U8;
  cudaError_t e = cudaMalloc((void **)&ix->vec, n * rowbytes);
  if (e == cudaSuccess && rowbytes != dim) e = cudaMemsetAsync(ix->vec, 0, n * rowbytes, h->stream);
  if (e == cudaSuccess) e = cudaMemcpy2DAsync(ix->vec, rowbytes, data, stride, dim, n, cudaMemcpyDefault, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) {
    if (ix->vec) cudaFree(ix->vec);
    delete ix;
    return set_err(h, B200ANN_ERR_CUDA, std::string("index_create_u8: ") + cudaGetErrorString(e));
  }
  *out = ix;
  return 0;
}

static void free_graph(b200ann_index *ix) {
  if (ix->graph) cudaFree(ix->graph);
  ix->graph = nullptr;
}
static void free_hnsw(b200ann_index *ix) {
  if (ix->links0) cudaFree(ix->links0);
  if (ix->levels) cudaFree(ix->levels);
  if (ix->upper_ptr) cudaFree(ix->upper_ptr);
  if (ix->upper_links) cudaFree(ix->upper_links);
  if (ix->labels) cudaFree(ix->labels);
  if (ix->upper_owner) cudaFree(ix->upper_owner);
  ix->upper_owner = nullptr;
  ix->cap_links0 = ix->cap_levels = ix->cap_uptr = ix->cap_upper = ix->cap_owner = ix->cap_labels = 0;
  ix->links0 = ix->upper_ptr = ix->upper_links = nullptr;
  ix->levels = nullptr; ix->labels = nullptr; ix->has_hnsw = false;
}

int b200ann_index_destroy(b200ann_index_t ix) {
  if (!ix) return 0;
  cudaSetDevice(ix->h->device);
  cudaStreamSynchronize(ix->h->stream);
  free_graph(ix);
  free_hnsw(ix);
  knn_tc_free(ix);
  if (ix->vec) cudaFree(ix->vec);
  delete ix;
  return 0;
}

size_t b200ann_index_size(b200ann_index_t ix) { return ix ? ix->n : 0; }
size_t b200ann_index_dim(b200ann_index_t ix) { return ix ? ix->dim : 0; }

int b200ann_index_set_graph(b200ann_index_t ix, const uint32_t *graph, size_t gstride, uint32_t ep) {
  if (!ix) return B200ANN_ERR_ARG;
  b200ann_handle *h = ix->h;
  if (!graph || gstride == 0 || ep >= ix->n) return set_err(h, B200ANN_ERR_ARG, "set_graph: bad graph/gstride/ep");
  CU_CHECK(h, cudaSetDevice(h->device));
  free_graph(ix);
  size_t gs = (gstride + 7) / 8 * 8;  // 32-byte rows
  InArg<uint32_t> in;
  CU_CHECK(h, in.stage(graph, ix->n * gstride, h->stream));
  CU_CHECK(h, cudaMalloc((void **)&ix->graph, ix->n * gs * sizeof(uint32_t)));
  DevBuf<uint32_t> w;
  CU_CHECK(h, w.alloc(1, h->stream));
  CU_CHECK(h, cudaMemsetAsync(w.p, 0, 4, h->stream));
  regraph_kernel<<<(unsigned)((ix->n + 127) / 128), 128, 0, h->stream>>>(in.d, gstride, ix->graph, gs, ix->n, w.p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  CU_CHECK(h, cudaMemcpyAsync(&ix->width, w.p, 4, cudaMemcpyDeviceToHost, h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  ix->gstride = gs; ix->ep = ep;
  return 0;
}

int b200ann_index_get_graph(b200ann_index_t ix, uint32_t *graph, size_t gstride_out, uint32_t *ep, uint32_t *width) {
  if (!ix) return B200ANN_ERR_ARG;
  b200ann_handle *h = ix->h;
  if (!ix->graph) return set_err(h, B200ANN_ERR_STATE, "get_graph: index has no graph");
  if (ep) *ep = ix->ep;
  if (width) *width = ix->width;
  if (!graph) return 0;
  if (gstride_out < ix->width) return set_err(h, B200ANN_ERR_ARG, "get_graph: gstride_out < width");
  CU_CHECK(h, cudaSetDevice(h->device));
  OutArg<uint32_t> out;
  CU_CHECK(h, out.stage(graph, ix->n * gstride_out, h->stream));
  regraph_kernel<<<(unsigned)((ix->n + 127) / 128), 128, 0, h->stream>>>(ix->graph, ix->gstride, out.d, gstride_out, ix->n, nullptr);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  CU_CHECK(h, out.finish(h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  return 0;
}

int b200ann_index_set_hnsw(b200ann_index_t ix, const uint32_t *links0, size_t maxM0, const int32_t *levels,
                           const uint32_t *upper_ptr, const uint32_t *upper_links, size_t upper_rows, size_t maxM,
                           const uint64_t *labels, uint32_t ep, int maxlevel) {
  if (!ix) return B200ANN_ERR_ARG;
  b200ann_handle *h = ix->h;
  if (!links0 || maxM0 == 0 || !levels || !upper_ptr || ep >= ix->n || maxlevel < 0 || (maxlevel > 0 && (!upper_links || maxM == 0)))
    return set_err(h, B200ANN_ERR_ARG, "set_hnsw: bad arguments");
  CU_CHECK(h, cudaSetDevice(h->device));
  free_hnsw(ix);
  size_t n = ix->n;
  CU_CHECK(h, upload(&ix->links0, links0, n * maxM0, h->stream));
  CU_CHECK(h, upload(&ix->levels, levels, n, h->stream));
  CU_CHECK(h, upload(&ix->upper_ptr, upper_ptr, n, h->stream));
  if (upper_rows) CU_CHECK(h, upload(&ix->upper_links, upper_links, upper_rows * maxM, h->stream));
  if (labels) CU_CHECK(h, upload(&ix->labels, labels, n, h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  ix->maxM0 = maxM0; ix->maxM = maxM ? maxM : 1; ix->upper_rows = upper_rows;
  ix->hnsw_ep = ep; ix->maxlevel = maxlevel; ix->has_hnsw = true;
  ix->hnsw_M = ix->maxM; ix->hnsw_efc = 0;
  ix->level_rng.seed(100);  // a loaded index draws new levels from a fresh generator (hnswalg.h:78-86 load ctor, random_seed = 100)
  return 0;
}

int b200ann_hnsw_build(b200ann_handle_t h, b200ann_index_t ix, size_t M, size_t efc, uint32_t seed, size_t ratio,
                       size_t bmax, const uint64_t *labels, uint64_t *info) {
  if (!h || !ix) return B200ANN_ERR_ARG;
  CU_CHECK(h, cudaSetDevice(h->device));
  free_hnsw(ix);
  CallTimer t(h);
  int rc = hnsw_build_launch(h, ix, M, efc, seed, ratio, bmax, labels, info);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_hnsw_insert_batch(b200ann_handle_t h, b200ann_index_t ix, const float *data, size_t stride, const uint64_t *labels,
                              size_t nb, size_t ef_construction, uint64_t *info) {
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!data || stride < ix->dim) return set_err(h, B200ANN_ERR_ARG, "hnsw_insert_batch: bad data / stride");
  if (nb == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  if (ef_construction) ix->hnsw_efc = ef_construction > ix->maxM ? ef_construction : ix->maxM;
  // a u8 index takes unsigned-byte rows, stride in bytes
  const size_t esz = ix_u8(ix) ? 1 : 4;
  InArg<uint8_t> din;
  CU_CHECK(h, din.stage((const uint8_t *)data, nb * stride * esz, h->stream));
  CallTimer t(h);
  int rc = hnsw_insert_launch(h, ix, din.d, stride * esz, nb, labels, 0, 0, info);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_index_set_labels(b200ann_index_t ix, const uint32_t *ids, const uint64_t *labels, size_t nb) {
  if (!ix) return B200ANN_ERR_ARG;
  b200ann_handle *h = ix->h;
  if (!ids || !labels) return set_err(h, B200ANN_ERR_ARG, "index_set_labels: need ids and labels");
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "index_set_labels: index has no HNSW graph");
  if (nb == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  std::vector<uint32_t> hid(nb);
  CU_CHECK(h, cudaMemcpy(hid.data(), ids, nb * 4, cudaMemcpyDefault));
  for (size_t i = 0; i < nb; i++)
    if (hid[i] >= ix->n) return set_err(h, B200ANN_ERR_ARG, "index_set_labels: an id is not a row of the index");
  if (!ix->labels) {  // label == row number so far
    CU_CHECK(h, cudaMalloc((void **)&ix->labels, ix->n * 8));
    ix->cap_labels = ix->n;
    iota_labels_kernel<<<(unsigned)((ix->n + 255) / 256), 256, 0, h->stream>>>(ix->labels, ix->n);
    h->launches++;
  }
  InArg<uint32_t> iin; InArg<uint64_t> lin;
  CU_CHECK(h, iin.stage(ids, nb, h->stream));
  CU_CHECK(h, lin.stage(labels, nb, h->stream));
  set_labels_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, h->stream>>>(ix->labels, iin.d, lin.d, nb);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  return 0;
}

int b200ann_hnsw_update_batch(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *ids, const float *data, size_t stride,
                              size_t nb, uint64_t *info) {
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ids || !data || stride < ix->dim) return set_err(h, B200ANN_ERR_ARG, "hnsw_update_batch: bad ids / data / stride");
  if (nb == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  std::vector<uint32_t> hid(nb);
  CU_CHECK(h, cudaMemcpy(hid.data(), ids, nb * 4, cudaMemcpyDefault));
  {
    std::vector<uint32_t> srt(hid);
    std::sort(srt.begin(), srt.end());
    if (srt.back() >= ix->n) return set_err(h, B200ANN_ERR_ARG, "hnsw_update_batch: an id is not a row of the index");
    for (size_t i = 1; i < nb; i++)
      if (srt[i] == srt[i - 1]) return set_err(h, B200ANN_ERR_ARG, "hnsw_update_batch: ids must be distinct within one call");
  }
  const size_t esz = ix_u8(ix) ? 1 : 4;
  InArg<uint8_t> din; InArg<uint32_t> iin;
  CU_CHECK(h, din.stage((const uint8_t *)data, nb * stride * esz, h->stream));
  CU_CHECK(h, iin.stage(ids, nb, h->stream));
  CallTimer t(h);
  // Like the build's batches, one pass moves at most 1/32 of the index: the elements of a pass do not see each other's
  // new positions, and 10 % of an index moved in one pass measured 0.011 recall@10 below a rebuild. A larger call is
  // a sequence of passes, each a call of its own as far as the result goes.
  const size_t sub = std::max<size_t>(1, ix->n / 32);
  uint64_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int rc = 0;
  for (size_t s = 0; s < nb && rc == 0; s += sub) {
    uint64_t one[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    rc = hnsw_update_launch(h, ix, iin.d + s, std::min(sub, nb - s), din.d + s * stride * esz, stride * esz, one);
    for (int i = 0; i < 5; i++) acc[i] += one[i];
    acc[5] = one[5]; acc[6] = one[6]; acc[7] = std::max(acc[7], one[7]);
  }
  t.stop();
  if (rc) return rc;
  if (info) memcpy(info, acc, sizeof acc);
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_index_hnsw_shape(b200ann_index_t ix, uint64_t *out5) {
  if (!ix || !out5) return B200ANN_ERR_ARG;
  if (!ix->has_hnsw) return set_err(ix->h, B200ANN_ERR_STATE, "index has no HNSW graph (set_hnsw / hnsw_build first)");
  out5[0] = ix->maxM0; out5[1] = ix->maxM; out5[2] = ix->upper_rows; out5[3] = (uint64_t)ix->maxlevel; out5[4] = ix->hnsw_ep;
  return 0;
}

int b200ann_index_get_hnsw(b200ann_index_t ix, uint32_t *links0, int32_t *levels, uint32_t *upper_ptr,
                           uint32_t *upper_links, uint64_t *labels) {
  if (!ix) return B200ANN_ERR_ARG;
  b200ann_handle *h = ix->h;
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "index has no HNSW graph (set_hnsw / hnsw_build first)");
  CU_CHECK(h, cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  if (links0) CU_CHECK(h, cudaMemcpyAsync(links0, ix->links0, ix->n * ix->maxM0 * 4, cudaMemcpyDefault, st));
  if (levels) CU_CHECK(h, cudaMemcpyAsync(levels, ix->levels, ix->n * 4, cudaMemcpyDefault, st));
  if (upper_ptr) CU_CHECK(h, cudaMemcpyAsync(upper_ptr, ix->upper_ptr, ix->n * 4, cudaMemcpyDefault, st));
  if (upper_links && ix->upper_rows)
    CU_CHECK(h, cudaMemcpyAsync(upper_links, ix->upper_links, ix->upper_rows * ix->maxM * 4, cudaMemcpyDefault, st));
  if (labels) {
    if (ix->labels) CU_CHECK(h, cudaMemcpyAsync(labels, ix->labels, ix->n * 8, cudaMemcpyDefault, st));
    else {
      std::vector<uint64_t> ident(ix->n);
      for (size_t i = 0; i < ix->n; i++) ident[i] = i;
      CU_CHECK(h, cudaMemcpyAsync(labels, ident.data(), ix->n * 8, cudaMemcpyDefault, st));
      CU_CHECK(h, cudaStreamSynchronize(st));
    }
  }
  CU_CHECK(h, cudaStreamSynchronize(st));
  return 0;
}

// tensor cores when the problem is big enough to fill them (or when forced), else the fp32 SIMT kernel
static int knn_dispatch(b200ann_handle *h, b200ann_index *ix, const float *q, size_t nq, size_t qstride, size_t k,
                        bool exclude_self, uint32_t *ids, float *dists, uint32_t self_base = 0) {
  bool tcp = false;
  if (h->knn_mode == 2) {
    if (!knn_tc_supported(ix, nq, k)) return set_err(h, B200ANN_ERR_ARG, "knn: tensor path forced but unsupported (k > 1024, dim > 4096 or no driver entry point)");
    tcp = true;
  } else if (h->knn_mode == 0) {
    tcp = knn_tc_supported(ix, nq, k) && ix->n >= 4096 && (double)nq * (double)ix->n >= (double)(1 << 24);
  }
  memset(h->knn_stats, 0, sizeof h->knn_stats);
  if (tcp) return launch_knn_tc(h, ix, q, nq, qstride, k, exclude_self, ids, dists, self_base);
  return launch_bf_knn(h, ix, q, nq, qstride, k, exclude_self, ids, dists, nullptr, self_base);
}

int b200ann_set_async(b200ann_handle_t h, int on) {
  if (!h) return B200ANN_ERR_ARG;
  h->async_mode = on != 0;
  return 0;
}

int b200ann_set_knn_mode(b200ann_handle_t h, int mode, int slack) {
  if (!h) return B200ANN_ERR_ARG;
  if (mode < 0 || mode > 2) return set_err(h, B200ANN_ERR_ARG, "set_knn_mode: mode must be 0 (auto), 1 (fp32 SIMT) or 2 (tensor cores)");
  h->knn_mode = mode; h->knn_slack = slack;
  return 0;
}

int b200ann_knn_last_stats(b200ann_handle_t h, uint64_t *out8) {
  if (!h || !out8) return B200ANN_ERR_ARG;
  memcpy(out8, h->knn_stats, sizeof h->knn_stats);
  return 0;
}

int b200ann_bf_knn(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride, size_t k,
                   uint32_t *ids, float *dists) {
  FLOAT_ONLY(h, ix, "bf_knn");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!q || !ids || k == 0 || qstride < ix->dim) return set_err(h, B200ANN_ERR_ARG, "bf_knn: bad q/ids/k/qstride");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<float> qin; OutArg<uint32_t> oi; OutArg<float> od;
  CU_CHECK(h, qin.stage(q, nq * qstride, h->stream));
  CU_CHECK(h, oi.stage(ids, nq * k, h->stream));
  CU_CHECK(h, od.stage(dists, nq * k, h->stream));
  CallTimer t(h);
  int rc = knn_dispatch(h, ix, qin.d, nq, qstride, k, false, oi.d, od.d);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_bf_knn_u8(b200ann_handle_t h, const uint8_t *base, size_t n, size_t dim, size_t bstride, const uint8_t *q,
                      size_t nq, size_t qstride, size_t k, uint32_t *ids, int32_t *dists) {
  if (!h) return B200ANN_ERR_ARG;
  if (!base || !q || !ids || k == 0 || dim == 0 || bstride < dim || qstride < dim)
    return set_err(h, B200ANN_ERR_ARG, "bf_knn_u8: bad base/q/ids/k/dim/stride");
  if (dim > 32768) return set_err(h, B200ANN_ERR_ARG, "bf_knn_u8: dim > 32768 (the reference's int distance can overflow)");
  if (n >= 0xFFFFFFFFull || nq >= 0xFFFFFFFFull) return set_err(h, B200ANN_ERR_ARG, "bf_knn_u8: n / nq must fit 32 bits");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<uint8_t> bin, qin; OutArg<uint32_t> oi; OutArg<int32_t> od;
  CU_CHECK(h, bin.stage(base, n ? (n - 1) * bstride + dim : 0, h->stream));
  CU_CHECK(h, qin.stage(q, (nq - 1) * qstride + dim, h->stream));
  CU_CHECK(h, oi.stage(ids, nq * k, h->stream));
  CU_CHECK(h, od.stage(dists, nq * k, h->stream));
  CallTimer t(h);
  int rc = launch_bf_knn_u8(h, bin.d, n, dim, bstride, qin.d, nq, qstride, k, oi.d, od.d);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_upload_u8(b200ann_handle_t h, const uint8_t *src, size_t bytes, uint8_t **out_dev) {
  if (!h || !out_dev || (!src && bytes)) return B200ANN_ERR_ARG;
  CU_CHECK(h, cudaSetDevice(h->device));
  uint8_t *d = nullptr;
  CU_CHECK(h, cudaMalloc((void **)&d, bytes ? bytes : 1));
  cudaError_t e = cudaMemcpyAsync(d, src, bytes, cudaMemcpyHostToDevice, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) { cudaFree(d); CU_CHECK(h, e); }
  *out_dev = d;
  return 0;
}

int b200ann_free_device(void *dev) {
  return cudaFree(dev) == cudaSuccess ? 0 : B200ANN_ERR_CUDA;
}

int b200ann_knn_graph(b200ann_handle_t h, b200ann_index_t ix, size_t K, uint32_t *knn) {
  FLOAT_ONLY(h, ix, "knn_graph");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!knn || K == 0 || K >= ix->n) return set_err(h, B200ANN_ERR_ARG, "knn_graph: need 0 < K < n");
  CU_CHECK(h, cudaSetDevice(h->device));
  OutArg<uint32_t> oi;
  CU_CHECK(h, oi.stage(knn, ix->n * K, h->stream));
  CallTimer t(h);
  int rc = knn_dispatch(h, ix, ix->vec, ix->n, ix->stride, K, true, oi.d, nullptr);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_knn_graph_rows(b200ann_handle_t h, b200ann_index_t ix, size_t row0, size_t nrows, size_t K, uint32_t *knn) {
  FLOAT_ONLY(h, ix, "knn_graph_rows");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!knn || K == 0 || K >= ix->n || row0 + nrows > ix->n) return set_err(h, B200ANN_ERR_ARG, "knn_graph_rows: need 0 < K < n and row0 + nrows <= n");
  if (nrows == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  OutArg<uint32_t> oi;
  CU_CHECK(h, oi.stage(knn, nrows * K, h->stream));
  CallTimer t(h);
  int rc = knn_dispatch(h, ix, ix->vec + row0 * ix->stride, nrows, ix->stride, K, true, oi.d, nullptr, (uint32_t)row0);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_nndescent(b200ann_handle_t h, b200ann_index_t ix, size_t K, size_t L, size_t iter, size_t S, size_t R,
                      uint32_t seed, uint32_t *knn, float *dists, uint64_t *info) {
  FLOAT_ONLY(h, ix, "nndescent");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!knn) return set_err(h, B200ANN_ERR_ARG, "nndescent: knn output is NULL");
  CU_CHECK(h, cudaSetDevice(h->device));
  OutArg<uint32_t> oi; OutArg<float> od;
  CU_CHECK(h, oi.stage(knn, ix->n * K, h->stream));
  CU_CHECK(h, od.stage(dists, ix->n * K, h->stream));
  CallTimer t(h);
  int rc = nndescent_launch(h, ix, K, L, iter, S, R, seed, oi.d, od.d, info);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_kdtree_graph(b200ann_handle_t h, b200ann_index_t ix, size_t K, size_t nTrees, size_t mLevel, uint32_t seed, uint32_t *knn,
                         float *dists, uint64_t *info) {
  FLOAT_ONLY(h, ix, "kdtree_graph");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!knn) return set_err(h, B200ANN_ERR_ARG, "kdtree_graph: knn output is NULL");
  CU_CHECK(h, cudaSetDevice(h->device));
  OutArg<uint32_t> oi; OutArg<float> od;
  CU_CHECK(h, oi.stage(knn, ix->n * K, h->stream));
  CU_CHECK(h, od.stage(dists, ix->n * K, h->stream));
  CallTimer t(h);
  int rc = kdtree_launch(h, ix, K, nTrees, mLevel, seed, oi.d, od.d, info);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_nndescent_refine(b200ann_handle_t h, b200ann_index_t ix, size_t K, size_t L, size_t iter, size_t S, size_t R,
                             uint32_t seed, const uint32_t *init_graph, size_t init_k, uint32_t *knn, float *dists, uint64_t *info) {
  FLOAT_ONLY(h, ix, "nndescent_refine");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!knn || !init_graph || init_k == 0) return set_err(h, B200ANN_ERR_ARG, "nndescent_refine: need knn, init_graph, init_k > 0");
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<uint32_t> gin; OutArg<uint32_t> oi; OutArg<float> od;
  CU_CHECK(h, gin.stage(init_graph, ix->n * init_k, h->stream));
  CU_CHECK(h, oi.stage(knn, ix->n * K, h->stream));
  CU_CHECK(h, od.stage(dists, ix->n * K, h->stream));
  CallTimer t(h);
  int rc = nndescent_launch(h, ix, K, L, iter, S, R, seed, oi.d, od.d, info, gin.d, init_k);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_nsg_search(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride, size_t K,
                       size_t L, uint32_t seed, uint32_t *ids, float *dists, uint32_t *stats) {
  FLOAT_ONLY(h, ix, "nsg_search");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ix->graph) return set_err(h, B200ANN_ERR_STATE, "nsg_search: index has no graph (set_graph / nsg_build first)");
  if (!q || !ids || K == 0 || L < K || L > ix->n || qstride < ix->dim)
    return set_err(h, B200ANN_ERR_ARG, "nsg_search: need q, ids, 0 < K <= L <= n, qstride >= dim");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<float> qin; OutArg<uint32_t> oi; OutArg<float> od; OutArg<uint32_t> os;
  CU_CHECK(h, qin.stage(q, nq * qstride, h->stream));
  CU_CHECK(h, oi.stage(ids, nq * K, h->stream));
  CU_CHECK(h, od.stage(dists, nq * K, h->stream));
  CU_CHECK(h, os.stage(stats, nq * 2, h->stream));
  CallTimer t(h);
  int rc = launch_nsg_search(h, ix, qin.d, nq, qstride, K, L, seed, oi.d, od.d, os.d);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, os.finish(h->stream));
  if (h->async_mode && !oi.host && !od.host && !os.host && is_device_ptr(q)) return 0;  // stream-ordered, no host sync
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_nsg_search_clustered(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride,
                                 size_t K, size_t L, uint32_t seed, const uint32_t *ep, const uint32_t *base,
                                 const uint32_t *size, uint32_t *ids, float *dists, uint32_t *stats) {
  FLOAT_ONLY(h, ix, "nsg_search_clustered");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ix->graph) return set_err(h, B200ANN_ERR_STATE, "nsg_search_clustered: index has no graph (set_graph first)");
  if (!q || !ids || !ep || !base || !size || K == 0 || L < K || qstride < ix->dim)
    return set_err(h, B200ANN_ERR_ARG, "nsg_search_clustered: need q, ids, ep, base, size, 0 < K <= L, qstride >= dim");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<float> qin; InArg<uint32_t> iep, ib, is; OutArg<uint32_t> oi; OutArg<float> od; OutArg<uint32_t> os;
  CU_CHECK(h, qin.stage(q, nq * qstride, h->stream));
  CU_CHECK(h, iep.stage(ep, nq, h->stream));
  CU_CHECK(h, ib.stage(base, nq, h->stream));
  CU_CHECK(h, is.stage(size, nq, h->stream));
  CU_CHECK(h, oi.stage(ids, nq * K, h->stream));
  CU_CHECK(h, od.stage(dists, nq * K, h->stream));
  CU_CHECK(h, os.stage(stats, nq * 2, h->stream));
  CallTimer t(h);
  int rc = launch_nsg_search_clustered(h, ix, qin.d, nq, qstride, K, L, seed, iep.d, ib.d, is.d, oi.d, od.d, os.d);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, os.finish(h->stream));
  CU_CHECK(h, t.sync());
  return 0;
}

static int search_from_common(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride,
                              size_t K, size_t L, uint32_t seed, const uint32_t *entry, bool hybrid, uint32_t *ids,
                              float *dists, uint32_t *entry_out, uint32_t *stats, const char *who) {
  FLOAT_ONLY(h, ix, "nsg_search_from / hybrid_search");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ix->graph) return set_err(h, B200ANN_ERR_STATE, std::string(who) + ": index has no NSG graph (set_graph / nsg_build first)");
  if (hybrid && !ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, std::string(who) + ": index has no HNSW levels (set_hnsw first)");
  if (!q || !ids || K == 0 || L < K || L >= ix->n || qstride < ix->dim || (!hybrid && !entry))
    return set_err(h, B200ANN_ERR_ARG, std::string(who) + ": need q, ids, 0 < K <= L < n, qstride >= dim");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<float> qin; InArg<uint32_t> ein; OutArg<uint32_t> oi; OutArg<float> od; OutArg<uint32_t> os; OutArg<uint32_t> oe;
  CU_CHECK(h, qin.stage(q, nq * qstride, h->stream));
  if (!hybrid) CU_CHECK(h, ein.stage(entry, nq, h->stream));
  CU_CHECK(h, oi.stage(ids, nq * K, h->stream));
  CU_CHECK(h, od.stage(dists, nq * K, h->stream));
  CU_CHECK(h, os.stage(stats, nq * 2, h->stream));
  CU_CHECK(h, oe.stage(hybrid ? entry_out : nullptr, nq, h->stream));
  CallTimer t(h);
  int rc = launch_nsg_search_from(h, ix, qin.d, nq, qstride, K, L, seed, hybrid ? nullptr : ein.d, oe.d, oi.d, od.d, os.d);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, oi.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, os.finish(h->stream));
  CU_CHECK(h, oe.finish(h->stream));
  if (h->async_mode && !oi.host && !od.host && !os.host && !oe.host && is_device_ptr(q) && (hybrid || is_device_ptr(entry)))
    return 0;  // stream-ordered, no host sync
  CU_CHECK(h, t.sync());
  return 0;
}

int b200ann_nsg_search_from(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride, size_t K,
                            size_t L, uint32_t seed, const uint32_t *entry, uint32_t *ids, float *dists,
                            uint32_t *stats) {
  return search_from_common(h, ix, q, nq, qstride, K, L, seed, entry, false, ids, dists, nullptr, stats, "nsg_search_from");
}

int b200ann_hybrid_search(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride, size_t K,
                          size_t L, uint32_t seed, uint32_t *ids, float *dists, uint32_t *entry_out, uint32_t *stats) {
  return search_from_common(h, ix, q, nq, qstride, K, L, seed, nullptr, true, ids, dists, entry_out, stats, "hybrid_search");
}

int b200ann_hnsw_search(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride, size_t k,
                        size_t ef, uint64_t *labels, float *dists, uint32_t *stats) {
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "hnsw_search: index has no HNSW graph (set_hnsw first)");
  if (!q || !labels || !dists || k == 0 || qstride < ix->dim)
    return set_err(h, B200ANN_ERR_ARG, "hnsw_search: need q, labels, dists, k > 0, qstride >= dim");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  QueryArg qin; OutArg<uint64_t> ol; OutArg<float> od; OutArg<uint32_t> os;
  CU_CHECK(h, qin.stage(ix, q, nq, qstride, h->stream));
  CU_CHECK(h, ol.stage(labels, nq * k, h->stream));
  CU_CHECK(h, od.stage(dists, nq * k, h->stream));
  CU_CHECK(h, os.stage(stats, nq * 2, h->stream));
  CallTimer t(h);
  int rc = launch_hnsw_search(h, ix, qin.d, nq, qin.qstride, k, ef, ol.d, od.d, os.d);
  if (!rc) rc = u8_fix_pad(h, ix, od.d, nq * k);
  t.stop();
  if (rc) return rc;
  CU_CHECK(h, ol.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, os.finish(h->stream));
  CU_CHECK(h, t.sync());
  // fewer than k results for some query -> error like bindings.cpp:646-648 (rows stay padded)
  if (!ol.host) return 0;
  for (size_t i = 0; i < nq; i++)
    if (labels[i * k + k - 1] == ~(uint64_t)0)
      return set_err(h, B200ANN_ERR_FEWER, "hnsw_search: fewer than k results for some query; raise ef or lower k");
  return 0;
}

// shared by the filtered and the epsilon entry points: two passes (shared-memory heaps, then global-memory heaps for the
// queries the first pass could not prove exact). allow == NULL: every element may be returned.
static int filtered_common(b200ann_handle_t h, b200ann_index_t ix, const char *who, const float *q, size_t nq, size_t qstride,
                           size_t k, size_t ef, const uint32_t *allow, uint64_t *labels, float *dists, uint32_t *stats,
                           uint8_t *inexact, float eps, size_t min_cand, uint32_t *counts, bool internal_ids,
                           const uint32_t *doc = nullptr, size_t efcoll = 0, size_t docs_keep = 0) {
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<uint32_t> docin;
  CU_CHECK(h, docin.stage(doc, doc ? ix->n : 0, h->stream));
  QueryArg qin; InArg<uint32_t> ain; OutArg<uint64_t> ol; OutArg<float> od; OutArg<uint32_t> os, oc;
  DevBuf<uint8_t> flags;
  DevBuf<uint32_t> ones;
  const size_t awords = (ix->n + 31) / 32;
  CU_CHECK(h, qin.stage(ix, q, nq, qstride, h->stream));
  const uint32_t *adev;
  if (allow) {
    CU_CHECK(h, ain.stage(allow, awords, h->stream));
    adev = ain.d;
  } else {
    CU_CHECK(h, ones.alloc(awords, h->stream));
    CU_CHECK(h, cudaMemsetAsync(ones.p, 0xff, awords * 4, h->stream));
    adev = ones.p;
  }
  CU_CHECK(h, ol.stage(labels, nq * k, h->stream));
  CU_CHECK(h, od.stage(dists, nq * k, h->stream));
  CU_CHECK(h, os.stage(stats, nq * 2, h->stream));
  CU_CHECK(h, oc.stage(counts, nq, h->stream));
  CU_CHECK(h, flags.alloc(nq, h->stream));
  std::vector<uint8_t> hf(nq);
  CallTimer t(h);
  size_t left = 0;
  for (int pass = 0; pass < 2; pass++) {
    int rc = launch_hnsw_search_filtered(h, ix, qin.d, nq, qin.qstride, k, ef, adev, 0, pass == 1, pass ? flags.p : nullptr,
                                         ol.d, od.d, os.d, flags.p, nullptr, eps, min_cand, oc.d, internal_ids, docin.d, efcoll,
                                         docs_keep);
    if (rc) return rc;
    CU_CHECK(h, cudaMemcpyAsync(hf.data(), flags.p, nq, cudaMemcpyDeviceToHost, h->stream));
    CU_CHECK(h, cudaStreamSynchronize(h->stream));
    left = 0;
    for (size_t i = 0; i < nq; i++) left += hf[i] != 0;
    if (!left) break;  // pass 1 searches only the queries pass 0 reported
  }
  { int rc = u8_fix_pad(h, ix, od.d, nq * k); if (rc) return rc; }
  t.stop();
  CU_CHECK(h, ol.finish(h->stream));
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, os.finish(h->stream));
  CU_CHECK(h, oc.finish(h->stream));
  CU_CHECK(h, t.sync());
  if (inexact) std::memcpy(inexact, hf.data(), nq);
  else if (left) {
    std::string m = std::string(who) + (doc ? ": the result list of some query needs more than max_results slots (or its candidate set "
                                              "overflowed); raise max_results"
                                            : ": the filter is too selective for the graph walk on some queries "
                                              "(candidate set overflow); pass `inexact` to see which, or scan the allowed rows with bf_knn");
    return set_err(h, B200ANN_ERR_CAPACITY, m.c_str());
  }
  return 0;
}

int b200ann_hnsw_search_filtered(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride, size_t k,
                                 size_t ef, const uint32_t *allow, uint64_t *labels, float *dists, uint32_t *stats,
                                 uint8_t *inexact) {
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "hnsw_search_filtered: index has no HNSW graph (set_hnsw / hnsw_build first)");
  if (!q || !allow || !labels || !dists || k == 0 || qstride < ix->dim)
    return set_err(h, B200ANN_ERR_ARG, "hnsw_search_filtered: need q, allow, labels, dists, k > 0, qstride >= dim");
  if (inexact && is_device_ptr(inexact)) return set_err(h, B200ANN_ERR_ARG, "hnsw_search_filtered: inexact must be host memory");
  if (nq == 0) return 0;
  return filtered_common(h, ix, "hnsw_search_filtered", q, nq, qstride, k, ef, allow, labels, dists, stats, inexact, INFINITY, 0,
                         nullptr, false);
}

int b200ann_hnsw_search_epsilon(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride, float epsilon,
                                size_t min_num_candidates, size_t max_num_candidates, const uint32_t *allow, int internal_ids,
                                uint64_t *labels, float *dists, uint32_t *counts, uint32_t *stats) {
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "hnsw_search_epsilon: index has no HNSW graph (set_hnsw / hnsw_build first)");
  if (!q || !labels || !dists || !counts || max_num_candidates == 0 || qstride < ix->dim)
    return set_err(h, B200ANN_ERR_ARG, "hnsw_search_epsilon: need q, labels, dists, counts, max_num_candidates > 0, qstride >= dim");
  if (min_num_candidates > max_num_candidates)  // the reference asserts this (stop_condition.h:227)
    return set_err(h, B200ANN_ERR_ARG, "hnsw_search_epsilon: min_num_candidates > max_num_candidates");
  if (epsilon != epsilon) return set_err(h, B200ANN_ERR_ARG, "hnsw_search_epsilon: epsilon is NaN");
  if (nq == 0) return 0;
  return filtered_common(h, ix, "hnsw_search_epsilon", q, nq, qstride, max_num_candidates, max_num_candidates, allow, labels, dists,
                         stats, nullptr, epsilon, min_num_candidates, counts, internal_ids != 0);
}

int b200ann_hnsw_search_multivector(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride,
                                    const uint32_t *doc_ids, size_t num_docs_to_search, size_t ef_collection, size_t max_results,
                                    const uint32_t *allow, int internal_ids, uint64_t *labels, float *dists, uint32_t *counts,
                                    uint32_t *stats) {
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "hnsw_search_multivector: index has no HNSW graph (set_hnsw / hnsw_build first)");
  if (!q || !doc_ids || !labels || !dists || !counts || max_results == 0 || num_docs_to_search == 0 || qstride < ix->dim)
    return set_err(h, B200ANN_ERR_ARG, "hnsw_search_multivector: need q, doc_ids, labels, dists, counts, num_docs_to_search > 0, max_results > 0, qstride >= dim");
  if (ef_collection < num_docs_to_search) ef_collection = num_docs_to_search;  // stop_condition.h:164
  if (nq == 0) return 0;
  return filtered_common(h, ix, "hnsw_search_multivector", q, nq, qstride, max_results, max_results, allow, labels, dists, stats, nullptr,
                         INFINITY, 0, counts, internal_ids != 0, doc_ids, ef_collection, num_docs_to_search);
}

}  // extern "C"

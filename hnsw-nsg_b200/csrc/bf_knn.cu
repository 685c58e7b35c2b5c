// Exact brute-force kNN, fp32 SIMT tiles with a fused running top-k.
//
// Replaces hnswlib::BruteforceSearch::searchKnn (hnswlib/hnswlib/bruteforce.h:106-135) and
// IndexGraph::generate_control_set (CNNS/efanna_graph/src/index_graph.cpp:138-153), batched over
// queries; with exclude_self it yields the exact kNN graph that IndexGraph::Build (:265-294)
// approximates. Result = the k smallest under (distance, id), ties to the smaller id, ascending.
//
// CTA = 256 threads = 64 queries x 128 base rows per tile, 4x8 register tile per thread, DK=16 dims
// per shared-memory stage. Distances are accumulated sequentially over the dimension with fmaf
// (oracle mode ORC_*_SEQ_FMA restates this order). Each warp owns 8 queries outright, so the running
// top-k needs no block-level synchronisation: candidates below the query's current k-th
// (distance,id) go to a 32-entry shared-memory pending buffer; a full buffer is bitonic-sorted across
// the warp and merged into the query's sorted k-list in global memory (ping-pong between two lists).
#include <mutex>
#include "common.cuh"

namespace {

constexpr int QT = 64, BT = 128, DK = 16, NT = 256, PEND = 32;
constexpr int QS = QT + 4, BS = BT + 4;  // padded smem row strides (floats), multiples of 4

struct BfParams {
  const float *base; size_t bstride; uint32_t n; uint32_t dim;
  const float *q; size_t qstride; uint32_t nq;
  uint32_t k; int ip; int exclude_self;
  const uint32_t *self_ids;  // optional: the id query i must skip (exclude_self with gathered queries)
  uint32_t self_base;        // added to the id to skip (queries = a block of base rows starting at self_base)
  uint32_t splits;          // base range splits (grid.y)
  float *list_d[2]; uint32_t *list_i[2];  // [splits, nq, k] ping-pong lists
  uint8_t *cur;             // [splits, nq] which list is current at exit
};

__device__ __forceinline__ bool lex_less(float d0, uint32_t i0, float d1, uint32_t i1) {
  return d0 < d1 || (d0 == d1 && i0 < i1);
}

// merge the (sorted ascending across lanes) pending entries of one query into its k-list
__device__ __forceinline__ void merge_pending(float pd, uint32_t pi, const volatile float *src_d,
                                              const volatile uint32_t *src_i, float *dst_d, uint32_t *dst_i,
                                              uint32_t k, int lane, float &thr_d, uint32_t &thr_i) {
  // bitonic sort of 32 (d,id) pairs across the warp
#pragma unroll
  for (int sz = 2; sz <= 32; sz <<= 1) {
#pragma unroll
    for (int st = sz >> 1; st >= 1; st >>= 1) {
      float od = __shfl_xor_sync(0xffffffffu, pd, st);
      uint32_t oi = __shfl_xor_sync(0xffffffffu, pi, st);
      bool keep_min = (((lane & sz) == 0) == ((lane & st) == 0));
      bool take = keep_min ? lex_less(od, oi, pd, pi) : lex_less(pd, pi, od, oi);
      if (take) { pd = od; pi = oi; }
    }
  }
  // list elements: new position = j + #{pending < list[j]}
  for (uint32_t j0 = 0; j0 < k; j0 += 32) {
    uint32_t j = j0 + lane;
    float ld = INFINITY; uint32_t li = INVALID_ID;
    if (j < k) { ld = src_d[j]; li = src_i[j]; }
    uint32_t cnt = 0;
#pragma unroll
    for (int s = 0; s < 32; s++) {
      float sd = __shfl_sync(0xffffffffu, pd, s);
      uint32_t si = __shfl_sync(0xffffffffu, pi, s);
      cnt += lex_less(sd, si, ld, li) ? 1u : 0u;
    }
    if (j < k && j + cnt < k) { dst_d[j + cnt] = ld; dst_i[j + cnt] = li; }
  }
  // pending elements: new position = lane + #{list < pending}  (binary search; keys are unique)
  if (pi != INVALID_ID) {
    uint32_t lo = 0, hi = k;
    while (lo < hi) {
      uint32_t mid = (lo + hi) >> 1;
      if (lex_less(src_d[mid], src_i[mid], pd, pi)) lo = mid + 1; else hi = mid;
    }
    uint32_t pos = lo + lane;
    if (pos < k) { dst_d[pos] = pd; dst_i[pos] = pi; }
  }
  __syncwarp();
  thr_d = ((const volatile float *)dst_d)[k - 1];
  thr_i = ((const volatile uint32_t *)dst_i)[k - 1];
}

// running top-k state of the CTA's 64 queries (each owned by one warp)
struct TopkSmem {
  float pend_d[QT][PEND]; uint32_t pend_i[QT][PEND];
  uint32_t pend_n[QT]; float thr_d[QT]; uint32_t thr_i[QT]; uint8_t curbuf[QT];
};

// Merge query ql's pending entries into its k-list. NOT inlined on purpose: the insertion pass is fully unrolled over
// the 4x8 register tile and tests two queries per step; 64 inlined copies of the bitonic sort + merge made that pass
// instruction-fetch bound (measured on the u8 twin of this kernel, bf_knn_u8.cu: 1.6x - 2x end to end).
__device__ __noinline__ void flush_query(TopkSmem *st, float *l0d, float *l1d, uint32_t *l0i, uint32_t *l1i, size_t off,
                                         uint32_t k, int ql, int lane) {
  uint32_t cnt = st->pend_n[ql];
  if (cnt == 0) return;
  float pd = (uint32_t)lane < cnt ? st->pend_d[ql][lane] : INFINITY;
  uint32_t pi = (uint32_t)lane < cnt ? st->pend_i[ql][lane] : INVALID_ID;
  int c = st->curbuf[ql];
  float td; uint32_t ti;
  merge_pending(pd, pi, (c ? l1d : l0d) + off, (c ? l1i : l0i) + off, (c ? l0d : l1d) + off, (c ? l0i : l1i) + off, k,
                lane, td, ti);
  if (lane == 0) { st->curbuf[ql] = c ^ 1; st->pend_n[ql] = 0; st->thr_d[ql] = td; st->thr_i[ql] = ti; }
  __syncwarp();
}

__global__ void __launch_bounds__(NT, 2) bf_knn_kernel(const BfParams p) {
  __shared__ __align__(16) float Qs[DK][QS];
  __shared__ __align__(16) float Bs[DK][BS];
  __shared__ TopkSmem st;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ty = tid >> 4, tx = tid & 15;
  const uint32_t q0 = blockIdx.x * QT;
  const uint32_t split = blockIdx.y;
  const uint32_t per = ((p.n + p.splits - 1) / p.splits + BT - 1) / BT * BT;
  const uint32_t b_begin = split * per;
  const uint32_t b_end = min(p.n, b_begin + per);
  const size_t lbase = ((size_t)split * p.nq) * p.k;

  for (int i = tid; i < QT; i += NT) {
    st.pend_n[i] = 0; st.thr_d[i] = INFINITY; st.thr_i[i] = INVALID_ID; st.curbuf[i] = 0;
  }
  // lists start as (+inf, INVALID)
  for (uint32_t qq = 0; qq < QT; qq++) {
    uint32_t qg = q0 + qq;
    if (qg >= p.nq) break;
    for (uint32_t j = tid; j < p.k; j += NT) {
      p.list_d[0][lbase + (size_t)qg * p.k + j] = INFINITY;
      p.list_i[0][lbase + (size_t)qg * p.k + j] = INVALID_ID;
    }
  }
  __syncthreads();

  auto flush = [&](int ql) {  // warp-cooperative; ql is warp-uniform and owned by this warp
    flush_query(&st, p.list_d[0], p.list_d[1], p.list_i[0], p.list_i[1], lbase + (size_t)(q0 + ql) * p.k, p.k, ql, lane);
  };

  for (uint32_t b0 = b_begin; b0 < b_end; b0 += BT) {
    float acc[4][8];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 8; j++) acc[i][j] = 0.f;

    for (uint32_t d0 = 0; d0 < p.dim; d0 += DK) {
      // stage Q (64 x 16) and B (128 x 16) transposed into shared memory, zero padded
      {
        int r = tid >> 2, dc = (tid & 3) * 4;
        uint32_t qg = q0 + r;
#pragma unroll
        for (int j = 0; j < 4; j++) {
          uint32_t d = d0 + dc + j;
          Qs[dc + j][r] = (qg < p.nq && d < p.dim) ? __ldg(p.q + (size_t)qg * p.qstride + d) : 0.f;
        }
#pragma unroll
        for (int h = 0; h < 2; h++) {
          int rb = r + h * 64;
          uint32_t bg = b0 + rb;
          if (bg < b_end && d0 + dc + 3 < p.dim + 0u && ((p.bstride & 3) == 0)) {
            float4 v = __ldg(reinterpret_cast<const float4 *>(p.base + (size_t)bg * p.bstride + d0 + dc));
            Bs[dc + 0][rb] = v.x; Bs[dc + 1][rb] = v.y; Bs[dc + 2][rb] = v.z; Bs[dc + 3][rb] = v.w;
          } else {
#pragma unroll
            for (int j = 0; j < 4; j++) {
              uint32_t d = d0 + dc + j;
              Bs[dc + j][rb] = (bg < b_end && d < p.dim) ? __ldg(p.base + (size_t)bg * p.bstride + d) : 0.f;
            }
          }
        }
      }
      __syncthreads();
#pragma unroll
      for (int d = 0; d < DK; d++) {
        float4 qv = *reinterpret_cast<const float4 *>(&Qs[d][ty * 4]);
        float4 bA = *reinterpret_cast<const float4 *>(&Bs[d][tx * 4]);
        float4 bB = *reinterpret_cast<const float4 *>(&Bs[d][64 + tx * 4]);
        const float qa[4] = {qv.x, qv.y, qv.z, qv.w};
        const float ba[8] = {bA.x, bA.y, bA.z, bA.w, bB.x, bB.y, bB.z, bB.w};
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 8; j++) {
            if (p.ip) acc[i][j] = fmaf(qa[i], ba[j], acc[i][j]);
            else { float df = qa[i] - ba[j]; acc[i][j] = fmaf(df, df, acc[i][j]); }
          }
      }
      __syncthreads();
    }

    // ---- running top-k: this warp owns queries ty*4+i of its two ty values
    uint32_t self_id[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const uint32_t qg = q0 + ty * 4 + i;
      self_id[i] = ((p.self_ids && qg < p.nq) ? __ldg(p.self_ids + qg) : qg) + p.self_base;
    }
    // distances in place; a tile in which no lane of the warp beats its query's current k-th (thresholds only
    // tighten) skips the insertion pass altogether
    uint32_t mask = 0;  // bit i*8+j: this lane's tile entry (i,j) beats its query's current k-th
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int ql = ty * 4 + i;
      const uint32_t qg = q0 + ql;
      const float td = st.thr_d[ql]; const uint32_t ti = st.thr_i[ql];
#pragma unroll
      for (int j = 0; j < 8; j++) {
        uint32_t bg = b0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
        float d = p.ip ? 1.0f - acc[i][j] : acc[i][j];
        acc[i][j] = d;
        if (qg < p.nq && bg < b_end && !(p.exclude_self && bg == self_id[i]) && lex_less(d, bg, td, ti)) mask |= 1u << (i * 8 + j);
      }
    }
    // only the (i,j) steps in which some lane of the warp has an entry to insert are executed (warp-uniform test):
    // the CTA barrier of the next tile waits for the slowest warp, so the cost of this pass is paid by all 8 warps
    const uint32_t umask = __reduce_or_sync(0xffffffffu, mask);
    if (umask == 0) continue;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int ql = ty * 4 + i;
      const uint32_t qg = q0 + ql;
#pragma unroll
      for (int j = 0; j < 8; j++) {
        if (!((umask >> (i * 8 + j)) & 1u)) continue;
        // make room: at most 16 lanes push to one query per step
        int qa = (warp * 2) * 4 + i, qb = (warp * 2 + 1) * 4 + i;
        if (st.pend_n[qa] > PEND - 16) flush(qa);
        if (st.pend_n[qb] > PEND - 16) flush(qb);
        uint32_t bg = b0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
        float d = acc[i][j];
        bool pass = qg < p.nq && bg < b_end && !(p.exclude_self && bg == self_id[i]) &&
                    lex_less(d, bg, st.thr_d[ql], st.thr_i[ql]);
        if (pass) {
          uint32_t s = atomicAdd(&st.pend_n[ql], 1u);
          st.pend_d[ql][s] = d; st.pend_i[ql][s] = bg;
        }
        __syncwarp();
      }
    }
  }
  // final flush of this warp's 8 queries
  for (int i = 0; i < 8; i++) {
    int ql = warp * 8 + i;
    if (q0 + ql < p.nq) flush(ql);
  }
  __syncwarp();
  for (int i = 0; i < 8; i++) {
    int ql = warp * 8 + i;
    if (q0 + ql < p.nq && lane == 0) p.cur[(size_t)split * p.nq + q0 + ql] = st.curbuf[ql];
  }
}

// pick the current list of every (split, query), merge the splits: out = k smallest of the union.
__global__ void bf_finalize_kernel(const BfParams p, uint32_t *out_i, float *out_d) {
  const uint32_t qg = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (qg >= p.nq) return;
  const uint32_t total = p.splits * p.k;
  auto get = [&](uint32_t e, float &d, uint32_t &i) {
    uint32_t s = e / p.k, j = e % p.k;
    int c = p.cur[(size_t)s * p.nq + qg];
    size_t off = ((size_t)s * p.nq + qg) * p.k + j;
    d = p.list_d[c][off]; i = p.list_i[c][off];
  };
  for (uint32_t e = lane; e < total; e += 32) {
    float d; uint32_t i;
    get(e, d, i);
    uint32_t s = e / p.k, j = e % p.k;
    // rank = own index within its sorted list + number of smaller entries in the other lists
    uint32_t rank = j;
    for (uint32_t s2 = 0; s2 < p.splits; s2++) {
      if (s2 == s) continue;
      uint32_t lo = 0, hi = p.k;
      while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        float d2; uint32_t i2;
        get(s2 * p.k + mid, d2, i2);
        // INVALID entries (+inf, INVALID) are never "less" than anything valid; among themselves
        // order by split so that ranks stay unique
        bool less = (i2 == INVALID_ID && i == INVALID_ID) ? (s2 < s) : lex_less(d2, i2, d, i);
        if (less) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    if (rank < p.k) {
      out_i[(size_t)qg * p.k + rank] = i;
      if (out_d) out_d[(size_t)qg * p.k + rank] = d;
    }
  }
}

}  // namespace

int launch_bf_knn(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                  size_t k, bool exclude_self, uint32_t *ids, float *dists, const uint32_t *self_ids, uint32_t self_base) {
  if (nq == 0) return 0;
  BfParams p = {};
  p.base = ix->vec; p.bstride = ix->stride; p.n = (uint32_t)ix->n; p.dim = (uint32_t)ix->dim;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.k = (uint32_t)k;
  p.ip = ix->metric == B200ANN_IP; p.exclude_self = exclude_self; p.self_ids = self_ids; p.self_base = self_base;
  uint32_t qtiles = (uint32_t)((nq + QT - 1) / QT);
  // 2 CTAs are resident per SM. A grid slightly above a whole number of waves wastes most of the last one, and every
  // split pays its own top-k warm-up: pick the split count with the least time per CTA of useful work (bf_knn_u8.cu).
  const uint32_t slots = 2u * (uint32_t)h->sm_count;
  uint32_t cap = (uint32_t)((ix->n + 4 * BT - 1) / (4 * BT));
  if (cap > 64) cap = 64;
  if (cap < 1) cap = 1;
  uint32_t splits = 1; double best_cost = 1e300;
  for (uint32_t s = 1; s <= cap; s++) {
    double ctas = (double)qtiles * s, waves = (double)((qtiles * (uint64_t)s + slots - 1) / slots);
    double cost = waves / ctas * (1.0 + 0.02 * s);
    if (cost < best_cost * 0.999) { best_cost = cost; splits = s; }
  }
  while (splits > 1 && (size_t)splits * nq * k * 16 > ((size_t)1 << 30)) splits--;  // <= 1 GiB of partial lists
  p.splits = splits;
  size_t cnt = (size_t)splits * nq * k;
  DevBuf<float> ld0, ld1; DevBuf<uint32_t> li0, li1; DevBuf<uint8_t> cur;
  CU_CHECK(h, ld0.alloc(cnt, h->stream)); CU_CHECK(h, ld1.alloc(cnt, h->stream));
  CU_CHECK(h, li0.alloc(cnt, h->stream)); CU_CHECK(h, li1.alloc(cnt, h->stream));
  CU_CHECK(h, cur.alloc((size_t)splits * nq, h->stream));
  p.list_d[0] = ld0.p; p.list_d[1] = ld1.p; p.list_i[0] = li0.p; p.list_i[1] = li1.p; p.cur = cur.p;
  dim3 grid(qtiles, splits);
  bf_knn_kernel<<<grid, NT, 0, h->stream>>>(p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  bf_finalize_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(p, ids, dists);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// b200ann_bf_knn_noalloc: the SIMT kernel pair above for callers that cannot allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// The per-split partial lists and the "current list" flags, which launch_bf_knn takes from the stream
// ordered allocator, live in ONE file-scope device array of B200ANN_NOALLOC_SCRATCH_BYTES (16 MiB) that every GPU
// loading the library holds until exit. Every call shares it and finds the previous call's contents in it: the part
// a call uses is cleared first, and calls that overlap on the GPU (two streams, two handles, two host threads) race.
namespace {
constexpr size_t NOALLOC_BYTES = B200ANN_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char bf_noalloc_scratch[NOALLOC_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex noalloc_mu;
unsigned char *noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of bf_noalloc_scratch on device i (0 = not asked yet)

// In host code `bf_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address a kernel
// can use comes from cudaGetSymbolAddress, asked once per device with that device current.
int noalloc_scratch(b200ann_handle *h, unsigned char **out) {
  if (h->device < 0 || h->device >= NOALLOC_MAX_DEVICES) return set_err(h, B200ANN_ERR_ARG, "bf_knn_noalloc: device index >= 64");
  std::lock_guard<std::mutex> lock(noalloc_mu);
  if (!noalloc_ptr[h->device]) {
    int cur = -1;
    CU_CHECK(h, cudaGetDevice(&cur));
    if (cur != h->device) return set_err(h, B200ANN_ERR_STATE, "bf_knn_noalloc: the handle's device is not current");
    unsigned char *p = nullptr;
    CU_CHECK(h, cudaGetSymbolAddress((void **)&p, bf_noalloc_scratch));
    if (!p) return set_err(h, B200ANN_ERR_CUDA, "bf_knn_noalloc: cudaGetSymbolAddress returned a null address");
    noalloc_ptr[h->device] = p;
  }
  *out = noalloc_ptr[h->device];
  return 0;
}
}  // namespace

extern "C" int b200ann_bf_knn_noalloc(b200ann_handle_t h, b200ann_index_t ix, const float *q, size_t nq, size_t qstride,
                                      size_t k, uint32_t *ids, float *dists) {
  FLOAT_ONLY(h, ix, "bf_knn_noalloc");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!q || !ids || k == 0 || qstride < ix->dim) return set_err(h, B200ANN_ERR_ARG, "bf_knn_noalloc: bad q/ids/k/qstride");
  if (nq == 0) return 0;
  if (nq >= 0xFFFFFFFFull || k >= 0xFFFFFFFFull) return set_err(h, B200ANN_ERR_ARG, "bf_knn_noalloc: nq / k must fit 32 bits");
  CU_CHECK(h, cudaSetDevice(h->device));
  // host buffers would need staging copies, i.e. allocations
  if (!is_device_ptr(q) || !is_device_ptr(ids) || (dists && !is_device_ptr(dists)))
    return set_err(h, B200ANN_ERR_ARG, "bf_knn_noalloc: q, ids and dists must be device memory (host buffers need staging allocations; use b200ann_bf_knn)");
  // scratch per split: two (distance, id) list pairs of nq*k entries + nq flag bytes (rounded to 16)
  const size_t per_split = nq * k * 16 + (nq + 15) / 16 * 16;
  if (per_split > NOALLOC_BYTES) {
    char b[256];
    snprintf(b, sizeof b, "bf_knn_noalloc: nq * k * 16 + nq = %zu bytes of partial lists exceed the %zu-byte static scratch "
             "(lower nq or k, or use b200ann_bf_knn)", per_split, NOALLOC_BYTES);
    return set_err(h, B200ANN_ERR_ARG, b);
  }
  unsigned char *scratch = nullptr;
  int rc = noalloc_scratch(h, &scratch);
  if (rc) return rc;

  BfParams p = {};
  p.base = ix->vec; p.bstride = ix->stride; p.n = (uint32_t)ix->n; p.dim = (uint32_t)ix->dim;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.k = (uint32_t)k;
  p.ip = ix->metric == B200ANN_IP; p.exclude_self = 0; p.self_ids = nullptr; p.self_base = 0;
  const uint32_t qtiles = (uint32_t)((nq + QT - 1) / QT);
  const uint32_t slots = 2u * (uint32_t)h->sm_count;  // launch_bf_knn's split rule, then as many splits as the array holds
  uint32_t cap = (uint32_t)((ix->n + 4 * BT - 1) / (4 * BT));
  if (cap > 64) cap = 64;
  if (cap < 1) cap = 1;
  uint32_t splits = 1; double best_cost = 1e300;
  for (uint32_t s = 1; s <= cap; s++) {
    double ctas = (double)qtiles * s, waves = (double)((qtiles * (uint64_t)s + slots - 1) / slots);
    double cost = waves / ctas * (1.0 + 0.02 * s);
    if (cost < best_cost * 0.999) { best_cost = cost; splits = s; }
  }
  while (splits > 1 && (size_t)splits * per_split > NOALLOC_BYTES) splits--;
  p.splits = splits;
  const size_t cnt = (size_t)splits * nq * k;  // entries per list (4-byte entries: every section is aligned for its type)
  p.list_d[0] = reinterpret_cast<float *>(scratch);
  p.list_d[1] = p.list_d[0] + cnt;
  p.list_i[0] = reinterpret_cast<uint32_t *>(p.list_d[1] + cnt);
  p.list_i[1] = p.list_i[0] + cnt;
  p.cur = reinterpret_cast<uint8_t *>(p.list_i[1] + cnt);
  // the array still holds the last call's lists and flags: clear what this call uses
  CU_CHECK(h, cudaMemsetAsync(scratch, 0, cnt * 16 + (size_t)splits * nq, h->stream));
  dim3 grid(qtiles, splits);
  bf_knn_kernel<<<grid, NT, 0, h->stream>>>(p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  bf_finalize_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(p, ids, dists);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  CU_CHECK(h, cudaStreamIsCapturing(h->stream, &cs));
  if (cs == cudaStreamCaptureStatusNone && !h->async_mode) CU_CHECK(h, cudaStreamSynchronize(h->stream));
  return 0;
}

// Exact brute-force kNN over unsigned-byte vectors with integer squared-L2 distances.
//
// Replaces hnswlib::BruteforceSearch<int>::searchKnn (hnswlib/hnswlib/bruteforce.h:106-135) over
// hnswlib::L2SpaceI (space_l2.h:255-323: L2SqrI4x / L2SqrI, int accumulation of (a-b)^2 over u8), batched
// over queries. Integer arithmetic: every distance equals the reference's, so ids AND distances are
// bit-exact; result = the k smallest under (distance, id), ties to the smaller id, ascending.
//
// Same tiling as bf_knn.cu (CTA = 256 threads = 64 queries x 128 base rows, 4x8 register tile, warp-owned
// running top-k with a 32-entry pending buffer per query), but a shared-memory stage holds 64 dimensions per
// row packed four to a 32-bit word and the tile loop is ONE integer instruction per 4 dimensions of a pair:
// __dp4a (IDP.4A.U8.U8) accumulates the byte products a.b in u32; the distance is then
// |a|^2 + |b|^2 - 2 a.b with the row norms from a one-pass pre-kernel (u8_norms_kernel) - all integers, so
// the identity is exact and equals the reference's (a-b)^2 sum bit for bit (a first version with
// __vabsdiffu4 + __dp4a on the differences needed two instructions). Rows are fetched with 16-byte loads when
// base and stride are 16-byte aligned; the tail of a row is zero filled (zero bytes add nothing to products
// or norms). HBM traffic = n*dim bytes per 64-query tile; the kernel is bound by the integer pipe.
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int QT = 64, BT = 128, DK = 32, NT = 256, PEND = 32;  // DK words = 128 dimensions per stage
constexpr int QS = QT + 4, BS = BT + 4;  // padded smem row strides (dkeys), multiples of 4

typedef uint32_t dkey;                 // integer squared distance; DINF marks an empty slot
constexpr dkey DINF = 0x7FFFFFFFu;      // > 255^2 * 32768, the largest distance the entry point admits

struct BfParams {
  const uint8_t *base; size_t bstride; uint32_t n; uint32_t dim;  // strides in bytes
  const uint8_t *q; size_t qstride; uint32_t nq;
  uint32_t k; int bvec, qvec;  // rows 16-byte aligned -> uint4 loads
  const uint32_t *bnorm, *qnorm;  // sum of squares per row (u8_norms_kernel)
  uint32_t splits;          // base range splits (grid.y)
  dkey *list_d[2]; uint32_t *list_i[2];  // [splits, nq, k] ping-pong lists
  uint8_t *cur;             // [splits, nq] which list is current at exit
  dkey *gthr;               // [nq] smallest k-th distance any split of the query has reached so far
};

// 16 bytes of a row starting at byte c0, zero past the row end (zero bytes add 0 to the distance)
__device__ __forceinline__ uint4 load16(const uint8_t *row, uint32_t c0, uint32_t dim, bool vec) {
  if (vec && c0 + 16 <= dim) return __ldg(reinterpret_cast<const uint4 *>(row + c0));
  auto word = [&](uint32_t c) {
    uint32_t w = 0;
#pragma unroll
    for (int j = 0; j < 4; j++)
      if (c + j < dim) w |= (uint32_t)__ldg(row + c + j) << (8 * j);
    return w;
  };
  return make_uint4(word(c0), word(c0 + 4), word(c0 + 8), word(c0 + 12));
}

__device__ __forceinline__ bool lex_less(dkey d0, uint32_t i0, dkey d1, uint32_t i1) {
  return d0 < d1 || (d0 == d1 && i0 < i1);
}

// merge the (sorted ascending across lanes) pending entries of one query into its k-list
__device__ __forceinline__ void merge_pending(dkey pd, uint32_t pi, const volatile dkey *src_d,
                                              const volatile uint32_t *src_i, dkey *dst_d, uint32_t *dst_i,
                                              uint32_t k, int lane, dkey &thr_d, uint32_t &thr_i) {
  // bitonic sort of 32 (d,id) pairs across the warp
#pragma unroll
  for (int sz = 2; sz <= 32; sz <<= 1) {
#pragma unroll
    for (int st = sz >> 1; st >= 1; st >>= 1) {
      dkey od = __shfl_xor_sync(0xffffffffu, pd, st);
      uint32_t oi = __shfl_xor_sync(0xffffffffu, pi, st);
      bool keep_min = (((lane & sz) == 0) == ((lane & st) == 0));
      bool take = keep_min ? lex_less(od, oi, pd, pi) : lex_less(pd, pi, od, oi);
      if (take) { pd = od; pi = oi; }
    }
  }
  // list elements: new position = j + #{pending < list[j]}
  for (uint32_t j0 = 0; j0 < k; j0 += 32) {
    uint32_t j = j0 + lane;
    dkey ld = DINF; uint32_t li = INVALID_ID;
    if (j < k) { ld = src_d[j]; li = src_i[j]; }
    uint32_t cnt = 0;
#pragma unroll
    for (int s = 0; s < 32; s++) {
      dkey sd = __shfl_sync(0xffffffffu, pd, s);
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
  thr_d = ((const volatile dkey *)dst_d)[k - 1];
  thr_i = ((const volatile uint32_t *)dst_i)[k - 1];
}

// sum of squares of every row, one warp per row
__global__ void u8_norms_kernel(const uint8_t *rows, size_t stride, uint32_t n, uint32_t dim, int vec, uint32_t *out) {
  const uint32_t r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= n) return;
  const uint8_t *row = rows + (size_t)r * stride;
  uint32_t acc = 0;
  for (uint32_t c = lane * 16; c < dim; c += 512) {
    uint4 v = load16(row, c, dim, vec);
    acc = __dp4a(v.x, v.x, acc); acc = __dp4a(v.y, v.y, acc); acc = __dp4a(v.z, v.z, acc); acc = __dp4a(v.w, v.w, acc);
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) out[r] = acc;
}

// running top-k state of the CTA's 64 queries (each owned by one warp)
struct TopkSmem {
  dkey pend_d[QT][PEND]; uint32_t pend_i[QT][PEND];
  uint32_t pend_n[QT]; dkey thr_d[QT]; uint32_t thr_i[QT]; uint8_t curbuf[QT];
  dkey gth[QT];  // copy of BfParams::gthr for the CTA's queries, refreshed every few tiles
};

// Merge query ql's pending entries into its k-list. NOT inlined on purpose: the insertion pass below is fully
// unrolled over the 4x8 register tile and tests two queries per step, and 64 inlined copies of the bitonic sort +
// merge made the pass instruction-fetch bound (1M x 128: 3x the cost of the tile's arithmetic).
__device__ __noinline__ void flush_query(TopkSmem *st, dkey *l0d, dkey *l1d, uint32_t *l0i, uint32_t *l1i, size_t off,
                                         uint32_t k, int ql, int lane, dkey *gthr_q) {
  uint32_t cnt = st->pend_n[ql];
  if (cnt == 0) return;
  dkey pd = (uint32_t)lane < cnt ? st->pend_d[ql][lane] : DINF;
  uint32_t pi = (uint32_t)lane < cnt ? st->pend_i[ql][lane] : INVALID_ID;
  int c = st->curbuf[ql];
  dkey td; uint32_t ti;
  merge_pending(pd, pi, (c ? l1d : l0d) + off, (c ? l1i : l0i) + off, (c ? l0d : l1d) + off, (c ? l0i : l1i) + off, k,
                lane, td, ti);
  if (lane == 0) {
    st->curbuf[ql] = c ^ 1; st->pend_n[ql] = 0; st->thr_d[ql] = td; st->thr_i[ql] = ti;
    if (td < st->gth[ql]) { st->gth[ql] = td; atomicMin(gthr_q, td); }  // a full list: publish its k-th distance
  }
  __syncwarp();
}

__global__ void __launch_bounds__(NT, 2) bf_knn_u8_kernel(const BfParams p) {
  __shared__ __align__(16) uint32_t Qs[DK][QS];
  __shared__ __align__(16) uint32_t Bs[DK][BS];
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
    st.gth[i] = DINF; st.pend_n[i] = 0; st.thr_d[i] = DINF; st.thr_i[i] = INVALID_ID; st.curbuf[i] = 0;
  }
  // lists start as (+inf, INVALID)
  for (uint32_t qq = 0; qq < QT; qq++) {
    uint32_t qg = q0 + qq;
    if (qg >= p.nq) break;
    for (uint32_t j = tid; j < p.k; j += NT) {
      p.list_d[0][lbase + (size_t)qg * p.k + j] = DINF;
      p.list_i[0][lbase + (size_t)qg * p.k + j] = INVALID_ID;
    }
  }
  __syncthreads();

  auto flush = [&](int ql) {  // warp-cooperative; ql is warp-uniform and owned by this warp
    flush_query(&st, p.list_d[0], p.list_d[1], p.list_i[0], p.list_i[1], lbase + (size_t)(q0 + ql) * p.k, p.k, ql, lane,
                p.gthr + (q0 + ql));
  };

  // Register double buffer: the global loads of stage s+1 are issued before the arithmetic of stage s, so their
  // latency hides behind 32 x 32 dp4a per thread instead of sitting in front of the CTA barrier (ncu of the
  // single-buffered form: barrier = top stall, issue slots 30 % busy).
  uint4 pq[2], pb[2][2];
  auto fetch = [&](uint32_t fb0, uint32_t fc0) {  // 4 threads per row, two 16-byte chunks each
#pragma unroll
    for (int half = 0; half < 2; half++) {
      const int r = tid >> 2, wc = ((tid & 3) + 4 * half) * 4;
      const uint32_t qg = q0 + r;
      pq[half] = make_uint4(0, 0, 0, 0);
      if (qg < p.nq) pq[half] = load16(p.q + (size_t)qg * p.qstride, fc0 + wc * 4, p.dim, p.qvec);
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const uint32_t bg = fb0 + r + h * 64;
        pb[half][h] = make_uint4(0, 0, 0, 0);
        if (bg < b_end) pb[half][h] = load16(p.base + (size_t)bg * p.bstride, fc0 + wc * 4, p.dim, p.bvec);
      }
    }
  };
  if (b_begin < b_end) fetch(b_begin, 0);

  for (uint32_t b0 = b_begin; b0 < b_end; b0 += BT) {
    dkey acc[4][8];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 8; j++) acc[i][j] = 0u;

    for (uint32_t c0 = 0; c0 < p.dim; c0 += DK * 4) {
      // stage Q (64 rows x 128 bytes) and B (128 rows x 128 bytes) as packed u32 words, word-major
#pragma unroll
      for (int half = 0; half < 2; half++) {
        const int r = tid >> 2, wc = ((tid & 3) + 4 * half) * 4;
        const uint4 v = pq[half];
        Qs[wc + 0][r] = v.x; Qs[wc + 1][r] = v.y; Qs[wc + 2][r] = v.z; Qs[wc + 3][r] = v.w;
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int rb = r + h * 64;
          const uint4 w = pb[half][h];
          Bs[wc + 0][rb] = w.x; Bs[wc + 1][rb] = w.y; Bs[wc + 2][rb] = w.z; Bs[wc + 3][rb] = w.w;
        }
      }
      __syncthreads();
      {
        uint32_t nc0 = c0 + DK * 4, nb0 = b0;
        if (nc0 >= p.dim) { nc0 = 0; nb0 = b0 + BT; }
        if (nb0 < b_end) fetch(nb0, nc0);
      }
#pragma unroll
      for (int d = 0; d < DK; d++) {
        uint4 qv = *reinterpret_cast<const uint4 *>(&Qs[d][ty * 4]);
        uint4 bA = *reinterpret_cast<const uint4 *>(&Bs[d][tx * 4]);
        uint4 bB = *reinterpret_cast<const uint4 *>(&Bs[d][64 + tx * 4]);
        const uint32_t qa[4] = {qv.x, qv.y, qv.z, qv.w};
        const uint32_t ba[8] = {bA.x, bA.y, bA.z, bA.w, bB.x, bB.y, bB.z, bB.w};
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 8; j++) {
            acc[i][j] = __dp4a(qa[i], ba[j], acc[i][j]);  // sum of 4 byte products
          }
      }
      __syncthreads();
    }

    // ---- running top-k: this warp owns queries ty*4+i of its two ty values
    uint32_t qn[4], bn[8];
#pragma unroll
    for (int i = 0; i < 4; i++) { uint32_t qg = q0 + ty * 4 + i; qn[i] = qg < p.nq ? __ldg(p.qnorm + qg) : 0u; }
#pragma unroll
    for (int j = 0; j < 8; j++) {
      uint32_t bg = b0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      bn[j] = bg < b_end ? __ldg(p.bnorm + bg) : 0u;
    }
    // distances in place; a tile in which no lane of the warp beats its query's current k-th (thresholds only
    // tighten) skips the insertion pass altogether
    // Every split of a query keeps its own list, but an entry farther than the k-th distance ANY split has reached
    // cannot be among the k nearest overall (equal distances still go through: the id decides), so the splits share
    // that bound through global memory and only the first ones pay the full list warm-up. The merged result is the
    // same exact top-k; which non-members a split happens to keep is the only thing that varies from run to run.
    if ((((b0 - b_begin) / BT) & 3u) == 0) {
      if (lane < 8) {
        const int ql = warp * 8 + lane;
        if (q0 + ql < p.nq) {
          dkey g = *((volatile dkey *)(p.gthr + q0 + ql));
          if (g < st.gth[ql]) st.gth[ql] = g;
        }
      }
      __syncwarp();
    }
    uint32_t mask = 0;  // bit i*8+j: this lane's tile entry (i,j) beats its query's current k-th
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int ql = ty * 4 + i;
      const uint32_t qg = q0 + ql;
      const dkey td = st.thr_d[ql]; const uint32_t ti = st.thr_i[ql];
      const dkey tg = st.gth[ql];
#pragma unroll
      for (int j = 0; j < 8; j++) {
        uint32_t bg = b0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
        // |a-b|^2 = |a|^2 + |b|^2 - 2 a.b, exact in u32 (2 a.b <= |a|^2 + |b|^2 <= 2 * 255^2 * 32768 < 2^32)
        dkey d = qn[i] + bn[j] - 2u * acc[i][j];
        acc[i][j] = d;
        if (qg < p.nq && bg < b_end && d <= tg && lex_less(d, bg, td, ti)) mask |= 1u << (i * 8 + j);
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
        dkey d = acc[i][j];
        bool pass = qg < p.nq && bg < b_end && d <= st.gth[ql] && lex_less(d, bg, st.thr_d[ql], st.thr_i[ql]);
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
__global__ void bf_u8_finalize_kernel(const BfParams p, uint32_t *out_i, dkey *out_d) {
  const uint32_t qg = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (qg >= p.nq) return;
  const uint32_t total = p.splits * p.k;
  auto get = [&](uint32_t e, dkey &d, uint32_t &i) {
    uint32_t s = e / p.k, j = e % p.k;
    int c = p.cur[(size_t)s * p.nq + qg];
    size_t off = ((size_t)s * p.nq + qg) * p.k + j;
    d = p.list_d[c][off]; i = p.list_i[c][off];
  };
  for (uint32_t e = lane; e < total; e += 32) {
    dkey d; uint32_t i;
    get(e, d, i);
    uint32_t s = e / p.k, j = e % p.k;
    // rank = own index within its sorted list + number of smaller entries in the other lists
    uint32_t rank = j;
    for (uint32_t s2 = 0; s2 < p.splits; s2++) {
      if (s2 == s) continue;
      uint32_t lo = 0, hi = p.k;
      while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        dkey d2; uint32_t i2;
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

int launch_bf_knn_u8(b200ann_handle *h, const uint8_t *base, size_t n, size_t dim, size_t bstride, const uint8_t *q,
                     size_t nq, size_t qstride, size_t k, uint32_t *ids, int32_t *dists) {
  if (nq == 0) return 0;
  BfParams p = {};
  p.base = base; p.bstride = bstride; p.n = (uint32_t)n; p.dim = (uint32_t)dim;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.k = (uint32_t)k;
  p.bvec = ((uintptr_t)base % 16 == 0) && (bstride % 16 == 0);
  p.qvec = ((uintptr_t)q % 16 == 0) && (qstride % 16 == 0);
  uint32_t qtiles = (uint32_t)((nq + QT - 1) / QT);
  // 2 CTAs are resident per SM (127 registers x 256 threads). A grid slightly above a whole number of waves wastes
  // most of the last one (1M x 128, 10k queries: 157 query tiles x 2 splits = 314 CTAs on 296 slots = 2 waves), and
  // every split pays its own top-k warm-up: pick the split count with the least time per CTA of useful work.
  const uint32_t slots = 2u * (uint32_t)h->sm_count;
  uint32_t cap = (uint32_t)((n + 4 * BT - 1) / (4 * BT));
  if (cap > 64) cap = 64;
  if (cap < 1) cap = 1;
  uint32_t splits = 1; double best_cost = 1e300;
  for (uint32_t s = 1; s <= cap; s++) {
    double ctas = (double)qtiles * s, waves = (double)((qtiles * (uint64_t)s + slots - 1) / slots);
    double cost = waves / ctas * (1.0 + 0.02 * s);
    if (cost < best_cost * 0.999) { best_cost = cost; splits = s; }
  }
  if (const char *e = getenv("B200ANN_U8_SPLITS")) {  // A/B knob for the heuristic above
    int v = atoi(e);
    if (v >= 1 && (uint32_t)v <= cap) splits = (uint32_t)v;
  }
  while (splits > 1 && (size_t)splits * nq * k * 16 > ((size_t)1 << 30)) splits--;  // <= 1 GiB of partial lists
  p.splits = splits;
  size_t cnt = (size_t)splits * nq * k;
  DevBuf<dkey> ld0, ld1; DevBuf<uint32_t> li0, li1; DevBuf<uint8_t> cur;
  CU_CHECK(h, ld0.alloc(cnt, h->stream)); CU_CHECK(h, ld1.alloc(cnt, h->stream));
  CU_CHECK(h, li0.alloc(cnt, h->stream)); CU_CHECK(h, li1.alloc(cnt, h->stream));
  CU_CHECK(h, cur.alloc((size_t)splits * nq, h->stream));
  DevBuf<dkey> gthr;
  CU_CHECK(h, gthr.alloc(nq, h->stream));
  CU_CHECK(h, cudaMemsetAsync(gthr.p, 0x7F, nq * sizeof(dkey), h->stream));  // 0x7F7F7F7F > any distance, <= DINF
  p.gthr = gthr.p;
  p.list_d[0] = ld0.p; p.list_d[1] = ld1.p; p.list_i[0] = li0.p; p.list_i[1] = li1.p; p.cur = cur.p;
  DevBuf<uint32_t> bnorm, qnorm;
  CU_CHECK(h, bnorm.alloc(n, h->stream)); CU_CHECK(h, qnorm.alloc(nq, h->stream));
  p.bnorm = bnorm.p; p.qnorm = qnorm.p;
  if (n) { u8_norms_kernel<<<(unsigned)((n + 7) / 8), 256, 0, h->stream>>>(base, bstride, p.n, p.dim, p.bvec, bnorm.p); h->launches++; }
  u8_norms_kernel<<<(unsigned)((nq + 7) / 8), 256, 0, h->stream>>>(q, qstride, p.nq, p.dim, p.qvec, qnorm.p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  dim3 grid(qtiles, splits);
  bf_knn_u8_kernel<<<grid, NT, 0, h->stream>>>(p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  bf_u8_finalize_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(p, ids, reinterpret_cast<dkey *>(dists));
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

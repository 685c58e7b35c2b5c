// Filtered / deleted-element HNSW search: hnswlib::HierarchicalNSW::searchKnn(query, k, isIdAllowed) when a filter
// functor is given or elements are marked deleted (hnswalg.h:1305-1313 -> searchBaseLayerST<false>, :309-440, no
// stop condition). The host functor becomes a device bitmap over INTERNAL ids (bit = allowed and not deleted).
//
// One warp per query, persistent over the batch. Both of the reference's heaps live in shared memory: top_candidates
// (allowed elements only, <= ef) as an ascending (distance,id) array, candidate_set (every considered element) as a
// binary min-heap on (distance,id) maintained by lane 0 (O(log) per push / pop; a selective filter makes it hold
// tens of thousands of entries, where shifting a sorted array costs 50x more). Ties are broken by id in both, so
// the pop order and the eviction order are those of oracle.c's sorted arrays.
// candidate_set is unbounded in the reference (a selective filter makes the walk consider ~ef/selectivity elements
// before top_candidates fills). The first pass holds `ccap` entries per query: the farthest entry is dropped when the
// heap is full, and the query is reported INEXACT as soon as the reference could have expanded a dropped entry
// (oracle.c hnsw_base_pools_filtered states the rule) - a query that is not reported is the reference's answer bit
// for bit. Reported queries are searched again (retry pass) with the heap in global memory and room for all n
// elements, which cannot overflow. The visited set is exact (the reference's per-query tag array,
// visited_list_pool.h:10-31): one bitmap of n bits per resident warp in global memory, cleared per query.
#include <string.h>
#include "beam.cuh"

using namespace beam;

namespace {

constexpr int FW = 4;  // warps per block

struct FilterParams {
  SearchParams s;
  const uint32_t *allow;   // [ceil(n/32)] bit i = internal id i may be returned
  uint32_t ccap;           // candidate_set capacity (entries)
  uint32_t *visited;       // [total_warps, vwords]
  uint32_t vwords;         // multiple of 4
  uint32_t total_warps;
  uint32_t wpb;            // warps per block
  float *gheap_d;          // retry pass: candidate_set heaps in global memory, [total_warps, ccap] (NULL: shared memory)
  uint32_t *gheap_i;
  uint8_t *inexact;        // [nq]
  const uint8_t *only;     // NULL, or [nq]: search only the queries with only[q] != 0 (retry pass)
  // EpsilonSearchStopCondition (stop_condition.h:219-275) with max_num_candidates = ef: eps = +inf, min_cand = 0
  // switches it off (the plain filtered search). counts [nq] = results left by filter_results, or NULL.
  float eps;
  uint32_t min_cand;
  uint32_t *counts;
  // MultiVectorSearchStopCondition (stop_condition.h:146-215): doc [n] = document id of every element (NULL = off).
  // The result set is bounded by its number of DISTINCT documents (efcoll), not by its size; afterwards the farthest
  // results are dropped until docs_keep documents are left. s.L = capacity of the result list (an overflow is reported
  // through `inexact`).
  const uint32_t *doc;
  uint32_t efcoll, docs_keep;
};

__host__ __device__ inline uint32_t fsmem_words(const FilterParams &p) {
  return smem_words(p.s.pool_cap, p.s.cand_cap, 0) + (p.gheap_d ? 0u : 2 * p.ccap) + (p.doc ? p.s.pool_cap : 0u);
}

// number of entries of a[lo, lo+n) ordered before (d,id)
__device__ __forceinline__ uint32_t count_before(const float *ad, const uint32_t *ai, uint32_t lo, uint32_t n, float d,
                                                 uint32_t id, int lane) {
  uint32_t pos = 0;
  for (uint32_t j0 = 0; j0 < n; j0 += 32) {
    uint32_t j = j0 + lane;
    bool in = j < n;
    float pd = in ? ad[lo + j] : 0.f;
    uint32_t pi = in ? ai[lo + j] : 0u;
    bool before = in && (pd < d || (pd == d && pi < id));
    pos += __popc(__ballot_sync(0xffffffffu, before));
  }
  return pos;
}
// entries [from, to) move up by one slot
__device__ __forceinline__ void shift_up(float *ad, uint32_t *ai, int from, int to, int lane) {
  if (to <= from) return;
  for (int jb = from + ((to - 1 - from) / 32) * 32; jb >= from; jb -= 32) {
    int j = jb + lane;
    bool mv = j < to;
    float td = 0.f; uint32_t ti = 0;
    if (mv) { td = ad[j]; ti = ai[j]; }
    __syncwarp();
    if (mv) { ad[j + 1] = td; ai[j + 1] = ti; }
    __syncwarp();
  }
}
__device__ __forceinline__ void shift_up1(uint32_t *a, int from, int to, int lane) {
  if (to <= from) return;
  for (int jb = from + ((to - 1 - from) / 32) * 32; jb >= from; jb -= 32) {
    int j = jb + lane;
    bool mv = j < to;
    uint32_t t = 0;
    if (mv) t = a[j];
    __syncwarp();
    if (mv) a[j + 1] = t;
    __syncwarp();
  }
}
// is document D among a[0, n)?
__device__ __forceinline__ bool has_doc(const uint32_t *a, uint32_t n, uint32_t D, int lane) {
  bool f = false;
  for (uint32_t j0 = 0; j0 < n && !f; j0 += 32) {
    uint32_t j = j0 + lane;
    f = __any_sync(0xffffffffu, j < n && a[j] == D);
  }
  return f;
}
__device__ __forceinline__ bool pless(float da, uint32_t ia, float db, uint32_t ib) {
  return da < db || (da == db && ia < ib);
}
// binary min-heap on (d,id); called by ONE lane
__device__ __forceinline__ void heap_sift_up(float *hd, uint32_t *hi, uint32_t i, float d, uint32_t id) {
  while (i > 0) {
    uint32_t par = (i - 1) >> 1;
    float pd = hd[par]; uint32_t pi = hi[par];
    if (!pless(d, id, pd, pi)) break;
    hd[i] = pd; hi[i] = pi;
    i = par;
  }
  hd[i] = d; hi[i] = id;
}
// removes the root of a heap of n entries (n >= 1)
__device__ __forceinline__ void heap_pop(float *hd, uint32_t *hi, uint32_t n) {
  n--;
  float d = hd[n]; uint32_t id = hi[n];
  uint32_t i = 0;
  for (;;) {
    uint32_t l = 2 * i + 1;
    if (l >= n) break;
    float cdv = hd[l]; uint32_t civ = hi[l]; uint32_t c = l;
    if (l + 1 < n) {
      float rd = hd[l + 1]; uint32_t ri = hi[l + 1];
      if (pless(rd, ri, cdv, civ)) { cdv = rd; civ = ri; c = l + 1; }
    }
    if (!pless(cdv, civ, d, id)) break;
    hd[i] = cdv; hi[i] = civ;
    i = c;
  }
  if (n) { hd[i] = d; hi[i] = id; }
}
// slot of the largest (d,id) of a heap of n entries: it is a leaf; warp-wide scan of the leaves
__device__ __forceinline__ uint32_t heap_argmax(const float *hd, const uint32_t *hi, uint32_t n, int lane) {
  float bd = -INFINITY; uint32_t bi = 0, bs = 0xffffffffu;
  for (uint32_t j = n / 2 + lane; j < n; j += 32) {
    float d = hd[j]; uint32_t id = hi[j];
    if (bs == 0xffffffffu || pless(bd, bi, d, id)) { bd = d; bi = id; bs = j; }
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    float od = __shfl_xor_sync(0xffffffffu, bd, off);
    uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, off);
    uint32_t os = __shfl_xor_sync(0xffffffffu, bs, off);
    if (os != 0xffffffffu && (bs == 0xffffffffu || pless(bd, bi, od, oi))) { bd = od; bi = oi; bs = os; }
  }
  return bs;
}

template <int TEAM, int NCH, int IP>
__global__ void __launch_bounds__(FW * 32) hnsw_filtered_kernel(const FilterParams fp) {
  extern __shared__ uint32_t smem[];
  const SearchParams &p = fp.s;
  const int warp = threadIdx.x >> 5;
  const uint32_t gw = blockIdx.x * fp.wpb + warp;
  if (gw >= fp.total_warps) return;
  uint32_t *base = smem + (size_t)warp * fsmem_words(fp);
  Beam<TEAM, NCH, IP, true> b(p, base);
  float *cd = reinterpret_cast<float *>(base + smem_words(p.pool_cap, p.cand_cap, 0));
  uint32_t *ci = base + smem_words(p.pool_cap, p.cand_cap, 0) + fp.ccap;
  if (fp.gheap_d) {  // retry pass: the heap lives in global memory (L2), sized so that it cannot overflow
    cd = fp.gheap_d + (size_t)gw * fp.ccap;
    ci = fp.gheap_i + (size_t)gw * fp.ccap;
  }
  uint32_t *vis = fp.visited + (size_t)gw * fp.vwords;
  const int lane = b.lane;
  const uint32_t ef = p.L;
  const bool docm = fp.doc != nullptr;
  uint32_t *pool_doc = base + smem_words(p.pool_cap, p.cand_cap, 0) + (fp.gheap_d ? 0u : 2 * fp.ccap);  // doc mode only

  for (uint32_t qi = gw; qi < p.nq; qi += fp.total_warps) {
    if (fp.only && !fp.only[qi]) continue;
    b.begin(p.q + (size_t)qi * p.qstride);
    {  // empty visited set
      uint4 *v4 = reinterpret_cast<uint4 *>(vis);
      for (uint32_t i = lane; i < fp.vwords / 4; i += 32) v4[i] = make_uint4(0, 0, 0, 0);
    }
    float curdist;
    const uint32_t ep = b.descend_upper(p.ep, curdist);
    uint32_t nt = 0, nc = 0, ndocs = 0;
    float LB, dmin = INFINITY;
    bool bad = false;
    const bool ep_ok = (__ldg(fp.allow + (ep >> 5)) >> (ep & 31)) & 1u;
    if (ep_ok) {  // :325-333 (the reference evaluates the entry point again; same value)
      LB = curdist;
      if (lane == 0) { b.pool_d[0] = curdist; b.pool_i[0] = ep; cd[0] = curdist; ci[0] = ep; }
      nt = 1;
      if (docm) { if (lane == 0) pool_doc[0] = __ldg(fp.doc + ep); ndocs = 1; }
    } else {      // :334-337: queued at FLT_MAX, never evaluated by the base layer
      LB = 3.402823466e+38f;
      if (lane == 0) { cd[0] = LB; ci[0] = ep; }
    }
    nc = 1;
    if (lane == 0) vis[ep >> 5] = 1u << (ep & 31);
    __syncwarp();

    for (;;) {
      const float c0d = nc ? cd[0] : INFINITY;
      // "the result set is full": ef elements, or (multi-vector) efcoll distinct documents (stop_condition.h:186-199)
      bool full = docm ? ndocs == fp.efcoll : nt == ef;
      if (dmin < INFINITY && (nc == 0 || dmin <= c0d) && !(dmin > LB && full) &&
          !(dmin > fp.eps && nt >= fp.min_cand)) bad = true;
      if (nc == 0) break;
      if (c0d > LB && full) break;  // :355, stop_condition.h:243 / :186
      if (c0d > fp.eps && nt >= fp.min_cand) break;  // stop_condition.h:247: out of the epsilon ball, enough seen
      const uint32_t node = ci[0];
      __syncwarp();
      if (lane == 0) heap_pop(cd, ci, nc);
      nc--;
      b.n_hops++;
      // unvisited neighbours, in row order, marked visited (:385-386)
      const uint32_t *row = p.graph + (size_t)node * p.gstride;
      uint32_t m = 0;
      __syncwarp();
      for (uint32_t j0 = 0; j0 < p.gstride; j0 += 32) {
        uint32_t j = j0 + lane;
        uint32_t id = j < p.gstride ? __ldg(row + j) : INVALID_ID;
        bool valid = id != INVALID_ID;
        if (valid) {
          uint32_t bit = 1u << (id & 31);
          valid = !(atomicOr(vis + (id >> 5), bit) & bit);
        }
        unsigned mask = __ballot_sync(0xffffffffu, valid);
        if (valid) b.cand_i[m + __popc(mask & ((1u << lane) - 1))] = id;
        m += __popc(mask);
        if (__any_sync(0xffffffffu, id == INVALID_ID)) break;
      }
      __syncwarp();
      b.eval(m);
      for (uint32_t c0 = 0; c0 < m; c0 += 32) {
        uint32_t c = c0 + lane;
        float d = c < m ? b.cand_d[c] : INFINITY;
        full = docm ? ndocs == fp.efcoll : nt == ef;
        bool pass = c < m && (!full || d < LB);
        unsigned mask = __ballot_sync(0xffffffffu, pass);
        while (mask) {
          int bsel = __ffs(mask) - 1;
          mask &= mask - 1;
          float dd = __shfl_sync(0xffffffffu, d, bsel);
          uint32_t id = b.cand_i[c0 + bsel];
          full = docm ? ndocs == fp.efcoll : nt == ef;
          if (!(!full || dd < LB)) continue;  // :395 with the live bound
          // candidate_set.emplace (:398)
          if (nc == fp.ccap) {  // full: the farthest of (entries, new) goes
            uint32_t slot = heap_argmax(cd, ci, nc, lane);
            float ld = cd[slot];
            uint32_t li = ci[slot];
            float yd = dd;
            __syncwarp();
            if (pless(dd, id, ld, li)) {
              yd = ld;
              if (lane == 0) heap_sift_up(cd, ci, slot, dd, id);  // a leaf got smaller: it can only move up
            }
            if (!(full && yd > LB) && yd < dmin) dmin = yd;
          } else {
            if (lane == 0) heap_sift_up(cd, ci, nc, dd, id);
            nc++;
          }
          __syncwarp();
          // top_candidates.emplace for allowed elements (:406-408), pop the farthest beyond ef (:413-424)
          if ((__ldg(fp.allow + (id >> 5)) >> (id & 31)) & 1u) {
            if (!docm) {
              if (nt == ef) nt--;  // the new element is closer than the farthest one (dd < LB)
              uint32_t before = count_before(b.pool_d, b.pool_i, 0, nt, dd, id, lane);
              shift_up(b.pool_d, b.pool_i, (int)before, (int)nt, lane);
              if (lane == 0) { b.pool_d[before] = dd; b.pool_i[before] = id; }
              nt++;
              __syncwarp();
            } else {
              // add_point_to_result (stop_condition.h:167-174), then drop the farthest results while more than efcoll
              // documents are held (should_remove_extra / remove_point_from_result, :176-183, :201-204)
              const uint32_t D = __ldg(fp.doc + id);
              uint32_t before = count_before(b.pool_d, b.pool_i, 0, nt, dd, id, lane);
              if (nt == ef) {          // the list is out of slots: the farthest of (list, new) cannot be kept
                bad = true;
                if (before == nt) before = 0xffffffffu;
                else { const uint32_t Dl = pool_doc[nt - 1]; nt--; if (!has_doc(pool_doc, nt, Dl, lane)) ndocs--; }
              }
              if (before != 0xffffffffu) {
                if (!has_doc(pool_doc, nt, D, lane)) ndocs++;
                shift_up(b.pool_d, b.pool_i, (int)before, (int)nt, lane);
                shift_up1(pool_doc, (int)before, (int)nt, lane);
                if (lane == 0) { b.pool_d[before] = dd; b.pool_i[before] = id; pool_doc[before] = D; }
                nt++;
                __syncwarp();
                while (ndocs > fp.efcoll) {
                  const uint32_t Dl = pool_doc[nt - 1];
                  nt--;
                  if (!has_doc(pool_doc, nt, Dl, lane)) ndocs--;
                }
              }
            }
          }
          if (nt) LB = b.pool_d[nt - 1];  // :426-427
        }
      }
    }

    // filter_results (stop_condition.h:265-272): the list is ascending, so dropping the tail beyond epsilon leaves
    // the entries with distance <= epsilon (NaN-free data); nt <= ef = max_num_candidates already
    if (docm) {  // multi-vector filter_results (:206-219): the farthest results go until docs_keep documents are left
      while (ndocs > fp.docs_keep && nt > 0) {
        const uint32_t Dl = pool_doc[nt - 1];
        nt--;
        if (!has_doc(pool_doc, nt, Dl, lane)) ndocs--;
      }
    }
    uint32_t nres = 0;
    for (uint32_t i0 = 0; i0 < nt; i0 += 32) {
      uint32_t i = i0 + lane;
      nres += __popc(__ballot_sync(0xffffffffu, i < nt && !(b.pool_d[i] > fp.eps)));
    }
    for (uint32_t i = lane; i < p.K; i += 32) {
      bool ok = i < nres;
      uint32_t id = ok ? b.pool_i[i] : INVALID_ID;
      p.labels_out[(size_t)qi * p.K + i] = ok ? (p.labels ? __ldg(p.labels + id) : (uint64_t)id) : ~(uint64_t)0;
      if (p.dists) p.dists[(size_t)qi * p.K + i] = ok ? b.pool_d[i] : INFINITY;
    }
    if (lane == 0) {
      fp.inexact[qi] = bad ? 1 : 0;
      if (fp.counts) fp.counts[qi] = nres < p.K ? nres : p.K;
      if (p.stats) { p.stats[2 * (size_t)qi] = b.n_dist; p.stats[2 * (size_t)qi + 1] = b.n_hops; }
    }
    __syncwarp();
  }
}

template <int TEAM, int NCH, int IP>
int launch_ft(b200ann_handle *h, FilterParams &fp) {
  size_t smem = (size_t)fp.wpb * fsmem_words(fp) * 4;
  auto kern = hnsw_filtered_kernel<TEAM, NCH, IP>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((fp.total_warps + fp.wpb - 1) / fp.wpb), block(fp.wpb * 32);
  kern<<<grid, block, smem, h->stream>>>(fp);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

template <int IP>
int launch_fdim(b200ann_handle *h, FilterParams &fp) {
  int nch = fp.s.nch;
  if (nch <= 8) return launch_ft<8, 1, IP>(h, fp);
  if (nch <= 16) return launch_ft<8, 2, IP>(h, fp);
  if (nch <= 32) return launch_ft<8, 4, IP>(h, fp);
  if (nch <= 64) return launch_ft<8, 8, IP>(h, fp);
  if (nch <= 128) return launch_ft<32, 4, IP>(h, fp);
  if (nch <= 256) return launch_ft<32, 8, IP>(h, fp);
  if (nch <= 512) return launch_ft<32, 16, IP>(h, fp);
  return set_err(h, B200ANN_ERR_ARG, "hnsw_search_filtered: dim > 2048 (8192 for byte rows) is not supported");
}

}  // namespace

// big = false: candidate_set heaps of min(ccap ? ccap : 2048, what fits) entries in shared memory. big = true (retry
// pass): the heaps move to global memory with up to n entries each (n entries cannot overflow: a query then is always
// exact), as many resident warps as 16 GiB of heaps (a quarter of the free memory at most) and 2 GiB of visited bitmaps allow. only = NULL or a device mask of
// the queries to search (may alias inexact). All pointers are device pointers.
int launch_hnsw_search_filtered(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                                size_t k, size_t ef, const uint32_t *allow_bits, size_t ccap, bool big,
                                const uint8_t *only, uint64_t *labels, float *dists, uint32_t *stats, uint8_t *inexact,
                                size_t *ccap_used, float eps, size_t min_cand, uint32_t *counts, bool internal_ids,
                                const uint32_t *doc, size_t efcoll, size_t docs_keep) {
  FilterParams fp = {};
  fp.doc = doc; fp.efcoll = (uint32_t)efcoll; fp.docs_keep = (uint32_t)docs_keep;
  fp.eps = eps; fp.min_cand = (uint32_t)min_cand; fp.counts = counts;
  if (ix_u8(ix) && eps < INFINITY)  // int distances travel as bit patterns (dist.cuh): d > eps <=> bits(d) > bits(floor(eps))
  {
    const int e = (int)fminf(floorf(eps), 2139095039.f);
    fp.eps = -1.0f;
    if (eps >= 0.f) memcpy(&fp.eps, &e, 4);
  }
  SearchParams &p = fp.s;
  // epsilon mode (counts given): k is only the row stride of the outputs, the walk keeps ef = max_num_candidates
  // results; more than n + 1 can never be held, and with n + 1 slots "nt == ef" stays false like in the reference
  size_t L = ef > k ? ef : k;
  if (counts) L = ef < ix->n + 1 ? ef : ix->n + 1;
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = ix_nch(ix);
  p.graph = ix->links0; p.gstride = (uint32_t)ix->maxM0; p.ep = ix->hnsw_ep;
  p.upper_links = ix->upper_links; p.upper_ptr = ix->upper_ptr; p.maxM = (uint32_t)ix->maxM;
  p.maxlevel = ix->maxlevel; p.labels = internal_ids ? nullptr : ix->labels;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.dim = (uint32_t)ix->dim;
  p.K = (uint32_t)k; p.L = (uint32_t)L; p.seed = 0; p.hbits = 0;
  size_t cc = ix->maxM0 > ix->maxM ? ix->maxM0 : ix->maxM;
  p.cand_cap = round_up((uint32_t)cc, 32);
  p.pool_cap = round_up((uint32_t)L + 1, 32);
  p.labels_out = labels; p.dists = dists; p.stats = stats;
  fp.allow = allow_bits; fp.inexact = inexact; fp.only = only;
  fp.vwords = round_up((uint32_t)((ix->n + 31) / 32), 4);
  const size_t fixed = smem_words(p.pool_cap, p.cand_cap, 0) + (doc ? p.pool_cap : 0u);
  // warps per block: FW, fewer when one warp's result list is large (an epsilon search with max_num_candidates in the
  // thousands), so that the list plus a candidate heap of >= 512 entries still fits the block's 227 KB
  fp.wpb = FW;
  while (fp.wpb > 1 && fixed + 2 * 512 > 227 * 1024 / 4 / fp.wpb) fp.wpb >>= 1;
  const size_t budget = 227 * 1024 / 4 / fp.wpb;
  if (fixed + 2 * 64 > budget)
    return set_err(h, B200ANN_ERR_ARG, "hnsw_search_filtered / epsilon: ef (max_num_candidates) / degree need more than 227 KB "
                                       "shared memory per query (about 28000 results)");
  size_t warps;
  DevBuf<float> gd; DevBuf<uint32_t> gi;
  if (!big) {
    size_t most = (budget - fixed) / 2 / 32 * 32;
    if (ccap == 0) ccap = 2048;
    if (ccap > most) ccap = most;
    fp.ccap = (uint32_t)ccap;
    size_t per_sm = 227 * 1024 / ((size_t)fp.wpb * fsmem_words(fp) * 4);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 8) per_sm = 8;
    warps = (size_t)h->sm_count * per_sm * fp.wpb;
  } else {
    ccap = ix->n;  // every element can be queued at most once (the visited set is exact)
    warps = (size_t)h->sm_count * 6 * fp.wpb;
    size_t free_b = 0, total_b = 0;
    size_t heap_budget = (size_t)16 << 30;  // up to 16 GiB of heaps, at most a quarter of what is free
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && free_b / 4 < heap_budget) heap_budget = free_b / 4;
    size_t by_heap = heap_budget / (ccap * 8);
    if (by_heap < 1) by_heap = 1;
    if (warps > by_heap) warps = by_heap;
  }
  size_t by_mem = ((size_t)2 << 30) / ((size_t)fp.vwords * 4);
  if (by_mem < 1) by_mem = 1;
  if (warps > by_mem) warps = by_mem;
  if (warps > nq) warps = nq;
  fp.total_warps = (uint32_t)warps;
  fp.ccap = (uint32_t)ccap;
  if (ccap_used) *ccap_used = ccap;
  if (big) {
    CU_CHECK(h, gd.alloc(warps * ccap, h->stream));
    CU_CHECK(h, gi.alloc(warps * ccap, h->stream));
    fp.gheap_d = gd.p; fp.gheap_i = gi.p;
  }
  DevBuf<uint32_t> vis;
  CU_CHECK(h, vis.alloc((size_t)fp.total_warps * fp.vwords, h->stream));
  fp.visited = vis.p;
  if (ix_u8(ix)) return launch_fdim<MET_U8>(h, fp);
  return ix->metric == B200ANN_IP ? launch_fdim<MET_IP>(h, fp) : launch_fdim<MET_L2>(h, fp);
}

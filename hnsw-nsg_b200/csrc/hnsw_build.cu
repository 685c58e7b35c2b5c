// HNSW construction on the GPU.
//
// Replaces the addPoint loop of hnswlib::HierarchicalNSW (hnswlib/hnswlib/hnswalg.h:1153-1267: level draw
// :206-210, greedy descent :1214-1239, searchBaseLayer :225-305, mutuallyConnectNewElement :506-630 with
// getNeighborsByHeuristic2 :443-483), as the drivers run it over a whole dataset
// (hnswlib/tests/cpp/sift_1m.cpp:321-336, python_bindings/bindings.cpp:287-303).
//
// The reference inserts one point at a time under per-node mutexes. Here points are inserted in batches
// ("batch form", restated serially by oracle/oracle.c orc_hnsw_build variant 1 and compared link for link):
//   schedule   batch = min(bmax, max(1, inserted / ratio)) consecutive rows (ratio at most 8 below min(2048, n/16) rows); a row whose level exceeds the
//              current top level is a batch of its own (it becomes the new entry point, :1262-1265).
//   insert     one warp per new point, on the graph as it was BEFORE the batch: greedy descent to the point's
//              level, then per level the ef_construction beam (beam.cuh, single-pool form) and the heuristic
//              (occlude_scan) pick <= M links; the point's own rows are written (farthest first, the order the
//              reference pops its max-heap in, :514-517) and every link is recorded as a reverse edge.
//   reverse    reverse edges are grouped per (neighbour, level) row; one warp per touched row appends them in
//              ascending id when they fit in Mmax (M0 = 2M on level 0), else re-runs the heuristic over
//              existing + new candidates sorted by (distance, id) (:581-618 does it one edge at a time).
// Levels come from the same std::default_random_engine / uniform_real_distribution<double> sequence as the
// reference (this file is compiled against the same libstdc++), so level assignment is identical to it.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <random>
#include <vector>
#include "beam.cuh"
#include "topbuf.cuh"

using namespace beam;

namespace {

constexpr int WARPS = 4;

struct InsertParams {
  SearchParams sp;          // graph = links0 (gstride = maxM0), upper_*, ep = entry point, maxlevel, L = ef_construction
  uint32_t s, e;            // batch rows [s, e)
  uint32_t M, nlev;         // links per new point; levels per point in the edge buffer
  const int32_t *levels;
  uint32_t *links0, *upper_links;   // writable views of sp.graph / sp.upper_links
  uint32_t *rev_row, *rev_src; float *rev_d;   // [(e-s) * nlev * M]
  uint32_t *rcount, *touched, *counters;       // [n + rows], [n + rows], {ntouched, maxgroup, cursor}
  unsigned long long *stats;                   // {beam distance evaluations, hops, heuristic evaluations}
  // update mode (hnsw_update_launch, phase B = repairConnectionsForUpdate, hnswalg.h:1073-1140): the batch is
  // upd_ids[s..e) - elements that are IN the graph; the point itself is filtered out of its candidates (:1112-1118)
  // and its new rows go to a staging area (other warps are walking the rows it would overwrite)
  const uint32_t *upd_ids;                     // NULL = insertion of rows [s, e)
  uint32_t *own_stage, *own_ridx;              // [(e-s) * nlev, gstride], [(e-s) * nlev] (row index or INVALID)
};

__host__ __device__ inline uint32_t insert_smem_words(const InsertParams &ip) {
  return smem_words(ip.sp.pool_cap, ip.sp.cand_cap, ip.sp.hbits) + 2 * ip.sp.gstride;
}

template <int TEAM, int NCH, int IP>
__global__ void __launch_bounds__(WARPS * 32) hnsw_insert_kernel(const InsertParams ip) {
  extern __shared__ uint32_t smem[];
  const SearchParams &p = ip.sp;
  const int warp = threadIdx.x >> 5;
  const uint32_t pos = blockIdx.x * WARPS + warp;  // position in the batch
  if (ip.s + pos >= ip.e) return;
  const uint32_t pt = ip.upd_ids ? ip.upd_ids[ip.s + pos] : ip.s + pos;
  uint32_t *base = smem + (size_t)warp * insert_smem_words(ip);
  typedef Beam<TEAM, NCH, IP, true, false, (NCH <= 4 ? 4 : 2)> B;  // 4 rows per team in flight: one warp's hops are the critical path
  B b(p, base);
  uint32_t *res_i = base + smem_words(p.pool_cap, p.cand_cap, p.hbits);
  float *res_d = reinterpret_cast<float *>(res_i + p.gstride);
  NoRec rec;
  const float *qv = p.vec + (size_t)pt * p.stride;
#ifdef BEAM_PROF
  const long long tk0 = clock64();
  long long t_heur = 0;
#endif
  b.begin(qv);
  const int curlevel = ip.levels[pt];
  uint32_t cur = p.ep;
  if (curlevel < p.maxlevel) {  // :1214-1239
    float cd;
    cur = b.descend_upper(p.ep, cd, curlevel);
  }
  uint32_t n_heur = 0;
  for (int level = curlevel < p.maxlevel ? curlevel : p.maxlevel; level >= 0; --level) {
    // ---- searchBaseLayer(cur, point, level)
    b.lvl = level;
    for (uint32_t i = b.lane; i <= b.hmask; i += 32) b.tab[i] = INVALID_ID;
    __syncwarp();
    if (b.lane == 0) b.cand_i[0] = cur;
    b.eval(1);
    if (b.lane == 0) { b.pool_d[0] = b.cand_d[0]; b.pool_i[0] = cur | FLAG; }
    b.test_and_set(cur);
    b.np = 1;
    __syncwarp();
    b.expand_all(rec);
    if (ip.upd_ids) {  // the element finds itself (distance 0 to its own new vector): drop it from the candidates
      int at = -1;
      for (uint32_t j0 = 0; j0 < b.np && at < 0; j0 += 32) {
        uint32_t j = j0 + b.lane;
        unsigned mk = __ballot_sync(0xffffffffu, j < b.np && (b.pool_i[j] & ~FLAG) == pt);
        if (mk) at = (int)j0 + __ffs(mk) - 1;
      }
      if (at >= 0) {
        for (uint32_t j0 = (uint32_t)at; j0 + 1 < b.np; j0 += 32) {
          uint32_t j = j0 + b.lane;
          float td = 0.f; uint32_t ti = 0;
          bool mv = j + 1 < b.np;
          if (mv) { td = b.pool_d[j + 1]; ti = b.pool_i[j + 1]; }
          __syncwarp();
          if (mv) { b.pool_d[j] = td; b.pool_i[j] = ti; }
          __syncwarp();
        }
        b.np--;
      }
      if (b.np == 0) continue;  // nothing but itself was reachable: the rows stay as they are (:1122)
    }
    // ---- getNeighborsByHeuristic2(top_candidates, M): fewer than M -> all kept
    uint32_t nres = 0;
    if (b.np < ip.M) {
      for (uint32_t i = b.lane; i < b.np; i += 32) { res_i[i] = b.pool_i[i] & ~FLAG; res_d[i] = b.pool_d[i]; }
      nres = b.np;
      __syncwarp();
    } else {
#ifdef BEAM_PROF
      const long long th0 = clock64();
#endif
      occlude_scan(b, b.pool_d, b.pool_i, 0, b.np, ip.M, res_i, res_d, nres, n_heur);
      b.load_query(qv);  // occlude_scan used the query registers for the candidate rows
#ifdef BEAM_PROF
      t_heur += clock64() - th0;
#endif
    }
    __syncwarp();
    // ---- own row, farthest first; reverse edges
    uint32_t *row; uint32_t rl;
    if (level == 0) { row = ip.links0 + (size_t)pt * p.gstride; rl = p.gstride; }
    else { row = ip.upper_links + ((size_t)p.upper_ptr[pt] + (level - 1)) * p.maxM; rl = p.maxM; }
    if (ip.upd_ids) {
      row = ip.own_stage + ((size_t)pos * ip.nlev + level) * p.gstride;
      if (b.lane == 0) ip.own_ridx[(size_t)pos * ip.nlev + level] = level == 0 ? pt : p.n + p.upper_ptr[pt] + (level - 1);
    }
    for (uint32_t i = b.lane; i < rl; i += 32) row[i] = i < nres ? res_i[nres - 1 - i] : INVALID_ID;
    const size_t eb = ((size_t)pos * ip.nlev + level) * ip.M;
    for (uint32_t i = b.lane; i < nres; i += 32) {
      uint32_t sn = res_i[i];
      uint32_t ridx = level == 0 ? sn : p.n + p.upper_ptr[sn] + (level - 1);
      ip.rev_row[eb + i] = ridx; ip.rev_src[eb + i] = pt; ip.rev_d[eb + i] = res_d[i];
      uint32_t c = atomicAdd(&ip.rcount[ridx], 1u);
      if (c == 0) ip.touched[atomicAdd(&ip.counters[0], 1u)] = ridx;
      atomicMax(&ip.counters[1], c + 1);
    }
    cur = res_i[0];  // closest selected = next entry point (:520)
    __syncwarp();
  }
  if (b.lane == 0) {
    atomicAdd(ip.stats + 0, (unsigned long long)b.n_dist);
    atomicAdd(ip.stats + 1, (unsigned long long)b.n_hops);
    atomicAdd(ip.stats + 2, (unsigned long long)n_heur);
#ifdef BEAM_PROF
    atomicAdd(ip.stats + 4, (unsigned long long)(clock64() - tk0));
    atomicAdd(ip.stats + 5, (unsigned long long)b.tp[3]);
    atomicAdd(ip.stats + 6, (unsigned long long)b.tp[0]);
    atomicAdd(ip.stats + 7, (unsigned long long)b.tp[1]);
    atomicAdd(ip.stats + 8, (unsigned long long)b.tp[2]);
    atomicAdd(ip.stats + 9, (unsigned long long)t_heur);
    atomicAdd(ip.stats + 10, (unsigned long long)b.tp[4]);
    atomicAdd(ip.stats + 11, (unsigned long long)b.tp[5]);
#endif
  }
}

// offsets of the touched rows' groups in the grouped edge arrays (any order: groups are sorted when used)
__global__ void rev_offsets_kernel(const uint32_t *touched, const uint32_t *counters, const uint32_t *rcount, uint32_t *roff,
                                   uint32_t *cursor) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= counters[0]) return;
  uint32_t r = touched[t];
  roff[r] = atomicAdd(cursor, rcount[r]);
}
__global__ void rev_group_kernel(const uint32_t *rev_row, const uint32_t *rev_src, const float *rev_d, size_t total,
                                 const uint32_t *roff, uint32_t *rfill, uint32_t *g_src, float *g_d) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= total) return;
  uint32_t r = rev_row[e];
  if (r == INVALID_ID) return;
  uint32_t pos = roff[r] + atomicAdd(&rfill[r], 1u);
  g_src[pos] = rev_src[e];
  g_d[pos] = rev_d[e];
}

struct ReverseParams {
  SearchParams sp;          // vec/stride/nch/dim (+ pool_cap/cand_cap/hbits for the smem carve-up); gstride = maxM0
  uint32_t *links0, *upper_links;
  const uint32_t *row_owner;        // [rows]: node that owns upper row r
  const uint32_t *touched, *counters, *roff, *g_src; const float *g_d;
  uint32_t *rcount, *rfill;
  uint32_t mcap;                    // candidates a warp can hold: Mmax0 + largest group
  unsigned long long *stats;
  int update;                       // 1: a source already present in the row is skipped (isUpdate, hnswalg.h:548-560)
};
__host__ __device__ inline uint32_t reverse_smem_words(const ReverseParams &rp) {
  return smem_words(rp.sp.pool_cap, rp.sp.cand_cap, rp.sp.hbits) + 4 * rp.mcap + 2 * rp.sp.gstride;
}

template <int TEAM, int NCH, int IP>
__global__ void __launch_bounds__(WARPS * 32) hnsw_reverse_kernel(const ReverseParams rp) {
  extern __shared__ uint32_t smem[];
  const SearchParams &p = rp.sp;
  const int warp = threadIdx.x >> 5;
  const uint32_t t = blockIdx.x * WARPS + warp;
  if (t >= rp.counters[0]) return;
  uint32_t *base = smem + (size_t)warp * reverse_smem_words(rp);
  typedef Beam<TEAM, NCH, IP, true> B;
  B b(p, base);
  uint32_t *ui = base + smem_words(p.pool_cap, p.cand_cap, p.hbits);   // unsorted ids / distances
  float *ud = reinterpret_cast<float *>(ui + rp.mcap);
  uint32_t *si = ui + 2 * rp.mcap;                                     // sorted
  float *sd = reinterpret_cast<float *>(si + rp.mcap);
  uint32_t *res_i = si + 2 * rp.mcap;
  float *res_d = reinterpret_cast<float *>(res_i + p.gstride);
  const int lane = b.lane;

  const uint32_t ridx = rp.touched[t];
  uint32_t node, cap; uint32_t *row;
  if (ridx < p.n) { node = ridx; cap = p.gstride; row = rp.links0 + (size_t)node * p.gstride; }
  else { node = rp.row_owner[ridx - p.n]; cap = p.maxM; row = rp.upper_links + (size_t)(ridx - p.n) * p.maxM; }
  const uint32_t k = rp.rcount[ridx], off = rp.roff[ridx];
  // existing links (rows are padded with INVALID)
  uint32_t c = 0;
  for (uint32_t j0 = 0; j0 < cap; j0 += 32) {
    uint32_t j = j0 + lane;
    uint32_t id = j < cap ? row[j] : INVALID_ID;
    c += __popc(__ballot_sync(0xffffffffu, id != INVALID_ID));
    if (j < cap) b.cand_i[j] = id;
  }
  __syncwarp();
  // new sources -> ui/ud[c ..): all of the group, or (update) those the row does not hold yet
  uint32_t kn = 0;
  for (uint32_t i0 = 0; i0 < k; i0 += 32) {
    uint32_t i = i0 + lane;
    uint32_t src = i < k ? rp.g_src[off + i] : INVALID_ID;
    bool ok = i < k;
    if (ok && rp.update)
      for (uint32_t j = 0; j < c; j++) ok &= b.cand_i[j] != src;
    unsigned mk = __ballot_sync(0xffffffffu, ok);
    if (ok) { uint32_t at = c + kn + __popc(mk & ((1u << lane) - 1)); ui[at] = src; ud[at] = rp.g_d[off + i]; }
    kn += __popc(mk);
  }
  __syncwarp();
  uint32_t n_heur = 0;
  if (c + kn <= cap) {
    // append in ascending id (the order a one-thread addPoint loop would append them in)
    for (uint32_t i = lane; i < kn; i += 32) {
      uint32_t id = ui[c + i], r = 0;
      for (uint32_t j = 0; j < kn; j++) r += ui[c + j] < id;
      row[c + r] = id;
    }
  } else {
    // candidates = existing (distance from this node) + new (distance already known), sorted by (distance,id)
    b.load_query(p.vec + (size_t)node * p.stride);
    b.n_dist = 0;
    b.eval(c);
    for (uint32_t i = lane; i < c; i += 32) { ui[i] = b.cand_i[i]; ud[i] = b.cand_d[i]; }
    const uint32_t m = c + kn;
    __syncwarp();
    for (uint32_t i = lane; i < m; i += 32) {
      float di = ud[i]; uint32_t ii = ui[i], r = 0;
      for (uint32_t j = 0; j < m; j++) {
        float dj = ud[j]; uint32_t ij = ui[j];
        r += (dj < di || (dj == di && ij < ii)) ? 1u : 0u;
      }
      sd[r] = di; si[r] = ii;
    }
    __syncwarp();
    uint32_t nres = 0;
    occlude_scan(b, sd, si, 0, m, cap, res_i, res_d, nres, n_heur);
    __syncwarp();
    for (uint32_t i = lane; i < cap; i += 32) row[i] = i < nres ? res_i[nres - 1 - i] : INVALID_ID;
    n_heur += c;
  }
  if (lane == 0) {
    rp.rcount[ridx] = 0; rp.rfill[ridx] = 0;
    if (n_heur) atomicAdd(rp.stats + 2, (unsigned long long)n_heur);
  }
}

template <int TEAM, int NCH, int IP>
int launch_insert_t(b200ann_handle *h, const InsertParams &ip) {
  size_t smem = (size_t)WARPS * insert_smem_words(ip) * 4;
  if (smem > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "hnsw_build: ef_construction / M need more than 227 KB shared memory per block");
  auto kern = hnsw_insert_kernel<TEAM, NCH, IP>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(ip.e - ip.s + WARPS - 1) / WARPS, WARPS * 32, smem, h->stream>>>(ip);
  h->launches++;
  return 0;
}
template <int TEAM, int NCH, int IP>
int launch_reverse_t(b200ann_handle *h, const ReverseParams &rp, uint32_t ntouched) {
  size_t smem = (size_t)WARPS * reverse_smem_words(rp) * 4;
  if (smem > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "hnsw_build: a reverse-link group needs more than 227 KB shared memory (lower bmax)");
  auto kern = hnsw_reverse_kernel<TEAM, NCH, IP>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(ntouched + WARPS - 1) / WARPS, WARPS * 32, smem, h->stream>>>(rp);
  h->launches++;
  return 0;
}

#define HB_NCH(FN, MET, ...)                                  \
  do {                                                        \
    if (nch <= 8) return FN<8, 1, MET>(__VA_ARGS__);          \
    if (nch <= 16) return FN<8, 2, MET>(__VA_ARGS__);         \
    if (nch <= 32) return FN<8, 4, MET>(__VA_ARGS__);         \
    if (nch <= 64) return FN<8, 8, MET>(__VA_ARGS__);         \
    if (nch <= 128) return FN<32, 4, MET>(__VA_ARGS__);       \
    if (nch <= 256) return FN<32, 8, MET>(__VA_ARGS__);       \
    return FN<32, 16, MET>(__VA_ARGS__);                      \
  } while (0)
// met: MET_L2 / MET_IP / MET_U8 (dist.cuh); nch = 16-byte chunks per row (4 floats or 16 bytes each)
#define HB_DISPATCH(FN, ...)                                                                                  \
  do {                                                                                                        \
    if (nch > 512) return set_err(h, B200ANN_ERR_ARG, "dim > 2048 (8192 for byte rows) is not supported");    \
    if (met == MET_IP) HB_NCH(FN, MET_IP, __VA_ARGS__);                                                       \
    else if (met == MET_U8) HB_NCH(FN, MET_U8, __VA_ARGS__);                                                  \
    else HB_NCH(FN, MET_L2, __VA_ARGS__);                                                                     \
  } while (0)

int launch_insert(b200ann_handle *h, const InsertParams &ip, int met) {
  int nch = ip.sp.nch;
  HB_DISPATCH(launch_insert_t, h, ip);
}
int launch_reverse(b200ann_handle *h, const ReverseParams &rp, uint32_t ntouched, int met) {
  int nch = rp.sp.nch;
  HB_DISPATCH(launch_reverse_t, h, rp, ntouched);
}

// ---------------------------------------------------------------- updatePoint, phase A (hnswalg.h:993-1070)
// The one-hop neighbours of an updated element get NEW lists: the ef_construction closest of sCand = {element} +
// one-hop + two-hop neighbours (minus the neighbour itself), through the heuristic with Mmax. Batch form (oracle.c
// orc_hnsw_update variant 1): every list is computed on the graph as it was before the call and staged; a row
// claimed by several updated elements is computed for the last of them.
struct UpdParams {
  SearchParams sp;               // vec/stride/nch/dim, graph = links0 (gstride = maxM0), upper_*, maxM; L = ef_construction
  const uint32_t *ids; uint32_t nb, nlev;
  const int32_t *levels;
  uint32_t cap2;                 // collector capacity (power of two >= 2 * ef_construction)
  uint32_t *claim;               // [n + rows] zero-initialised
  uint32_t *stage, *stage_ridx;  // [nb * nlev * maxM0, maxM0], [nb * nlev * maxM0]
  unsigned long long *stats;     // [2] += heuristic evaluations; [3] += rows rewritten
};
__device__ __forceinline__ const uint32_t *upd_row(const SearchParams &p, uint32_t node, int layer, uint32_t &cap) {
  if (layer == 0) { cap = p.gstride; return p.graph + (size_t)node * p.gstride; }
  cap = p.maxM;
  return p.upper_links + ((size_t)p.upper_ptr[node] + (layer - 1)) * p.maxM;
}
// work item w = (position, layer, slot) of the updated elements' lists
__global__ void upd_claim_kernel(const UpdParams up) {
  const SearchParams &p = up.sp;
  size_t w = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (w >= (size_t)up.nb * up.nlev * p.gstride) return;
  const uint32_t slot = (uint32_t)(w % p.gstride), layer = (uint32_t)((w / p.gstride) % up.nlev), pos = (uint32_t)(w / ((size_t)p.gstride * up.nlev));
  const uint32_t u = up.ids[pos];
  if ((int)layer > up.levels[u]) return;
  uint32_t cap;
  const uint32_t *row = upd_row(p, u, (int)layer, cap);
  if (slot >= cap) return;
  const uint32_t neigh = row[slot];
  if (neigh == INVALID_ID) return;
  const uint32_t ridx = layer == 0 ? neigh : p.n + p.upper_ptr[neigh] + (layer - 1);
  atomicMax(&up.claim[ridx], pos + 1);
}
__host__ __device__ inline uint32_t upd_smem_words(const UpdParams &up) {
  return smem_words(up.sp.pool_cap, up.sp.cand_cap, up.sp.hbits) + 2 * up.cap2 + 2 * up.sp.gstride;
}
template <int TEAM, int NCH, int IP>
__global__ void __launch_bounds__(WARPS * 32) hnsw_update_neighbors_kernel(const UpdParams up) {
  extern __shared__ uint32_t smem[];
  const SearchParams &p = up.sp;
  const int warp = threadIdx.x >> 5;
  const size_t w = (size_t)blockIdx.x * WARPS + warp;
  if (w >= (size_t)up.nb * up.nlev * p.gstride) return;
  const uint32_t slot = (uint32_t)(w % p.gstride), layer = (uint32_t)((w / p.gstride) % up.nlev), pos = (uint32_t)(w / ((size_t)p.gstride * up.nlev));
  const uint32_t u = up.ids[pos];
  if ((int)layer > up.levels[u]) return;
  uint32_t cap;
  const uint32_t *r1 = upd_row(p, u, (int)layer, cap);
  if (slot >= cap) return;
  const uint32_t neigh = r1[slot];
  if (neigh == INVALID_ID) return;
  const uint32_t ridx = layer == 0 ? neigh : p.n + p.upper_ptr[neigh] + (layer - 1);
  if (up.claim[ridx] != pos + 1) return;

  uint32_t *base = smem + (size_t)warp * upd_smem_words(up);
  typedef Beam<TEAM, NCH, IP, true> B;
  B b(p, base);
  TopBuf tb;
  tb.d = reinterpret_cast<float *>(base + smem_words(p.pool_cap, p.cand_cap, p.hbits));
  tb.i = base + smem_words(p.pool_cap, p.cand_cap, p.hbits) + up.cap2;
  tb.cap2 = up.cap2; tb.keep = p.L; tb.lane = b.lane;
  tb.reset();
  uint32_t *res_i = base + smem_words(p.pool_cap, p.cand_cap, p.hbits) + 2 * up.cap2;
  float *res_d = reinterpret_cast<float *>(res_i + p.gstride);
  const int lane = b.lane;
  b.load_query(p.vec + (size_t)neigh * p.stride);
  b.n_dist = 0;
  // chunk -1: the element and its one-hop list; chunk a: the list of one-hop neighbour a (two-hop)
  for (int a = -1; a < (int)cap; a++) {
    const uint32_t *row = r1;
    if (a >= 0) {
      const uint32_t wn = r1[a];
      if (wn == INVALID_ID) break;
      uint32_t c2;
      row = upd_row(p, wn, (int)layer, c2);
    }
    uint32_t m = 0;
    __syncwarp();
    if (a < 0 && u != neigh) { if (lane == 0) b.cand_i[0] = u; m = 1; }
    for (uint32_t j0 = 0; j0 < cap; j0 += 32) {
      uint32_t j = j0 + lane;
      uint32_t id = j < cap ? row[j] : INVALID_ID;
      bool ok = id != INVALID_ID && id != neigh;
      unsigned mk = __ballot_sync(0xffffffffu, ok);
      if (ok) b.cand_i[m + __popc(mk & ((1u << lane) - 1))] = id;
      m += __popc(mk);
    }
    __syncwarp();
    b.eval(m);
    tb.push_batch(b.cand_i, b.cand_d, m, 0.f, 0u, false);
  }
  tb.compact();
  uint32_t nres = 0, n_heur = 0;
  if (tb.cnt < cap) {   // getNeighborsByHeuristic2: fewer than Mmax -> all kept
    for (uint32_t i = lane; i < tb.cnt; i += 32) { res_i[i] = tb.i[i]; res_d[i] = tb.d[i]; }
    nres = tb.cnt;
  } else {
    occlude_scan(b, tb.d, tb.i, 0, tb.cnt, cap, res_i, res_d, nres, n_heur);
  }
  __syncwarp();
  uint32_t *out = up.stage + w * p.gstride;
  for (uint32_t i = lane; i < p.gstride; i += 32) out[i] = i < nres ? res_i[nres - 1 - i] : INVALID_ID;
  if (lane == 0) {
    up.stage_ridx[w] = ridx;
    atomicAdd(up.stats + 0, (unsigned long long)b.n_dist);
    atomicAdd(up.stats + 2, (unsigned long long)n_heur);
    atomicAdd(up.stats + 3, 1ull);
  }
}
// staged rows -> the graph: item t holds the new content of row ridx[t] (INVALID = nothing staged)
__global__ void apply_rows_kernel(const uint32_t *stage, const uint32_t *ridx, size_t items, uint32_t sstride, uint32_t *links0,
                                  uint32_t *upper_links, uint32_t n, uint32_t maxM0, uint32_t maxM) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= items * sstride) return;
  const size_t t = e / sstride; const uint32_t j = (uint32_t)(e % sstride);
  const uint32_t r = ridx[t];
  if (r == INVALID_ID) return;
  if (r < n) { if (j < maxM0) links0[(size_t)r * maxM0 + j] = stage[e]; }
  else if (j < maxM) upper_links[(size_t)(r - n) * maxM + j] = stage[e];
}
// new vectors of the updated elements: row ids[i] of the index := row i of src (rb payload bytes; the padding stays 0)
__global__ void scatter_rows_kernel(const unsigned char *src, size_t sstride_bytes, const uint32_t *ids, size_t nb, unsigned char *vec,
                                    size_t vstride_bytes, size_t rb) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= nb * rb) return;
  const size_t i = e / rb, c = e % rb;
  vec[(size_t)ids[i] * vstride_bytes + c] = src[i * sstride_bytes + c];
}
template <int TEAM, int NCH, int IP>
int launch_updn_t(b200ann_handle *h, const UpdParams &up) {
  size_t smem = (size_t)WARPS * upd_smem_words(up) * 4;
  if (smem > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "hnsw_update: ef_construction / M need more than 227 KB shared memory per block");
  auto kern = hnsw_update_neighbors_kernel<TEAM, NCH, IP>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const size_t items = (size_t)up.nb * up.nlev * up.sp.gstride;
  kern<<<(unsigned)((items + WARPS - 1) / WARPS), WARPS * 32, smem, h->stream>>>(up);
  h->launches++;
  return 0;
}

int launch_updn(b200ann_handle *h, const UpdParams &up, int met) {
  int nch = up.sp.nch;
  HB_DISPATCH(launch_updn_t, h, up);
}

__global__ void fill_u32_kernel(uint32_t *p, size_t count, uint32_t v) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (; i < count; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

}  // namespace

// HierarchicalNSW::getRandomLevel (hnswalg.h:206-210): the next `count` draws of the index's level generator
// (std::default_random_engine seeded once, :122 / the load ctor; the sequence continues across calls like the
// reference's member generator does across addPoint calls)
static void draw_levels(b200ann_index *ix, size_t count, size_t M, std::vector<int32_t> &levels) {
  const double mult = 1 / std::log(1.0 * (double)M);          // :114
  levels.resize(count);
  for (size_t i = 0; i < count; i++) {
    std::uniform_real_distribution<double> distribution(0.0, 1.0);
    double r = -std::log(distribution(ix->level_rng)) * mult;
    levels[i] = (int)r;
  }
}

// undo of a batch whose reverse pass cannot run (see hnsw_run_batches): forget the reverse-edge counts it recorded
__global__ void rev_reset_kernel(const uint32_t *touched, uint32_t ntouched, uint32_t *rcount) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < ntouched) rcount[touched[t]] = 0;
}

__global__ void iota_u64_kernel(uint64_t *p, size_t from, size_t to) {
  size_t i = from + blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < to) p[i] = i;
}

// grow a device array to hold `need` elements, keeping the first `keep` (capacity doubles: appends are amortised)
template <typename T>
static cudaError_t grow(T **p, size_t keep, size_t need, size_t *cap, cudaStream_t st) {
  if (need <= *cap && *p) return cudaSuccess;
  size_t nc = *cap * 2 > need ? *cap * 2 : need;
  T *q = nullptr;
  cudaError_t e = cudaMalloc((void **)&q, (nc ? nc : 1) * sizeof(T));
  if (e != cudaSuccess) return e;
  if (*p && keep) e = cudaMemcpyAsync(q, *p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (*p) cudaFree(*p);
  *p = q; *cap = nc;
  return e;
}

// The insertion beam's visited set: 2^11 words per warp (measured on 1M x 128, M=16, efC=200: 2^10 1.68 s, 2^11 1.64 s,
// 2^12 2.10 s, 2^13 3.2 s - a smaller table re-evaluates ~5 % more rows but doubles the resident warps), holding
// 16-bit tags instead of 32-bit ids when the row count allows: twice the slots in the same words, so a beam that
// visits 3-4k rows no longer overfills its table (every test then ran the whole probe sequence and forgot an entry).
// What the set forgets only costs re-evaluations: a re-evaluated row is in the pool already or fails the pool's
// threshold again, so the links do not depend on the table.
static void insert_visited_set(const b200ann_handle *h, SearchParams &sp, size_t n) {
  sp.hbits = (h->link_visited_bits && h->link_visited_bits < 11) ? h->link_visited_bits : 11;
  if (const char *e = getenv("B200ANN_HNSW_HBITS")) sp.hbits = std::min(13, std::max(6, atoi(e)));   // A/B runs
  sp.nbits = 1;
  while (sp.nbits < 32 && (1ull << sp.nbits) < (unsigned long long)n) sp.nbits++;
  if (sp.nbits < sp.hbits + 1) sp.nbits = sp.hbits + 1;
  sp.vis16 = sp.nbits - (sp.hbits + 1) <= 12 ? 1 : 0;
  if (const char *e = getenv("B200ANN_HNSW_VIS16")) sp.vis16 = sp.vis16 && atoi(e);   // A/B runs
}

// Inserts rows [s0, n) of the index (vectors, levels, upper_ptr and the INVALID-filled link rows are in place) into
// the HNSW graph held by the index, batch by batch; updates the entry point / top level. levels = host levels of
// rows [s0, n). stats4 += {beam evaluations, hops, heuristic evaluations, 0}.
static int hnsw_run_batches(b200ann_handle *h, b200ann_index *ix, size_t s0, const std::vector<int32_t> &lv, size_t M, size_t efc,
                            size_t ratio, size_t bmax, unsigned long long *stats4, uint64_t *batches_out, uint64_t *biggest_out) {
  const size_t n = ix->n, rows = ix->upper_rows, maxM0 = ix->maxM0, maxM = ix->maxM;
  cudaStream_t st = h->stream;
  auto level_of = [&](size_t i) { return lv[i - s0]; };
  int toplevel = ix->maxlevel;
  for (size_t i = s0; i < n; i++) toplevel = std::max(toplevel, (int)level_of(i));
  std::vector<size_t> uprow(n - s0 + 1);   // first upper row of point i (the arrays were extended before this call)
  {
    size_t tot = 0;
    for (size_t i = s0; i < n; i++) tot += (size_t)level_of(i);
    size_t acc = rows - tot;
    for (size_t i = s0; i < n; i++) { uprow[i - s0] = acc; acc += (size_t)level_of(i); }
    uprow[n - s0] = acc;
  }
  const uint32_t nlev = (uint32_t)toplevel + 1;
  if (bmax > n - s0) bmax = n - s0 ? n - s0 : 1;
  DevBuf<uint32_t> rstate, touched, counters, rev_row, rev_src, g_src;
  DevBuf<float> rev_d, g_d;
  DevBuf<unsigned long long> stats;
  const size_t nr = n + rows, ecap = bmax * nlev * M;
  CU_CHECK(h, rstate.alloc(3 * nr, st));          // rcount | rfill | roff
  CU_CHECK(h, touched.alloc(nr, st));
  CU_CHECK(h, counters.alloc(4, st));
  CU_CHECK(h, rev_row.alloc(ecap, st)); CU_CHECK(h, rev_src.alloc(ecap, st)); CU_CHECK(h, rev_d.alloc(ecap, st));
  CU_CHECK(h, g_src.alloc(ecap, st)); CU_CHECK(h, g_d.alloc(ecap, st));
  CU_CHECK(h, stats.alloc(16, st));
  CU_CHECK(h, cudaMemsetAsync(rstate.p, 0, 3 * nr * 4, st));
  CU_CHECK(h, cudaMemsetAsync(stats.p, 0, 16 * sizeof(unsigned long long), st));
  uint32_t *rcount = rstate.p, *rfill = rstate.p + nr, *roff = rstate.p + 2 * nr;

  const int ipm = ix->metric == B200ANN_IP ? MET_IP : (ix_u8(ix) ? MET_U8 : MET_L2);
  InsertParams ip = {};
  ip.sp.vec = ix->vec; ip.sp.stride = ix->stride; ip.sp.n = (uint32_t)n; ip.sp.nch = ix_nch(ix);
  ip.sp.dim = (uint32_t)ix->dim;
  ip.sp.graph = ix->links0; ip.sp.gstride = (uint32_t)maxM0;
  ip.sp.upper_links = ix->upper_links; ip.sp.upper_ptr = ix->upper_ptr; ip.sp.maxM = (uint32_t)maxM;
  ip.sp.L = (uint32_t)efc; ip.sp.K = (uint32_t)M;
  insert_visited_set(h, ip.sp, n);
  ip.sp.pool_cap = round_up((uint32_t)efc + 1, 32);
  ip.sp.cand_cap = round_up((uint32_t)maxM0, 32);
  ip.M = (uint32_t)M; ip.nlev = nlev; ip.levels = ix->levels; ip.links0 = ix->links0; ip.upper_links = ix->upper_links;
  ip.rev_row = rev_row.p; ip.rev_src = rev_src.p; ip.rev_d = rev_d.p;
  ip.rcount = rcount; ip.touched = touched.p; ip.counters = counters.p; ip.stats = stats.p;
  ReverseParams rp = {};
  rp.sp = ip.sp; rp.sp.pool_cap = 32; rp.sp.hbits = 5; rp.sp.vis16 = 0;
  rp.links0 = ix->links0; rp.upper_links = ix->upper_links; rp.row_owner = ix->upper_owner;
  rp.touched = touched.p; rp.counters = counters.p; rp.roff = roff; rp.g_src = g_src.p; rp.g_d = g_d.p;
  rp.rcount = rcount; rp.rfill = rfill; rp.stats = stats.p;

  uint32_t ep = ix->hnsw_ep;
  int maxlevel = ix->maxlevel;
  uint64_t batches = 0, biggest = 0;
  size_t s = s0;
  int rc = 0;
  if (s == 0) {  // first element: entry point, nothing to link (:1255-1258)
    ep = 0; maxlevel = level_of(0); s = 1; batches = 1;
  }
  // a warp of the reverse pass holds one touched row's old links + ALL reverse edges the batch adds to it in shared
  // memory; a hub (duplicate-heavy data, inner product) can collect more than fits. Such a batch is undone - nothing
  // but the new points' own rows was written - and re-run at half the size.
  size_t max_group = (227 * 1024 / (WARPS * 4) - smem_words(32, round_up((uint32_t)maxM0, 32), 5) - 2 * maxM0) / 4 - maxM0 - 32;
  if (const char *e = getenv("B200ANN_HNSW_MAX_GROUP")) max_group = std::min<size_t>(max_group, (size_t)atol(e));  // tests
  size_t bcap = bmax;
  // The first min(2048, n/16) rows go in batches of up to 1/8 of the graph instead of 1/ratio: they are a small share of any
  // index worth building here and nearly all their links are rewritten by the rows that follow, while at 1/32 they
  // are 160 of a 50k-row build's 290 latency-bound batches. Measured against the reference-built graph
  // (scripts/hnsw_ratio_vs_ref.py, recall@10 at ef 10 - 100): still within 0.005 everywhere (worst -0.0037; 4096
  // rows at 1/8 or 2048 rows at 1/4 are not: -0.0054, -0.0086); 50k rows 0.183 -> 0.136 s.
  // An index of fewer than 32k rows takes n/16 rows instead (6000 x 32 uniform with 2048 early rows: -0.0055).
  size_t early_rows = std::min<size_t>(2048, n / 16), early_ratio = 8;
  if (const char *e = getenv("B200ANN_HNSW_EARLY")) { early_rows = (size_t)atol(e); if (const char *c = strchr(e, ',')) early_ratio = (size_t)atol(c + 1); }  // A/B runs
  while (s < n && rc == 0) {
    size_t e;
    if (level_of(s) > maxlevel) e = s + 1;
    else {
      size_t r_eff = ratio;
      if (early_rows && s < early_rows) r_eff = std::min(ratio, early_ratio);
      size_t b = std::min(std::min(bmax, bcap), std::max<size_t>(1, s / r_eff));
      e = std::min(n, s + b);
      for (size_t i = s + 1; i < e; i++)
        if (level_of(i) > maxlevel) { e = i; break; }
    }
    batches++;
    const size_t total = (e - s) * nlev * M;
    cudaMemsetAsync(rev_row.p, 0xFF, total * 4, st);
    cudaMemsetAsync(counters.p, 0, 4 * 4, st);
    ip.s = (uint32_t)s; ip.e = (uint32_t)e; ip.sp.ep = ep; ip.sp.maxlevel = maxlevel;
    rc = launch_insert(h, ip, ipm);
    if (rc) break;
    uint32_t hc[4];
    cudaMemcpyAsync(hc, counters.p, sizeof hc, cudaMemcpyDeviceToHost, st);
    cudaError_t ce = cudaStreamSynchronize(st);
    if (ce != cudaSuccess) { rc = set_err(h, B200ANN_ERR_CUDA, std::string("hnsw insert: ") + cudaGetErrorString(ce)); break; }
    if (hc[0] && hc[1] > max_group && e - s > 1) {
      rev_reset_kernel<<<(hc[0] + 255) / 256, 256, 0, st>>>(touched.p, hc[0], rcount);
      fill_u32_kernel<<<h->sm_count, 256, 0, st>>>(ix->links0 + s * maxM0, (e - s) * maxM0, INVALID_ID);
      const size_t ur0 = uprow[s - s0], ur1 = uprow[e - s0];   // the upper rows of the batch's points are contiguous
      if (ur1 > ur0) fill_u32_kernel<<<h->sm_count, 256, 0, st>>>(ix->upper_links + ur0 * maxM, (ur1 - ur0) * maxM, INVALID_ID);
      h->launches += 2;
      bcap = (e - s) / 2;
      batches--;
      continue;
    }
    if (hc[0]) {
      biggest = std::max<uint64_t>(biggest, hc[1]);
      rev_offsets_kernel<<<(hc[0] + 255) / 256, 256, 0, st>>>(touched.p, counters.p, rcount, roff, counters.p + 2);
      rev_group_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(rev_row.p, rev_src.p, rev_d.p, total, roff, rfill, g_src.p, g_d.p);
      h->launches += 2;
      rp.mcap = round_up((uint32_t)maxM0 + hc[1], 32);
      rc = launch_reverse(h, rp, hc[0], ipm);
      if (rc) break;
    }
    for (size_t i = s; i < e; i++)
      if (level_of(i) > maxlevel) { maxlevel = level_of(i); ep = (uint32_t)i; }   // :1262-1265
    s = e;
    if (bcap < bmax) bcap = std::min(bmax, bcap * 2);
  }
  unsigned long long hs[4] = {0, 0, 0, 0};
  if (rc == 0) {
    cudaMemcpyAsync(hs, stats.p, sizeof hs, cudaMemcpyDeviceToHost, st);
    cudaError_t ce = cudaStreamSynchronize(st);
#ifdef BEAM_PROF
    {
      unsigned long long ps[16];
      cudaMemcpy(ps, stats.p, sizeof ps, cudaMemcpyDeviceToHost);
      const double pts = (double)(n - s0);
      fprintf(stderr, "HB_PROF rows %zu: cycles/row total %.0f | expand_all %.0f (load_row %.0f eval %.0f insert %.0f) | heuristic %.0f | hops/row %.1f evals/row %.0f | spec hits/row %.1f scan cycles %.0f\n",
              n - s0, ps[4] / pts, ps[5] / pts, ps[6] / pts, ps[7] / pts, ps[8] / pts, ps[9] / pts, ps[1] / pts, ps[0] / pts, ps[10] / pts, ps[11] / pts);
    }
#endif
    if (ce == cudaSuccess) ce = cudaGetLastError();
    if (ce != cudaSuccess) rc = set_err(h, B200ANN_ERR_CUDA, std::string("hnsw insert: ") + cudaGetErrorString(ce));
  }
  if (rc) return rc;
  ix->hnsw_ep = ep; ix->maxlevel = maxlevel;
  for (int i = 0; i < 4; i++) stats4[i] += hs[i];
  *batches_out += batches;
  *biggest_out = std::max<uint64_t>(*biggest_out, biggest);
  return 0;
}

// rows [n_old, n) get their levels, upper rows (INVALID-filled) and owners; the arrays grow as needed
static int hnsw_extend_arrays(b200ann_handle *h, b200ann_index *ix, size_t n_old, const std::vector<int32_t> &lv) {
  cudaStream_t st = h->stream;
  const size_t n = ix->n, maxM0 = ix->maxM0, maxM = ix->maxM;
  std::vector<uint32_t> uptr(n - n_old), owner;
  size_t rows = ix->upper_rows;
  for (size_t i = n_old; i < n; i++) {
    uptr[i - n_old] = (uint32_t)rows;
    for (int l = 0; l < lv[i - n_old]; l++) owner.push_back((uint32_t)i);
    rows += (size_t)lv[i - n_old];
  }
  const size_t rows_old = ix->upper_rows;
  size_t c;
  c = ix->cap_links0; CU_CHECK(h, grow(&ix->links0, n_old * maxM0, n * maxM0, &c, st)); ix->cap_links0 = c;
  c = ix->cap_levels; CU_CHECK(h, grow(&ix->levels, n_old, n, &c, st)); ix->cap_levels = c;
  c = ix->cap_uptr; CU_CHECK(h, grow(&ix->upper_ptr, n_old, n, &c, st)); ix->cap_uptr = c;
  c = ix->cap_upper; CU_CHECK(h, grow(&ix->upper_links, rows_old * maxM, (rows ? rows : 1) * maxM, &c, st)); ix->cap_upper = c;
  c = ix->cap_owner; CU_CHECK(h, grow(&ix->upper_owner, rows_old, rows ? rows : 1, &c, st)); ix->cap_owner = c;
  CU_CHECK(h, cudaMemcpyAsync(ix->levels + n_old, lv.data(), (n - n_old) * 4, cudaMemcpyHostToDevice, st));
  CU_CHECK(h, cudaMemcpyAsync(ix->upper_ptr + n_old, uptr.data(), (n - n_old) * 4, cudaMemcpyHostToDevice, st));
  if (rows > rows_old) CU_CHECK(h, cudaMemcpyAsync(ix->upper_owner + rows_old, owner.data(), (rows - rows_old) * 4, cudaMemcpyHostToDevice, st));
  fill_u32_kernel<<<h->sm_count * 8, 256, 0, st>>>(ix->links0 + n_old * maxM0, (n - n_old) * maxM0, INVALID_ID);
  if (rows > rows_old) fill_u32_kernel<<<h->sm_count * 8, 256, 0, st>>>(ix->upper_links + rows_old * maxM, (rows - rows_old) * maxM, INVALID_ID);
  CU_CHECK(h, cudaStreamSynchronize(st));   // the host staging vectors go out of scope
  ix->upper_rows = rows;
  return 0;
}

// Builds the HNSW graph of the index's rows (inserted in row order) and installs it in the index.
// info8 = {beam distance evaluations, expanded nodes, heuristic distance evaluations, batches, upper rows,
// maxlevel, entry point, largest reverse group}.
int hnsw_build_launch(b200ann_handle *h, b200ann_index *ix, size_t M, size_t efc, uint32_t seed, size_t ratio, size_t bmax,
                      const uint64_t *labels_host_or_dev, uint64_t *info8) {
  const size_t n = ix->n;
  if (M < 2 || M > 128 || n >= (1ull << 31))
    return set_err(h, B200ANN_ERR_ARG, "hnsw_build: need 2 <= M <= 128");
  if (efc < M) efc = M;                                        // hnswalg.h:107
  if (ratio == 0) ratio = 32;
  if (bmax == 0) bmax = 16384;
  cudaStream_t st = h->stream;
  // the caller freed any previous HNSW graph: start from empty arrays
  ix->maxM = M; ix->maxM0 = 2 * M; ix->upper_rows = 0; ix->hnsw_ep = 0; ix->maxlevel = -1;
  ix->cap_links0 = ix->cap_levels = ix->cap_uptr = ix->cap_upper = ix->cap_owner = ix->cap_labels = 0;
  ix->level_rng.seed(seed); ix->hnsw_M = M; ix->hnsw_efc = efc;
  std::vector<int32_t> levels;
  draw_levels(ix, n, M, levels);
  int rc = hnsw_extend_arrays(h, ix, 0, levels);
  if (rc) return rc;
  if (labels_host_or_dev) {
    CU_CHECK(h, cudaMalloc((void **)&ix->labels, n * 8));
    ix->cap_labels = n;
    CU_CHECK(h, cudaMemcpyAsync(ix->labels, labels_host_or_dev, n * 8, cudaMemcpyDefault, st));
  }
  unsigned long long hs[4] = {0, 0, 0, 0};
  uint64_t batches = 0, biggest = 0;
  rc = hnsw_run_batches(h, ix, 0, levels, M, efc, ratio, bmax, hs, &batches, &biggest);
  if (rc) return rc;
  ix->has_hnsw = true;
  if (info8) {
    info8[0] = hs[0]; info8[1] = hs[1]; info8[2] = hs[2]; info8[3] = batches; info8[4] = ix->upper_rows;
    info8[5] = (uint64_t)ix->maxlevel; info8[6] = ix->hnsw_ep; info8[7] = biggest;
  }
  return 0;
}

// Appends nb rows to the index and links them into its HNSW graph in place: what nb addPoint calls do to a live
// hnswlib index (hnswalg.h:1153-1267). data = nb rows dstride_bytes apart (floats, or bytes for a u8 index), device memory. The level generator continues where the
// build (or the previous append) left it. Any NSG graph / tensor-core cache of the index is dropped (their row set
// changed). info8 as hnsw_build_launch (counters of this call).
int hnsw_insert_launch(b200ann_handle *h, b200ann_index *ix, const void *data, size_t dstride_bytes, size_t nb,
                       const uint64_t *labels_host_or_dev, size_t ratio, size_t bmax, uint64_t *info8) {
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "hnsw_insert_batch: index has no HNSW graph (hnsw_build / set_hnsw first)");
  if (ix->n + nb >= (1ull << 31)) return set_err(h, B200ANN_ERR_ARG, "hnsw_insert_batch: too many rows");
  if (ratio == 0) ratio = 32;
  if (bmax == 0) bmax = 16384;
  cudaStream_t st = h->stream;
  const size_t n_old = ix->n, n = n_old + nb;
  const size_t M = ix->hnsw_M ? ix->hnsw_M : ix->maxM;
  const size_t efc = ix->hnsw_efc ? ix->hnsw_efc : 200;
  // vectors
  {
    size_t c = ix->cap_vec ? ix->cap_vec : n_old * ix->stride;
    CU_CHECK(h, grow(&ix->vec, n_old * ix->stride, n * ix->stride, &c, st));
    ix->cap_vec = c;
    const size_t rb = ix_row_bytes(ix);
    if (ix->stride * 4 != rb) CU_CHECK(h, cudaMemsetAsync(ix->vec + n_old * ix->stride, 0, nb * ix->stride * 4, st));
    CU_CHECK(h, cudaMemcpy2DAsync(ix->vec + n_old * ix->stride, ix->stride * 4, data, dstride_bytes, rb, nb, cudaMemcpyDefault, st));
  }
  if (ix->graph) { cudaFree(ix->graph); ix->graph = nullptr; }
  knn_tc_free(ix);
  // labels: an index without a label array has label == row number
  if (labels_host_or_dev || ix->labels) {
    const bool had = ix->labels != nullptr;
    size_t c = ix->cap_labels ? ix->cap_labels : (had ? n_old : 0);
    CU_CHECK(h, grow(&ix->labels, had ? n_old : 0, n, &c, st));
    ix->cap_labels = c;
    if (!had) iota_u64_kernel<<<(unsigned)((n_old + 255) / 256), 256, 0, st>>>(ix->labels, 0, n_old);
    if (labels_host_or_dev) CU_CHECK(h, cudaMemcpyAsync(ix->labels + n_old, labels_host_or_dev, nb * 8, cudaMemcpyDefault, st));
    else iota_u64_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, st>>>(ix->labels, n_old, n);
  }
  // capacities of arrays installed by b200ann_index_set_hnsw (exact fit)
  if (!ix->cap_links0) { ix->cap_links0 = n_old * ix->maxM0; ix->cap_levels = n_old; ix->cap_uptr = n_old;
                         ix->cap_upper = (ix->upper_rows ? ix->upper_rows : 1) * ix->maxM; }
  if (!ix->upper_owner) {  // set_hnsw path: rebuild the owner of every upper row from the levels
    std::vector<int32_t> lv(n_old);
    CU_CHECK(h, cudaMemcpy(lv.data(), ix->levels, n_old * 4, cudaMemcpyDeviceToHost));
    std::vector<uint32_t> up(n_old);
    CU_CHECK(h, cudaMemcpy(up.data(), ix->upper_ptr, n_old * 4, cudaMemcpyDeviceToHost));
    std::vector<uint32_t> owner(ix->upper_rows ? ix->upper_rows : 1, 0);
    for (size_t i = 0; i < n_old; i++)
      for (int l = 0; l < lv[i]; l++) owner[up[i] + l] = (uint32_t)i;
    CU_CHECK(h, cudaMalloc((void **)&ix->upper_owner, owner.size() * 4));
    ix->cap_owner = owner.size();
    CU_CHECK(h, cudaMemcpy(ix->upper_owner, owner.data(), owner.size() * 4, cudaMemcpyHostToDevice));
  }
  ix->n = n;
  std::vector<int32_t> levels;
  draw_levels(ix, nb, M, levels);
  int rc = hnsw_extend_arrays(h, ix, n_old, levels);
  if (rc) { ix->n = n_old; return rc; }
  unsigned long long hs[4] = {0, 0, 0, 0};
  uint64_t batches = 0, biggest = 0;
  rc = hnsw_run_batches(h, ix, n_old, levels, M, efc, ratio, bmax, hs, &batches, &biggest);
  if (rc) return rc;
  if (info8) {
    info8[0] = hs[0]; info8[1] = hs[1]; info8[2] = hs[2]; info8[3] = batches; info8[4] = ix->upper_rows;
    info8[5] = (uint64_t)ix->maxlevel; info8[6] = ix->hnsw_ep; info8[7] = biggest;
  }
  return 0;
}

// updatePoint for nb elements of a built index (hnswalg.h:993-1140; what addPoint does for a label that exists):
// ids [nb] = DISTINCT internal ids, data = their new vectors (nb rows dstride_bytes apart), both device memory.
// Batch form (oracle.c orc_hnsw_update variant 1): the vectors are replaced; phase A gives every one-hop neighbour of
// an updated element a new list computed on the graph as it was; phase B re-links the elements themselves like an
// insertion into the graph phase A left (own rows staged, reverse links grouped, present sources skipped).
// Levels, entry point and labels do not change. info8 = {beam evaluations, hops, heuristic evaluations, chunks,
// rows rewritten by phase A, maxlevel, entry point, largest reverse group}.
int hnsw_update_launch(b200ann_handle *h, b200ann_index *ix, const uint32_t *ids, size_t nb, const void *data, size_t dstride_bytes,
                       uint64_t *info8) {
  if (!ix->has_hnsw) return set_err(h, B200ANN_ERR_STATE, "hnsw_update_batch: index has no HNSW graph (hnsw_build / set_hnsw first)");
  cudaStream_t st = h->stream;
  const size_t n = ix->n, rows = ix->upper_rows, maxM0 = ix->maxM0, maxM = ix->maxM;
  const size_t M = ix->hnsw_M ? ix->hnsw_M : maxM;
  size_t efc = ix->hnsw_efc ? ix->hnsw_efc : 200;
  if (efc < M) efc = M;
  if (n == 1) {  // :999-1001: a single element that is its own entry point: only the vector changes
    scatter_rows_kernel<<<1, 256, 0, st>>>((const unsigned char *)data, dstride_bytes, ids, 1, (unsigned char *)ix->vec, ix->stride * 4, ix_row_bytes(ix));
    h->launches++;
    CU_CHECK(h, cudaStreamSynchronize(st));
    return 0;
  }
  if (!ix->upper_owner && rows) {  // a graph installed by b200ann_index_set_hnsw: the owner of every upper row
    std::vector<int32_t> lv(n);
    std::vector<uint32_t> up(n), owner(rows);
    CU_CHECK(h, cudaMemcpy(lv.data(), ix->levels, n * 4, cudaMemcpyDeviceToHost));
    CU_CHECK(h, cudaMemcpy(up.data(), ix->upper_ptr, n * 4, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++)
      for (int l = 0; l < lv[i]; l++) owner[up[i] + l] = (uint32_t)i;
    CU_CHECK(h, cudaMalloc((void **)&ix->upper_owner, rows * 4));
    ix->cap_owner = rows;
    CU_CHECK(h, cudaMemcpy(ix->upper_owner, owner.data(), rows * 4, cudaMemcpyHostToDevice));
  }
  const uint32_t nlev = (uint32_t)ix->maxlevel + 1;
  const size_t rb = ix_row_bytes(ix);
  {
    const size_t tot = nb * rb;
    scatter_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>((const unsigned char *)data, dstride_bytes, ids, nb,
                                                                     (unsigned char *)ix->vec, ix->stride * 4, rb);
    h->launches++;
  }
  if (ix->graph) { cudaFree(ix->graph); ix->graph = nullptr; }
  knn_tc_free(ix);
  const int met = ix->metric == B200ANN_IP ? MET_IP : (ix_u8(ix) ? MET_U8 : MET_L2);
  const size_t nr = n + rows;
  size_t chunk = std::min<size_t>(nb, 8192);
  while (chunk > 64 && chunk * nlev * maxM0 * maxM0 * 4 > ((size_t)1 << 30)) chunk /= 2;  // <= 1 GiB of staged phase-A rows
  DevBuf<uint32_t> rstate, touched, counters, rev_row, rev_src, g_src, claim, stage, stage_ridx, own_stage, own_ridx;
  DevBuf<float> rev_d, g_d;
  DevBuf<unsigned long long> stats;
  const size_t ecap = chunk * nlev * M, items = chunk * nlev * maxM0, oitems = chunk * nlev;
  CU_CHECK(h, rstate.alloc(3 * nr, st));
  CU_CHECK(h, touched.alloc(nr, st));
  CU_CHECK(h, counters.alloc(4, st));
  CU_CHECK(h, rev_row.alloc(ecap, st)); CU_CHECK(h, rev_src.alloc(ecap, st)); CU_CHECK(h, rev_d.alloc(ecap, st));
  CU_CHECK(h, g_src.alloc(ecap, st)); CU_CHECK(h, g_d.alloc(ecap, st));
  CU_CHECK(h, claim.alloc(nr, st));
  CU_CHECK(h, stage.alloc(items * maxM0, st)); CU_CHECK(h, stage_ridx.alloc(items, st));
  CU_CHECK(h, own_stage.alloc(oitems * maxM0, st)); CU_CHECK(h, own_ridx.alloc(oitems, st));
  CU_CHECK(h, stats.alloc(16, st));
  CU_CHECK(h, cudaMemsetAsync(rstate.p, 0, 3 * nr * 4, st));
  CU_CHECK(h, cudaMemsetAsync(claim.p, 0, nr * 4, st));
  CU_CHECK(h, cudaMemsetAsync(stats.p, 0, 16 * sizeof(unsigned long long), st));
  uint32_t *rcount = rstate.p, *rfill = rstate.p + nr, *roff = rstate.p + 2 * nr;

  SearchParams sp = {};
  sp.vec = ix->vec; sp.stride = ix->stride; sp.n = (uint32_t)n; sp.nch = ix_nch(ix); sp.dim = (uint32_t)ix->dim;
  sp.graph = ix->links0; sp.gstride = (uint32_t)maxM0;
  sp.upper_links = ix->upper_links; sp.upper_ptr = ix->upper_ptr; sp.maxM = (uint32_t)maxM;
  sp.L = (uint32_t)efc; sp.K = (uint32_t)M; sp.ep = ix->hnsw_ep; sp.maxlevel = ix->maxlevel;
  UpdParams up = {};
  up.sp = sp; up.sp.pool_cap = 32; up.sp.hbits = 5; up.sp.cand_cap = round_up((uint32_t)maxM0 + 1, 32);
  up.nlev = nlev; up.levels = ix->levels;
  up.cap2 = 64; while (up.cap2 < 2 * efc) up.cap2 <<= 1;
  up.claim = claim.p; up.stage = stage.p; up.stage_ridx = stage_ridx.p; up.stats = stats.p;
  InsertParams ip = {};
  ip.sp = sp;
  insert_visited_set(h, ip.sp, n);
  ip.sp.pool_cap = round_up((uint32_t)efc + 1, 32);
  ip.sp.cand_cap = round_up((uint32_t)maxM0, 32);
  ip.M = (uint32_t)M; ip.nlev = nlev; ip.levels = ix->levels; ip.links0 = ix->links0; ip.upper_links = ix->upper_links;
  ip.rev_row = rev_row.p; ip.rev_src = rev_src.p; ip.rev_d = rev_d.p;
  ip.rcount = rcount; ip.touched = touched.p; ip.counters = counters.p; ip.stats = stats.p;
  ip.upd_ids = ids; ip.own_stage = own_stage.p; ip.own_ridx = own_ridx.p;
  ReverseParams rp = {};
  rp.sp = ip.sp; rp.sp.pool_cap = 32; rp.sp.hbits = 5; rp.sp.vis16 = 0;
  rp.links0 = ix->links0; rp.upper_links = ix->upper_links; rp.row_owner = ix->upper_owner;
  rp.touched = touched.p; rp.counters = counters.p; rp.roff = roff; rp.g_src = g_src.p; rp.g_d = g_d.p;
  rp.rcount = rcount; rp.rfill = rfill; rp.stats = stats.p; rp.update = 1;
  const size_t max_group = (227 * 1024 / (WARPS * 4) - smem_words(32, round_up((uint32_t)maxM0, 32), 5) - 2 * maxM0) / 4 - maxM0 - 32;

  uint64_t chunks = 0, biggest = 0;
  for (size_t s = 0; s < nb; s += chunk) {
    const size_t e = std::min(nb, s + chunk), cn = e - s;
    chunks++;
    // ---- phase A
    up.ids = ids + s; up.nb = (uint32_t)cn;
    const size_t it = cn * nlev * maxM0;
    CU_CHECK(h, cudaMemsetAsync(stage_ridx.p, 0xFF, it * 4, st));
    upd_claim_kernel<<<(unsigned)((it + 255) / 256), 256, 0, st>>>(up);
    h->launches++;
    int rc = launch_updn(h, up, met);
    if (rc) return rc;
    apply_rows_kernel<<<(unsigned)((it * maxM0 + 255) / 256), 256, 0, st>>>(stage.p, stage_ridx.p, it, (uint32_t)maxM0, ix->links0,
                                                                        ix->upper_links, (uint32_t)n, (uint32_t)maxM0, (uint32_t)maxM);
    h->launches++;
    CU_CHECK(h, cudaMemsetAsync(claim.p, 0, nr * 4, st));
    // ---- phase B
    const size_t total = cn * nlev * M;
    CU_CHECK(h, cudaMemsetAsync(rev_row.p, 0xFF, total * 4, st));
    CU_CHECK(h, cudaMemsetAsync(counters.p, 0, 4 * 4, st));
    CU_CHECK(h, cudaMemsetAsync(own_ridx.p, 0xFF, cn * nlev * 4, st));
    ip.s = (uint32_t)s; ip.e = (uint32_t)e;
    rc = launch_insert(h, ip, met);
    if (rc) return rc;
    apply_rows_kernel<<<(unsigned)((cn * nlev * maxM0 + 255) / 256), 256, 0, st>>>(own_stage.p, own_ridx.p, cn * nlev, (uint32_t)maxM0,
                                                                               ix->links0, ix->upper_links, (uint32_t)n,
                                                                               (uint32_t)maxM0, (uint32_t)maxM);
    h->launches++;
    uint32_t hc[4];
    CU_CHECK(h, cudaMemcpyAsync(hc, counters.p, sizeof hc, cudaMemcpyDeviceToHost, st));
    CU_CHECK(h, cudaStreamSynchronize(st));
    if (hc[0]) {
      if (hc[1] > max_group)
        return set_err(h, B200ANN_ERR_ARG, "hnsw_update_batch: one element collects more reverse links than a warp can hold; update in smaller batches");
      biggest = std::max<uint64_t>(biggest, hc[1]);
      rev_offsets_kernel<<<(hc[0] + 255) / 256, 256, 0, st>>>(touched.p, counters.p, rcount, roff, counters.p + 2);
      rev_group_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(rev_row.p, rev_src.p, rev_d.p, total, roff, rfill, g_src.p, g_d.p);
      h->launches += 2;
      rp.mcap = round_up((uint32_t)maxM0 + hc[1], 32);
      rc = launch_reverse(h, rp, hc[0], met);
      if (rc) return rc;
    }
  }
  unsigned long long hs[4] = {0, 0, 0, 0};
  CU_CHECK(h, cudaMemcpyAsync(hs, stats.p, sizeof hs, cudaMemcpyDeviceToHost, st));
  CU_CHECK(h, cudaStreamSynchronize(st));
  CU_CHECK(h, cudaGetLastError());
  if (info8) {
    info8[0] = hs[0]; info8[1] = hs[1]; info8[2] = hs[2]; info8[3] = chunks; info8[4] = hs[3];
    info8[5] = (uint64_t)ix->maxlevel; info8[6] = ix->hnsw_ep; info8[7] = biggest;
  }
  return 0;
}

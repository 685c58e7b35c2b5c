// Batched graph search kernels: one warp per query over the beam core in beam.cuh.
//   b200ann_nsg_search  <- efanna2e::IndexNSG::Search / SearchWithOptGraph (index_nsg.cpp:506-657)
//   b200ann_hnsw_search <- hnswlib::HierarchicalNSW::searchKnn (hnswalg.h:1270-1324)
//   b200ann_hybrid_search   <- hnsw_nsg::HNSW_NSG::searchKnn (hnsw_nsg/include/index_hnsw_nsg.h:107-151)
//   b200ann_nsg_search_from <- IndexNSG::SearchFromEnterpoint (hnsw_nsg/src/index_nsg.cpp:703-782)
#include "beam.cuh"

using namespace beam;

int launch_search2_pr0(b200ann_handle *h, const SearchParams &p, bool ip, bool hnsw);
int launch_search2_pr1(b200ann_handle *h, const SearchParams &p, bool ip, bool hnsw);
int launch_search2_pr2(b200ann_handle *h, const SearchParams &p, bool ip, bool hnsw);

namespace {

constexpr int WARPS = 4;

// Persistent form (p.work != NULL): the grid is one resident wave and every WARP pulls its next query from an
// atomic counter, so a warp whose query ends early starts another one instead of idling until the slowest warp of
// its block (and the last partial wave) is done. Results do not depend on which warp ran which query.
template <int TEAM, int NCH, int IP, bool HNSW>
__global__ void __launch_bounds__(WARPS * 32, (NCH <= 4 ? 6 : 1)) beam_search_kernel(const SearchParams p) {
  extern __shared__ uint32_t smem[];
  const int warp = threadIdx.x >> 5;
  Beam<TEAM, NCH, IP, HNSW> b(p, smem + (size_t)warp * smem_words(p.pool_cap, p.cand_cap, p.hbits));
  NoRec rec;
  for (bool first = true;; first = false) {
    uint32_t qi;
    if (p.work) {
      qi = 0;
      if (b.lane == 0) qi = atomicAdd(p.work, 1u);
      qi = __shfl_sync(0xffffffffu, qi, 0);
    } else {
      if (!first) break;
      qi = blockIdx.x * WARPS + warp;
    }
    if (qi >= p.nq) break;
    if (p.order) qi = __ldg(p.order + qi);
    b.begin(p.q + (size_t)qi * p.qstride);
    if (HNSW) {
      b.descend_hnsw();
    } else if (p.mode == 0) {
      b.seed_nsg(p.qidx_base + qi, rec);
    } else if (p.mode == 4) {
      const uint32_t size = __ldg(p.cl_size + qi);
      if (size == 0) {  // query not routed to any cluster in this pass
        for (uint32_t i = b.lane; i < p.K; i += 32) {
          p.ids[(size_t)qi * p.K + i] = INVALID_ID;
          if (p.dists) p.dists[(size_t)qi * p.K + i] = INFINITY;
        }
        if (p.stats && b.lane == 0) { p.stats[2 * (size_t)qi] = 0; p.stats[2 * (size_t)qi + 1] = 0; }
        continue;
      }
      b.seed_nsg_local(p.qidx_base + qi, __ldg(p.entry + qi), __ldg(p.cl_base + qi), size);
    } else {
      uint32_t entry;
      if (p.mode == 2) {  // hybrid: HNSW levels >= 1, external label of the node reached = NSG entry
        float cd;
        uint32_t cur = b.descend_upper(p.ep_hnsw, cd);
        entry = p.labels ? (uint32_t)__ldg(p.labels + cur) : cur;
        if (p.entry_out && b.lane == 0) p.entry_out[qi] = entry;
      } else {
        entry = __ldg(p.entry + qi);
      }
      if (entry >= p.n) entry = p.ep;  // the reference would index out of bounds; start from the navigating node
      b.seed_from(entry, p.qidx_base + qi);
    }
    b.expand_all(rec);

    // ---- results
    for (uint32_t i = b.lane; i < p.K; i += 32) {
      bool ok = i < b.np;
      uint32_t id = ok ? (b.pool_i[i] & ~FLAG) : INVALID_ID;
      float d = ok ? b.pool_d[i] : INFINITY;
      if (HNSW) {
        p.labels_out[(size_t)qi * p.K + i] = ok ? (p.labels ? __ldg(p.labels + id) : (uint64_t)id) : ~(uint64_t)0;
      } else {
        p.ids[(size_t)qi * p.K + i] = id + (ok ? p.id_base : 0u);
      }
      if (p.dists) p.dists[(size_t)qi * p.K + i] = d;
    }
    if (p.stats && b.lane == 0) {
      p.stats[2 * (size_t)qi] = b.n_dist;
      p.stats[2 * (size_t)qi + 1] = b.n_hops;
    }
    __syncwarp();
  }
}

template <int TEAM, int NCH, int IP, bool HNSW>
int launch_t(b200ann_handle *h, const SearchParams &p0) {
  SearchParams p = p0;
  size_t smem = (size_t)WARPS * smem_words(p.pool_cap, p.cand_cap, p.hbits) * 4;
  if (smem > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "search: L / degree / visited_bits need more than 227 KB shared memory per block");
  auto kern = beam_search_kernel<TEAM, NCH, IP, HNSW>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  size_t blocks = (p.nq + WARPS - 1) / WARPS;
  p.work = nullptr;
  if (h->persistent) {
    int occ = 0;
    CU_CHECK(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WARPS * 32, smem));
    if (h->search_ctas_per_sm > 0 && h->search_ctas_per_sm < occ) occ = h->search_ctas_per_sm;
    size_t wave = (size_t)(occ > 0 ? occ : 1) * (size_t)h->sm_count;
    if (blocks > wave) {  // more than one wave of blocks: pull queries from a queue instead
      if (!h->work) CU_CHECK(h, cudaMalloc((void **)&h->work, 2 * sizeof(uint32_t)));
      CU_CHECK(h, cudaMemsetAsync(h->work, 0, sizeof(uint32_t), h->stream));
      p.work = h->work;
      blocks = wave;
    }
  }
  dim3 grid((unsigned)blocks), block(WARPS * 32);
  kern<<<grid, block, smem, h->stream>>>(p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

template <int IP, bool HNSW>
int launch_dim(b200ann_handle *h, const SearchParams &p0) {
  SearchParams p = p0;
  p.order = (h->query_order && h->query_order_n == p.nq) ? h->query_order : nullptr;
  p.prefetch = h->search_prefetch;
  p.id_base = h->id_base;
  if (const char *e = getenv("B200ANN_MERGE_MIN")) p.merge_min = (uint32_t)atoi(e);   // A/B runs
  p.nbits = 1;
  while (p.nbits < 32 && (1ull << p.nbits) < (unsigned long long)p.n) p.nbits++;
  if (p.nbits < p.hbits + 1) p.nbits = p.hbits + 1;
  p.vis16 = (h->visited_compact && p.nbits - (p.hbits + 1) <= 12) ? 1 : 0;
  // version 0 (default): v2 where its register pool applies (L <= 64), v1 beyond (measured, 1M x 128 NSG, L = 100:
  // 4.33 ms v1 against 5.10 ms v2 - the shared-memory pool form of v2 has the smaller visited table)
  if (h->search_version != 1 && IP != MET_U8 && (h->search_version == 2 || p.L <= 64)) {  // byte rows (MET_U8) run on the v1 kernel; v2: TMA row ring + register pool (beam2.cuh); B200ANN_SEARCH_VERSION=1 keeps v1 for A/B runs
    if (p.L <= 32) return launch_search2_pr1(h, p, IP == MET_IP, HNSW);
    if (p.L <= 64) return launch_search2_pr2(h, p, IP == MET_IP, HNSW);
    return launch_search2_pr0(h, p, IP == MET_IP, HNSW);
  }
  int nch = p.nch;
  if (nch <= 8) return launch_t<8, 1, IP, HNSW>(h, p);
  if (nch <= 16) return launch_t<8, 2, IP, HNSW>(h, p);
  if (nch <= 32) return launch_t<8, 4, IP, HNSW>(h, p);
  if (nch <= 64) return launch_t<8, 8, IP, HNSW>(h, p);
  if (nch <= 128) return launch_t<32, 4, IP, HNSW>(h, p);
  if (nch <= 256) return launch_t<32, 8, IP, HNSW>(h, p);
  if (nch <= 512) return launch_t<32, 16, IP, HNSW>(h, p);
  return set_err(h, B200ANN_ERR_ARG, "search: dim > 2048 (8192 for byte rows) is not supported");
}

}  // namespace

// ---------------------------------------------------------------- random-row gather probe
// What HBM delivers for the access pattern of graph search itself: whole rows of the index at pseudo-random
// positions, read with the search kernels' own loads (a team of 8 / 32 lanes per row, 16 bytes per lane and load,
// `inflight` rows per team in flight), no dependence between rows. The measured rate is the ceiling a beam search
// over this index can reach; bench.py reports it next to the streaming-copy peak.
namespace {
template <int TEAM, int NCH, int INFLIGHT>
__global__ void __launch_bounds__(256) gather_probe_kernel(const float *vec, size_t stride, uint32_t n, int nch,
                                                           uint64_t rows_total, uint32_t seed, float *sink) {
  constexpr int TPW = 32 / TEAM;
  const int lane = threadIdx.x & 31, t = lane % TEAM, g = lane / TEAM;
  const uint64_t team_id = ((uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * TPW + g;
  const uint64_t teams = (uint64_t)gridDim.x * (blockDim.x >> 5) * TPW;
  float acc = 0.f;
  for (uint64_t r0 = team_id * INFLIGHT; r0 < rows_total; r0 += teams * INFLIGHT) {
    float4 v[INFLIGHT][NCH];
#pragma unroll
    for (int i = 0; i < INFLIGHT; i++) {
      uint32_t row = seed_hash(seed, (uint32_t)(r0 + i), (uint32_t)((r0 + i) >> 32)) % n;
      const float4 *rp = reinterpret_cast<const float4 *>(vec + (size_t)row * stride);
#pragma unroll
      for (int c = 0; c < NCH; c++) {
        int ch = t + c * TEAM;
        v[i][c] = ch < nch ? ld_row16(rp + ch) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int i = 0; i < INFLIGHT; i++)
#pragma unroll
      for (int c = 0; c < NCH; c++) acc += v[i][c].x + v[i][c].y + v[i][c].z + v[i][c].w;
  }
  if (acc == 1.2345e-30f) sink[0] = acc;  // keeps the loads alive
}

template <int TEAM, int NCH>
int probe_launch(b200ann_handle *h, const b200ann_index *ix, uint64_t rows, int inflight, int wpb, int bpsm,
                 uint32_t seed, float *sink) {
  dim3 grid((unsigned)(h->sm_count * bpsm)), block((unsigned)(wpb * 32));
  const int nch = (int)((ix->dim + 3) / 4);
#define PROBE_CASE(F)                                                                                              \
  case F:                                                                                                          \
    gather_probe_kernel<TEAM, NCH, F><<<grid, block, 0, h->stream>>>(ix->vec, ix->stride, (uint32_t)ix->n, nch, rows, seed, sink); \
    break;
  switch (inflight) {
    PROBE_CASE(1) PROBE_CASE(2) PROBE_CASE(4)
    default: return set_err(h, B200ANN_ERR_ARG, "probe_gather: rows in flight per team must be 1, 2 or 4");
  }
#undef PROBE_CASE
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}
}  // namespace

int launch_gather_probe(b200ann_handle *h, const b200ann_index *ix, uint64_t rows, int inflight, int wpb, int bpsm,
                        uint32_t seed, float *sink) {
  const int nch = (int)((ix->dim + 3) / 4);
  if (wpb < 1 || wpb > 8 || bpsm < 1 || bpsm > 32) return set_err(h, B200ANN_ERR_ARG, "probe_gather: warps per block in [1,8], blocks per SM in [1,32]");
  if (nch <= 8) return probe_launch<8, 1>(h, ix, rows, inflight, wpb, bpsm, seed, sink);
  if (nch <= 16) return probe_launch<8, 2>(h, ix, rows, inflight, wpb, bpsm, seed, sink);
  if (nch <= 32) return probe_launch<8, 4>(h, ix, rows, inflight, wpb, bpsm, seed, sink);
  if (nch <= 64) return probe_launch<8, 8>(h, ix, rows, inflight, wpb, bpsm, seed, sink);
  if (nch <= 128) return probe_launch<32, 4>(h, ix, rows, inflight, wpb, bpsm, seed, sink);
  if (nch <= 256) return probe_launch<32, 8>(h, ix, rows, inflight, wpb, bpsm, seed, sink);
  return set_err(h, B200ANN_ERR_ARG, "probe_gather: dim > 1024 is not supported");
}

int launch_nsg_search(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                      size_t K, size_t L, uint32_t seed, uint32_t *ids, float *dists, uint32_t *stats,
                      uint32_t qidx_base) {
  SearchParams p = {};
  p.qidx_base = qidx_base;
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = (int)((ix->dim + 3) / 4);
  p.graph = ix->graph; p.gstride = (uint32_t)ix->gstride; p.ep = ix->ep;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.dim = (uint32_t)ix->dim;
  p.K = (uint32_t)K; p.L = (uint32_t)L; p.seed = seed; p.hbits = h->visited_bits;
  p.cand_cap = round_up((uint32_t)(ix->gstride > L ? ix->gstride : L), 32);
  p.pool_cap = round_up((uint32_t)L + 1, 32);
  p.ids = ids; p.dists = dists; p.stats = stats;
  return ix->metric == B200ANN_IP ? launch_dim<true, false>(h, p) : launch_dim<false, false>(h, p);
}

// mode 2 (entry == nullptr): hybrid; mode 3: explicit entry points
int launch_nsg_search_from(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                           size_t K, size_t L, uint32_t seed, const uint32_t *entry, uint32_t *entry_out, uint32_t *ids,
                           float *dists, uint32_t *stats) {
  SearchParams p = {};
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = (int)((ix->dim + 3) / 4);
  p.graph = ix->graph; p.gstride = (uint32_t)ix->gstride; p.ep = ix->ep;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.dim = (uint32_t)ix->dim;
  p.K = (uint32_t)K; p.L = (uint32_t)L; p.seed = seed; p.hbits = h->visited_bits;
  size_t cc = ix->gstride > L ? ix->gstride : L;
  if (entry) {
    p.mode = 3; p.entry = entry;
  } else {
    p.mode = 2; p.entry_out = entry_out;
    p.upper_links = ix->upper_links; p.upper_ptr = ix->upper_ptr; p.maxM = (uint32_t)ix->maxM;
    p.maxlevel = ix->maxlevel; p.labels = ix->labels; p.ep_hnsw = ix->hnsw_ep;
    if (ix->maxM > cc) cc = ix->maxM;
  }
  p.cand_cap = round_up((uint32_t)cc, 32);
  p.pool_cap = round_up((uint32_t)L + 1, 32);
  p.ids = ids; p.dists = dists; p.stats = stats;
  return ix->metric == B200ANN_IP ? launch_dim<true, false>(h, p) : launch_dim<false, false>(h, p);
}

// mode 4: every query searches its own cluster of a block-diagonal graph
int launch_nsg_search_clustered(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                                size_t K, size_t L, uint32_t seed, const uint32_t *ep, const uint32_t *base,
                                const uint32_t *size, uint32_t *ids, float *dists, uint32_t *stats) {
  SearchParams p = {};
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = (int)((ix->dim + 3) / 4);
  p.graph = ix->graph; p.gstride = (uint32_t)ix->gstride; p.ep = ix->ep;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.dim = (uint32_t)ix->dim;
  p.K = (uint32_t)K; p.L = (uint32_t)L; p.seed = seed; p.hbits = h->visited_bits;
  p.mode = 4; p.entry = ep; p.cl_base = base; p.cl_size = size;
  p.cand_cap = round_up((uint32_t)(ix->gstride > L ? ix->gstride : L), 32);
  p.pool_cap = round_up((uint32_t)L + 1, 32);
  p.ids = ids; p.dists = dists; p.stats = stats;
  return ix->metric == B200ANN_IP ? launch_dim<true, false>(h, p) : launch_dim<false, false>(h, p);
}

int launch_hnsw_search(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                       size_t k, size_t ef, uint64_t *labels, float *dists, uint32_t *stats) {
  SearchParams p = {};
  size_t L = ef > k ? ef : k;  // hnswalg.h:1309 std::max(ef_, k)
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = ix_nch(ix);
  p.graph = ix->links0; p.gstride = (uint32_t)ix->maxM0; p.ep = ix->hnsw_ep;
  p.upper_links = ix->upper_links; p.upper_ptr = ix->upper_ptr; p.maxM = (uint32_t)ix->maxM;
  p.maxlevel = ix->maxlevel; p.labels = ix->labels;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.dim = (uint32_t)ix->dim;  // u8 index: q = byte rows, qstride in 4-byte units
  p.K = (uint32_t)k; p.L = (uint32_t)L; p.seed = 0; p.hbits = h->visited_bits;
  size_t cc = ix->maxM0 > ix->maxM ? ix->maxM0 : ix->maxM;
  p.cand_cap = round_up((uint32_t)cc, 32);
  p.pool_cap = round_up((uint32_t)L + 1, 32);
  p.labels_out = labels; p.dists = dists; p.stats = stats;
  if (ix_u8(ix)) return launch_dim<MET_U8, true>(h, p);
  return ix->metric == B200ANN_IP ? launch_dim<MET_IP, true>(h, p) : launch_dim<MET_L2, true>(h, p);
}

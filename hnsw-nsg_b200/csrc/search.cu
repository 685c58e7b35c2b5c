// Batched graph beam search: one warp per query.
//
// One kernel core, two front-ends:
//   NSG  — efanna2e::IndexNSG::Search / SearchWithOptGraph (CNNS/src/nsg/index_nsg.cpp:506-568, :570-657):
//          sorted pool of L, seeds = first L out-neighbours of ep + pseudo-random fill, expand the first
//          unexpanded pool entry, InsertIntoPool (CNNS/include/nsg/neighbor.h:107-135), k/nk rewind.
//   HNSW — hnswlib::HierarchicalNSW::searchKnn (hnswlib/hnswlib/hnswalg.h:1270-1324): greedy descent on
//          levels >= 1, then searchBaseLayerST<bare_bone> (:309-440) in its single-pool form (one
//          ascending (distance,id) pool of <= ef entries with an "unexpanded" flag; see oracle.c
//          hnsw_base_pool for why it equals the reference's two heaps).
//
// Per-warp shared memory: the pool (distance + id|flag), the compacted candidate list of the node
// being expanded, and a direct-mapped "forgetful" visited filter (u32 ids, 2^hbits slots). The
// filter never reports an unvisited node as visited; when it forgets a node the node's distance is
// recomputed and the pool's (distance,id) duplicate check / the monotone admission threshold reject
// it again, so results are identical to an exact visited set (the reference's N-bit bitset,
// index_nsg.cpp:512, or u16[N] epoch array, visited_list_pool.h:10-31) without N bits per query.
//
// Neighbour vectors are gathered with 16-byte loads, a team of 8 (dim <= 256) or 32 lanes per row,
// two rows per team in flight (dist.cuh).
#include "common.cuh"
#include "dist.cuh"

namespace {

constexpr int WARPS = 4;
constexpr uint32_t FLAG = 0x80000000u;
constexpr int BIGPOS = 0x7fffffff;

struct SearchParams {
  const float *vec; size_t stride; uint32_t n; int nch;
  const uint32_t *graph; uint32_t gstride; uint32_t ep;
  const uint32_t *upper_links; const uint32_t *upper_ptr; uint32_t maxM; int maxlevel;
  const uint64_t *labels;
  const float *q; size_t qstride; uint32_t nq; uint32_t dim;
  uint32_t K, L, seed; int hbits;
  uint32_t cand_cap, pool_cap;  // words
  uint32_t *ids; uint64_t *labels_out; float *dists; uint32_t *stats;
};

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
// stateless replacement for the reference's rand() seed fill (index_nsg.cpp:521-527)
__device__ __forceinline__ uint32_t seed_hash(uint32_t seed, uint32_t query, uint32_t counter) {
  return mix32(mix32(seed ^ (query * 0x9E3779B9U)) + counter * 0x85EBCA6BU);
}

template <int TEAM, int NCH, bool IP, bool HNSW>
struct Beam {
  static constexpr int TPW = 32 / TEAM;
  const SearchParams &p;
  float *pool_d; uint32_t *pool_i;   // [pool_cap]
  uint32_t *cand_i; float *cand_d;   // [cand_cap]
  uint32_t *tab; uint32_t hshift;
  float4 qreg[NCH];
  int lane, t, g;
  uint32_t np, cap;
  uint32_t n_dist, n_hops;

  __device__ __forceinline__ Beam(const SearchParams &pp) : p(pp) {}

  __device__ __forceinline__ uint32_t slot(uint32_t id) const { return (id * 0x9E3779B1u) >> hshift; }

  // distances of cand_i[0..m) -> cand_d[0..m)
  __device__ __forceinline__ void eval(uint32_t m) {
    __syncwarp();
    for (uint32_t c0 = 0; c0 < m; c0 += 2 * TPW) {
      uint32_t i0 = c0 + g, i1 = c0 + TPW + g;
      const float4 *r0 = nullptr, *r1 = nullptr;
      if (i0 < m) r0 = reinterpret_cast<const float4 *>(p.vec + (size_t)cand_i[i0] * p.stride);
      if (i1 < m) r1 = reinterpret_cast<const float4 *>(p.vec + (size_t)cand_i[i1] * p.stride);
      float d0, d1;
      team_dist2<TEAM, NCH, IP>(r0, r1, qreg, t, p.nch, d0, d1);
      if (t == 0) {
        if (i0 < m) cand_d[i0] = d0;
        if (i1 < m) cand_d[i1] = d1;
      }
    }
    n_dist += m;
    __syncwarp();
  }

  // copy an adjacency row into cand_i; with `filter`, drop ids the visited filter remembers.
  __device__ __forceinline__ uint32_t load_row(const uint32_t *row, uint32_t rowlen, uint32_t limit, bool filter) {
    uint32_t m = 0;
    __syncwarp();
    for (uint32_t j0 = 0; j0 < rowlen; j0 += 32) {
      uint32_t j = j0 + lane;
      uint32_t id = (j < rowlen && j < limit) ? __ldg(row + j) : INVALID_ID;
      bool valid = id != INVALID_ID;
      if (valid && filter) {
        uint32_t s = slot(id);
        if (tab[s] == id) valid = false;
        else tab[s] = id;
      }
      unsigned mask = __ballot_sync(0xffffffffu, valid);
      if (valid) cand_i[m + __popc(mask & ((1u << lane) - 1))] = id;
      m += __popc(mask);
      if (__any_sync(0xffffffffu, id == INVALID_ID)) break;
    }
    __syncwarp();
    return m;
  }

  // sorted insert. NSG: after every entry with distance <= d (InsertIntoPool, neighbor.h:107-135);
  // HNSW: (distance,id) lexicographic. Returns the position, or BIGPOS for a duplicate.
  __device__ __forceinline__ int insert(float d, uint32_t id) {
    int pos = 0;
    bool dup = false;
    for (uint32_t j0 = 0; j0 < np; j0 += 32) {
      uint32_t j = j0 + lane;
      bool in = j < np;
      float pdj = in ? pool_d[j] : 0.f;
      uint32_t pij = in ? (pool_i[j] & ~FLAG) : INVALID_ID;
      bool before = in && (HNSW ? (pdj < d || (pdj == d && pij < id)) : (pdj <= d));
      dup |= in && pdj == d && pij == id;
      pos += __popc(__ballot_sync(0xffffffffu, before));
    }
    if (__any_sync(0xffffffffu, dup)) return BIGPOS;
    int end = (np < cap) ? (int)np : (int)cap - 1;  // entries [pos,end) move up by one
    if (end > pos) {
      for (int jb = pos + ((end - 1 - pos) / 32) * 32; jb >= pos; jb -= 32) {
        int j = jb + lane;
        bool mv = j < end;
        float td = 0.f; uint32_t ti = 0;
        if (mv) { td = pool_d[j]; ti = pool_i[j]; }
        __syncwarp();
        if (mv) { pool_d[j + 1] = td; pool_i[j + 1] = ti; }
        __syncwarp();
      }
    }
    if (lane == 0) { pool_d[pos] = d; pool_i[pos] = id | FLAG; }
    if (np < cap) np++;
    __syncwarp();
    return pos;
  }

  __device__ __forceinline__ void run(uint32_t qi) {
    // ---- query -> registers (any alignment / stride)
    const float *qv = p.q + (size_t)qi * p.qstride;
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      uint32_t b = (uint32_t)(t + c * TEAM) * 4;
      qreg[c].x = b + 0 < p.dim ? __ldg(qv + b + 0) : 0.f;
      qreg[c].y = b + 1 < p.dim ? __ldg(qv + b + 1) : 0.f;
      qreg[c].z = b + 2 < p.dim ? __ldg(qv + b + 2) : 0.f;
      qreg[c].w = b + 3 < p.dim ? __ldg(qv + b + 3) : 0.f;
    }
    for (uint32_t i = lane; i < (1u << p.hbits); i += 32) tab[i] = INVALID_ID;
    n_dist = 0; n_hops = 0;
    cap = p.L;
    __syncwarp();

    if (HNSW) {
      // ---- upper layers: greedy descent (hnswalg.h:1275-1303)
      uint32_t cur = p.ep;
      if (lane == 0) cand_i[0] = cur;
      eval(1);
      float curdist = cand_d[0];
      for (int level = p.maxlevel; level > 0; level--) {
        bool changed = true;
        while (changed) {
          changed = false;
          const uint32_t *row = p.upper_links + ((size_t)__ldg(p.upper_ptr + cur) + (level - 1)) * p.maxM;
          uint32_t m = load_row(row, p.maxM, p.maxM, false);
          n_hops++;
          eval(m);
          // sequential "if (d < curdist)" over the list == first index of the minimum, if below curdist
          float bd = INFINITY; uint32_t bi = INVALID_ID;
          for (uint32_t c = lane; c < m; c += 32) {
            float d = cand_d[c];
            if (d < bd) { bd = d; bi = c; }
          }
#pragma unroll
          for (int off = 16; off >= 1; off >>= 1) {
            float od = __shfl_xor_sync(0xffffffffu, bd, off);
            uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (od < bd || (od == bd && oi < bi)) { bd = od; bi = oi; }
          }
          if (bi != INVALID_ID && bd < curdist) { curdist = bd; cur = cand_i[bi]; changed = true; }
        }
      }
      // ---- base layer starts from one entry (hnswalg.h:323-340)
      if (lane == 0) { pool_d[0] = curdist; pool_i[0] = cur | FLAG; tab[slot(cur)] = cur; }
      np = 1;
      __syncwarp();
    } else {
      // ---- seeds: first L out-neighbours of ep (index_nsg.cpp:516-519) ...
      uint32_t m = load_row(p.graph + (size_t)p.ep * p.gstride, p.gstride, p.L, false);
      // ... then pseudo-random ids not yet chosen (:521-527)
      uint32_t ctr = 0;
      while (m < p.L) {
        uint32_t id = seed_hash(p.seed, qi, ctr++) % p.n;
        bool hit = false;
        for (uint32_t j = lane; j < m; j += 32) hit |= cand_i[j] == id;
        if (!__any_sync(0xffffffffu, hit)) {
          if (lane == 0) cand_i[m] = id;
          m++;
        }
        __syncwarp();
      }
      eval(m);
      // std::sort(retset) (:538) as a rank sort on (distance,id); all seeds unexpanded
      for (uint32_t i = lane; i < m; i += 32) {
        float di = cand_d[i]; uint32_t ii = cand_i[i];
        uint32_t r = 0;
        for (uint32_t j = 0; j < m; j++) {
          float dj = cand_d[j]; uint32_t ij = cand_i[j];
          // position breaks (distance,id) ties: a reference-built row can list an id twice
          // (InterInsert leaves stale tail entries, index_nsg.cpp:399-404) and Search keeps both seeds
          r += (dj < di || (dj == di && (ij < ii || (ij == ii && j < i)))) ? 1u : 0u;
        }
        pool_d[r] = di; pool_i[r] = ii | FLAG;
        tab[slot(ii)] = ii;
      }
      np = m;
      __syncwarp();
    }

    // ---- main loop (index_nsg.cpp:539-564 / hnswalg.h:342-436)
    uint32_t k = 0;
    while (k < np) {
      int nk = (int)np;
      uint32_t e = pool_i[k];
      if (e & FLAG) {
        __syncwarp();
        if (lane == 0) pool_i[k] = e & ~FLAG;
        uint32_t node = e & ~FLAG;
        n_hops++;
        uint32_t m = load_row(p.graph + (size_t)node * p.gstride, p.gstride, p.gstride, true);
        eval(m);
        for (uint32_t c0 = 0; c0 < m; c0 += 32) {
          uint32_t c = c0 + lane;
          float d = c < m ? cand_d[c] : INFINITY;
          float last = pool_d[np - 1];
          bool pass = c < m && ((HNSW && np < cap) || d < last);
          unsigned mask = __ballot_sync(0xffffffffu, pass);
          while (mask) {
            int b = __ffs(mask) - 1;
            mask &= mask - 1;
            float dd = __shfl_sync(0xffffffffu, d, b);
            uint32_t id = cand_i[c0 + b];
            last = pool_d[np - 1];
            if (!((HNSW && np < cap) || dd < last)) continue;  // threshold moved since the vote
            int r = insert(dd, id);
            if (r < nk) nk = r;
          }
        }
      }
      k = ((uint32_t)nk <= k) ? (uint32_t)nk : k + 1;
    }
    __syncwarp();

    // ---- results
    for (uint32_t i = lane; i < p.K; i += 32) {
      bool ok = i < np;
      uint32_t id = ok ? (pool_i[i] & ~FLAG) : INVALID_ID;
      float d = ok ? pool_d[i] : INFINITY;
      if (HNSW) {
        p.labels_out[(size_t)qi * p.K + i] = ok ? (p.labels ? __ldg(p.labels + id) : (uint64_t)id) : ~(uint64_t)0;
      } else {
        p.ids[(size_t)qi * p.K + i] = id;
      }
      if (p.dists) p.dists[(size_t)qi * p.K + i] = d;
    }
    if (p.stats && lane == 0) {
      p.stats[2 * (size_t)qi] = n_dist;
      p.stats[2 * (size_t)qi + 1] = n_hops;
    }
  }
};

template <int TEAM, int NCH, bool IP, bool HNSW>
__global__ void __launch_bounds__(WARPS * 32) beam_search_kernel(const SearchParams p) {
  extern __shared__ uint32_t smem[];
  const int warp = threadIdx.x >> 5;
  const uint32_t qi = blockIdx.x * WARPS + warp;
  if (qi >= p.nq) return;
  const uint32_t per_warp = 2 * p.pool_cap + 2 * p.cand_cap + (1u << p.hbits);
  uint32_t *base = smem + (size_t)warp * per_warp;
  Beam<TEAM, NCH, IP, HNSW> b(p);
  b.pool_d = reinterpret_cast<float *>(base);
  b.pool_i = base + p.pool_cap;
  b.cand_i = base + 2 * p.pool_cap;
  b.cand_d = reinterpret_cast<float *>(base + 2 * p.pool_cap + p.cand_cap);
  b.tab = base + 2 * p.pool_cap + 2 * p.cand_cap;
  b.hshift = 32 - p.hbits;
  b.lane = threadIdx.x & 31;
  b.t = b.lane % TEAM;
  b.g = b.lane / TEAM;
  b.run(qi);
}

template <int TEAM, int NCH, bool IP, bool HNSW>
int launch_t(b200ann_handle *h, const SearchParams &p) {
  const uint32_t per_warp = 2 * p.pool_cap + 2 * p.cand_cap + (1u << p.hbits);
  size_t smem = (size_t)WARPS * per_warp * 4;
  if (smem > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "search: L / degree / visited_bits need more than 227 KB shared memory per block");
  auto kern = beam_search_kernel<TEAM, NCH, IP, HNSW>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((p.nq + WARPS - 1) / WARPS), block(WARPS * 32);
  kern<<<grid, block, smem, h->stream>>>(p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

template <bool IP, bool HNSW>
int launch_dim(b200ann_handle *h, const SearchParams &p) {
  int nch = p.nch;
  if (nch <= 8) return launch_t<8, 1, IP, HNSW>(h, p);
  if (nch <= 16) return launch_t<8, 2, IP, HNSW>(h, p);
  if (nch <= 32) return launch_t<8, 4, IP, HNSW>(h, p);
  if (nch <= 64) return launch_t<8, 8, IP, HNSW>(h, p);
  if (nch <= 128) return launch_t<32, 4, IP, HNSW>(h, p);
  if (nch <= 256) return launch_t<32, 8, IP, HNSW>(h, p);
  if (nch <= 512) return launch_t<32, 16, IP, HNSW>(h, p);
  return set_err(h, B200ANN_ERR_ARG, "search: dim > 2048 is not supported");
}

inline uint32_t round_up(uint32_t x, uint32_t m) { return (x + m - 1) / m * m; }

}  // namespace

int launch_nsg_search(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                      size_t K, size_t L, uint32_t seed, uint32_t *ids, float *dists, uint32_t *stats) {
  SearchParams p = {};
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = (int)((ix->dim + 3) / 4);
  p.graph = ix->graph; p.gstride = (uint32_t)ix->gstride; p.ep = ix->ep;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.dim = (uint32_t)ix->dim;
  p.K = (uint32_t)K; p.L = (uint32_t)L; p.seed = seed; p.hbits = h->visited_bits;
  p.cand_cap = round_up((uint32_t)(ix->gstride > L ? ix->gstride : L), 32);
  p.pool_cap = round_up((uint32_t)L + 1, 32);
  p.ids = ids; p.dists = dists; p.stats = stats;
  return ix->metric == B200ANN_IP ? launch_dim<true, false>(h, p) : launch_dim<false, false>(h, p);
}

int launch_hnsw_search(b200ann_handle *h, const b200ann_index *ix, const float *q, size_t nq, size_t qstride,
                       size_t k, size_t ef, uint64_t *labels, float *dists, uint32_t *stats) {
  SearchParams p = {};
  size_t L = ef > k ? ef : k;  // hnswalg.h:1309 std::max(ef_, k)
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = (int)((ix->dim + 3) / 4);
  p.graph = ix->links0; p.gstride = (uint32_t)ix->maxM0; p.ep = ix->hnsw_ep;
  p.upper_links = ix->upper_links; p.upper_ptr = ix->upper_ptr; p.maxM = (uint32_t)ix->maxM;
  p.maxlevel = ix->maxlevel; p.labels = ix->labels;
  p.q = q; p.qstride = qstride; p.nq = (uint32_t)nq; p.dim = (uint32_t)ix->dim;
  p.K = (uint32_t)k; p.L = (uint32_t)L; p.seed = 0; p.hbits = h->visited_bits;
  size_t cc = ix->maxM0 > ix->maxM ? ix->maxM0 : ix->maxM;
  p.cand_cap = round_up((uint32_t)cc, 32);
  p.pool_cap = round_up((uint32_t)L + 1, 32);
  p.labels_out = labels; p.dists = dists; p.stats = stats;
  return ix->metric == B200ANN_IP ? launch_dim<true, true>(h, p) : launch_dim<false, true>(h, p);
}

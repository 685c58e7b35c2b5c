// Warp-per-query beam search core shared by the search kernels (search.cu) and the NSG build
// kernels (nsg_build.cu).
//
// One core, two front-ends:
//   NSG  — efanna2e::IndexNSG::Search / SearchWithOptGraph (CNNS/src/nsg/index_nsg.cpp:506-568, :570-657)
//          and IndexNSG::get_neighbors (:150-285): sorted pool of L, seeds = first L out-neighbours of
//          ep + pseudo-random fill, expand the first unexpanded pool entry, InsertIntoPool
//          (CNNS/include/nsg/neighbor.h:107-135), k/nk rewind.
//   HNSW — hnswlib::HierarchicalNSW::searchKnn (hnswlib/hnswlib/hnswalg.h:1270-1324): greedy descent on
//          levels >= 1, then searchBaseLayerST<bare_bone> (:309-440) in its single-pool form (one
//          ascending (distance,id) pool of <= ef entries with an "unexpanded" flag; see oracle.c
//          hnsw_base_pool for why it equals the reference's two heaps).
//
// Per-warp shared memory: the pool (distance + id|flag), the compacted candidate list of the node
// being expanded, and the visited set: an open-addressing hash set of u32 ids (2^hbits slots, linear
// probing, <= PROBES probes). It never reports an unvisited node as visited. When a probe sequence
// is full the home slot is overwritten ("forgets" one id); a forgotten node's distance is recomputed
// and the pool's (distance,id) duplicate check / the monotone admission threshold reject it again,
// so results are identical to an exact visited set (the reference's N-bit bitset, index_nsg.cpp:512,
// or u16[N] epoch array, visited_list_pool.h:10-31) without N bits per query.
//
// Neighbour vectors are gathered with 16-byte loads, a team of 8 (dim <= 256) or 32 lanes per row,
// two rows per team in flight (dist.cuh).
//
// `Rec` is a warp-uniform callback rec(cand_i, cand_d, m, is_seed) invoked after every batch of
// distance evaluations; get_neighbors' `fullset` (:250,:273) is built through it.
#pragma once
#include "common.cuh"
#include "dist.cuh"

namespace beam {

constexpr uint32_t FLAG = 0x80000000u;
constexpr int BIGPOS = 0x7fffffff;
constexpr int PROBES = 8;

struct SearchParams {
  const float *vec; size_t stride; uint32_t n; int nch;
  const uint32_t *graph; uint32_t gstride; uint32_t ep;
  const uint32_t *upper_links; const uint32_t *upper_ptr; uint32_t maxM; int maxlevel;
  const uint64_t *labels;
  const float *q; size_t qstride; uint32_t nq; uint32_t dim;
  uint32_t K, L, seed; int hbits;
  uint32_t qidx_base;           // query index fed to the seed hash = qidx_base + query
  uint32_t id_base;             // added to every valid result id (row shards: local -> global row numbers)
  int mode;                     // non-HNSW kernels: 0 = Search from ep, 2 = hybrid (HNSW descent, then
                                // SearchFromEnterpoint), 3 = SearchFromEnterpoint from entry[query],
                                // 4 = Search inside one cluster per query (block-diagonal graph):
                                //     entry[q] = the cluster's navigating node, cl_base/cl_size = its rows
  const uint32_t *cl_base, *cl_size;  // mode 4: [nq]; cl_size 0 = the query is skipped
  uint32_t ep_hnsw;             // mode 2: enterpoint_node_ of the HNSW levels
  const uint32_t *entry;        // mode 3: [nq]
  uint32_t *entry_out;          // mode 2: [nq] or NULL
  uint32_t cand_cap, pool_cap;  // words
  uint32_t *ids; uint64_t *labels_out; float *dists; uint32_t *stats;
  uint32_t *work;               // search kernels: queue counter (NULL = query index from the launch grid)
  const uint32_t *order;        // search kernels: work item w -> query order[w] (NULL = identity); scheduling only
  int vis16;                    // visited set entries: 0 = u32 ids (2^hbits slots), 1 = u16 tags (2^(hbits+1) slots)
  int nbits;                    // vis16: ids are hashed as nbits-bit numbers (n <= 2^nbits)
  int prefetch;                 // 1: expand_all prefetches the adjacency row of the next unexpanded pool entry
  uint32_t merge_min;           // HNSW pools below this many entries insert one candidate at a time (A/B knob, 0: never)
  uint32_t nst;                 // search core v2 (beam2.cuh): stages of the per-warp row ring
  uint32_t q1;                  // search core v2: work items [0, q1) run one per warp, [q1, nq) one per block; work[0..1] = the two queues
};

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
// stateless replacement for the reference's rand() seed fill (index_nsg.cpp:521-527)
__device__ __forceinline__ uint32_t seed_hash(uint32_t seed, uint32_t query, uint32_t counter) {
  return mix32(mix32(seed ^ (query * 0x9E3779B9U)) + counter * 0x85EBCA6BU);
}

struct NoRec {
  __device__ __forceinline__ void operator()(const uint32_t *, const float *, uint32_t, bool) const {}
};

// words of shared memory one warp needs
__host__ __device__ inline uint32_t smem_words(uint32_t pool_cap, uint32_t cand_cap, int hbits) {
  return 2 * pool_cap + 2 * cand_cap + (1u << hbits);
}

// COHERENT: adjacency rows are read with ordinary (L2-coherent) loads instead of the read-only path,
// for kernels that append edges while they search (tree_grow).
// ROWS: candidate rows a team evaluates per round of eval() (2, or 4 for latency-bound kernels with spare registers).
template <int TEAM, int NCH, int IP, bool HNSW, bool COHERENT = false, int ROWS = 2>
struct Beam {
  static constexpr int TPW = 32 / TEAM;
  static constexpr int TEAMSZ = TEAM, NCHUNK = NCH;
  static constexpr int IS_IP = IP;  // the metric: MET_L2 / MET_IP / MET_U8 (dist.cuh)
  const SearchParams &p;
  float *pool_d; uint32_t *pool_i;   // [pool_cap]
  uint32_t *cand_i; float *cand_d;   // [cand_cap]
  uint32_t *tab; uint32_t hshift, hmask;
  uint32_t nmask, rbits, rmask, smask;  // vis16 geometry
  float4 qreg[NCH];
  int lane, t, g;
  uint32_t np, cap;
  uint32_t n_dist, n_hops;
  uint32_t never = INVALID_ID;  // a node load_row(filter) always drops, even after the visited set forgot it
  int lvl = 0;  // HNSW only: the layer expand_all walks (0 = links0 rows, l >= 1 = the node's upper row l-1)

  __device__ __forceinline__ Beam(const SearchParams &pp, uint32_t *smem_base) : p(pp) {
    pool_d = reinterpret_cast<float *>(smem_base);
    pool_i = smem_base + p.pool_cap;
    cand_i = smem_base + 2 * p.pool_cap;
    cand_d = reinterpret_cast<float *>(smem_base + 2 * p.pool_cap + p.cand_cap);
    tab = smem_base + 2 * p.pool_cap + 2 * p.cand_cap;
    hshift = 32 - p.hbits;
    hmask = (1u << p.hbits) - 1;
    nmask = rbits = rmask = smask = 0;
    if (p.vis16) {
      nmask = p.nbits >= 32 ? 0xFFFFFFFFu : ((1u << p.nbits) - 1);
      rbits = (uint32_t)(p.nbits - (p.hbits + 1));
      rmask = (1u << rbits) - 1;
      smask = (2u << p.hbits) - 1;
    }
    lane = threadIdx.x & 31;
    t = lane % TEAM;
    g = lane / TEAM;
  }

  __device__ __forceinline__ uint32_t slot(uint32_t id) const { return (id * 0x9E3779B1u) >> hshift; }

  // visited set: true if `id` was already present, else inserts it
  __device__ __forceinline__ bool test_and_set(uint32_t id) {
    if (p.vis16) {
      const uint32_t x = (id * 0x9E3779B1u) & nmask;
      const uint32_t s = x >> rbits;
      const uint32_t tag = (x & rmask) << 3;
      unsigned short *t16 = reinterpret_cast<unsigned short *>(tab);
#pragma unroll 1
      for (uint32_t pr = 0; pr < (uint32_t)PROBES; pr++) {
        const unsigned short e = (unsigned short)(tag | pr);
        unsigned short old = atomicCAS(&t16[(s + pr) & smask], (unsigned short)0xFFFFu, e);
        if (old == 0xFFFFu) return false;
        if (old == e) return true;
      }
      t16[s] = (unsigned short)tag;  // probe sequence full: forget the home slot's entry
      return false;
    }
    uint32_t s = slot(id);
#pragma unroll 1
    for (int pr = 0; pr < PROBES; pr++) {
      uint32_t old = atomicCAS(&tab[(s + pr) & hmask], INVALID_ID, id);
      if (old == INVALID_ID) return false;
      if (old == id) return true;
    }
    tab[s] = id;  // probe sequence full: forget the home slot's id
    return false;
  }

  // The same test for a whole adjacency row (lane = entry, all 32 lanes call it): plain loads, and a new id claims
  // its slot by store + read-back - the table belongs to this warp alone, so of the lanes that store to one slot in
  // the same round exactly one reads its own entry back and the others probe on. Shared-memory atomics cost the
  // insertion beam ~1500 cycles per hop (32 tests, most of them for ids already in the set). Two lanes with the
  // SAME id would both report "new": the caller passes each id once (load_row).
  __device__ __forceinline__ bool test_and_set_row(uint32_t id, bool valid) {
    bool seen = false, done = !valid;
    if (p.vis16) {
      const uint32_t x = (id * 0x9E3779B1u) & nmask;
      const uint32_t s = x >> rbits;
      const uint32_t tag = (x & rmask) << 3;
      volatile unsigned short *t16 = reinterpret_cast<volatile unsigned short *>(tab);  // no store -> load forwarding
#pragma unroll 1
      for (uint32_t pr = 0; pr < (uint32_t)PROBES; pr++) {
        const uint32_t sl = (s + pr) & smask;
        const unsigned short me = (unsigned short)(tag | pr);
        const unsigned short e = done ? (unsigned short)0 : t16[sl];
        if (!done && e == me) { seen = true; done = true; }
        const bool claim = !done && e == 0xFFFFu;
        if (claim) t16[sl] = me;
        __syncwarp();
        if (claim && t16[sl] == me) done = true;
        if (__all_sync(0xffffffffu, done)) break;
      }
      if (!done) t16[s] = (unsigned short)tag;  // probe sequence full: forget the home slot's entry
      __syncwarp();
      return seen;
    }
    const uint32_t s = slot(id);
    volatile uint32_t *vt = tab;
#pragma unroll 1
    for (int pr = 0; pr < PROBES; pr++) {
      const uint32_t sl = (s + pr) & hmask;
      const uint32_t e = done ? 0u : vt[sl];
      if (!done && e == id) { seen = true; done = true; }
      const bool claim = !done && e == INVALID_ID;
      if (claim) vt[sl] = id;
      __syncwarp();
      if (claim && vt[sl] == id) done = true;
      if (__all_sync(0xffffffffu, done)) break;
    }
    if (!done) vt[s] = id;
    __syncwarp();
    return seen;
  }

  __device__ __forceinline__ void load_query(const float *qv) {
    if constexpr (IP == MET_U8) {  // qv = p.dim unsigned bytes (any alignment); bytes past dim read as 0
      const unsigned char *qb = reinterpret_cast<const unsigned char *>(qv);
#pragma unroll
      for (int c = 0; c < NCH; c++) {
        const uint32_t b = (uint32_t)(t + c * TEAM) * 16;
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          w[j] = 0;
#pragma unroll
          for (int k = 0; k < 4; k++) {
            const uint32_t i = b + 4 * j + k;
            if (i < p.dim) w[j] |= (uint32_t)__ldg(qb + i) << (8 * k);
          }
        }
        qreg[c] = make_float4(__uint_as_float(w[0]), __uint_as_float(w[1]), __uint_as_float(w[2]), __uint_as_float(w[3]));
      }
    } else {
#pragma unroll
      for (int c = 0; c < NCH; c++) {
        uint32_t b = (uint32_t)(t + c * TEAM) * 4;
        qreg[c].x = b + 0 < p.dim ? __ldg(qv + b + 0) : 0.f;
        qreg[c].y = b + 1 < p.dim ? __ldg(qv + b + 1) : 0.f;
        qreg[c].z = b + 2 < p.dim ? __ldg(qv + b + 2) : 0.f;
        qreg[c].w = b + 3 < p.dim ? __ldg(qv + b + 3) : 0.f;
      }
    }
  }

  // distances of cand_i[0..m) -> cand_d[0..m)
  __device__ __forceinline__ void eval(uint32_t m) {
    __syncwarp();
    if constexpr (ROWS > 2) {
      for (uint32_t c0 = 0; c0 < m; c0 += ROWS * TPW) {
        const float4 *r[ROWS];
        float d[ROWS];
#pragma unroll
        for (int k = 0; k < ROWS; k++) {
          const uint32_t i = c0 + k * TPW + g;
          r[k] = i < m ? reinterpret_cast<const float4 *>(p.vec + (size_t)cand_i[i] * p.stride) : nullptr;
        }
        team_dist_rows<TEAM, NCH, IP, ROWS>(r, qreg, t, p.nch, d);
        if (t == 0) {
#pragma unroll
          for (int k = 0; k < ROWS; k++) {
            const uint32_t i = c0 + k * TPW + g;
            if (i < m) cand_d[i] = d[k];
          }
        }
      }
    } else {
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
    }
    n_dist += m;
    __syncwarp();
  }

  __device__ __forceinline__ uint32_t ld_adj(const uint32_t *a) const {
    if (COHERENT) return __ldcg(a);
    return __ldg(a);
  }

  // copy an adjacency row into cand_i; with `filter`, drop ids already in the visited set.
  // have_pre: the caller already holds the row's first 32 entries (pre = entry `lane`, see expand_all)
  __device__ __forceinline__ uint32_t load_row(const uint32_t *row, uint32_t rowlen, uint32_t limit, bool filter,
                                               bool have_pre = false, uint32_t pre = INVALID_ID) {
    uint32_t m = 0;
    __syncwarp();
    for (uint32_t j0 = 0; j0 < rowlen; j0 += 32) {
      uint32_t j = j0 + lane;
      uint32_t id = (j < rowlen && j < limit) ? ((have_pre && j0 == 0) ? pre : ld_adj(row + j)) : INVALID_ID;
      bool valid = id != INVALID_ID;
      if (filter) {
        valid = valid && id != never;
        // a row listing a node twice: the first copy stands (test_and_set_row would call both new)
        const unsigned same = __match_any_sync(0xffffffffu, id);   // every lane takes part: no short-circuit
        valid = valid && (uint32_t)__ffs(same) - 1u == (uint32_t)lane;
        valid = !test_and_set_row(id, valid) && valid;
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

  // HNSW pools of up to 32 * MCH entries: one chunk of candidates (one per lane) merged into the pool in ONE pass,
  // with the result the sequential loop above produces candidate by candidate (lane order = arrival order):
  //  * candidate c is accepted iff  #{pool: dist <= d_c} + #{earlier candidates that are no duplicates: dist <= d_c}
  //    < cap. (Sequential rule: accepted iff the pool is not full or d_c < the pool's largest distance, i.e. iff fewer
  //    than cap of the current entries have dist <= d_c; the current entries are the cap smallest of pool + accepted,
  //    and a candidate rejected earlier with dist <= d_c implies c is rejected too, so "accepted earlier" can be
  //    replaced by "earlier".) A candidate equal to a pool entry (the visited set forgot it) changes nothing.
  //  * the new pool is the cap (distance, id)-smallest of pool + accepted: an accepted candidate lands at
  //    (#pool entries before it) + (#accepted before it); the pool entries fill the other positions in order.
  // Every lane searches its own candidate (three levels of independent probes: strides 32, 4, 1); the pool sits in
  // registers (entry i * 32 + lane) and every entry moves up by the number of accepted candidates before it. A
  // chunk costs about the same however many candidates it inserts (the sequential form re-scans and shifts the
  // pool per candidate: 3500 cycles per hop at ef_construction 200, where 7-10 of a hop's ~13 evaluated rows pass).
  // Returns the smallest position written (BIGPOS if none).
  static constexpr int MCH = 8;
  __device__ __forceinline__ int merge_chunk(float d, uint32_t id, bool valid) {
    const uint32_t np0 = np;
    const bool pass = valid && (np0 < cap || d < pool_d[np0 - 1]);
    const unsigned mask = __ballot_sync(0xffffffffu, pass);
    if (!mask) return BIGPOS;
    float pd[MCH]; uint32_t pi[MCH];
    const int nchk = (int)((np0 + 31) >> 5);
#pragma unroll
    for (int i = 0; i < MCH; i++) {
      const uint32_t j = (uint32_t)i * 32u + lane;
      const bool in = i < nchk && j < np0;
      pd[i] = in ? pool_d[j] : INFINITY;
      pi[i] = in ? pool_i[j] : 0u;
    }
    uint32_t lex = 0, le = 0;   // pool entries (distance, id)-before mine / with distance <= mine
    bool dup = false;
    if (pass) {
      uint32_t c = 0;           // entries with distance < d: whole 32-blocks, then 4-blocks, then single entries
#pragma unroll
      for (int t = 1; t < MCH; t++) c += ((uint32_t)t * 32u - 1u < np0 && pool_d[t * 32 - 1] < d) ? 1u : 0u;
      uint32_t base = c * 32u;
      c = 0;
#pragma unroll
      for (int t = 1; t < 8; t++) c += (base + (uint32_t)t * 4u - 1u < np0 && pool_d[base + t * 4 - 1] < d) ? 1u : 0u;
      base += c * 4u;
      c = 0;
#pragma unroll
      for (int t = 0; t < 3; t++) c += (base + (uint32_t)t < np0 && pool_d[base + t] < d) ? 1u : 0u;
      lex = le = base + c;
      if (lex < np0 && pool_d[lex] == d) {  // ties in distance: order by id, find a duplicate
        while (lex < np0 && pool_d[lex] == d && (pool_i[lex] & ~FLAG) < id) lex++;
        dup = lex < np0 && pool_d[lex] == d && (pool_i[lex] & ~FLAG) == id;
        while (le < np0 && pool_d[le] == d) le++;
      }
    }
    __syncwarp();
    const unsigned live = __ballot_sync(0xffffffffu, pass && !dup);
    uint32_t earlier = 0;
    for (unsigned mm = live; mm; mm &= mm - 1) {
      const int b = __ffs(mm) - 1;
      const float db = __shfl_sync(0xffffffffu, d, b);
      earlier += (b < lane && db <= d) ? 1u : 0u;
    }
    const bool acc = pass && !dup && le + earlier < cap;
    const unsigned amask = __ballot_sync(0xffffffffu, acc);
    if (!amask) return BIGPOS;
    uint32_t before = 0, sh[MCH];
#pragma unroll
    for (int i = 0; i < MCH; i++) sh[i] = 0;
    for (unsigned mm = amask; mm; mm &= mm - 1) {
      const int b = __ffs(mm) - 1;
      const float db = __shfl_sync(0xffffffffu, d, b);
      const uint32_t idb = __shfl_sync(0xffffffffu, id, b);
      const uint32_t lxb = __shfl_sync(0xffffffffu, lex, b);
      before += (db < d || (db == d && idb < id)) ? 1u : 0u;
#pragma unroll
      for (int i = 0; i < MCH; i++) sh[i] += (lxb <= (uint32_t)i * 32u + lane) ? 1u : 0u;
    }
    const uint32_t at = lex + before;
    const int first = (int)__reduce_min_sync(0xffffffffu, acc ? at : (uint32_t)BIGPOS);
    __syncwarp();
#pragma unroll
    for (int i = 0; i < MCH; i++) {
      const uint32_t j = (uint32_t)i * 32u + lane;
      if (i < nchk && j < np0 && sh[i] && j + sh[i] < cap) { pool_d[j + sh[i]] = pd[i]; pool_i[j + sh[i]] = pi[i]; }
    }
    if (acc && at < cap) { pool_d[at] = d; pool_i[at] = id | FLAG; }
    np = min(cap, np0 + (uint32_t)__popc(amask));
    __syncwarp();
    return first;
  }

  // query -> registers (any alignment / stride), empty visited set
  __device__ __forceinline__ void begin(const float *qv) {
    load_query(qv);
    for (uint32_t i = lane; i <= hmask; i += 32) tab[i] = INVALID_ID;
    n_dist = 0; n_hops = 0;
    cap = p.L;
    np = 0;
    __syncwarp();
  }

  // HNSW upper layers: greedy descent (hnswalg.h:1275-1303); the base layer then starts from the
  // one entry it ends on (:323-340)
  __device__ __forceinline__ void descend_hnsw() {
    float curdist;
    uint32_t cur = descend_upper(p.ep, curdist);
    if (lane == 0) { pool_d[0] = curdist; pool_i[0] = cur | FLAG; }
    test_and_set(cur);  // every lane offers the same id: one wins, the rest see it present
    np = 1;
    __syncwarp();
  }

  // the greedy walk itself (also the first phase of HNSW_NSG::searchKnn,
  // hnsw_nsg/include/index_hnsw_nsg.h:111-143): returns the node it ends on
  __device__ __forceinline__ uint32_t descend_upper(uint32_t start, float &curdist, int stop_level = 0) {
    uint32_t cur = start;
    if (lane == 0) cand_i[0] = cur;
    eval(1);
    curdist = cand_d[0];
    for (int level = p.maxlevel; level > stop_level; level--) {
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
    __syncwarp();
    return cur;
  }

  // warp-uniform visited test: lane 0 probes, everyone gets the answer
  __device__ __forceinline__ bool test_and_set_uniform(uint32_t id) {
    int was = 0;
    if (lane == 0) was = test_and_set(id) ? 1 : 0;
    return __shfl_sync(0xffffffffu, was, 0) != 0;
  }

  // rank sort of cand_i/cand_d[0..m) into the pool: std::sort(retset) as (distance,id,position)
  __device__ __forceinline__ void seeds_to_pool(uint32_t m, bool flag_ids) {
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
      if (flag_ids) test_and_set(ii);
    }
    np = m;
    __syncwarp();
  }

  // seeds of IndexNSG::SearchFromEnterpoint (hnsw_nsg/src/index_nsg.cpp:703-749), quirks included:
  // the entry point is flagged but never evaluated (:711-712); the neighbours-of-seeds pass indexes
  // each seed's row with the running seed count and leaves id 0 in the slot of an already flagged
  // neighbour (:720-728; init_ids is zero-initialised, :707); slots holding that 0 are evaluated
  // without being flagged (:741-747).
  __device__ __forceinline__ void seed_from(uint32_t ep, uint32_t qidx) {
    never = ep;  // flagged but never evaluated (:711-712): must stay excluded when the hash set forgets it
    test_and_set_uniform(ep);
    uint32_t m = load_row(p.graph + (size_t)ep * p.gstride, p.gstride, p.L, false);
    for (uint32_t j = lane; j < m; j += 32) test_and_set(cand_i[j]);
    __syncwarp();
    const uint32_t cur = m;
    uint32_t lim = p.L < p.gstride ? p.L : p.gstride;
    for (uint32_t i = 0; i < cur && m < p.L; i++) {
      const uint32_t *row = p.graph + (size_t)cand_i[i] * p.gstride;
      bool ended = false;
      while (m < lim && !ended) {
        uint32_t j = m + lane;
        uint32_t v = j < lim ? ld_adj(row + j) : INVALID_ID;
        unsigned inv = __ballot_sync(0xffffffffu, v == INVALID_ID);
        uint32_t cnt = inv ? (uint32_t)(__ffs(inv) - 1) : 32u;  // entries of this chunk before the row end
        if ((uint32_t)lane < cnt) cand_i[j] = test_and_set(v) ? 0u : v;
        m += cnt;
        ended = inv != 0;
        __syncwarp();
      }
    }
    uint32_t ctr = 0;
    while (m < p.L) {
      uint32_t id = seed_hash(p.seed, qidx, ctr++) % p.n;
      if (!test_and_set_uniform(id)) {
        if (lane == 0) cand_i[m] = id;
        m++;
      }
    }
    __syncwarp();
    eval(m);
    seeds_to_pool(m, false);
  }

  // NSG seeds: first L out-neighbours of ep (index_nsg.cpp:516-519), then pseudo-random ids not yet
  // chosen (:521-527); distances; std::sort(retset) (:538) as a rank sort on (distance,id,position)
  template <class Rec>
  __device__ __forceinline__ void seed_nsg(uint32_t qidx, Rec &rec) {
    uint32_t m = load_row(p.graph + (size_t)p.ep * p.gstride, p.gstride, p.L, false);
    uint32_t ctr = 0;
    while (m < p.L) {
      uint32_t id = seed_hash(p.seed, qidx, ctr++) % p.n;
      bool hit = false;
      for (uint32_t j = lane; j < m; j += 32) hit |= cand_i[j] == id;
      if (!__any_sync(0xffffffffu, hit)) {
        if (lane == 0) cand_i[m] = id;
        m++;
      }
      __syncwarp();
    }
    eval(m);
    rec(cand_i, cand_d, m, true);
    seeds_to_pool(m, true);
  }

  // seed_nsg for a query confined to the rows [base, base + size) of a block-diagonal graph (one NSG per cluster,
  // ids global): the per-cluster IndexNSG::Search of CNNS/tests/cluster_hnsw_nsg_search.cpp:203-207. The random
  // fill draws inside the cluster (rand() % nd_ there is the cluster's own nd_); a cluster smaller than L is
  // searched with a pool of its size.
  __device__ __forceinline__ void seed_nsg_local(uint32_t qidx, uint32_t ep, uint32_t base, uint32_t size) {
    const uint32_t Lq = p.L < size ? p.L : size;
    uint32_t m = load_row(p.graph + (size_t)ep * p.gstride, p.gstride, Lq, false);
    uint32_t ctr = 0;
    while (m < Lq) {
      uint32_t id = base + seed_hash(p.seed, qidx, ctr++) % size;
      bool hit = false;
      for (uint32_t j = lane; j < m; j += 32) hit |= cand_i[j] == id;
      if (!__any_sync(0xffffffffu, hit)) {
        if (lane == 0) cand_i[m] = id;
        m++;
      }
      __syncwarp();
    }
    eval(m);
    seeds_to_pool(m, true);
    cap = Lq;
  }

  // Speculative prefetch (the reference's _mm_prefetch of the next rows, index_nsg.cpp:627-632, hnswalg.h:370-383):
  // while node pool[k] is expanded, the adjacency row of the next unexpanded pool entry - the node the loop turns to
  // next unless a closer candidate arrives - is pulled into L2, so its load later costs an L2 hit instead of a DRAM
  // round trip on the hop's critical path. Touches no state: results cannot change.
  __device__ __forceinline__ void prefetch_next_row(uint32_t k) {
    uint32_t nxt = INVALID_ID;
    for (uint32_t j0 = k + 1; j0 < np; j0 += 32) {
      uint32_t j = j0 + lane;
      uint32_t e = j < np ? pool_i[j] : 0u;
      unsigned m = __ballot_sync(0xffffffffu, (e & FLAG) != 0);
      if (m) { nxt = __shfl_sync(0xffffffffu, e, __ffs(m) - 1) & ~FLAG; break; }
    }
    if (nxt != INVALID_ID) {
      const uint32_t *row = p.graph + (size_t)nxt * p.gstride;
      if ((uint32_t)lane * 32u < p.gstride) asm volatile("prefetch.global.L2 [%0];" ::"l"(row + lane * 32));
    }
  }

  // main loop (index_nsg.cpp:539-564 / hnswalg.h:342-436)
#ifdef BEAM_PROF
  long long tp[6] = {0, 0, 0, 0, 0, 0};   // load_row, eval, insert, total of expand_all, speculation hits, spec scan
#define BP_T(x) const long long x = clock64()
#define BP_ADD(i, a, b) tp[i] += (b) - (a)
#else
#define BP_T(x)
#define BP_ADD(i, a, b)
#endif
  template <class Rec>
  __device__ __forceinline__ void expand_all(Rec &rec) {
    uint32_t k = 0;
    // HNSW level 0: while a node's neighbours are evaluated, the adjacency row of the node the loop most likely
    // turns to next (the next unexpanded pool entry) is loaded into a register per lane; when the guess holds - no
    // closer candidate arrived - the next hop starts without its dependent row load. A wrong guess is dropped.
    uint32_t spec_node = INVALID_ID, spec_val = INVALID_ID;
    BP_T(t_begin);
    while (k < np) {
      {  // skip to the first unexpanded entry at or after k, 32 entries per step
        uint32_t kk = np;
        for (uint32_t j0 = k & ~31u; j0 < np; j0 += 32) {
          const uint32_t j = j0 + lane;
          const unsigned mf = __ballot_sync(0xffffffffu, j >= k && j < np && (pool_i[j] & FLAG));
          if (mf) { kk = j0 + __ffs(mf) - 1; break; }
        }
        if (kk >= np) break;
        k = kk;
      }
      int nk = (int)np;
      uint32_t e = pool_i[k];
      if (e & FLAG) {
        __syncwarp();
        if (lane == 0) pool_i[k] = e & ~FLAG;
        uint32_t node = e & ~FLAG;
        n_hops++;
        if (p.prefetch && !(HNSW && lvl > 0)) prefetch_next_row(k);
        const uint32_t *row = p.graph + (size_t)node * p.gstride;
        uint32_t rl = p.gstride;
        if (HNSW && lvl > 0) {
          row = p.upper_links + ((size_t)__ldg(p.upper_ptr + node) + (lvl - 1)) * p.maxM;
          rl = p.maxM;
        }
        BP_T(t0);
        const bool hit = HNSW && !COHERENT && lvl == 0 && node == spec_node;
        uint32_t m = load_row(row, rl, rl, true, hit, spec_val);
#ifdef BEAM_PROF
        tp[4] += hit;
        const long long ts0 = clock64();
#endif
        if (HNSW && !COHERENT && lvl == 0) {
          spec_node = INVALID_ID;
          for (uint32_t j0 = (k + 1) & ~31u; j0 < np; j0 += 32) {
            const uint32_t j = j0 + lane;
            const uint32_t pe = (j > k && j < np) ? pool_i[j] : 0u;
            const unsigned mf = __ballot_sync(0xffffffffu, (pe & FLAG) != 0);
            if (mf) { spec_node = __shfl_sync(0xffffffffu, pe, __ffs(mf) - 1) & ~FLAG; break; }
          }
          if (spec_node != INVALID_ID)
            spec_val = (uint32_t)lane < p.gstride ? ld_adj(p.graph + (size_t)spec_node * p.gstride + lane) : INVALID_ID;
        }
#ifdef BEAM_PROF
        tp[5] += clock64() - ts0;
#endif
        BP_T(t1);
        eval(m);
        BP_T(t2);
        BP_ADD(0, t0, t1); BP_ADD(1, t1, t2);
        rec(cand_i, cand_d, m, false);
        for (uint32_t c0 = 0; c0 < m; c0 += 32) {
          uint32_t c = c0 + lane;
          float d = c < m ? cand_d[c] : INFINITY;
          if (HNSW && cap <= 32u * MCH && cap >= p.merge_min) {
            const int r = merge_chunk(d, c < m ? cand_i[c] : INVALID_ID, c < m);
            if (r < nk) nk = r;
            continue;
          }
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
        BP_T(t3);
        BP_ADD(2, t2, t3);
      }
      k = ((uint32_t)nk <= k) ? (uint32_t)nk : k + 1;
    }
    __syncwarp();
    BP_T(t_end);
    BP_ADD(3, t_begin, t_end);
  }
};

// ---------------------------------------------------------------- occlusion scan
// The rule shared by sync_prune (index_nsg.cpp:328-345), InterInsert (:386-404) and hnswlib's
// getNeighborsByHeuristic2 (hnswalg.h:443-483): scan sorted candidates
// [start,end), keep p iff no kept r has r.id == p.id or d(r,p) < d(q,p); stop at R kept.
// Kept rows are compared 2*TPW at a time; the first kept entry must be placed by the caller or by
// passing nres == 0 (then the first candidate is kept unconditionally, like `result.push_back(pool[start])`).
template <class B>
__device__ __forceinline__ void occlude_scan(B &b, const float *sd, const uint32_t *si, uint32_t start,
                                             uint32_t end, uint32_t R, uint32_t *res_i, float *res_d,
                                             uint32_t &nres, uint32_t &n_eval) {
  constexpr int TPW = B::TPW;
  for (uint32_t s = start; s < end && nres < R; s++) {
    const uint32_t pid = si[s];
    const float pd = sd[s];
    bool occ = false;
    if (nres > 0) {
      const float4 *prow = reinterpret_cast<const float4 *>(b.p.vec + (size_t)pid * b.p.stride);
#pragma unroll
      for (int c = 0; c < B::NCHUNK; c++) {
        int ch = b.t + c * B::TEAMSZ;
        b.qreg[c] = ch < b.p.nch ? ld_row16_ca(prow + ch) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      for (uint32_t t0 = 0; t0 < nres && !occ; t0 += 2 * TPW) {
        uint32_t i0 = t0 + b.g, i1 = t0 + TPW + b.g;
        const float4 *r0 = nullptr, *r1 = nullptr;
        uint32_t id0 = INVALID_ID, id1 = INVALID_ID;
        if (i0 < nres) { id0 = res_i[i0]; r0 = reinterpret_cast<const float4 *>(b.p.vec + (size_t)id0 * b.p.stride); }
        if (i1 < nres) { id1 = res_i[i1]; r1 = reinterpret_cast<const float4 *>(b.p.vec + (size_t)id1 * b.p.stride); }
        float d0, d1;
        team_dist2<B::TEAMSZ, B::NCHUNK, B::IS_IP, true>(r0, r1, b.qreg, b.t, b.p.nch, d0, d1);
        bool o = (i0 < nres && (id0 == pid || d0 < pd)) || (i1 < nres && (id1 == pid || d1 < pd));
        occ = __any_sync(0xffffffffu, o);
        n_eval += min(nres - t0, (uint32_t)(2 * TPW));
      }
    }
    if (!occ) {
      __syncwarp();
      if (b.lane == 0) { res_i[nres] = pid; res_d[nres] = pd; }
      nres++;
      __syncwarp();
    }
  }
}

// dispatch on the row length: nch = ceil(dim/4) float4 chunks -> (TEAM, NCH)
#define BEAM_DISPATCH_NCH(nch, CALL)                 \
  do {                                               \
    if ((nch) <= 8) { CALL(8, 1); }                  \
    else if ((nch) <= 16) { CALL(8, 2); }            \
    else if ((nch) <= 32) { CALL(8, 4); }            \
    else if ((nch) <= 64) { CALL(8, 8); }            \
    else if ((nch) <= 128) { CALL(32, 4); }          \
    else if ((nch) <= 256) { CALL(32, 8); }          \
    else if ((nch) <= 512) { CALL(32, 16); }         \
  } while (0)

inline uint32_t round_up(uint32_t x, uint32_t m) { return (x + m - 1) / m * m; }

}  // namespace beam

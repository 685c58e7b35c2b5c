// NSG construction on the GPU: efanna2e::IndexNSG::Build (CNNS/src/nsg/index_nsg.cpp:465-504).
//
//   init_graph  (:287-303)  centroid_kernel (sequential fp32 sum per dimension, the reference's order; rows staged by loader warps)
//                           + one beam search for the centroid (search.cu)
//   Link        (:426-457)  link_kernel: warp per node. get_neighbors (:218-285) = the beam core over
//                           the kNN graph with every evaluated node recorded; sync_prune (:305-355) =
//                           sort + occlusion scan by the same warp.
//   InterInsert (:357-424)  order-independent batch form (oracle.c orc_nsg_interinsert variant 1):
//                           reverse lists by destination (count / scan / fill), then a warp per
//                           destination merges own + reverse candidates and re-prunes on overflow.
//   compaction  (:473-487)  cut_to_graph_kernel
//   tree_grow   (:684-764)  level-synchronous BFS from the navigating node, then one persistent CTA
//                           that attaches each remaining component exactly like findroot (:712-749).
//
// get_neighbors' `fullset` only matters up to its C smallest entries (sync_prune scans
// pool[0..C), :328), so a warp keeps a 2C-entry buffer in shared memory and compacts it (bitonic
// sort on (distance,id), drop duplicates, keep C) whenever it fills; entries above the current
// C-th smallest are dropped on arrival. Duplicates can only come from the forgetful visited set and
// carry identical (distance,id), so the compacted list equals the reference's sorted pool prefix.
#include "beam.cuh"
#include "topbuf.cuh"

using namespace beam;

namespace {

constexpr int WARPS = 4;
constexpr int LINK_MAXW = 8;  // link_kernel: warps per block are chosen at launch (shared-memory occupancy)

struct TopRec {  // Beam callback: every evaluated node goes to the collector (Link's get_neighbors)
  TopBuf &tb;
  __device__ __forceinline__ void operator()(const uint32_t *ci, const float *cd, uint32_t m, bool) {
    tb.push_batch(ci, cd, m, 0.f, 0u, false);
  }
};

// ---------------------------------------------------------------- init_graph: centroid
// center[j] = (sum over rows in index order of x[i][j]) / n: the fp32 additions of one dimension happen in the
// reference's order (:290-297), so every dimension is one dependent chain of n additions - 4 cycles each, 2 ms at
// 1M rows, whatever reads the rows. A block owns 8 dimensions (32 contiguous bytes of every row): eight loader
// warps stage tiles of 512 rows in shared memory (16-byte loads, double-buffered) while eight lanes of warp 0 walk
// the previous tile in row order. One thread per dimension reading global memory itself (round 1) kept one SM's
// 16 KB in flight: 46 ms at 1M x 128.
constexpr int CT_ROWS = 512;
__global__ void __launch_bounds__(288) centroid_kernel(const float *vec, size_t stride, uint32_t n, uint32_t dim, float *center) {
  __shared__ float4 tile[2][CT_ROWS][2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t d0 = blockIdx.x * 8;
  const uint32_t ntiles = (n + CT_ROWS - 1) / CT_ROWS;
  auto load = [&](uint32_t t, int buf) {   // loader threads (warps 1-8): 4 x 16 bytes each per tile
    const int lt = (int)threadIdx.x - 32;
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int idx = lt + u * 256;
      const uint32_t row = t * CT_ROWS + (uint32_t)(idx >> 1), c = d0 + (uint32_t)(idx & 1) * 4u;
      v[u] = (row < n && c < stride) ? __ldg(reinterpret_cast<const float4 *>(vec + (size_t)row * stride + c))
                                     : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int idx = lt + u * 256;
      tile[buf][idx >> 1][idx & 1] = v[u];
    }
  };
  if (warp > 0) load(0, 0);
  __syncthreads();
  float acc = 0.f;
  for (uint32_t t = 0; t < ntiles; t++) {
    const int buf = (int)(t & 1u);
    if (warp > 0) {
      if (t + 1 < ntiles) load(t + 1, buf ^ 1);
    } else if (lane < 8) {
      const float *tp = reinterpret_cast<const float *>(&tile[buf][0][0]) + lane;
      const uint32_t rows = min((uint32_t)CT_ROWS, n - t * CT_ROWS);
#pragma unroll 8
      for (uint32_t r = 0; r < rows; r++) acc += tp[r * 8];
    }
    __syncthreads();
  }
  if (warp == 0 && lane < 8 && d0 + lane < dim) center[d0 + lane] = acc / (float)n;
}

// ---------------------------------------------------------------- Link: get_neighbors + sync_prune
struct LinkParams {
  SearchParams sp;      // graph = kNN graph, ep = navigating node, L
  uint32_t R, C, cap2;  // cap2: collector capacity (power of two >= 2C)
  uint32_t *cut_ids; float *cut_d;  // [n,R]
  uint32_t *stats;      // [n,3] or null: beam evaluations, expanded nodes, prune evaluations
};

__host__ __device__ inline uint32_t link_smem_words(const LinkParams &lp) {
  return smem_words(lp.sp.pool_cap, lp.sp.cand_cap, lp.sp.hbits) + 2 * lp.cap2;
}

template <int TEAM, int NCH, bool IP>
__global__ void __launch_bounds__(LINK_MAXW * 32, 1) link_kernel(const LinkParams lp) {
  extern __shared__ uint32_t smem[];
  const SearchParams &p = lp.sp;
  const int warp = threadIdx.x >> 5;
  const uint32_t node = blockIdx.x * (blockDim.x >> 5) + warp;
  if (node >= p.n) return;
  uint32_t *base = smem + (size_t)warp * link_smem_words(lp);
  typedef Beam<TEAM, NCH, IP, false> B;
  B b(p, base);
  TopBuf tb;
  tb.d = reinterpret_cast<float *>(base + smem_words(p.pool_cap, p.cand_cap, p.hbits));
  tb.i = base + smem_words(p.pool_cap, p.cand_cap, p.hbits) + lp.cap2;
  tb.cap2 = lp.cap2; tb.keep = lp.C; tb.lane = b.lane;
  tb.reset();
  TopRec rec{tb};

  b.begin(p.vec + (size_t)node * p.stride);
  b.seed_nsg(node, rec);
  b.expand_all(rec);
  const uint32_t beam_evals = b.n_dist;
  // sync_prune :314-321 — the node's own kNN list, minus what the search already visited
  uint32_t m = b.load_row(p.graph + (size_t)node * p.gstride, p.gstride, p.gstride, true);
  b.eval(m);
  rec(b.cand_i, b.cand_d, m, false);
  tb.compact();

  // sync_prune :323-345 on the sorted, duplicate-free pool prefix; results live in the (now free)
  // candidate arrays
  uint32_t *res_i = b.cand_i;
  float *res_d = b.cand_d;
  uint32_t nres = 0, n_prune = m;
  uint32_t start = (tb.cnt > 0 && tb.i[0] == node) ? 1u : 0u;
  occlude_scan(b, tb.d, tb.i, start, tb.cnt, lp.R, res_i, res_d, nres, n_prune);
  __syncwarp();
  for (uint32_t t = b.lane; t < lp.R; t += 32) {
    lp.cut_ids[(size_t)node * lp.R + t] = t < nres ? res_i[t] : INVALID_ID;
    lp.cut_d[(size_t)node * lp.R + t] = t < nres ? res_d[t] : -1.0f;
  }
  if (lp.stats && b.lane == 0) {
    lp.stats[3 * (size_t)node] = beam_evals;
    lp.stats[3 * (size_t)node + 1] = b.n_hops;
    lp.stats[3 * (size_t)node + 2] = n_prune;
  }
}

// ---------------------------------------------------------------- InterInsert, batch form
__global__ void rev_count_kernel(const uint32_t *cut_ids, const float *cut_d, size_t total, uint32_t *cnt) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= total) return;
  if (cut_d[e] != -1.0f) atomicAdd(&cnt[cut_ids[e]], 1u);
}

// exclusive scan of cnt[0..n) -> off[0..n], one block (n is at most a few 10^7: memory-trivial)
__global__ void scan_kernel(const uint32_t *cnt, uint32_t n, uint32_t *off) {
  __shared__ uint32_t part[1024];
  const uint32_t tid = threadIdx.x, nt = blockDim.x;
  const uint32_t per = (n + nt - 1) / nt;
  const uint32_t b = min(n, tid * per), e = min(n, b + per);
  uint32_t s = 0;
  for (uint32_t i = b; i < e; i++) s += cnt[i];
  part[tid] = s;
  __syncthreads();
  if (tid == 0) {
    uint32_t run = 0;
    for (uint32_t t = 0; t < nt; t++) { uint32_t v = part[t]; part[t] = run; run += v; }
    off[n] = run;
  }
  __syncthreads();
  s = part[tid];
  for (uint32_t i = b; i < e; i++) { off[i] = s; s += cnt[i]; }
}

__global__ void rev_fill_kernel(const uint32_t *cut_ids, const float *cut_d, size_t total, uint32_t R,
                                const uint32_t *off, uint32_t *fill, uint32_t *rev_src, float *rev_d) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= total) return;
  float d = cut_d[e];
  if (d == -1.0f) return;
  uint32_t des = cut_ids[e];
  uint32_t pos = off[des] + atomicAdd(&fill[des], 1u);
  rev_src[pos] = (uint32_t)(e / R);
  rev_d[pos] = d;
}

struct InterParams {
  SearchParams sp;  // vec/stride/nch/dim only (+ pool_cap/cand_cap/hbits for the smem carve-up)
  uint32_t R, keep, cap2;
  const uint32_t *cut_ids; const float *cut_d;     // snapshot rows [n,R]
  const uint32_t *off; const uint32_t *rev_src; const float *rev_d;
  uint32_t *new_ids; float *new_d;                 // [n,R]
  unsigned long long *n_eval;
};

// words per warp: own row (2R) + result (2R) + small scratch for the fits-in-R case (2R) + collector
__host__ __device__ inline uint32_t inter_tmp_cap(uint32_t R) { return R > 32 ? R : 32; }
__host__ __device__ inline uint32_t inter_smem_words(const InterParams &ip) {
  return 4 * ip.R + 2 * inter_tmp_cap(ip.R) + 2 * ip.cap2;
}

template <int TEAM, int NCH, bool IP>
__global__ void __launch_bounds__(WARPS * 32) interinsert_kernel(const InterParams ip) {
  extern __shared__ uint32_t smem[];
  const SearchParams &p = ip.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t des = blockIdx.x * WARPS + warp;
  if (des >= p.n) return;
  const uint32_t R = ip.R;
  uint32_t *base = smem + (size_t)warp * inter_smem_words(ip);
  uint32_t *own_i = base; float *own_d = reinterpret_cast<float *>(base + R);
  uint32_t *res_i = base + 2 * R; float *res_d = reinterpret_cast<float *>(base + 3 * R);
  uint32_t *tmp_i = base + 4 * R; float *tmp_d = reinterpret_cast<float *>(base + 4 * R + inter_tmp_cap(R));
  TopBuf tb;
  tb.d = reinterpret_cast<float *>(base + 4 * R + 2 * inter_tmp_cap(R));
  tb.i = base + 4 * R + 2 * inter_tmp_cap(R) + ip.cap2;
  tb.cap2 = ip.cap2; tb.keep = ip.keep; tb.lane = lane;

  // own sync_prune row
  uint32_t own = 0;
  for (uint32_t j0 = 0; j0 < R; j0 += 32) {
    uint32_t j = j0 + lane;
    float d = j < R ? ip.cut_d[(size_t)des * R + j] : -1.0f;
    uint32_t id = j < R ? ip.cut_ids[(size_t)des * R + j] : INVALID_ID;
    bool ok = d != -1.0f;
    if (ok) { own_i[j] = id; own_d[j] = d; }
    unsigned mask = __ballot_sync(0xffffffffu, ok);
    own += __popc(mask);
    if (mask != 0xffffffffu) break;
  }
  __syncwarp();
  const uint32_t rb = ip.off[des], re = ip.off[des + 1];
  auto is_dup = [&](uint32_t src) {
    bool dup = false;
    for (uint32_t j = 0; j < own; j++) dup |= own_i[j] == src;
    return dup;
  };
  // how many reverse candidates are new?
  uint32_t nrev = 0;
  for (uint32_t e0 = rb; e0 < re; e0 += 32) {
    uint32_t e = e0 + lane;
    bool ok = e < re && !is_dup(ip.rev_src[e]);
    nrev += __popc(__ballot_sync(0xffffffffu, ok));
  }
  uint32_t nres = 0;
  const uint32_t *out_i; const float *out_d;
  if (own + nrev <= R) {
    // fits: own row, then the reverse candidates in ascending source id (arrival order of the
    // serial reference loop, :406-415)
    uint32_t m = 0;
    for (uint32_t e0 = rb; e0 < re; e0 += 32) {
      uint32_t e = e0 + lane;
      uint32_t src = e < re ? ip.rev_src[e] : INVALID_ID;
      bool ok = e < re && !is_dup(src);
      unsigned mask = __ballot_sync(0xffffffffu, ok);
      if (ok) { uint32_t pos = m + __popc(mask & ((1u << lane) - 1)); tmp_i[pos] = src; tmp_d[pos] = ip.rev_d[e]; }
      m += __popc(mask);
    }
    __syncwarp();
    for (uint32_t a = lane; a < m; a += 32) {
      uint32_t ia = tmp_i[a], r = 0;
      for (uint32_t c = 0; c < m; c++) r += tmp_i[c] < ia ? 1u : 0u;
      own_i[own + r] = ia; own_d[own + r] = tmp_d[a];
    }
    __syncwarp();
    nres = own + m;
    out_i = own_i; out_d = own_d;
  } else {
    // overflow: all candidates ascending by (distance,id), occlusion scan until R are kept
    // (:386-404). The collector hands them out `keep` at a time, smallest first.
    typedef Beam<TEAM, NCH, IP, false> B;
    B b(p, base);  // only qreg / lane / team geometry are used; its smem carve-up is not touched
    b.lane = lane; b.t = lane % TEAM; b.g = lane / TEAM;
    float lo_d = -INFINITY; uint32_t lo_i = 0; bool use_lo = false;
    uint32_t n_eval = 0;
    for (;;) {
      tb.reset();
      tb.push_batch(own_i, own_d, own, lo_d, lo_i, use_lo);
      for (uint32_t e0 = rb; e0 < re; e0 += 32) {
        uint32_t e = e0 + lane;
        uint32_t src = e < re ? ip.rev_src[e] : INVALID_ID;
        float d = e < re ? ip.rev_d[e] : INFINITY;
        bool ok = e < re && !is_dup(src);
        // stage this chunk through tmp so that push_batch sees a dense list
        unsigned mask = __ballot_sync(0xffffffffu, ok);
        if (ok) { uint32_t pos = __popc(mask & ((1u << lane) - 1)); tmp_i[pos] = src; tmp_d[pos] = d; }
        __syncwarp();
        tb.push_batch(tmp_i, tmp_d, __popc(mask), lo_d, lo_i, use_lo);
      }
      tb.compact();
      occlude_scan(b, tb.d, tb.i, 0, tb.cnt, R, res_i, res_d, nres, n_eval);
      if (nres >= R || tb.cnt < tb.keep) break;
      lo_d = tb.d[tb.cnt - 1]; lo_i = tb.i[tb.cnt - 1]; use_lo = true;
      __syncwarp();
    }
    if (lane == 0 && ip.n_eval) atomicAdd(ip.n_eval, (unsigned long long)n_eval);
    out_i = res_i; out_d = res_d;
  }
  __syncwarp();
  for (uint32_t t = lane; t < R; t += 32) {
    ip.new_ids[(size_t)des * R + t] = t < nres ? out_i[t] : INVALID_ID;
    ip.new_d[(size_t)des * R + t] = t < nres ? out_d[t] : -1.0f;
  }
}

// ---------------------------------------------------------------- compaction (:473-487)
__global__ void cut_to_graph_kernel(const uint32_t *cut_ids, const float *cut_d, uint32_t n, uint32_t R,
                                    uint32_t *graph, uint32_t gs, uint32_t *deg) {
  uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  uint32_t k = 0;
  while (k < R && cut_d[(size_t)r * R + k] != -1.0f) k++;
  for (uint32_t c = 0; c < gs; c++) graph[(size_t)r * gs + c] = c < k ? cut_ids[(size_t)r * R + c] : INVALID_ID;
  deg[r] = k;
}

// ---------------------------------------------------------------- tree_grow (:684-764)
// reach[] = 1 for nodes reachable from the navigating node. Level-synchronous BFS: a warp per
// frontier node scans its row, newly reached ids are appended to the next frontier.
__global__ void bfs_level_kernel(const uint32_t *graph, uint32_t gs, const uint32_t *fin, uint32_t nin,
                                 uint32_t *fout, uint32_t *nout, uint32_t *reach) {
  const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= nin) return;
  const uint32_t *row = graph + (size_t)fin[w] * gs;
  for (uint32_t j = lane; j < gs; j += 32) {
    uint32_t id = row[j];
    if (id == INVALID_ID) break;
    if (atomicExch(&reach[id], 1u) == 0u) fout[atomicAdd(nout, 1u)] = id;
  }
}

struct GrowParams {
  SearchParams sp;      // graph = the NSG adjacency being repaired (read coherently), ep, L
  uint32_t *graph_rw;   // same buffer, writable
  uint32_t *deg;        // [n] current row lengths
  uint32_t *reach;      // [n]
  uint32_t *queue;      // [n] BFS scratch
  uint32_t reached;     // nodes reached by the initial BFS
  uint32_t *added;      // out: number of edges added
};

// Beam callback for findroot (:727-737): the nearest evaluated node (ascending (distance,id), seeds
// excluded like the first get_neighbors overload, :150-216) that is already reached and has a free
// slot in its row.
struct RootRec {
  const uint32_t *reach; const uint32_t *deg; uint32_t gs;
  float best_d; uint32_t best_i; int lane;
  __device__ __forceinline__ void operator()(const uint32_t *ci, const float *cd, uint32_t m, bool seeds) {
    if (seeds) return;
    for (uint32_t c = lane; c < m; c += 32) {
      uint32_t id = ci[c];
      float d = cd[c];
      if (lex_less(d, id, best_d, best_i) && __ldcg(reach + id) != 0u && __ldcg(deg + id) < gs) { best_d = d; best_i = id; }
    }
  }
  __device__ __forceinline__ void reduce() {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      float od = __shfl_xor_sync(0xffffffffu, best_d, off);
      uint32_t oi = __shfl_xor_sync(0xffffffffu, best_i, off);
      if (lex_less(od, oi, best_d, best_i)) { best_d = od; best_i = oi; }
    }
  }
};

constexpr int GROW_THREADS = 256;

// One persistent CTA: while some node is unreached { id = first unreached in index order; warp 0
// beam-searches the current graph for vec[id] and picks the parent; edge parent -> id is appended;
// everything reachable from id is marked by a block-wide BFS }.
template <int TEAM, int NCH, bool IP>
__global__ void __launch_bounds__(GROW_THREADS) grow_kernel(const GrowParams gp) {
  extern __shared__ uint32_t smem[];
  __shared__ uint32_t s_scan, s_id, s_head, s_tail, s_cnt, s_min[GROW_THREADS / 32];
  const SearchParams &p = gp.sp;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { s_scan = 0; s_cnt = gp.reached; }
  __syncthreads();
  uint32_t added = 0;
  while (s_cnt < p.n) {
    // ---- first unreached id >= s_scan
    for (;;) {
      uint32_t base = s_scan;
      uint32_t i = base + tid;
      uint32_t cand = (i < p.n && __ldcg(gp.reach + i) == 0u) ? i : INVALID_ID;
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, off));
      if (lane == 0) s_min[warp] = cand;
      __syncthreads();
      if (tid == 0) {
        uint32_t mn = INVALID_ID;
        for (int w = 0; w < GROW_THREADS / 32; w++) mn = min(mn, s_min[w]);
        s_id = mn;
        s_scan = mn != INVALID_ID ? mn : base + GROW_THREADS;
      }
      __syncthreads();
      if (s_id != INVALID_ID || s_scan >= p.n) break;
    }
    const uint32_t id = s_id;
    if (id == INVALID_ID) break;  // cannot happen while s_cnt < n
    // ---- findroot
    if (warp == 0) {
      typedef Beam<TEAM, NCH, IP, false, true> B;
      B b(p, smem);
      RootRec rec{gp.reach, gp.deg, p.gstride, INFINITY, INVALID_ID, (int)lane};
      b.begin(p.vec + (size_t)id * p.stride);
      b.seed_nsg(id, rec);
      b.expand_all(rec);
      rec.reduce();
      uint32_t root = rec.best_i;
      if (root == INVALID_ID) {  // :739-746, with the counter hash in place of rand()
        for (uint32_t t = 0;; t++) {
          uint32_t rid = seed_hash(p.seed, id, 0x10000u + t) % p.n;
          if (__ldcg(gp.reach + rid) != 0u && __ldcg(gp.deg + rid) < p.gstride) { root = rid; break; }
        }
      }
      if (lane == 0) {
        uint32_t dg = gp.deg[root];
        gp.graph_rw[(size_t)root * p.gstride + dg] = id;
        gp.deg[root] = dg + 1;
        gp.reach[id] = 1u;
        gp.queue[0] = id;
        s_head = 0; s_tail = 1;
        __threadfence();
      }
    }
    __syncthreads();
    added++;
    // ---- mark everything reachable from id
    uint32_t head = s_head, tail = s_tail;
    while (head < tail) {
      __syncthreads();
      for (uint32_t qx = head + tid; qx < tail; qx += GROW_THREADS) {
        const uint32_t *row = gp.graph_rw + (size_t)gp.queue[qx] * p.gstride;
        for (uint32_t j = 0; j < p.gstride; j++) {
          uint32_t nb = __ldcg(row + j);
          if (nb == INVALID_ID) break;
          if (atomicExch(&gp.reach[nb], 1u) == 0u) gp.queue[atomicAdd(&s_tail, 1u)] = nb;
        }
      }
      __syncthreads();
      head = tail;
      tail = s_tail;
    }
    __syncthreads();
    if (tid == 0) s_cnt += tail;  // tail = size of the newly reached component
    __threadfence();
    __syncthreads();
  }
  if (tid == 0 && gp.added) *gp.added = added;
}

__global__ void max_deg_kernel(const uint32_t *deg, uint32_t n, uint32_t *width) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t v = i < n ? deg[i] : 0u;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, off));
  if ((threadIdx.x & 31) == 0 && v) atomicMax(width, v);
}

}  // namespace

// ================================================================= host side
namespace {

uint32_t pow2_at_least(uint32_t x) { uint32_t p = 64; while (p < x) p <<= 1; return p; }

SearchParams base_params(const b200ann_index *ix) {
  SearchParams p = {};
  p.vec = ix->vec; p.stride = ix->stride; p.n = (uint32_t)ix->n; p.nch = (int)((ix->dim + 3) / 4);
  p.dim = (uint32_t)ix->dim;
  return p;
}

template <int TEAM, int NCH, bool IP>
int launch_link_t(b200ann_handle *h, const LinkParams &lp) {
  // the kernel is occupancy-bound by shared memory (pool + candidates + visited table + collector per warp): pick the
  // warps per block that leave the fewest bytes of the SM's 228 KB unused (each block also reserves 1 KB)
  const size_t per_warp = (size_t)link_smem_words(lp) * 4;
  if (per_warp > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "nsg_link: L / K / C need more than 227 KB shared memory per block");
  int W = 1; size_t best = 0;
  for (int w = 1; w <= LINK_MAXW; w++) {
    size_t blk = per_warp * w;
    if (blk > 227 * 1024) break;
    size_t resident = (228 * 1024 / (blk + 1024)) * w;
    if (resident > 48) resident = 48;  // beyond the register limit more warps per block only coarsen the grid
    if (resident > best || (resident == best && w <= 4)) { best = resident; W = w; }
  }
  size_t smem = per_warp * W;
  auto kern = link_kernel<TEAM, NCH, IP>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(lp.sp.n + W - 1) / W, W * 32, smem, h->stream>>>(lp);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

template <int TEAM, int NCH, bool IP>
int launch_inter_t(b200ann_handle *h, const InterParams &ip) {
  size_t smem = (size_t)WARPS * inter_smem_words(ip) * 4;
  if (smem > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "nsg_interinsert: R needs more than 227 KB shared memory per block");
  auto kern = interinsert_kernel<TEAM, NCH, IP>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(ip.sp.n + WARPS - 1) / WARPS, WARPS * 32, smem, h->stream>>>(ip);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

template <int TEAM, int NCH, bool IP>
int launch_grow_t(b200ann_handle *h, const GrowParams &gp) {
  size_t smem = (size_t)smem_words(gp.sp.pool_cap, gp.sp.cand_cap, gp.sp.hbits) * 4;
  auto kern = grow_kernel<TEAM, NCH, IP>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<1, GROW_THREADS, smem, h->stream>>>(gp);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

#define DISPATCH_METRIC(FN, ARG)                                                              \
  do {                                                                                        \
    if (nch > 512) return set_err(h, B200ANN_ERR_ARG, "dim > 2048 is not supported");         \
    if (ip_metric) { BEAM_DISPATCH_NCH(nch, FN##_IP); } else { BEAM_DISPATCH_NCH(nch, FN##_L2); } \
  } while (0)

}  // namespace

// ---- phase launchers (all asynchronous on h->stream unless they return a host value)
int nsg_link_launch(b200ann_handle *h, const b200ann_index *ix, const uint32_t *knn, size_t kstride, uint32_t ep,
                    size_t L, size_t R, size_t C, uint32_t seed, uint32_t *cut_ids, float *cut_d, uint32_t *stats) {
  LinkParams lp = {};
  lp.sp = base_params(ix);
  lp.sp.graph = knn; lp.sp.gstride = (uint32_t)kstride; lp.sp.ep = ep;
  lp.sp.L = (uint32_t)L; lp.sp.K = (uint32_t)L; lp.sp.seed = seed;
  lp.sp.hbits = h->link_visited_bits ? h->link_visited_bits : (ix->dim <= 128 ? 9 : 10);
  // 16-bit tags in the same words (twice the slots) when the row count allows: what the set forgets is re-evaluated
  // (24 % of the beam's rows with 2^9 32-bit slots at 1M), never lost - the candidate buffer drops duplicates
  lp.sp.nbits = 1;
  while (lp.sp.nbits < 32 && (1ull << lp.sp.nbits) < (unsigned long long)ix->n) lp.sp.nbits++;
  if (lp.sp.nbits < lp.sp.hbits + 1) lp.sp.nbits = lp.sp.hbits + 1;
  lp.sp.vis16 = lp.sp.nbits - (lp.sp.hbits + 1) <= 12 ? 1 : 0;
  if (const char *e = getenv("B200ANN_LINK_VIS16")) lp.sp.vis16 = lp.sp.vis16 && atoi(e);   // A/B runs
  size_t cc = kstride > L ? kstride : L;
  if (R > cc) cc = R;
  lp.sp.cand_cap = round_up((uint32_t)cc, 32);
  lp.sp.pool_cap = round_up((uint32_t)L + 1, 32);
  lp.R = (uint32_t)R; lp.C = (uint32_t)C; lp.cap2 = pow2_at_least(2 * (uint32_t)C);
  lp.cut_ids = cut_ids; lp.cut_d = cut_d; lp.stats = stats;
  const int nch = lp.sp.nch;
  const bool ip_metric = ix->metric == B200ANN_IP;
  int rc = 0;
#define LINK_IP(T, N) rc = launch_link_t<T, N, true>(h, lp)
#define LINK_L2(T, N) rc = launch_link_t<T, N, false>(h, lp)
  DISPATCH_METRIC(LINK, lp);
  return rc;
}

int nsg_interinsert_launch(b200ann_handle *h, const b200ann_index *ix, size_t R, uint32_t *cut_ids, float *cut_d,
                           unsigned long long *n_eval_dev) {
  const size_t n = ix->n, total = n * R;
  DevBuf<uint32_t> cnt, off, fill, rev_src, new_ids;
  DevBuf<float> rev_d, new_d;
  CU_CHECK(h, cnt.alloc(n, h->stream)); CU_CHECK(h, off.alloc(n + 1, h->stream)); CU_CHECK(h, fill.alloc(n, h->stream));
  CU_CHECK(h, rev_src.alloc(total, h->stream)); CU_CHECK(h, rev_d.alloc(total, h->stream));
  CU_CHECK(h, new_ids.alloc(total, h->stream)); CU_CHECK(h, new_d.alloc(total, h->stream));
  CU_CHECK(h, cudaMemsetAsync(cnt.p, 0, n * 4, h->stream));
  CU_CHECK(h, cudaMemsetAsync(fill.p, 0, n * 4, h->stream));
  const unsigned eb = (unsigned)((total + 255) / 256);
  rev_count_kernel<<<eb, 256, 0, h->stream>>>(cut_ids, cut_d, total, cnt.p);
  scan_kernel<<<1, 1024, 0, h->stream>>>(cnt.p, (uint32_t)n, off.p);
  rev_fill_kernel<<<eb, 256, 0, h->stream>>>(cut_ids, cut_d, total, (uint32_t)R, off.p, fill.p, rev_src.p, rev_d.p);
  h->launches += 3;
  CU_CHECK(h, cudaGetLastError());
  InterParams ip = {};
  ip.sp = base_params(ix);
  ip.R = (uint32_t)R; ip.keep = 512; ip.cap2 = 1024;
  ip.cut_ids = cut_ids; ip.cut_d = cut_d; ip.off = off.p; ip.rev_src = rev_src.p; ip.rev_d = rev_d.p;
  ip.new_ids = new_ids.p; ip.new_d = new_d.p; ip.n_eval = n_eval_dev;
  const int nch = ip.sp.nch;
  const bool ip_metric = ix->metric == B200ANN_IP;
  int rc = 0;
#define INTER_IP(T, N) rc = launch_inter_t<T, N, true>(h, ip)
#define INTER_L2(T, N) rc = launch_inter_t<T, N, false>(h, ip)
  DISPATCH_METRIC(INTER, ip);
  if (rc) return rc;
  CU_CHECK(h, cudaMemcpyAsync(cut_ids, new_ids.p, total * 4, cudaMemcpyDeviceToDevice, h->stream));
  CU_CHECK(h, cudaMemcpyAsync(cut_d, new_d.p, total * 4, cudaMemcpyDeviceToDevice, h->stream));
  return 0;
}

// cut graph -> ix->graph, then tree_grow on it. Synchronises (the BFS level loop reads counters).
int nsg_finish(b200ann_handle *h, b200ann_index *ix, const uint32_t *cut_ids, const float *cut_d, size_t R, uint32_t ep,
               size_t L, uint32_t seed, size_t slack, uint32_t *added_out) {
  const size_t n = ix->n;
  const size_t gs = (R + slack + 7) / 8 * 8;
  uint32_t *graph = nullptr;
  CU_CHECK(h, cudaMalloc((void **)&graph, n * gs * 4));
  DevBuf<uint32_t> deg, reach, fa, fb, ctr;
  cudaError_t e = cudaSuccess;
  if ((e = deg.alloc(n, h->stream)) != cudaSuccess || (e = reach.alloc(n, h->stream)) != cudaSuccess ||
      (e = fa.alloc(n, h->stream)) != cudaSuccess || (e = fb.alloc(n, h->stream)) != cudaSuccess ||
      (e = ctr.alloc(4, h->stream)) != cudaSuccess) {
    cudaFree(graph);
    return set_err(h, B200ANN_ERR_CUDA, std::string("nsg_finish: ") + cudaGetErrorString(e));
  }
  cut_to_graph_kernel<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>(cut_ids, cut_d, (uint32_t)n, (uint32_t)R, graph,
                                                                        (uint32_t)gs, deg.p);
  h->launches++;
  // BFS from ep
  cudaMemsetAsync(reach.p, 0, n * 4, h->stream);
  uint32_t one = 1;
  cudaMemcpyAsync(reach.p + ep, &one, 4, cudaMemcpyHostToDevice, h->stream);
  cudaMemcpyAsync(fa.p, &ep, 4, cudaMemcpyHostToDevice, h->stream);
  uint32_t nin = 1, reached = 1;
  uint32_t *fin = fa.p, *fout = fb.p;
  while (nin > 0) {
    cudaMemsetAsync(ctr.p, 0, 4, h->stream);
    bfs_level_kernel<<<(unsigned)(((size_t)nin * 32 + 255) / 256), 256, 0, h->stream>>>(graph, (uint32_t)gs, fin, nin, fout,
                                                                                     ctr.p, reach.p);
    h->launches++;
    cudaMemcpyAsync(&nin, ctr.p, 4, cudaMemcpyDeviceToHost, h->stream);
    e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) { cudaFree(graph); return set_err(h, B200ANN_ERR_CUDA, std::string("nsg_finish bfs: ") + cudaGetErrorString(e)); }
    reached += nin;
    uint32_t *t = fin; fin = fout; fout = t;
  }
  uint32_t added = 0;
  if (reached < n) {
    GrowParams gp = {};
    gp.sp = base_params(ix);
    gp.sp.graph = graph; gp.sp.gstride = (uint32_t)gs; gp.sp.ep = ep;
    gp.sp.L = (uint32_t)L; gp.sp.K = (uint32_t)L; gp.sp.seed = seed; gp.sp.hbits = h->visited_bits;
    gp.sp.cand_cap = round_up((uint32_t)(gs > L ? gs : L), 32);
    gp.sp.pool_cap = round_up((uint32_t)L + 1, 32);
    gp.graph_rw = graph; gp.deg = deg.p; gp.reach = reach.p; gp.queue = fa.p; gp.reached = reached;
    gp.added = ctr.p + 1;
    const int nch = gp.sp.nch;
    const bool ip_metric = ix->metric == B200ANN_IP;
    int rc = 0;
#define GROW_IP(T, N) rc = launch_grow_t<T, N, true>(h, gp)
#define GROW_L2(T, N) rc = launch_grow_t<T, N, false>(h, gp)
    DISPATCH_METRIC(GROW, gp);
    if (rc) { cudaFree(graph); return rc; }
    cudaMemcpyAsync(&added, ctr.p + 1, 4, cudaMemcpyDeviceToHost, h->stream);
  }
  uint32_t width = 0;
  cudaMemsetAsync(ctr.p + 2, 0, 4, h->stream);
  max_deg_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(deg.p, (uint32_t)n, ctr.p + 2);
  h->launches++;
  cudaMemcpyAsync(&width, ctr.p + 2, 4, cudaMemcpyDeviceToHost, h->stream);
  e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) { cudaFree(graph); return set_err(h, B200ANN_ERR_CUDA, std::string("nsg_finish: ") + cudaGetErrorString(e)); }
  if (ix->graph) cudaFree(ix->graph);
  ix->graph = graph; ix->gstride = gs; ix->ep = ep;
  ix->width = width > R ? width : (uint32_t)R;  // sync_prune sets width = R (:310), tree_grow only raises it
  if (added_out) *added_out = added;
  return 0;
}

int nsg_centroid_launch(b200ann_handle *h, const b200ann_index *ix, float *center) {
  centroid_kernel<<<(unsigned)((ix->dim + 7) / 8), 288, 0, h->stream>>>(ix->vec, ix->stride, (uint32_t)ix->n,
                                                                      (uint32_t)ix->dim, center);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

// ================================================================= C ABI (include/b200ann.h)
namespace {
// host restatement of beam::seed_hash (same constants) for the default ep0
inline uint32_t host_mix32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
inline uint32_t host_seed_hash(uint32_t seed, uint32_t query, uint32_t counter) {
  return host_mix32(host_mix32(seed ^ (query * 0x9E3779B9U)) + counter * 0x85EBCA6BU);
}

struct PhaseTimer {
  b200ann_handle *h;
  explicit PhaseTimer(b200ann_handle *hh) : h(hh) { cudaEventRecord(h->ev0, h->stream); }
  cudaError_t finish() {
    cudaEventRecord(h->ev1, h->stream);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e == cudaSuccess) cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return e;
  }
};

int check_build_args(b200ann_handle *h, b200ann_index *ix, const uint32_t *knn, size_t kstride, size_t L, size_t R,
                     size_t C) {
  if (!knn || kstride == 0) return set_err(h, B200ANN_ERR_ARG, "nsg: knn graph missing");
  if (L == 0 || L > ix->n) return set_err(h, B200ANN_ERR_ARG, "nsg: need 0 < L <= n");
  if (R == 0 || C == 0 || C > 8192) return set_err(h, B200ANN_ERR_ARG, "nsg: need R > 0 and 0 < C <= 8192");
  return 0;
}

// init_graph on device pointers; ep_out is a host value (synchronises)
int init_ep_dev(b200ann_handle *h, b200ann_index *ix, const uint32_t *knn_d, size_t kstride, size_t L, uint32_t seed,
                uint32_t ep0, uint32_t *ep_out, float *center_dev) {
  if (ep0 == INVALID_ID) ep0 = host_seed_hash(seed, 0xFFFFFFFFu, 0) % (uint32_t)ix->n;
  if (ep0 >= ix->n) return set_err(h, B200ANN_ERR_ARG, "nsg_init_ep: ep0 >= n");
  int rc = nsg_centroid_launch(h, ix, center_dev);
  if (rc) return rc;
  b200ann_index view = *ix;
  view.graph = const_cast<uint32_t *>(knn_d); view.gstride = kstride; view.ep = ep0;
  DevBuf<uint32_t> out;
  CU_CHECK(h, out.alloc(1, h->stream));
  rc = launch_nsg_search(h, &view, center_dev, 1, ix->dim, 1, L, seed, out.p, nullptr, nullptr, (uint32_t)ix->n);
  if (rc) return rc;
  CU_CHECK(h, cudaMemcpyAsync(ep_out, out.p, 4, cudaMemcpyDeviceToHost, h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  return 0;
}
}  // namespace


extern "C" {

int b200ann_set_link_visited_bits(b200ann_handle_t h, int bits) {
  if (!h) return B200ANN_ERR_ARG;
  if (bits != 0 && (bits < 5 || bits > 14)) return set_err(h, B200ANN_ERR_ARG, "link_visited_bits must be 0 (auto) or in [5,14]");
  h->link_visited_bits = bits;
  return 0;
}

int b200ann_nsg_init_ep(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *knn, size_t kstride, size_t L,
                        uint32_t seed, uint32_t ep0, uint32_t *ep_out, float *center_out) {
  FLOAT_ONLY(h, ix, "nsg_init_ep");
  if (!h || !ix || !ep_out) return B200ANN_ERR_ARG;
  int rc = check_build_args(h, ix, knn, kstride, L, 1, 1);
  if (rc) return rc;
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<uint32_t> kin;
  CU_CHECK(h, kin.stage(knn, ix->n * kstride, h->stream));
  DevBuf<float> center;
  CU_CHECK(h, center.alloc(ix->dim, h->stream));
  rc = init_ep_dev(h, ix, kin.d, kstride, L, seed, ep0, ep_out, center.p);
  if (rc) return rc;
  if (center_out) {
    CU_CHECK(h, cudaMemcpyAsync(center_out, center.p, ix->dim * 4, cudaMemcpyDefault, h->stream));
    CU_CHECK(h, cudaStreamSynchronize(h->stream));
  }
  return 0;
}

int b200ann_nsg_link(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *knn, size_t kstride, uint32_t ep, size_t L,
                     size_t R, size_t C, uint32_t seed, uint32_t *cut_ids, float *cut_d, uint32_t *stats) {
  FLOAT_ONLY(h, ix, "nsg_link");
  if (!h || !ix) return B200ANN_ERR_ARG;
  int rc = check_build_args(h, ix, knn, kstride, L, R, C);
  if (rc) return rc;
  if (!cut_ids || !cut_d || ep >= ix->n) return set_err(h, B200ANN_ERR_ARG, "nsg_link: need cut_ids, cut_d, ep < n");
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<uint32_t> kin; OutArg<uint32_t> oi; OutArg<float> od; OutArg<uint32_t> os;
  CU_CHECK(h, kin.stage(knn, ix->n * kstride, h->stream));
  CU_CHECK(h, oi.stage(cut_ids, ix->n * R, h->stream));
  CU_CHECK(h, od.stage(cut_d, ix->n * R, h->stream));
  CU_CHECK(h, os.stage(stats, ix->n * 3, h->stream));
  PhaseTimer t(h);
  rc = nsg_link_launch(h, ix, kin.d, kstride, ep, L, R, C, seed, oi.d, od.d, os.d);
  if (rc) return rc;
  CU_CHECK(h, t.finish());
  CU_CHECK(h, oi.finish(h->stream)); CU_CHECK(h, od.finish(h->stream)); CU_CHECK(h, os.finish(h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  return 0;
}

int b200ann_nsg_interinsert(b200ann_handle_t h, b200ann_index_t ix, size_t R, uint32_t *cut_ids, float *cut_d,
                            uint64_t *n_eval) {
  FLOAT_ONLY(h, ix, "nsg_interinsert");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!cut_ids || !cut_d || R == 0) return set_err(h, B200ANN_ERR_ARG, "nsg_interinsert: need cut_ids, cut_d, R > 0");
  CU_CHECK(h, cudaSetDevice(h->device));
  const size_t total = ix->n * R;
  const bool dev = is_device_ptr(cut_ids);
  DevBuf<uint32_t> bi; DevBuf<float> bd; DevBuf<unsigned long long> ne;
  uint32_t *ci = cut_ids; float *cd = cut_d;
  if (!dev) {
    CU_CHECK(h, bi.alloc(total, h->stream)); CU_CHECK(h, bd.alloc(total, h->stream));
    CU_CHECK(h, cudaMemcpyAsync(bi.p, cut_ids, total * 4, cudaMemcpyHostToDevice, h->stream));
    CU_CHECK(h, cudaMemcpyAsync(bd.p, cut_d, total * 4, cudaMemcpyHostToDevice, h->stream));
    ci = bi.p; cd = bd.p;
  }
  CU_CHECK(h, ne.alloc(1, h->stream));
  CU_CHECK(h, cudaMemsetAsync(ne.p, 0, 8, h->stream));
  PhaseTimer t(h);
  int rc = nsg_interinsert_launch(h, ix, R, ci, cd, ne.p);
  if (rc) return rc;
  CU_CHECK(h, t.finish());
  if (!dev) {
    CU_CHECK(h, cudaMemcpyAsync(cut_ids, ci, total * 4, cudaMemcpyDeviceToHost, h->stream));
    CU_CHECK(h, cudaMemcpyAsync(cut_d, cd, total * 4, cudaMemcpyDeviceToHost, h->stream));
  }
  unsigned long long v = 0;
  CU_CHECK(h, cudaMemcpyAsync(&v, ne.p, 8, cudaMemcpyDeviceToHost, h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  if (n_eval) *n_eval = v;
  return 0;
}

int b200ann_nsg_finish(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *cut_ids, const float *cut_d, size_t R,
                       uint32_t ep, size_t L, uint32_t seed, size_t slack, uint32_t *added) {
  FLOAT_ONLY(h, ix, "nsg_finish");
  if (!h || !ix) return B200ANN_ERR_ARG;
  if (!cut_ids || !cut_d || R == 0 || ep >= ix->n || L == 0 || L > ix->n)
    return set_err(h, B200ANN_ERR_ARG, "nsg_finish: need cut_ids, cut_d, R > 0, ep < n, 0 < L <= n");
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<uint32_t> ci; InArg<float> cd;
  CU_CHECK(h, ci.stage(cut_ids, ix->n * R, h->stream));
  CU_CHECK(h, cd.stage(cut_d, ix->n * R, h->stream));
  return nsg_finish(h, ix, ci.d, cd.d, R, ep, L, seed, slack, added);
}

int b200ann_nsg_build(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *knn, size_t kstride, size_t L, size_t R,
                      size_t C, uint32_t seed, uint32_t ep0, uint64_t *info) {
  FLOAT_ONLY(h, ix, "nsg_build");
  if (!h || !ix) return B200ANN_ERR_ARG;
  int rc = check_build_args(h, ix, knn, kstride, L, R, C);
  if (rc) return rc;
  CU_CHECK(h, cudaSetDevice(h->device));
  const size_t n = ix->n;
  InArg<uint32_t> kin;
  CU_CHECK(h, kin.stage(knn, n * kstride, h->stream));
  DevBuf<float> center, cut_d; DevBuf<uint32_t> cut_ids, stats; DevBuf<unsigned long long> ne;
  CU_CHECK(h, center.alloc(ix->dim, h->stream));
  CU_CHECK(h, cut_ids.alloc(n * R, h->stream)); CU_CHECK(h, cut_d.alloc(n * R, h->stream));
  CU_CHECK(h, stats.alloc(n * 3, h->stream));
  CU_CHECK(h, ne.alloc(1, h->stream));
  CU_CHECK(h, cudaMemsetAsync(ne.p, 0, 8, h->stream));
  cudaEvent_t e0, e1;
  CU_CHECK(h, cudaEventCreate(&e0)); CU_CHECK(h, cudaEventCreate(&e1));
  cudaEventRecord(e0, h->stream);
  uint32_t ep = 0, added = 0;
  rc = init_ep_dev(h, ix, kin.d, kstride, L, seed, ep0, &ep, center.p);
  if (!rc) rc = nsg_link_launch(h, ix, kin.d, kstride, ep, L, R, C, seed, cut_ids.p, cut_d.p, stats.p);
  if (!rc) rc = nsg_interinsert_launch(h, ix, R, cut_ids.p, cut_d.p, ne.p);
  if (!rc) rc = nsg_finish(h, ix, cut_ids.p, cut_d.p, R, ep, L, seed, 14, &added);
  cudaEventRecord(e1, h->stream);
  cudaError_t ce = cudaStreamSynchronize(h->stream);
  if (!rc && ce == cudaSuccess) cudaEventElapsedTime(&h->last_ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (rc) return rc;
  CU_CHECK(h, ce);
  if (info) {
    // column sums of stats on the host (n*12 bytes once per build)
    std::vector<uint32_t> st(n * 3);
    CU_CHECK(h, cudaMemcpy(st.data(), stats.p, n * 12, cudaMemcpyDeviceToHost));
    unsigned long long v = 0;
    CU_CHECK(h, cudaMemcpy(&v, ne.p, 8, cudaMemcpyDeviceToHost));
    uint64_t s0 = 0, s1 = 0, s2 = 0;
    for (size_t i = 0; i < n; i++) { s0 += st[3 * i]; s1 += st[3 * i + 1]; s2 += st[3 * i + 2]; }
    info[0] = ep; info[1] = ix->width; info[2] = added; info[3] = s0; info[4] = s1; info[5] = s2; info[6] = v; info[7] = 0;
  }
  return 0;
}

}  // extern "C"

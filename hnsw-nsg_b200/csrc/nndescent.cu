// NN-descent kNN-graph construction on the GPU.
//
// Replaces efanna2e::IndexGraph::Build with the IndexRandom initializer
// (CNNS/efanna_graph/src/index_graph.cpp:265-294 = InitializeGraph :175-201 + NNDescent :122-136 =
// iter x { join :22-33 (nhood::join / nhood::insert, efanna2e/neighbor.h:57-87), update :34-120 } + the
// top-K extraction :275-291; index_random.cpp:24-27, util.h:17-31 GenRandom).
//
// The reference is a lock-per-node, order-dependent CPU loop. Here every phase is a set function of its
// inputs ("batch form", restated serially by oracle/oracle.c orc_nndescent_batch and compared bit for bit):
//   pool        u64 [n, L] sorted ascending, key = distance bits << 32 | id << 1 | flag (L2 distances are
//               >= 0, so the raw bits order like the floats). A pool is the set of the <= L smallest keys it
//               has been offered (dedup by id); an offer is admitted only if distance <= the pool's largest
//               distance at the start of the join (nhood::insert's heap-top test, neighbor.h:59).
//   join        one CTA per node: the sampled rows are gathered into shared memory in 64 x 64 tiles and all
//               new x new (each unordered pair once) and new x old distances are computed in fp32, k
//               sequential with fmaf (oracle mode L2_SEQ_FMA); pairs that pass the destination's threshold
//               go through a shared-memory queue to warp-cooperative sorted inserts under a per-node lock.
//   update      M / new / old sampling exactly as :49-98; the reverse lists keep the R offers with the smallest
//               (hash(seed_it, node, source), source) instead of "first R then overwrite slot rand() % R"
//               (:78-82), through the same sorted-insert primitive; `pool.back()` (:76,:88) follows the
//               reference's one-thread schedule: the largest distance when the other node's id is larger,
//               the std::make_heap leaf when it is smaller (leaf_rank table).
#include <algorithm>
#include <vector>
#include "common.cuh"

namespace nnd {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
// same stream as beam::seed_hash / oracle orc_seed_hash
__device__ __forceinline__ uint32_t seed_hash(uint32_t seed, uint32_t query, uint32_t counter) {
  return mix32(mix32(seed ^ (query * 0x9E3779B9U)) + counter * 0x85EBCA6BU);
}

constexpr uint32_t SEED_POOL = 0x68E31DA4U;
constexpr int MAX_GEN = 128;   // 2S and S+1 must fit
constexpr uint64_t EMPTY = ~0ull;

struct State {
  const float *vec; size_t stride; uint32_t n, dim;
  uint32_t L, S, R, K;
  uint32_t cap_fnew, cap_fold;          // 2S, min(L, 2R)
  uint64_t *pool;                       // [n, L]
  uint32_t *psize, *M, *worst, *leafd;  // [n]
  uint32_t *fwd_new, *fwd_old;          // [n, cap_fnew], [n, cap_fold]
  uint32_t *n_fnew, *n_fold;            // [n]
  uint64_t *rnn_new, *rnn_old;          // [n, R]
  uint32_t *n_rnew, *n_rold;            // [n]
  uint32_t *lock_pool, *lock_rnew, *lock_rold;  // [n]
  const uint32_t *leaf_rank;            // [L+1]
  // lock-free join (two passes over the same tiles): offers are counted per destination, then written to the
  // destination's segment of one buffer, then merged by one warp per destination (merge_offers_kernel)
  uint32_t *ocnt, *ofill;               // [n] offers per destination this iteration / write cursors
  const unsigned long long *ooff;       // [n + 1] exclusive scan of ocnt
  uint64_t *obuf;                       // [ooff[n]] offer keys, grouped by destination
  ulonglong2 *rec;                      // MODE 3: (key, destination) records in arrival order, scattered by scatter_offers_kernel
  unsigned long long *rec_cursor;       // MODE 3: records written (counts past rec_cap too: the host re-runs the pass then)
  unsigned long long rec_cap;           // MODE 3: records the buffer holds
  const uint32_t *init_graph;           // [n, init_k] or NULL: RefineGraph (index_graph.cpp:203-262) - the pools start from these rows
  uint32_t init_k;
  unsigned long long *stats;            // {distance evaluations, offers queued, inserts done, next node}
};

// efanna2e::GenRandom (util.h:17-31) over the stateless stream seed_hash(seed, node, draw)
__device__ void gen_random(uint32_t seed, uint32_t node, uint32_t *addr, uint32_t size, uint32_t N) {
  uint32_t ctr = 0;
  for (uint32_t i = 0; i < size; ++i) {
    uint32_t v = seed_hash(seed, node, ctr++) % (N - size);
    uint32_t j = i;  // insertion sort == std::sort on integers
    while (j > 0 && addr[j - 1] > v) { addr[j] = addr[j - 1]; --j; }
    addr[j] = v;
  }
  for (uint32_t i = 1; i < size; ++i)
    if (addr[i] <= addr[i - 1]) addr[i] = addr[i - 1] + 1;
  uint32_t off = seed_hash(seed, node, ctr++) % N;
  for (uint32_t i = 0; i < size; ++i) addr[i] = (addr[i] + off) % N;
}

// sequential-over-k fmaf order (oracle ORC_L2_SEQ_FMA), one thread
__device__ float seq_l2(const float *a, const float *b, uint32_t dim) {
  float acc = 0.f;
  for (uint32_t k = 0; k < dim; ++k) { float t = a[k] - b[k]; acc = fmaf(t, t, acc); }
  return acc;
}

// InitializeGraph (:175-201): nn_new = 2S random ids (nhood ctor), pool = first S of S+1 random ids minus self
__global__ void init_kernel(State s, uint32_t seed) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.n) return;
  uint32_t tmp[MAX_GEN];
  gen_random(seed, i, tmp, 2 * s.S, s.n);
  for (uint32_t j = 0; j < 2 * s.S; ++j) s.fwd_new[(size_t)i * s.cap_fnew + j] = tmp[j];
  s.n_fnew[i] = 2 * s.S;
  s.n_fold[i] = 0; s.n_rnew[i] = 0; s.n_rold[i] = 0;
  s.M[i] = s.S;
  s.lock_pool[i] = 0; s.lock_rnew[i] = 0; s.lock_rold[i] = 0;
  uint64_t *pool = s.pool + (size_t)i * s.L;
  uint32_t sz = 0;
  unsigned long long evals = 0;
  if (sz == 0) {
    gen_random(seed ^ SEED_POOL, i, tmp, s.S + 1, s.n);
    for (uint32_t j = 0; j < s.S; ++j) {
      uint32_t id = tmp[j];
      if (id == i) continue;
      float d = seq_l2(s.vec + (size_t)i * s.stride, s.vec + (size_t)id * s.stride, s.dim);
      ++evals;
      uint64_t key = ((uint64_t)__float_as_uint(d) << 32) | (id << 1) | 1u;
      uint32_t p = sz;
      while (p > 0 && pool[p - 1] > key) { pool[p] = pool[p - 1]; --p; }
      pool[p] = key;
      ++sz;
    }
  }
  for (uint32_t j = sz; j < s.L; ++j) pool[j] = EMPTY;
  s.psize[i] = sz;
  s.worst[i] = (uint32_t)(pool[sz - 1] >> 32);
  atomicAdd(s.stats + 0, evals);
}

// RefineGraph's start (InitializeGraph_Refine, index_graph.cpp:203-232), one warp per node, run AFTER init_kernel: the
// pool becomes the node's row of the initializer's graph - self, out-of-range ids and duplicates dropped, every entry
// flagged new, the L smallest (distance, id) kept in ascending order. A row with no usable entry keeps the random start.
constexpr int RF_WARPS = 4, RF_MAXK = 512;
__global__ void __launch_bounds__(RF_WARPS * 32) init_refine_kernel(State s) {
  __shared__ uint64_t keys_s[RF_WARPS][RF_MAXK];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t i = blockIdx.x * RF_WARPS + warp;
  if (i >= s.n) return;
  uint64_t *keys = keys_s[warp];
  const uint32_t k = s.init_k;
  uint32_t evals = 0;
  for (uint32_t j0 = 0; j0 < k; j0 += 32) {
    const uint32_t j = j0 + lane;
    uint64_t key = EMPTY;
    if (j < k) {
      const uint32_t id = s.init_graph[(size_t)i * k + j];
      if (id != 0xFFFFFFFFu && id != i && id < s.n) {
        const float d = seq_l2(s.vec + (size_t)i * s.stride, s.vec + (size_t)id * s.stride, s.dim);
        key = ((uint64_t)__float_as_uint(d) << 32) | (id << 1) | 1u;
        ++evals;
      }
      keys[j] = key;
    }
  }
  __syncwarp();
  // duplicates: the same id twice gives the same key; only its first occurrence stays
  bool dupf[RF_MAXK / 32];
#pragma unroll
  for (int t = 0; t < RF_MAXK / 32; ++t) {
    const uint32_t j = t * 32 + lane;
    dupf[t] = false;
    if (j < k && keys[j] != EMPTY)
      for (uint32_t u = 0; u < j; ++u) dupf[t] |= keys[u] == keys[j];
  }
  __syncwarp();
#pragma unroll
  for (int t = 0; t < RF_MAXK / 32; ++t) {
    const uint32_t j = t * 32 + lane;
    if (j < k && dupf[t]) keys[j] = EMPTY;
  }
  __syncwarp();
  uint32_t cnt = 0;
  for (uint32_t j0 = 0; j0 < k; j0 += 32) {
    const uint32_t j = j0 + lane;
    cnt += __popc(__ballot_sync(0xffffffffu, j < k && keys[j] != EMPTY));
  }
  evals = __reduce_add_sync(0xffffffffu, evals);
  if (cnt == 0) return;   // the random start of init_kernel stays
  uint64_t *pool = s.pool + (size_t)i * s.L;
  for (uint32_t j = lane; j < s.L; j += 32) pool[j] = EMPTY;
  __syncwarp();
  for (uint32_t j = lane; j < k; j += 32) {
    const uint64_t key = keys[j];
    if (key == EMPTY) continue;
    uint32_t r = 0;
    for (uint32_t u = 0; u < k; ++u) r += keys[u] < key;   // EMPTY is the largest value: never counted
    if (r < s.L) pool[r] = key;
  }
  __syncwarp();
  if (lane == 0) {
    const uint32_t sz = cnt < s.L ? cnt : s.L;
    // the random start's evaluations were counted by init_kernel; the reference evaluates only the initializer's rows
    const unsigned long long old = s.psize[i];
    s.psize[i] = sz;
    s.worst[i] = (uint32_t)(pool[sz - 1] >> 32);
    atomicAdd(s.stats + 0, (unsigned long long)evals - old);
  }
}

// ---- warp-cooperative "keep the cap smallest keys, sorted" under a per-list lock -------------------------
// All list accesses go through L2 (.cg) because other SMs update the same lists.
template <int NT>
__device__ bool warp_topl_insert(uint64_t *list, uint32_t *size_p, uint32_t *lock, uint32_t cap, uint64_t key,
                                 bool dedup, int lane) {
  // racy prefilter: the last slot only ever decreases (EMPTY while the list is not full)
  if (__ldcg((const unsigned long long *)(list + cap - 1)) <= key) return false;
  if (lane == 0) {
    while (atomicCAS(lock, 0u, 1u) != 0u) __nanosleep(32);
    __threadfence();
  }
  __syncwarp();
  uint32_t size = __ldcg(size_p);
  uint64_t e[NT];
  const uint32_t id = (uint32_t)key >> 1;
  int less = 0;
  bool dup = false;
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    uint32_t idx = t * 32 + lane;
    e[t] = idx < size ? __ldcg((const unsigned long long *)(list + idx)) : EMPTY;
    if (idx < size) {
      less += e[t] < key;
      dup |= dedup && (((uint32_t)e[t] >> 1) == id);
    }
  }
  uint32_t pos = __reduce_add_sync(0xffffffffu, less);
  bool ok = !__any_sync(0xffffffffu, dup) && pos < cap;
  if (ok) {
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      uint32_t idx = t * 32 + lane;
      if (idx >= pos && idx < size && idx + 1 < cap) __stcg((unsigned long long *)(list + idx + 1), e[t]);
    }
    if (lane == 0) {
      __stcg((unsigned long long *)(list + pos), key);
      if (size < cap) __stcg(size_p, size + 1);
    }
  }
  __threadfence();
  __syncwarp();
  if (lane == 0) atomicExch(lock, 0u);
  return ok;
}

__device__ __forceinline__ bool topl_insert_dispatch(uint64_t *list, uint32_t *size_p, uint32_t *lock, uint32_t cap,
                                                     uint64_t key, bool dedup, int lane) {
  if (cap <= 64) return warp_topl_insert<2>(list, size_p, lock, cap, key, dedup, lane);
  if (cap <= 128) return warp_topl_insert<4>(list, size_p, lock, cap, key, dedup, lane);
  if (cap <= 256) return warp_topl_insert<8>(list, size_p, lock, cap, key, dedup, lane);
  return warp_topl_insert<16>(list, size_p, lock, cap, key, dedup, lane);
}

// ---- join (:22-33) ---------------------------------------------------------------------------------------
constexpr int JT = 256;          // threads per CTA
constexpr int TILE = 64;         // pairs tile TILE x TILE, 4 x 4 per thread
constexpr int KC = 64;           // floats of the dimension staged per pass
constexpr int ROWF = KC + 4;     // padded shared-memory row (bank shift of 4 per row, 16-byte aligned)
constexpr int DROW = TILE + 1;   // padded row of the tile's distance matrix (conflict-free column reads)
constexpr uint32_t NO_PAIR = 0x7f800000u;  // +inf: never <= a threshold (thresholds are finite)

__device__ __forceinline__ void load_rows(float *sm, const State &s, const uint32_t *ids, uint32_t pos0, uint32_t limit,
                                          uint32_t kc, uint32_t kw4, int tid) {
  // rows pos0 .. min(pos0 + TILE, limit): a warp covers one row's float4s (512 contiguous bytes at KC = 128),
  // four rows per warp in flight. Rows past `limit` are never read by a pair that counts.
  const int warp = tid >> 5, lane = tid & 31;
  const int rows = (int)(limit > pos0 ? (limit - pos0 < (uint32_t)TILE ? limit - pos0 : (uint32_t)TILE) : 0u);
  for (int r0 = warp * 4; r0 < rows; r0 += (JT / 32) * 4) {
    for (uint32_t c = lane; c < kw4; c += 32) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (r0 + u < rows)
          v[u] = __ldg(reinterpret_cast<const float4 *>(s.vec + (size_t)ids[pos0 + r0 + u] * s.stride + kc) + c);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (r0 + u < rows) reinterpret_cast<float4 *>(sm + (r0 + u) * ROWF)[c] = v[u];
    }
  }
}

#ifdef NND_PROF
__device__ unsigned long long g_prof[16];
#define MPROF(i) do { if (lane == 0) { long long now__ = clock64(); atomicAdd(&g_prof[i], (unsigned long long)(now__ - mt__)); mt__ = now__; } } while (0)
#define MPROF_START() long long mt__ = clock64()
#else
#define MPROF(i)
#define MPROF_START()
#endif
// All offers one join tile makes to ONE destination pool, merged under one lock acquisition: the pool is
// loaded once into registers (entry idx = t*32 + lane), every admitted candidate is inserted by a
// register shift (shuffles, no memory traffic), and the pool is written back once.
// c0/c1: this lane's two candidate distance bits (members mpos0 + lane and mpos0 + lane + 32 of the tile's
// other side), NO_PAIR where the pair does not exist. Returns the number of inserts; *new_last = the pool's
// largest distance bits afterwards when it is full (else 0xFFFFFFFF).
template <int NT>
__device__ uint32_t warp_pool_merge(uint64_t *list, uint32_t *size_p, uint32_t *lock, uint32_t cap, const uint32_t (&c)[4],
                                    const uint32_t (&src)[4], uint32_t thr, int lane, uint32_t *new_last) {
  // lane-major registers: lane holds entries [lane*NT, lane*NT + NT) -> an insert shifts inside the lanes and
  // passes ONE entry to the next lane (one 64-bit shuffle per candidate)
  uint32_t mk[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) mk[g] = __ballot_sync(0xffffffffu, c[g] <= thr);
  *new_last = 0xFFFFFFFFu;
  if (!(mk[0] | mk[1] | mk[2] | mk[3])) return 0;
  MPROF_START();
  if (lane == 0) {
    while (atomicCAS(lock, 0u, 1u) != 0u) __nanosleep(32);
    __threadfence();
  }
  __syncwarp();
  MPROF(0);
  uint32_t size = __ldcg(size_p);
  uint64_t e[NT];
  const uint32_t base = lane * NT;
  const bool vec = base + NT <= cap && (reinterpret_cast<uintptr_t>(list) & 15) == 0;   // 16-byte accesses
  if (vec) {                                    // slots past `size` hold EMPTY: no dependence on the size load
#pragma unroll
    for (int t = 0; t < NT; t += 2) {
      ulonglong2 v = __ldcg(reinterpret_cast<const ulonglong2 *>(list + base + t));
      e[t] = v.x; e[t + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int t = 0; t < NT; ++t) e[t] = base + t < cap ? __ldcg((const unsigned long long *)(list + base + t)) : EMPTY;
  }
  uint32_t inserted = 0;
#ifdef NND_PROF
  { unsigned long long sink = 0;
    for (int t = 0; t < NT; ++t) sink += e[t];
    if (sink == 0x1234567ull) g_prof[15] = 1; }
#endif
  MPROF(1);
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint32_t m = mk[g];
#pragma unroll 1
    while (m) {
      int srcl = __ffs(m) - 1;
      m &= m - 1;
      uint32_t db = __shfl_sync(0xffffffffu, c[g], srcl);
      uint32_t sid = __shfl_sync(0xffffffffu, src[g], srcl);
      uint64_t key = ((uint64_t)db << 32) | (sid << 1) | 1u;
      int less = 0;
      bool dup = false;
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        less += e[t] < key;                       // EMPTY slots are never < key
        dup |= ((uint32_t)e[t] >> 1) == sid && e[t] != EMPTY;
      }
      uint32_t pos = __reduce_add_sync(0xffffffffu, less);
      if (__any_sync(0xffffffffu, dup) || pos >= cap) continue;
      uint64_t carry = __shfl_up_sync(0xffffffffu, e[NT - 1], 1);   // lane-1's last entry (before the shift)
      if (base + NT > pos) {                      // this lane holds pos or entries after it
        const int p = pos > base ? (int)(pos - base) : -1;          // local insert slot; -1: whole lane shifts
#pragma unroll
        for (int t = NT - 1; t >= 1; --t) e[t] = t > p ? e[t - 1] : e[t];
        if (p < 0) e[0] = carry;
#pragma unroll
        for (int t = 0; t < NT; ++t)
          if (t == p) e[t] = key;
        if (pos == base) e[0] = key;
      }
#pragma unroll
      for (int t = 0; t < NT; ++t)
        if (base + t >= cap) e[t] = EMPTY;        // the entry pushed out of a full pool is dropped
      if (size < cap) ++size;
      ++inserted;
    }
  }
  MPROF(2);
  if (inserted) {
    if (vec) {
#pragma unroll
      for (int t = 0; t < NT; t += 2) {
        if (base + t < size) __stcg(reinterpret_cast<ulonglong2 *>(list + base + t), make_ulonglong2(e[t], e[t + 1]));
      }
    } else {
#pragma unroll
      for (int t = 0; t < NT; ++t)
        if (base + t < size) __stcg((unsigned long long *)(list + base + t), e[t]);
    }
    if (lane == 0) __stcg(size_p, size);
  }
  if (size == cap) {
    const uint32_t lidx = cap - 1;
    uint64_t le = EMPTY;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      uint64_t v = __shfl_sync(0xffffffffu, e[t], lidx / NT);
      if (t == (int)(lidx % NT)) le = v;
    }
    *new_last = (uint32_t)(le >> 32);
  }
  __threadfence();
  __syncwarp();
  MPROF(3);
  if (lane == 0) atomicExch(lock, 0u);
#ifdef NND_PROF
  if (lane == 0) atomicAdd(&g_prof[4], 1ull);
#endif
  return inserted;
}

__device__ __forceinline__ uint32_t pool_merge_dispatch(uint64_t *list, uint32_t *size_p, uint32_t *lock, uint32_t cap,
                                                     const uint32_t (&c)[4], const uint32_t (&src)[4], uint32_t thr,
                                                     int lane, uint32_t *new_last) {
  if (cap <= 64) return warp_pool_merge<2>(list, size_p, lock, cap, c, src, thr, lane, new_last);
  if (cap <= 128) return warp_pool_merge<4>(list, size_p, lock, cap, c, src, thr, lane, new_last);
  if (cap <= 256) return warp_pool_merge<8>(list, size_p, lock, cap, c, src, thr, lane, new_last);
  return warp_pool_merge<16>(list, size_p, lock, cap, c, src, thr, lane, new_last);
}

// one k-chunk of a tile: NI x NJ 16-row sub-blocks per thread, k sequential per pair (fmaf), no branches
template <int NI, int NJ>
__device__ __forceinline__ void tile_fma(const float *As, const float *Bs, float (&acc)[4][4], uint32_t kw4, int ty, int tx) {
  const float4 *ap = reinterpret_cast<const float4 *>(As + ty * ROWF);
  const float4 *bp = reinterpret_cast<const float4 *>(Bs + tx * ROWF);
#pragma unroll 2
  for (uint32_t k4 = 0; k4 < kw4; ++k4) {
    float4 a[NI], b[NJ];
#pragma unroll
    for (int i = 0; i < NI; ++i) a[i] = ap[i * 16 * (ROWF / 4) + k4];
#pragma unroll
    for (int j = 0; j < NJ; ++j) b[j] = bp[j * 16 * (ROWF / 4) + k4];
#pragma unroll
    for (int i = 0; i < NI; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float t = a[i].x - b[j].x; acc[i][j] = fmaf(t, t, acc[i][j]);
      }
#pragma unroll
    for (int i = 0; i < NI; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float t = a[i].y - b[j].y; acc[i][j] = fmaf(t, t, acc[i][j]);
      }
#pragma unroll
    for (int i = 0; i < NI; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float t = a[i].z - b[j].z; acc[i][j] = fmaf(t, t, acc[i][j]);
      }
#pragma unroll
    for (int i = 0; i < NI; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float t = a[i].w - b[j].w; acc[i][j] = fmaf(t, t, acc[i][j]);
      }
  }
}

// MODE 0: offers are merged into the pools on the spot under per-destination locks (round 1). MODE 1 / 2: the lock-free
// form, pass 1 counts the offers per destination, pass 2 writes them; the pools are not touched by either, so the
// admission threshold is exactly the pool's largest distance at the start of the join (oracle.c: db <= oldmax).
template <int MODE>
__global__ void __launch_bounds__(JT, 3) join_kernel(State s, uint32_t cn_max) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float *As = reinterpret_cast<float *>(smem_raw);
  float *Bs = As + TILE * ROWF;
  uint32_t *Dm = reinterpret_cast<uint32_t *>(Bs + TILE * ROWF);   // [TILE][DROW] distance bits of the tile
  uint32_t *ids = Dm + TILE * DROW;
  uint32_t *thr = ids + cn_max;
  __shared__ uint32_t sh_node, sh_nn, sh_no;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ty = tid >> 4, tx = tid & 15;
  const uint32_t cap_old = 2 * s.R;
  unsigned long long my_dist = 0, my_offer = 0, my_ins = 0;
#ifdef NND_PROF
  long long t_fetch = 0, t_tile = 0, t_drain = 0, t_gather = 0, t_mark;
  unsigned long long sum_nn = 0, sum_no = 0, sum_tiles = 0;
#define PROF_MARK() t_mark = clock64()
#define PROF_ADD(x) do { long long now__ = clock64(); x += now__ - t_mark; t_mark = now__; } while (0)
#else
#define PROF_MARK()
#define PROF_ADD(x)
#endif

  for (;;) {
    PROF_MARK();
    if (tid == 0) {
      uint32_t node = (uint32_t)atomicAdd(s.stats + 3, 1ull);
      sh_node = node;
      if (node < s.n) {
        uint32_t no = s.n_fold[node] + s.n_rold[node];
        sh_nn = s.n_fnew[node] + s.n_rnew[node];
        sh_no = no < cap_old ? no : cap_old;
      }
    }
    __syncthreads();
    const uint32_t node = sh_node;
    if (node >= s.n) break;
    const uint32_t nn = sh_nn, no = sh_no, ct = nn + no;
    {  // member ids: new = fwd_new ++ rnn_new, old = (fwd_old ++ rnn_old)[:2R]; thresholds
      const uint32_t nfn = s.n_fnew[node], nfo = s.n_fold[node];
      for (uint32_t p = tid; p < ct; p += JT) {
        uint32_t id;
        if (p < nn) id = p < nfn ? s.fwd_new[(size_t)node * s.cap_fnew + p] : (uint32_t)s.rnn_new[(size_t)node * s.R + (p - nfn)];
        else {
          uint32_t q = p - nn;
          id = q < nfo ? s.fwd_old[(size_t)node * s.cap_fold + q] : (uint32_t)s.rnn_old[(size_t)node * s.R + (q - nfo)];
        }
        ids[p] = id;
        uint32_t last = (uint32_t)(__ldcg((const unsigned long long *)(s.pool + (size_t)id * s.L + s.L - 1)) >> 32);
        uint32_t w = s.worst[id];
        thr[p] = last < w ? last : w;
      }
    }
    __syncthreads();
    PROF_ADD(t_fetch);
    const uint32_t a_tiles = (nn + TILE - 1) / TILE, b_tiles = (ct + TILE - 1) / TILE;
#ifdef NND_PROF
    sum_nn += nn; sum_no += no;
    for (uint32_t a_ = 0; a_ < a_tiles; ++a_) sum_tiles += b_tiles - a_;
#endif
    const uint32_t kchunks = ((uint32_t)s.stride + KC - 1) / KC;
    for (uint32_t at = 0; at < a_tiles; ++at) {
      if (kchunks == 1) load_rows(As, s, ids, at * TILE, nn, 0, (uint32_t)s.stride / 4, tid);
      for (uint32_t bt = at; bt < b_tiles; ++bt) {
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        // 16-row sub-blocks of the tile that hold members (short sample lists leave most of a tile empty)
        const int ni = (int)(((nn - at * TILE < (uint32_t)TILE ? nn - at * TILE : (uint32_t)TILE) + 15) / 16);
        const int nj = (int)(((ct - bt * TILE < (uint32_t)TILE ? ct - bt * TILE : (uint32_t)TILE) + 15) / 16);
        for (uint32_t kcn = 0; kcn < kchunks; ++kcn) {
          const uint32_t kc = kcn * KC;
          const uint32_t kw4 = (((uint32_t)s.stride - kc) < (uint32_t)KC ? ((uint32_t)s.stride - kc) : (uint32_t)KC) / 4;
          if (kchunks > 1) load_rows(As, s, ids, at * TILE, nn, kc, kw4, tid);
          load_rows(Bs, s, ids, bt * TILE, ct, kc, kw4, tid);
          __syncthreads();
#ifdef NND_PROF
          PROF_ADD(t_gather);
#endif
          // 16-row sub-blocks that hold members, rounded up to 1, 2 or 4 (short sample lists leave most of a tile empty)
          if (ni <= 1) {
            if (nj <= 1) tile_fma<1, 1>(As, Bs, acc, kw4, ty, tx);
            else if (nj <= 2) tile_fma<1, 2>(As, Bs, acc, kw4, ty, tx);
            else tile_fma<1, 4>(As, Bs, acc, kw4, ty, tx);
          } else if (ni <= 2) {
            if (nj <= 2) tile_fma<2, 2>(As, Bs, acc, kw4, ty, tx);
            else tile_fma<2, 4>(As, Bs, acc, kw4, ty, tx);
          } else {
            tile_fma<4, 4>(As, Bs, acc, kw4, ty, tx);
          }
          __syncthreads();
        }
        // ---- the tile's distance matrix: row = A member, column = B member; NO_PAIR where no pair exists
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t ra = ty + 16 * i, rb = tx + 16 * j;
            uint32_t pa = at * TILE + ra, pb = bt * TILE + rb;
            bool pair = pa < nn && pb < ct && (pb >= nn || pb > pa) && ids[pa] != ids[pb];
            my_dist += pair;
            Dm[ra * DROW + rb] = pair ? __float_as_uint(acc[i][j]) : NO_PAIR;
          }
        __syncthreads();
        PROF_ADD(t_tile);
        // ---- offers, one destination per warp step: the A members (rows of Dm) and the B members (columns).
        // On the diagonal tile member r is both row r and column r: one merge takes both.
        const bool diag = at == bt;
        for (uint32_t dsel = warp; dsel < (diag ? TILE : 2 * TILE); dsel += JT / 32) {
          const bool to_a = dsel < TILE;
          const uint32_t r = to_a ? dsel : dsel - TILE;
          const uint32_t pd = (to_a ? at : bt) * TILE + r;           // destination member position
          if (pd >= (to_a && !diag ? nn : ct)) continue;
          const uint32_t obase = (to_a ? bt : at) * TILE;            // the other side's first position
          uint32_t c[4], src[4];
          c[0] = c[1] = c[2] = c[3] = NO_PAIR;
          if (to_a) { c[0] = Dm[r * DROW + lane]; c[1] = Dm[r * DROW + lane + 32]; }
          if (!to_a || diag) { c[2] = Dm[lane * DROW + r]; c[3] = Dm[(lane + 32) * DROW + r]; }
          const uint32_t t = thr[pd];
          if (!__any_sync(0xffffffffu, c[0] <= t || c[1] <= t || c[2] <= t || c[3] <= t)) continue;
          const uint32_t p0 = obase + lane, p1 = obase + lane + 32;
          src[0] = c[0] != NO_PAIR ? ids[p0] : 0u; src[1] = c[1] != NO_PAIR ? ids[p1] : 0u;
          src[2] = c[2] != NO_PAIR ? ids[p0] : 0u; src[3] = c[3] != NO_PAIR ? ids[p1] : 0u;
          const uint32_t dst = ids[pd];
          my_offer += (c[0] <= t) + (c[1] <= t) + (c[2] <= t) + (c[3] <= t);
          if (MODE == 0) {
            uint32_t new_last;
            uint32_t ins = pool_merge_dispatch(s.pool + (size_t)dst * s.L, s.psize + dst, s.lock_pool + dst, s.L, c, src, t,
                                               lane, &new_last);
            if (lane == 0) {
              my_ins += ins;
              if (new_last < t) thr[pd] = new_last;                  // tighter threshold for the next tiles
            }
          } else {
            // this warp step's offers to dst: one atomic for all of them
            uint32_t rank[4], total = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const unsigned m = __ballot_sync(0xffffffffu, c[k] <= t);
              rank[k] = total + __popc(m & ((1u << lane) - 1));
              total += __popc(m);
            }
            uint32_t base = 0;
            if (MODE == 3) {  // one pass: count per destination and append (key, destination) records
              unsigned long long cur = 0;
              if (lane == 0) { atomicAdd(s.ocnt + dst, total); cur = atomicAdd(s.rec_cursor, (unsigned long long)total); }
              cur = __shfl_sync(0xffffffffu, cur, 0);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (c[k] <= t && cur + rank[k] < s.rec_cap) s.rec[cur + rank[k]] = make_ulonglong2(((uint64_t)c[k] << 32) | ((uint64_t)src[k] << 1) | 1ull, dst);
              continue;
            }
            if (lane == 0) base = atomicAdd((MODE == 1 ? s.ocnt : s.ofill) + dst, total);
            if (MODE == 2) {
              base = __shfl_sync(0xffffffffu, base, 0);
              uint64_t *seg = s.obuf + s.ooff[dst] + base;
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (c[k] <= t) seg[rank[k]] = ((uint64_t)c[k] << 32) | ((uint64_t)src[k] << 1) | 1ull;
            }
          }
        }
        __syncthreads();
        PROF_ADD(t_drain);
      }
    }
  }
  if (MODE != 2) {  // the second pass of the lock-free form repeats the first one's work
    if (my_dist) atomicAdd(s.stats + 0, my_dist);
    if (my_offer) atomicAdd(s.stats + 1, my_offer);
    if (my_ins) atomicAdd(s.stats + 2, my_ins);
  }
#ifdef NND_PROF
  if (tid == 0) {
    atomicAdd(s.stats + 8, sum_nn); atomicAdd(s.stats + 9, sum_no); atomicAdd(s.stats + 10, sum_tiles);
    atomicAdd(s.stats + 7, (unsigned long long)t_gather);
    atomicAdd(s.stats + 4, (unsigned long long)t_fetch);
    atomicAdd(s.stats + 5, (unsigned long long)t_tile);
    atomicAdd(s.stats + 6, (unsigned long long)t_drain);
  }
#endif
}

// ---- lock-free join, step 3: one warp per destination merges its offers into its pool -------------------------
// pool' = the L smallest keys of (pool + offers) with one entry per id; a pool entry wins over an offer of the same
// id (same distance, its flag bit sorts first) - nhood::insert rejects an id that is already there (neighbor.h:61-66).
constexpr int MG_WARPS = 4, MG_CAP = 1024, MG_CHUNK = 512;
__device__ __forceinline__ void warp_bitonic_u64(uint64_t *a, uint32_t ns, int lane) {
  for (uint32_t k = 2; k <= ns; k <<= 1)
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t x = lane; x < (ns >> 1); x += 32) {
        const uint32_t lo = ((x & ~(j - 1)) << 1) | (x & (j - 1)), hi = lo | j;
        const uint64_t vl = a[lo], vh = a[hi];
        const bool up = (lo & k) == 0;
        if (up ? vh < vl : vl < vh) { a[lo] = vh; a[hi] = vl; }
      }
      __syncwarp();
    }
}
__global__ void __launch_bounds__(MG_WARPS * 32) merge_offers_kernel(State s) {
  __shared__ uint64_t buf_s[MG_WARPS][MG_CAP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t d = blockIdx.x * MG_WARPS + warp;
  if (d >= s.n) return;
  const uint32_t m = s.ocnt[d];
  if (m == 0) return;
  uint64_t *buf = buf_s[warp];
  uint64_t *pool = s.pool + (size_t)d * s.L;
  uint32_t sz = s.psize[d];
  for (uint32_t j = lane; j < sz; j += 32) buf[j] = pool[j];
  __syncwarp();
  const uint64_t *off = s.obuf + s.ooff[d];
  for (uint32_t c0 = 0; c0 < m; c0 += MG_CHUNK) {
    // offers of this chunk that can still enter (below the current L-th key when the pool is full)
    const uint64_t bound = sz == s.L ? buf[s.L - 1] : EMPTY;
    uint32_t cnt = sz;
    for (uint32_t j0 = c0; j0 < c0 + MG_CHUNK && j0 < m; j0 += 32) {
      const uint32_t j = j0 + lane;
      const uint64_t key = (j < m && j < c0 + MG_CHUNK) ? off[j] : EMPTY;
      const bool keep = key < bound;
      const unsigned mk = __ballot_sync(0xffffffffu, keep);
      if (keep) buf[cnt + __popc(mk & ((1u << lane) - 1))] = key;
      cnt += __popc(mk);
    }
    if (cnt == sz) continue;
    uint32_t ns = 64;
    while (ns < cnt) ns <<= 1;
    for (uint32_t j = cnt + lane; j < ns; j += 32) buf[j] = EMPTY;
    __syncwarp();
    warp_bitonic_u64(buf, ns, lane);
    // one entry per id (the first of equal ids: the pool's own entry, flag 0, sorts before an offer), the L smallest
    uint32_t out = 0;
    for (uint32_t j0 = 0; j0 < cnt && out < s.L; j0 += 32) {
      const uint32_t j = j0 + lane;
      uint64_t key = EMPTY;
      bool ok = false;
      if (j < cnt) { key = buf[j]; ok = !(j > 0 && (buf[j - 1] >> 1) == (key >> 1)); }
      __syncwarp();
      const unsigned mk = __ballot_sync(0xffffffffu, ok);
      const uint32_t pos = out + __popc(mk & ((1u << lane) - 1));
      if (ok && pos < s.L) buf[pos] = key;
      out += __popc(mk);
      __syncwarp();
    }
    sz = out < s.L ? out : s.L;
  }
  for (uint32_t j = lane; j < s.L; j += 32) pool[j] = j < sz ? buf[j] : EMPTY;
  if (lane == 0) s.psize[d] = sz;
}
// upper bound of the pairs one join evaluates (each pair makes at most two offers): sum over the nodes' sample lists
__global__ void join_bound_kernel(State s, unsigned long long *out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long pairs = 0;
  if (i < s.n) {
    const unsigned long long nn = s.n_fnew[i] + s.n_rnew[i];
    unsigned long long no = s.n_fold[i] + s.n_rold[i];
    if (no > 2ull * s.R) no = 2ull * s.R;
    pairs = nn * (nn - 1) / 2 + nn * no;
  }
  for (int o = 16; o; o >>= 1) pairs += __shfl_xor_sync(0xffffffffu, pairs, o);
  if ((threadIdx.x & 31) == 0 && pairs) atomicAdd(out, pairs);
}
// records (key, destination) -> the destination's segment of obuf
__global__ void scatter_offers_kernel(State s, unsigned long long total) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const ulonglong2 r = s.rec[i];
    const uint32_t d = (uint32_t)r.y;
    s.obuf[s.ooff[d] + atomicAdd(s.ofill + d, 1u)] = r.x;
  }
}
// exclusive scan of cnt[0..n) into 64-bit offsets, one block
__global__ void scan_u32_u64_kernel(const uint32_t *cnt, uint32_t n, unsigned long long *off) {
  __shared__ unsigned long long part[1024];
  const uint32_t tid = threadIdx.x, nt = blockDim.x;
  const uint32_t per = (n + nt - 1) / nt;
  const uint32_t b = min(n, tid * per), e = min(n, b + per);
  unsigned long long sum = 0;
  for (uint32_t i = b; i < e; i++) sum += cnt[i];
  part[tid] = sum;
  __syncthreads();
  if (tid == 0) {
    unsigned long long run = 0;
    for (uint32_t t = 0; t < nt; t++) { unsigned long long v = part[t]; part[t] = run; run += v; }
    off[n] = run;
  }
  __syncthreads();
  sum = part[tid];
  for (uint32_t i = b; i < e; i++) { off[i] = sum; sum += cnt[i]; }
}

// ---- update (:34-120) ------------------------------------------------------------------------------------
// phase A (:49-64): M, and the two values `pool.back()` can take for this pool; resets the sample lists
__global__ void update_a_kernel(State s) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.n) return;
  const uint64_t *pool = s.pool + (size_t)i * s.L;
  uint32_t psize = s.psize[i];
  uint32_t maxl = s.M[i] + s.S < psize ? s.M[i] + s.S : psize;
  uint32_t c = 0, l = 0;
  while (l < maxl && c < s.S) {
    if (pool[l] & 1ull) ++c;
    ++l;
  }
  s.M[i] = l;
  s.worst[i] = (uint32_t)(pool[psize - 1] >> 32);
  s.leafd[i] = (uint32_t)(pool[s.leaf_rank[psize]] >> 32);
  s.n_fnew[i] = 0; s.n_fold[i] = 0; s.n_rnew[i] = 0; s.n_rold[i] = 0;
}

// the reverse lists start empty every iteration
__global__ void fill_u64_kernel(uint64_t *p, size_t count, uint64_t v) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (; i < count; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

// phase B (:66-98): one warp per node; forward lists in pool order, reverse offers through the sorted insert
__global__ void update_b_kernel(State s, uint32_t seed_new, uint32_t seed_old) {
  const int lane = threadIdx.x & 31;
  uint32_t nd = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (nd >= s.n) return;
  uint64_t *pool = s.pool + (size_t)nd * s.L;
  const uint32_t M = s.M[nd];
  uint32_t cnt_new = 0, cnt_old = 0;
  for (uint32_t base = 0; base < M; base += 32) {
    uint32_t l = base + lane;
    bool valid = l < M;
    uint64_t e = valid ? pool[l] : 0ull;
    uint32_t id = (uint32_t)e >> 1, db = (uint32_t)(e >> 32);
    bool isnew = valid && (e & 1ull);
    bool isold = valid && !(e & 1ull);
    uint32_t mnew = __ballot_sync(0xffffffffu, isnew), mold = __ballot_sync(0xffffffffu, isold);
    uint32_t below = (1u << lane) - 1u;
    if (isnew) {
      s.fwd_new[(size_t)nd * s.cap_fnew + cnt_new + __popc(mnew & below)] = id;
      pool[l] = e & ~1ull;
    }
    if (isold) {
      uint32_t slot = cnt_old + __popc(mold & below);
      if (slot < s.cap_fold) s.fwd_old[(size_t)nd * s.cap_fold + slot] = id;
    }
    cnt_new += __popc(mnew);
    cnt_old += __popc(mold);
    bool rev = valid && db > (id < nd ? s.leafd[id] : s.worst[id]);
    uint32_t mrev = __ballot_sync(0xffffffffu, rev);
    while (mrev) {
      int src = __ffs(mrev) - 1;
      mrev &= mrev - 1;
      uint32_t o = __shfl_sync(0xffffffffu, id, src);
      bool onew = (mnew >> src) & 1u;
      uint64_t key = ((uint64_t)seed_hash(onew ? seed_new : seed_old, o, nd) << 32) | nd;
      if (onew) topl_insert_dispatch(s.rnn_new + (size_t)o * s.R, s.n_rnew + o, s.lock_rnew + o, s.R, key, false, lane);
      else topl_insert_dispatch(s.rnn_old + (size_t)o * s.R, s.n_rold + o, s.lock_rold + o, s.R, key, false, lane);
    }
  }
  if (lane == 0) {
    s.n_fnew[nd] = cnt_new;
    s.n_fold[nd] = cnt_old < s.cap_fold ? cnt_old : s.cap_fold;
  }
}

// :275-291: the K best ids (and distances) of every pool
__global__ void extract_kernel(State s, uint32_t *knn, float *knn_d) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t total = (size_t)s.n * s.K;
  for (; i < total; i += (size_t)gridDim.x * blockDim.x) {
    uint32_t node = (uint32_t)(i / s.K), j = (uint32_t)(i % s.K);
    bool have = j < s.psize[node];
    uint64_t e = have ? s.pool[(size_t)node * s.L + j] : EMPTY;
    knn[i] = have ? (uint32_t)e >> 1 : INVALID_ID;
    if (knn_d) knn_d[i] = have ? __uint_as_float((uint32_t)(e >> 32)) : __int_as_float(0x7f800000);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// KD-tree initializer: efanna2e::IndexKDtree::Build (CNNS/efanna_graph/src/index_kdtree.cpp:279-435) in the batch form
// oracle.c orc_kdtree_graph restates (see there for the rules); equal to it bit for bit. Level-synchronous build of all
// trees at once: one warp per active node picks the cut (meanSplit / selectDivision), one CTA per active node does the
// stable three-way partition of its segment and creates the two children; then one warp per (tree, point) walks the
// point's path and, below the merge level, searches the sibling subtree whenever the point sits in the smaller child
// (mergeSubGraphs), offering every distance to both ends' K-lists through the locked sorted insert of the join above.
constexpr uint32_t KD_SEED = 0x4B445452U;
constexpr uint32_t KD_TNS = 10, KD_SAMPLE = 101;
constexpr int KD_RAND_DIM = 5;
struct KdState {
  const float *vec; size_t stride; uint32_t n, dim, K, T, ml, seed;
  uint32_t *ids, *tmp;                         // [T, n] the trees' point orders; partition scratch
  uint32_t *n_start, *n_end, *n_left, *n_depth; int *n_cutdim; float *n_cutval;   // nodes (roots = 0..T-1; left = INVALID: leaf)
  uint32_t *act_in, *act_out;                  // active nodes of this / the next level
  uint32_t *counters;                          // {nodes allocated, next active count, deepest level, 0}
  uint64_t *lists; uint32_t *lsize, *lock;     // [n, K], [n], [n]
  unsigned long long *stats;                   // {distance evaluations}
};
__global__ void kd_perm_kernel(KdState s, const uint32_t *ab) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= (size_t)s.T * s.n) return;
  const uint32_t t = (uint32_t)(e / s.n), j = (uint32_t)(e % s.n);
  s.ids[e] = (uint32_t)(((uint64_t)ab[2 * t] * j + ab[2 * t + 1]) % s.n);
}
// meanSplit (:17-52) + selectDivision (:110-133) for active node a: one warp, lane = dimensions lane, lane + 32, ...
__global__ void __launch_bounds__(128) kd_cut_kernel(KdState s, uint32_t n_act) {
  extern __shared__ float kd_sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t a = blockIdx.x * 4 + warp;
  if (a >= n_act) return;
  float *mean = kd_sm + (size_t)warp * 2 * s.dim, *var = mean + s.dim;
  const uint32_t x = s.act_in[a];
  const uint32_t start = s.n_start[x], count = s.n_end[x] - start;
  const uint32_t t = start / s.n;   // segments are positions in the [T, n] array
  const uint32_t *ids = s.ids + start;
  const uint32_t cnt = count < KD_SAMPLE ? count : KD_SAMPLE;
  const float div = 1.0f / (float)cnt;
  for (uint32_t k = lane; k < s.dim; k += 32) {
    float m = 0.f;
    for (uint32_t j = 0; j < cnt; ++j) m += s.vec[(size_t)ids[j] * s.stride + k];
    m *= div;
    float v = 0.f;
    for (uint32_t j = 0; j < cnt; ++j) { float d = s.vec[(size_t)ids[j] * s.stride + k] - m; v = __fadd_rn(v, __fmul_rn(d, d)); }  // no fma: the oracle's order
    mean[k] = m; var[k] = v;
  }
  __syncwarp();
  if (lane == 0) {
    uint32_t topind[KD_RAND_DIM];
    int num = 0;
    for (uint32_t i = 0; i < s.dim; ++i) {
      if (num < KD_RAND_DIM || var[i] > var[topind[num - 1]]) {
        if (num < KD_RAND_DIM) topind[num++] = i; else topind[num - 1] = i;
        int j = num - 1;
        while (j > 0 && var[topind[j]] > var[topind[j - 1]]) { uint32_t w = topind[j]; topind[j] = topind[j - 1]; topind[j - 1] = w; --j; }
      }
    }
    const uint32_t pick = seed_hash(s.seed ^ KD_SEED, t, (start - t * s.n) * 31u + count);
    const uint32_t cd = topind[pick % (uint32_t)num];
    s.n_cutdim[x] = (int)cd;
    s.n_cutval[x] = mean[cd];
  }
}
// stable three-way partition of active node a's segment (planeSplit's outcome, :69-108) + the split index (:54-62);
// creates the children and queues those that are still larger than a leaf
__global__ void __launch_bounds__(256) kd_partition_kernel(KdState s, uint32_t n_act) {
  __shared__ uint32_t wtot[8][3];
  __shared__ uint32_t tot[2];
  const uint32_t a = blockIdx.x;
  if (a >= n_act) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t x = s.act_in[a];
  const uint32_t start = s.n_start[x], count = s.n_end[x] - start;
  const int cd = s.n_cutdim[x];
  const float cv = s.n_cutval[x];
  uint32_t *ids = s.ids + start, *tmp = s.tmp + start;
  // pass 1: lim1 = #(< cutval), lim2 = #(<= cutval)
  uint32_t c1 = 0, c2 = 0;
  for (uint32_t j = tid; j < count; j += 256) {
    const float v = s.vec[(size_t)ids[j] * s.stride + cd];
    c1 += v < cv; c2 += v <= cv;
  }
  c1 = __reduce_add_sync(0xffffffffu, c1); c2 = __reduce_add_sync(0xffffffffu, c2);
  if (lane == 0) { wtot[warp][0] = c1; wtot[warp][1] = c2; }
  __syncthreads();
  if (tid == 0) {
    uint32_t l1 = 0, l2 = 0;
    for (int w = 0; w < 8; ++w) { l1 += wtot[w][0]; l2 += wtot[w][1]; }
    tot[0] = l1; tot[1] = l2;
  }
  __syncthreads();
  const uint32_t lim1 = tot[0], lim2 = tot[1];
  // pass 2: stable scatter, class 0 -> [0, lim1), 1 -> [lim1, lim2), 2 -> [lim2, count)
  uint32_t run[3] = {0, lim1, lim2};
  for (uint32_t j0 = 0; j0 < count; j0 += 256) {
    const uint32_t j = j0 + tid;
    int cls = -1;
    uint32_t id = 0;
    if (j < count) {
      id = ids[j];
      const float v = s.vec[(size_t)id * s.stride + cd];
      cls = v < cv ? 0 : (v == cv ? 1 : 2);
    }
    uint32_t wr = 0;
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const unsigned m = __ballot_sync(0xffffffffu, cls == c);
      if (cls == c) wr = __popc(m & ((1u << lane) - 1));
      if (lane == 0) wtot[warp][c] = __popc(m);
    }
    __syncthreads();
    if (cls >= 0) {
      uint32_t off = wr;
      for (int w = 0; w < warp; ++w) off += wtot[w][cls];
      tmp[run[cls] + off] = id;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c)
      for (int w = 0; w < 8; ++w) run[c] += wtot[w][c];
  }
  __syncthreads();
  for (uint32_t j = tid; j < count; j += 256) ids[j] = tmp[j];
  if (tid == 0) {
    uint32_t index;
    if (lim1 > count / 2) index = lim1; else if (lim2 < count / 2) index = lim2; else index = count / 2;
    if (lim1 == count || lim2 == 0) index = count / 2;
    const uint32_t L = atomicAdd(&s.counters[0], 2u);
    const uint32_t dep = s.n_depth[x] + 1;
    s.n_left[x] = L;
    s.n_start[L] = start; s.n_end[L] = start + index; s.n_left[L] = INVALID_ID; s.n_depth[L] = dep;
    s.n_start[L + 1] = start + index; s.n_end[L + 1] = start + count; s.n_left[L + 1] = INVALID_ID; s.n_depth[L + 1] = dep;
    atomicMax(&s.counters[2], dep);
    if (index > KD_TNS) s.act_out[atomicAdd(&s.counters[1], 1u)] = L;
    if (count - index > KD_TNS) s.act_out[atomicAdd(&s.counters[1], 1u)] = L + 1;
  }
}
// mergeSubGraphs (:221-275) from the point's side: warp w = (tree t, position pos of the tree's final order)
__global__ void __launch_bounds__(128) kd_merge_kernel(KdState s) {
  const int lane = threadIdx.x & 31;
  const size_t w = (size_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (w >= (size_t)s.T * s.n) return;
  const uint32_t t = (uint32_t)(w / s.n);
  const uint32_t gpos = (uint32_t)w;                 // position in the [T, n] array = what n_start / n_end hold
  const uint32_t p = s.ids[w];
  const float *pv = s.vec + (size_t)p * s.stride;
  uint32_t x = t, evals = 0;
  while (true) {
    const uint32_t L = s.n_left[x];
    if (L == INVALID_ID) break;
    const uint32_t R = L + 1;
    const uint32_t endL = s.n_end[L];
    const bool inL = gpos < endL;
    const uint32_t numL = endL - s.n_start[L], numR = s.n_end[R] - s.n_start[R];
    const bool searchers_left = numL < numR;
    if (s.n_depth[x] >= s.ml && inL == searchers_left) {
      uint32_t y = inL ? R : L;
      for (uint32_t yl = s.n_left[y]; yl != INVALID_ID; yl = s.n_left[y]) y = pv[s.n_cutdim[y]] < s.n_cutval[y] ? yl : yl + 1;
      const uint32_t ls = s.n_start[y], ln = s.n_end[y] - ls;   // a leaf: <= KD_TNS points
      uint32_t q = INVALID_ID;
      float d = 0.f;
      if ((uint32_t)lane < ln) {
        q = s.ids[ls + lane];
        d = seq_l2(s.vec + (size_t)q * s.stride, pv, s.dim);
      }
      evals += ln;
      for (uint32_t j = 0; j < ln; ++j) {
        const uint32_t qj = __shfl_sync(0xffffffffu, q, (int)j);
        const uint64_t db = (uint64_t)__float_as_uint(__shfl_sync(0xffffffffu, d, (int)j)) << 32;
        topl_insert_dispatch(s.lists + (size_t)qj * s.K, s.lsize + qj, s.lock + qj, s.K, db | ((uint64_t)p << 1), true, lane);
        topl_insert_dispatch(s.lists + (size_t)p * s.K, s.lsize + p, s.lock + p, s.K, db | ((uint64_t)qj << 1), true, lane);
      }
    }
    x = inL ? L : R;
  }
  if (lane == 0 && evals) atomicAdd(s.stats + 0, (unsigned long long)evals);
}
// rows ascending (distance, id); short rows filled with distinct hashed ids at distance +inf (:413-424)
__global__ void kd_extract_kernel(KdState s, uint32_t *knn, float *knn_d) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.n) return;
  uint32_t sz = s.lsize[i], ctr = 0;
  const uint64_t *l = s.lists + (size_t)i * s.K;
  uint32_t *row = knn + (size_t)i * s.K;
  for (uint32_t j = 0; j < sz; ++j) { row[j] = (uint32_t)l[j] >> 1; if (knn_d) knn_d[(size_t)i * s.K + j] = __uint_as_float((uint32_t)(l[j] >> 32)); }
  while (sz < s.K) {
    const uint32_t id = seed_hash(s.seed ^ KD_SEED, i, 1000u + ctr++) % s.n;
    bool dup = false;
    for (uint32_t j = 0; j < sz; ++j) dup |= row[j] == id;
    if (dup) continue;
    row[sz] = id;
    if (knn_d) knn_d[(size_t)i * s.K + sz] = __int_as_float(0x7f800000);
    ++sz;
  }
}

}  // namespace nnd

// Host driver. knn [n,K] / knn_d [n,K] are device pointers. info8 = {distance evaluations, offers queued,
// pool inserts, iterations, join ms, update ms, 0, 0}.
int nndescent_launch(b200ann_handle *h, const b200ann_index *ix, size_t K, size_t L, size_t iter, size_t S, size_t R,
                     uint32_t seed, uint32_t *knn, float *knn_d, uint64_t *info8, const uint32_t *init_graph, size_t init_k) {
  using namespace nnd;
  const size_t n = ix->n;
  if (ix->metric != B200ANN_L2)
    return set_err(h, B200ANN_ERR_ARG, "nndescent: the reference NN-descent is L2 only (efanna2e::Index always builds DistanceL2)");
  if (S < 2 || 2 * S > (size_t)MAX_GEN || S + 1 > L || K > L || K == 0 || R == 0 || L > 512 || R > 512 || n <= 2 * S + 1 ||
      n >= (1ull << 31))
    return set_err(h, B200ANN_ERR_ARG, "nndescent: need 2 <= S <= 64, S < L <= 512, 0 < K <= L, 0 < R <= 512, n > 2S+1");
  cudaStream_t st = h->stream;
  State s{};
  s.vec = ix->vec; s.stride = ix->stride; s.n = (uint32_t)n; s.dim = (uint32_t)ix->dim;
  s.L = (uint32_t)L; s.S = (uint32_t)S; s.R = (uint32_t)R; s.K = (uint32_t)K;
  s.cap_fnew = (uint32_t)(2 * S);
  s.cap_fold = (uint32_t)std::min(L, 2 * R);
  DevBuf<uint64_t> pool, rnew, rold;
  DevBuf<uint32_t> u32s, fnew, fold, leaf;
  DevBuf<unsigned long long> stats;
  CU_CHECK(h, pool.alloc(n * L, st));
  CU_CHECK(h, rnew.alloc(n * R, st));
  CU_CHECK(h, rold.alloc(n * R, st));
  CU_CHECK(h, u32s.alloc(n * 11, st));
  CU_CHECK(h, fnew.alloc(n * s.cap_fnew, st));
  CU_CHECK(h, fold.alloc(n * s.cap_fold, st));
  CU_CHECK(h, leaf.alloc(L + 1, st));
  CU_CHECK(h, stats.alloc(16, st));
  s.pool = pool.p; s.rnn_new = rnew.p; s.rnn_old = rold.p;
  uint32_t *u = u32s.p;
  s.psize = u; s.M = u + n; s.worst = u + 2 * n; s.leafd = u + 3 * n; s.n_fnew = u + 4 * n; s.n_fold = u + 5 * n;
  s.n_rnew = u + 6 * n; s.n_rold = u + 7 * n; s.lock_pool = u + 8 * n; s.lock_rnew = u + 9 * n; s.lock_rold = u + 10 * n;
  s.fwd_new = fnew.p; s.fwd_old = fold.p; s.stats = stats.p;
  // leaf_rank[p]: sorted rank of the element std::make_heap (the reference's own libstdc++ algorithm,
  // index_graph.cpp:97) leaves at position p-1 of an ascending array of p distinct values
  std::vector<uint32_t> lr(L + 1, 0), hp;
  for (size_t p = 1; p <= L; ++p) {
    hp.resize(p);
    for (size_t i = 0; i < p; ++i) hp[i] = (uint32_t)i;
    std::make_heap(hp.begin(), hp.end());
    lr[p] = hp[p - 1];
  }
  CU_CHECK(h, cudaMemcpyAsync(leaf.p, lr.data(), (L + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  s.leaf_rank = leaf.p;
  CU_CHECK(h, cudaMemsetAsync(stats.p, 0, 16 * sizeof(unsigned long long), st));
  const int TB = 128;
  const unsigned nb = (unsigned)((n + TB - 1) / TB);
  init_kernel<<<nb, TB, 0, st>>>(s, seed);
  h->launches++;
  if (init_graph) {
    if (init_k == 0 || init_k > L || init_k > (size_t)RF_MAXK)
      return set_err(h, B200ANN_ERR_ARG, "nndescent_refine: need 0 < init_k <= L (<= 512)");
    s.init_graph = init_graph; s.init_k = (uint32_t)init_k;
    init_refine_kernel<<<(unsigned)((n + RF_WARPS - 1) / RF_WARPS), RF_WARPS * 32, 0, st>>>(s);
    h->launches++;
  }
  const uint32_t cn_max = (uint32_t)(std::max(2 * S, S + R) + 2 * R);
  const size_t smem = 2 * (size_t)TILE * ROWF * sizeof(float) + (size_t)TILE * DROW * sizeof(uint32_t) + 2 * (size_t)cn_max * sizeof(uint32_t);
  CU_CHECK(h, cudaFuncSetAttribute(join_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CU_CHECK(h, cudaFuncSetAttribute(join_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CU_CHECK(h, cudaFuncSetAttribute(join_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CU_CHECK(h, cudaFuncSetAttribute(join_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  CU_CHECK(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, join_kernel<1>, JT, smem));
  // B200ANN_NND_LOCKFREE=0: the round-1 join (offers merged under per-destination locks), for A/B runs
  // 2 (default): one tile pass + record scatter; 1: two tile passes (count, then write), no record buffer
  int lockfree = 2;
  if (const char *e = getenv("B200ANN_NND_LOCKFREE")) lockfree = atoi(e);
  DevBuf<unsigned long long> cursors;
  CU_CHECK(h, cursors.alloc(2, st));
  DevBuf<ulonglong2> rec;      // offer records and per-destination segments, kept across the iterations
  DevBuf<uint64_t> obuf;
  unsigned long long rec_cap = 0, obuf_cap = 0;
  DevBuf<unsigned long long> stats_keep;
  CU_CHECK(h, stats_keep.alloc(3, st));
  DevBuf<uint32_t> ocnt;
  DevBuf<unsigned long long> ooff;
  CU_CHECK(h, ocnt.alloc(2 * n, st));
  CU_CHECK(h, ooff.alloc(n + 1, st));
  s.ocnt = ocnt.p; s.ofill = ocnt.p + n; s.ooff = ooff.p;
  if (occ < 1) return set_err(h, B200ANN_ERR_ARG, "nndescent: sample lists do not fit in shared memory (S, R too large)");
  const unsigned jgrid = (unsigned)std::min<size_t>((size_t)h->sm_count * occ, n);
  struct Events {  // destroyed on every return path
    cudaEvent_t e[3] = {nullptr, nullptr, nullptr};
    ~Events() { for (auto x : e) if (x) cudaEventDestroy(x); }
  } evs;
  for (auto &x : evs.e) CU_CHECK(h, cudaEventCreate(&x));
  cudaEvent_t e0 = evs.e[0], e1 = evs.e[1], e2 = evs.e[2];
  float join_ms = 0.f, upd_ms = 0.f;
  for (size_t it = 0; it < iter; ++it) {
    cudaMemsetAsync(stats.p + 3, 0, sizeof(unsigned long long), st);
    cudaEventRecord(e0, st);
    if (!lockfree) {
      join_kernel<0><<<jgrid, JT, smem, st>>>(s, cn_max);
      h->launches++;
    } else if (lockfree == 2) {
      // one tile pass: offers go to a record buffer sized by the pair bound, then to their destinations' segments
      cudaMemsetAsync(ocnt.p, 0, 2 * n * sizeof(uint32_t), st);
      cudaMemsetAsync(cursors.p, 0, 2 * sizeof(unsigned long long), st);
      join_bound_kernel<<<nb, TB, 0, st>>>(s, cursors.p + 1);
      unsigned long long bound = 0;
      CU_CHECK(h, cudaMemcpyAsync(&bound, cursors.p + 1, sizeof bound, cudaMemcpyDeviceToHost, st));
      CU_CHECK(h, cudaStreamSynchronize(st));
      // The record buffer is sized by the pair bound once (first iteration: nearly every pair is an offer); later
      // bounds overshoot the offers 3-6x, and growing a multi-GB buffer costs more than the join (0.9 s measured at
      // 1M rows), so later iterations keep the buffer and re-run the pass in the rare case the records overflow it -
      // the pass only reads the graph, so a re-run writes the same records.
      if (!rec_cap) {
        rec_cap = 2 * bound + 1;   // cannot overflow in this iteration; a later one that does re-runs its pass
        if (const char *e = getenv("B200ANN_NND_REC_CAP")) rec_cap = std::max(1ll, atoll(e));  // tests: force the re-run path
        CU_CHECK(h, rec.alloc(rec_cap, st));
      }
      s.rec = rec.p; s.rec_cursor = cursors.p; s.rec_cap = rec_cap;
      unsigned long long total = 0;
      for (;;) {
        cudaMemcpyAsync(stats_keep.p, stats.p, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToDevice, st);
        join_kernel<3><<<jgrid, JT, smem, st>>>(s, cn_max);
        scan_u32_u64_kernel<<<1, 1024, 0, st>>>(s.ocnt, (uint32_t)n, ooff.p);
        CU_CHECK(h, cudaMemcpyAsync(&total, ooff.p + n, sizeof total, cudaMemcpyDeviceToHost, st));
        CU_CHECK(h, cudaStreamSynchronize(st));
        h->launches += 2;
        if (total <= rec_cap) break;
        cudaFreeAsync(rec.p, st); rec.p = nullptr;
        rec_cap = total + total / 4;
        CU_CHECK(h, rec.alloc(rec_cap, st));
        s.rec = rec.p; s.rec_cap = rec_cap;
        cudaMemcpyAsync(stats.p, stats_keep.p, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToDevice, st);
        cudaMemsetAsync(ocnt.p, 0, 2 * n * sizeof(uint32_t), st);
        cudaMemsetAsync(cursors.p, 0, sizeof(unsigned long long), st);
        cudaMemsetAsync(stats.p + 3, 0, sizeof(unsigned long long), st);   // the tile pass's work counter
      }
      if (total + 1 > obuf_cap) {
        if (obuf.p) { cudaFreeAsync(obuf.p, st); obuf.p = nullptr; }
        obuf_cap = (total + 1) + (total + 1) / 4;
        CU_CHECK(h, obuf.alloc(obuf_cap, st));
      }
      s.obuf = obuf.p;
      if (total) scatter_offers_kernel<<<h->sm_count * 16, 256, 0, st>>>(s, total);
      merge_offers_kernel<<<(unsigned)((n + MG_WARPS - 1) / MG_WARPS), MG_WARPS * 32, 0, st>>>(s);
      h->launches += 3;
    } else {
      cudaMemsetAsync(ocnt.p, 0, 2 * n * sizeof(uint32_t), st);
      join_kernel<1><<<jgrid, JT, smem, st>>>(s, cn_max);
      scan_u32_u64_kernel<<<1, 1024, 0, st>>>(s.ocnt, (uint32_t)n, ooff.p);
      unsigned long long total = 0;
      CU_CHECK(h, cudaMemcpyAsync(&total, ooff.p + n, sizeof total, cudaMemcpyDeviceToHost, st));
      CU_CHECK(h, cudaStreamSynchronize(st));
      DevBuf<uint64_t> obuf;
      CU_CHECK(h, obuf.alloc(total ? total : 1, st));
      s.obuf = obuf.p;
      cudaMemsetAsync(stats.p + 3, 0, sizeof(unsigned long long), st);
      join_kernel<2><<<jgrid, JT, smem, st>>>(s, cn_max);
      merge_offers_kernel<<<(unsigned)((n + MG_WARPS - 1) / MG_WARPS), MG_WARPS * 32, 0, st>>>(s);
      h->launches += 4;
    }
    cudaEventRecord(e1, st);
    update_a_kernel<<<nb, TB, 0, st>>>(s);
    fill_u64_kernel<<<h->sm_count * 8, 256, 0, st>>>(rnew.p, n * R, EMPTY);
    fill_u64_kernel<<<h->sm_count * 8, 256, 0, st>>>(rold.p, n * R, EMPTY);
    const uint32_t seed_new = seed + (uint32_t)(2 * it + 1) * 0x9E3779B1U, seed_old = seed + (uint32_t)(2 * it + 2) * 0x9E3779B1U;
    update_b_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, st>>>(s, seed_new, seed_old);
    cudaEventRecord(e2, st);
    h->launches += 4;
    CU_CHECK(h, cudaEventSynchronize(e2));
    float a = 0.f, b = 0.f;
    cudaEventElapsedTime(&a, e0, e1); cudaEventElapsedTime(&b, e1, e2);
    join_ms += a; upd_ms += b;
#ifdef NND_PROF
    {
      unsigned long long ps[16];
      cudaMemcpy(ps, stats.p, sizeof ps, cudaMemcpyDeviceToHost);
      fprintf(stderr, "NND_PROF it %zu: join %.1f ms update %.1f ms | cum dist %llu offers %llu ins %llu | cum members new %llu old %llu tiles %llu | cyc fetch %llu tile %llu drain %llu\n",
              it, a, b, ps[0], ps[1], ps[2], ps[8], ps[9], ps[10], ps[4], ps[5], ps[6]);
    }
#endif
  }
  extract_kernel<<<h->sm_count * 8, 256, 0, st>>>(s, knn, knn_d);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  unsigned long long hs[16];
  CU_CHECK(h, cudaMemcpyAsync(hs, stats.p, sizeof hs, cudaMemcpyDeviceToHost, st));
  CU_CHECK(h, cudaStreamSynchronize(st));
  if (info8) {
    info8[0] = hs[0]; info8[1] = hs[1]; info8[2] = hs[2]; info8[3] = iter;
    info8[4] = (uint64_t)(join_ms * 1000.f); info8[5] = (uint64_t)(upd_ms * 1000.f); info8[6] = info8[7] = 0;
#ifdef NND_PROF
    unsigned long long gp[16];
    cudaMemcpyFromSymbol(gp, nnd::g_prof, sizeof gp);
    fprintf(stderr, "NND_PROF cycles(thread 0 of each CTA): fetch %llu gather %llu compute+Dm %llu drain %llu\n", hs[4], hs[7], hs[5], hs[6]);
    fprintf(stderr, "NND_PROF merges %llu: avg cycles lock %llu load %llu insert %llu writeback+fence %llu\n", gp[4], gp[0] / (gp[4] + 1),
            gp[1] / (gp[4] + 1), gp[2] / (gp[4] + 1), gp[3] / (gp[4] + 1));
#endif
  }
  return 0;
}

// KD-tree initializer (see the kernels above). knn [n,K] / knn_d device pointers. info8 = {distance evaluations, tree
// nodes, leaves, deepest level, levels run, 0, 0, 0}.
int kdtree_launch(b200ann_handle *h, const b200ann_index *ix, size_t K, size_t nTrees, size_t ml, uint32_t seed, uint32_t *knn,
                  float *knn_d, uint64_t *info8) {
  using namespace nnd;
  const size_t n = ix->n, T = nTrees;
  if (ix->metric != B200ANN_L2) return set_err(h, B200ANN_ERR_ARG, "kdtree_graph: L2 only (efanna2e::Index always builds DistanceL2)");
  if (n < 2 || K == 0 || K >= n || K > 512 || T == 0 || T > 64 || T * n >= (1ull << 31) || ix->dim > 4096)
    return set_err(h, B200ANN_ERR_ARG, "kdtree_graph: need 0 < K < n, K <= 512, 1 <= nTrees <= 64, nTrees * n < 2^31, dim <= 4096");
  cudaStream_t st = h->stream;
  KdState s{};
  s.vec = ix->vec; s.stride = ix->stride; s.n = (uint32_t)n; s.dim = (uint32_t)ix->dim; s.K = (uint32_t)K; s.T = (uint32_t)T;
  s.ml = (uint32_t)ml; s.seed = seed;
  const size_t capn = 2 * T * n + 8;
  DevBuf<uint32_t> ids, tmp, nodes, act, counters, lsz, ab;
  DevBuf<float> cutv;
  DevBuf<uint64_t> lists;
  DevBuf<unsigned long long> stats;
  CU_CHECK(h, ids.alloc(T * n, st)); CU_CHECK(h, tmp.alloc(T * n, st));
  CU_CHECK(h, nodes.alloc(5 * capn, st)); CU_CHECK(h, cutv.alloc(capn, st));
  CU_CHECK(h, act.alloc(2 * capn, st)); CU_CHECK(h, counters.alloc(4, st));
  CU_CHECK(h, lists.alloc(n * K, st)); CU_CHECK(h, lsz.alloc(2 * n, st));
  CU_CHECK(h, ab.alloc(2 * T, st)); CU_CHECK(h, stats.alloc(4, st));
  s.ids = ids.p; s.tmp = tmp.p;
  s.n_start = nodes.p; s.n_end = nodes.p + capn; s.n_left = nodes.p + 2 * capn; s.n_depth = nodes.p + 3 * capn;
  s.n_cutdim = reinterpret_cast<int *>(nodes.p + 4 * capn); s.n_cutval = cutv.p;
  s.counters = counters.p; s.lists = lists.p; s.lsize = lsz.p; s.lock = lsz.p + n; s.stats = stats.p;
  // tree t lists the points in the order (a_t * j + b_t) mod n, a_t coprime to n (oracle.c kd_mult)
  auto mix = [](uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; };
  auto hash = [&](uint32_t sd, uint32_t q, uint32_t c) { return mix(mix(sd ^ (q * 0x9E3779B9U)) + c * 0x85EBCA6BU); };
  auto gcd = [](uint32_t a, uint32_t b) { while (b) { uint32_t t = a % b; a = b; b = t; } return a; };
  std::vector<uint32_t> hab(2 * T), hn(5 * T), hact(T);
  for (uint32_t t = 0; t < T; ++t) {
    uint32_t a = hash(seed ^ KD_SEED, t, 0x7001u) % (uint32_t)n;
    if (a == 0) a = 1;
    while (gcd(a, (uint32_t)n) != 1) a = a + 1 == (uint32_t)n ? 1 : a + 1;
    hab[2 * t] = a; hab[2 * t + 1] = hash(seed ^ KD_SEED, t, 0x7002u) % (uint32_t)n;
  }
  CU_CHECK(h, cudaMemcpyAsync(ab.p, hab.data(), 2 * T * 4, cudaMemcpyHostToDevice, st));
  kd_perm_kernel<<<(unsigned)((T * n + 255) / 256), 256, 0, st>>>(s, ab.p);
  // roots: node t = the whole of tree t
  std::vector<uint32_t> rs(T), re(T), rl(T, INVALID_ID), rd(T, 0u);
  for (uint32_t t = 0; t < T; ++t) { rs[t] = (uint32_t)(t * n); re[t] = (uint32_t)((t + 1) * n); hact[t] = t; }
  CU_CHECK(h, cudaMemcpyAsync(s.n_start, rs.data(), T * 4, cudaMemcpyHostToDevice, st));
  CU_CHECK(h, cudaMemcpyAsync(s.n_end, re.data(), T * 4, cudaMemcpyHostToDevice, st));
  CU_CHECK(h, cudaMemcpyAsync(s.n_left, rl.data(), T * 4, cudaMemcpyHostToDevice, st));
  CU_CHECK(h, cudaMemcpyAsync(s.n_depth, rd.data(), T * 4, cudaMemcpyHostToDevice, st));
  uint32_t hc[4] = {(uint32_t)T, 0u, 0u, 0u};
  CU_CHECK(h, cudaMemcpyAsync(counters.p, hc, sizeof hc, cudaMemcpyHostToDevice, st));
  CU_CHECK(h, cudaMemcpyAsync(act.p, hact.data(), T * 4, cudaMemcpyHostToDevice, st));
  CU_CHECK(h, cudaMemsetAsync(stats.p, 0, 4 * sizeof(unsigned long long), st));
  CU_CHECK(h, cudaMemsetAsync(lsz.p, 0, 2 * n * 4, st));
  fill_u64_kernel<<<h->sm_count * 8, 256, 0, st>>>(lists.p, n * K, EMPTY);
  h->launches += 2;
  const size_t cut_smem = (size_t)4 * 2 * ix->dim * sizeof(float);
  CU_CHECK(h, cudaFuncSetAttribute(kd_cut_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cut_smem));
  uint32_t n_act = n > KD_TNS ? (uint32_t)T : 0u, levels = 0;
  uint32_t *ain = act.p, *aout = act.p + capn;
  while (n_act) {
    s.act_in = ain; s.act_out = aout;
    kd_cut_kernel<<<(n_act + 3) / 4, 128, cut_smem, st>>>(s, n_act);
    kd_partition_kernel<<<n_act, 256, 0, st>>>(s, n_act);
    h->launches += 2;
    CU_CHECK(h, cudaMemcpyAsync(hc, counters.p, sizeof hc, cudaMemcpyDeviceToHost, st));
    CU_CHECK(h, cudaStreamSynchronize(st));
    n_act = hc[1];
    CU_CHECK(h, cudaMemsetAsync(counters.p + 1, 0, 4, st));
    std::swap(ain, aout);
    if (++levels > 4096) return set_err(h, B200ANN_ERR_CUDA, "kdtree_graph: the tree build does not terminate");
  }
  kd_merge_kernel<<<(unsigned)((T * n + 3) / 4), 128, 0, st>>>(s);
  kd_extract_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(s, knn, knn_d);
  h->launches += 2;
  CU_CHECK(h, cudaGetLastError());
  unsigned long long hs[4];
  CU_CHECK(h, cudaMemcpyAsync(hs, stats.p, sizeof hs, cudaMemcpyDeviceToHost, st));
  CU_CHECK(h, cudaMemcpyAsync(hc, counters.p, sizeof hc, cudaMemcpyDeviceToHost, st));
  CU_CHECK(h, cudaStreamSynchronize(st));
  if (info8) {
    // every internal node has two children: nodes = T + 2 * internal, leaves = nodes - internal
    const uint64_t nn = hc[0], internal = (nn - T) / 2;
    info8[0] = hs[0]; info8[1] = nn; info8[2] = nn - internal; info8[3] = hc[2]; info8[4] = levels; info8[5] = info8[6] = info8[7] = 0;
  }
  return 0;
}

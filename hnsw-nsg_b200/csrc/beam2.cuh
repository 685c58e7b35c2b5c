// Search core v2 (search kernels only; the build kernels keep beam.cuh): the same algorithm, statement for
// statement, as beam.cuh's Beam - NSG Search / SearchFromEnterpoint (CNNS/src/nsg/index_nsg.cpp:506-568,
// hnsw_nsg/src/index_nsg.cpp:703-782) and hnswlib searchKnn (hnswlib/hnswlib/hnswalg.h:1270-1324, :309-440) - laid
// out for what the round-1 and round-2 profiles showed: the kernel is bound by the LATENCY of one query's dependent
// chain (adjacency row -> visited filter -> 5 round trips of 8 rows -> inserts), not by HBM (a random-row probe
// reads 7.5 TB/s where v1 reaches 4.2), and with one warp per query a 10k-query batch is only ~2.5 rounds of
// resident warps, so the last round runs the GPU a third empty.
//
//  1. One BLOCK of W = 4 warps per query. Warp 0 (the owner) runs the sequential part of the reference loop -
//     pool, visited set, candidate list - unchanged; every distance evaluation is a block-wide step: the owner
//     posts "evaluate cand_i[0..m)", the warps take the rows in groups of RPS, and the owner continues when all
//     distances are in cand_d. A hop's ~37 rows are one or two round trips instead of five, a query is ~3x
//     shorter, and a batch is ~9 rounds of resident blocks (the tail costs a few per cent instead of 30).
//  2. Neighbour rows arrive through the TMA bulk-copy engine instead of register loads: `cp.async.bulk`
//     HBM -> shared memory, one instruction per ROW, completion counted in bytes on an mbarrier, into a per-warp ring
//     of `nst` stages x RPS rows; the warp computes stage g from shared memory while the later stages are in flight.
//     Rows in flight cost no registers (v1: 32 per lane), which is what lets the kernel fit 64 registers and 32
//     warps per SM. (Measured against `cp.async` 16-byte copies, one warp-wide instruction per 512 bytes: the
//     per-lane address arithmetic made that form slower. The bulk copy takes its operands from uniform registers,
//     so a row per lane costs an elect / broadcast loop of 9 instructions per row.)
//     (BASELINE north_star: "overlap ... prefetch with TMA or cp.async"; the reference's own _mm_prefetch of the
//     neighbour rows before evaluating them: index_nsg.cpp:627-632, hnswalg.h:370-383.)
//  3. For L <= 64 the sorted pool lives in the owner's registers, one entry per lane and register (PR = 1 or 2):
//     InsertIntoPool (CNNS/include/nsg/neighbor.h:107-135) is a ballot rank + one shuffle-up instead of a
//     shared-memory scan and shift loop, and "first unexpanded entry" is one ballot.
//
// Distances are summed in exactly v1's order (team of TEAM lanes per row, lane t takes 16-byte chunks t, t+TEAM, ..,
// x,y,z,w with fmaf, xor butterfly), so results are bit-identical to v1 and to the oracle's L2_GPU order.
#pragma once
#include "beam.cuh"

namespace beam2 {
using namespace beam;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// one row: global -> shared through the bulk-copy (TMA) engine; bytes and both addresses are multiples of 16
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

constexpr int W2 = 4;                       // warps per block = warps per query
constexpr uint32_t CMD_EXIT = 0xFFFFFFFFu;  // block commands (word 0 of the block's shared memory); anything else = "evaluate m rows"
constexpr uint32_t CMD_QUERY = 0xFFFFFFFEu; // word 1 = query index: load it, clear your quarter of the visited table

// Per-block shared memory, in words: [cmd, qi, 0, 0] then one region per warp:
//   [ring | seeds' sort buffer][mbarriers nst][cand_i][cand_d][visited table][pool (shared-memory pool form only)]
// The register-pool forms need a shared pool only once per query, as the buffer the seeds are rank-sorted into; it
// aliases the warp's (then idle) ring. Every region is a multiple of 4 words, so every ring starts on a 16-byte
// boundary. In warp mode every warp works in its own region; in block mode the owner (warp 0) works in its region and
// the other warps read warp 0's candidate list and clear their share of warp 0's table.
__host__ __device__ inline uint32_t ring_region_words(uint32_t ring_words, uint32_t pool_cap, bool reg_pool) {
  uint32_t w = ring_words;
  if (reg_pool && w < 2 * pool_cap) w = 2 * pool_cap;
  return (w + 3u) & ~3u;
}
__host__ __device__ inline uint32_t warp_region_words(uint32_t pool_cap, uint32_t cand_cap, int hbits, uint32_t ring_words,
                                                      uint32_t nst, bool reg_pool) {
  uint32_t w = ring_region_words(ring_words, pool_cap, reg_pool) + 2 * nst + 2 * cand_cap + (1u << hbits) +
               (reg_pool ? 0u : 2 * pool_cap);
  return (w + 3u) & ~3u;
}
__host__ __device__ inline uint32_t smem_words2(uint32_t pool_cap, uint32_t cand_cap, int hbits, uint32_t ring_words,
                                                uint32_t nst, bool reg_pool) {
  return 4 + W2 * warp_region_words(pool_cap, cand_cap, hbits, ring_words, nst, reg_pool);
}

template <int TEAM, int NCH, bool IP, bool HNSW, int PR>
struct Beam2 {
  static constexpr int TPW = 32 / TEAM;
  static constexpr int RPS = 2 * TPW;  // rows per ring stage: two per team, v1's evaluation batch
  static constexpr unsigned FULL = 0xffffffffu;
  const SearchParams &p;
  volatile uint32_t *cmd;            // block command words
  const float *ring;                 // this warp's ring: [nst][RPS][stride] floats
  uint32_t ring_s, bar_s, rowbytes;  // shared-space addresses of the ring and of this warp's mbarrier 0; bytes per row
  uint32_t phases;                   // bit s: the parity stage s completes next
  float *pool_d; uint32_t *pool_i;   // [pool_cap]: the pool itself (PR == 0) or the seeds' sort buffer (PR > 0)
  uint32_t *cand_i; float *cand_d;   // [cand_cap]: this warp's candidate list
  uint32_t *cand0_i; float *cand0_d; // warp 0's: what a block-wide evaluation reads and writes
  uint32_t *tab, *tab0; uint32_t hshift, hmask;
  bool coop = false;                 // block mode: warp 0 owns the query, every evaluation is block-wide
  uint32_t nmask, rbits, rmask, smask;  // vis16 geometry
  float pd0 = INFINITY, pd1 = INFINITY; uint32_t pi0 = 0, pi1 = 0;  // register pool: entry e lives in lane e % 32 of register e / 32
  float4 qreg[NCH];
  int lane, t, g, warp;
  uint32_t np, cap;
  uint32_t n_dist, n_hops;
  uint32_t never = INVALID_ID;

  __device__ __forceinline__ Beam2(const SearchParams &pp, uint32_t *smem_base) : p(pp) {
    lane = threadIdx.x & 31;
    warp = threadIdx.x >> 5;
    t = lane % TEAM;
    g = lane / TEAM;
    const uint32_t region = ring_region_words(p.nst * RPS * (uint32_t)p.stride, p.pool_cap, PR > 0);
    const uint32_t per = warp_region_words(p.pool_cap, p.cand_cap, p.hbits, p.nst * RPS * (uint32_t)p.stride, p.nst, PR > 0);
    cmd = smem_base;
    uint32_t *mine = smem_base + 4 + warp * per, *w0 = smem_base + 4;
    ring = reinterpret_cast<const float *>(mine);
    ring_s = smem_u32(mine);
    bar_s = smem_u32(mine + region);
    rowbytes = (uint32_t)p.stride * 4u;
    cand_i = mine + region + 2 * p.nst;
    cand_d = reinterpret_cast<float *>(cand_i + p.cand_cap);
    tab = cand_i + 2 * p.cand_cap;
    cand0_i = w0 + region + 2 * p.nst;
    cand0_d = reinterpret_cast<float *>(cand0_i + p.cand_cap);
    tab0 = cand0_i + 2 * p.cand_cap;
    uint32_t *pl = PR > 0 ? mine : tab + (1u << p.hbits);  // register pool: the sort buffer aliases the ring
    pool_d = reinterpret_cast<float *>(pl);
    pool_i = pl + p.pool_cap;
    hshift = 32 - p.hbits;
    hmask = (1u << p.hbits) - 1;
    nmask = rbits = rmask = smask = 0;
    if (p.vis16) {
      nmask = p.nbits >= 32 ? 0xFFFFFFFFu : ((1u << p.nbits) - 1);
      rbits = (uint32_t)(p.nbits - (p.hbits + 1));
      rmask = (1u << rbits) - 1;
      smask = (2u << p.hbits) - 1;
    }
    phases = 0;
    if (lane == 0) {
      for (uint32_t s = 0; s < p.nst; s++) mbar_init(bar_s + 8 * s, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
  }

  // ------------------------------------------------------------ visited set (see beam.cuh)
  __device__ __forceinline__ uint32_t slot(uint32_t id) const { return (id * 0x9E3779B1u) >> hshift; }
  // For the lanes with `valid`: true if `id` was already in the set, else inserts it. Called by the whole warp.
  __device__ __forceinline__ bool visit(uint32_t id, bool valid) {
    bool seen = false;
    if (!p.vis16) {  // 32-bit entries: one atomicCAS per probe
      if (valid) {
        const uint32_t s = slot(id);
        bool done = false;
#pragma unroll 1
        for (int pr = 0; pr < PROBES && !done; pr++) {
          const uint32_t old = atomicCAS(&tab[(s + pr) & hmask], INVALID_ID, id);
          if (old == INVALID_ID) done = true;
          else if (old == id) { seen = true; done = true; }
        }
        if (!done) tab[s] = id;  // probe sequence full: forget the home slot's id
      }
      return seen;
    }
    // 16-bit entries (see beam.cuh for the id <-> (slot, entry) bijection), claimed without atomics - a 16-bit CAS on
    // shared memory is a 30-instruction emulation loop: every pending lane reads its probe slot; an empty slot is
    // claimed with a plain store and confirmed by reading it back after a warp barrier (of several lanes claiming one
    // slot one sticks, the others move on to their next probe, exactly what a failed CAS does). Lanes that offered the
    // SAME id all see their entry stick: the lowest lane inserted it, the others find it present.
    const uint32_t x = (id * 0x9E3779B1u) & nmask;
    const uint32_t s = x >> rbits;
    const uint32_t tag = (x & rmask) << 3;
    volatile unsigned short *t16 = reinterpret_cast<volatile unsigned short *>(tab);
    bool pending = valid;
    for (uint32_t pr = 0; __any_sync(FULL, pending); pr++) {
      const uint32_t idx = (s + pr) & smask;
      const unsigned short e = (unsigned short)(tag | pr);
      bool claimed = false;
      if (pending) {
        if (pr == (uint32_t)PROBES) {
          t16[s] = (unsigned short)tag;  // probe sequence full: forget the home slot's entry
          pending = false;
        } else {
          const unsigned short cur = t16[idx];
          if (cur == e) { seen = true; pending = false; }
          else if (cur == 0xFFFFu) { t16[idx] = e; claimed = true; }
        }
      }
      __syncwarp();
      const bool won = claimed && t16[idx] == e;
      const unsigned wm = __ballot_sync(FULL, won);
      if (won) {
        const unsigned twins = __match_any_sync(wm, idx);
        if (twins & ((1u << lane) - 1)) seen = true;
        pending = false;
      }
      __syncwarp();
    }
    return seen;
  }
  __device__ __forceinline__ bool visit_uniform(uint32_t id) {  // one id for the whole warp
    const int was = visit(id, lane == 0) ? 1 : 0;
    return __shfl_sync(FULL, was, 0) != 0;
  }

  __device__ __forceinline__ void load_query(const float *qv) {
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      uint32_t b = (uint32_t)(t + c * TEAM) * 4;
      qreg[c].x = b + 0 < p.dim ? __ldg(qv + b + 0) : 0.f;
      qreg[c].y = b + 1 < p.dim ? __ldg(qv + b + 1) : 0.f;
      qreg[c].z = b + 2 < p.dim ? __ldg(qv + b + 2) : 0.f;
      qreg[c].w = b + 3 < p.dim ? __ldg(qv + b + 3) : 0.f;
    }
  }

  // ------------------------------------------------------------ distances through the ring
  // rows ci[gi*RPS ..) -> stage st
  __device__ __forceinline__ void issue(const uint32_t *ci, uint32_t gi, uint32_t st, uint32_t m) {
    const uint32_t first = gi * RPS;
    const uint32_t cnt = m - first < (uint32_t)RPS ? m - first : (uint32_t)RPS;
    if (lane == 0) mbar_expect_tx(bar_s + 8 * st, cnt * rowbytes);
    __syncwarp();
    if ((uint32_t)lane < cnt)
      bulk_g2s(ring_s + (st * RPS + lane) * rowbytes, p.vec + (size_t)ci[first + lane] * p.stride, rowbytes, bar_s + 8 * st);
  }

  // distances of ci[0..m) -> cd[0..m) for the groups gi = first, first + step, ... (warp mode: every group; block
  // mode: first = warp, step = W2)
  __device__ __forceinline__ void eval_share(const uint32_t *ci, float *cd, uint32_t m, uint32_t first, uint32_t step) {
    const uint32_t G = (m + RPS - 1) / RPS;
    const uint32_t nst = p.nst;
    uint32_t nxt = first, st_i = 0;
    for (; st_i < nst && nxt < G; st_i++, nxt += step) issue(ci, nxt, st_i, m);
    uint32_t st = 0;
    for (uint32_t gi = first; gi < G; gi += step) {
      mbar_wait(bar_s + 8 * st, (phases >> st) & 1u);
      phases ^= 1u << st;
      const uint32_t i0 = gi * RPS + g, i1 = i0 + TPW;
      const float4 *r0 = reinterpret_cast<const float4 *>(ring + (size_t)(st * RPS + g) * p.stride);
      const float4 *r1 = reinterpret_cast<const float4 *>(ring + (size_t)(st * RPS + TPW + g) * p.stride);
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int c = 0; c < NCH; c++) {
        const int ch = t + c * TEAM;
        if (ch < p.nch) {
          const float4 v0 = r0[ch], v1 = r1[ch];
          a0 = acc4<IP>(a0, qreg[c], v0);
          a1 = acc4<IP>(a1, qreg[c], v1);
        }
      }
      a0 = team_reduce<TEAM>(a0);
      a1 = team_reduce<TEAM>(a1);
      if (t == 0) {
        if (i0 < m) cd[i0] = IP ? 1.0f - a0 : a0;
        if (i1 < m) cd[i1] = IP ? 1.0f - a1 : a1;
      }
      __syncwarp();  // every lane is done with stage st before it is refilled
      if (nxt < G) { issue(ci, nxt, st, m); nxt += step; }
      st = st + 1 == nst ? 0 : st + 1;
    }
  }

  // distances of cand_i[0..m) -> cand_d[0..m): by this warp (warp mode) or by the whole block (block mode, owner)
  __device__ __forceinline__ void eval(uint32_t m) {
    __syncwarp();
    if (coop) {
      if (lane == 0) cmd[0] = m;
      __syncthreads();  // A: the command and cand_i are visible to the block
      eval_share(cand0_i, cand0_d, m, 0, W2);
      __syncthreads();  // B: cand_d is complete
    } else {
      eval_share(cand_i, cand_d, m, 0, 1);
      __syncwarp();
    }
    n_dist += m;
  }

  // block mode, warps 1 .. W2-1: serve the owner's commands until it posts CMD_EXIT
  __device__ __forceinline__ void follow() {
    for (;;) {
      __syncthreads();  // A
      const uint32_t c = cmd[0];
      if (c == CMD_EXIT) break;
      if (c == CMD_QUERY) {
        load_query(p.q + (size_t)cmd[1] * p.qstride);
        for (uint32_t i = threadIdx.x; i <= hmask; i += W2 * 32) tab0[i] = INVALID_ID;
      } else {
        eval_share(cand0_i, cand0_d, c, warp, W2);
      }
      __syncthreads();  // B
    }
  }
  // block mode, owner: the queue is empty
  __device__ __forceinline__ void finish() {
    __syncwarp();
    if (lane == 0) cmd[0] = CMD_EXIT;
    __syncthreads();  // A
  }

  // copy an adjacency row into cand_i; with `filter`, drop ids already in the visited set.
  __device__ __forceinline__ uint32_t load_row(const uint32_t *row, uint32_t rowlen, uint32_t limit, bool filter) {
    uint32_t m = 0;
    __syncwarp();
    for (uint32_t j0 = 0; j0 < rowlen; j0 += 32) {
      uint32_t j = j0 + lane;
      uint32_t id = (j < rowlen && j < limit) ? __ldg(row + j) : INVALID_ID;
      bool valid = id != INVALID_ID;
      if (filter) {
        const bool chk = valid && id != never;
        valid = !visit(id, chk) && chk;
      }
      unsigned mask = __ballot_sync(FULL, valid);
      if (valid) cand_i[m + __popc(mask & ((1u << lane) - 1))] = id;
      m += __popc(mask);
      if (__any_sync(FULL, id == INVALID_ID)) break;
    }
    __syncwarp();
    return m;
  }

  // ------------------------------------------------------------ pool
  // Register form (PR = 1, 2): entry e lives in lane e % 32 of pd0/pi0 (e < 32) or pd1/pi1. Named scalars, not arrays:
  // an array whose element is picked by a run-time value ends up in local memory.
  __device__ __forceinline__ float pool_get_d(uint32_t e) const {
    if constexpr (PR == 0) {
      return pool_d[e];
    } else {
      float v = __shfl_sync(FULL, pd0, e & 31);
      if constexpr (PR == 2) {
        const float w = __shfl_sync(FULL, pd1, e & 31);
        v = e >= 32 ? w : v;
      }
      return v;
    }
  }
  __device__ __forceinline__ uint32_t pool_get_i(uint32_t e) const {
    if constexpr (PR == 0) {
      return pool_i[e];
    } else {
      uint32_t v = __shfl_sync(FULL, pi0, e & 31);
      if constexpr (PR == 2) {
        const uint32_t w = __shfl_sync(FULL, pi1, e & 31);
        v = e >= 32 ? w : v;
      }
      return v;
    }
  }
  __device__ __forceinline__ void pool_clear_flag(uint32_t e) {
    if constexpr (PR == 0) {
      __syncwarp();
      if (lane == 0) pool_i[e] &= ~FLAG;
      __syncwarp();
    } else {
      const bool mine = (e & 31) == (uint32_t)lane;
      pi0 &= ~((mine && e < 32) ? FLAG : 0u);
      if constexpr (PR == 2) pi1 &= ~((mine && e >= 32) ? FLAG : 0u);
    }
  }
  // index of the first unexpanded entry at or after `from`, np if there is none
  __device__ __forceinline__ uint32_t first_flagged(uint32_t from) const {
    if constexpr (PR == 0) {
      for (uint32_t j0 = from & ~31u; j0 < np; j0 += 32) {
        uint32_t j = j0 + lane;
        bool f = j >= from && j < np && (pool_i[j] & FLAG);
        unsigned m = __ballot_sync(FULL, f);
        if (m) return j0 + __ffs(m) - 1;
      }
      return np;
    } else {
      const uint32_t e0 = lane;
      unsigned m = __ballot_sync(FULL, e0 >= from && e0 < np && (pi0 & FLAG));
      if (m) return __ffs(m) - 1;
      if constexpr (PR == 2) {
        const uint32_t e1 = 32 + lane;
        m = __ballot_sync(FULL, e1 >= from && e1 < np && (pi1 & FLAG));
        if (m) return 32 + __ffs(m) - 1;
      }
      return np;
    }
  }
  // pool[0..m) <- the sort buffer (pool_d / pool_i in shared memory)
  __device__ __forceinline__ void pool_from_smem(uint32_t m) {
    np = m;
    __syncwarp();
    if constexpr (PR > 0) {
      pd0 = (uint32_t)lane < m ? pool_d[lane] : INFINITY;
      pi0 = (uint32_t)lane < m ? pool_i[lane] : 0u;
      if constexpr (PR == 2) {
        pd1 = (uint32_t)lane + 32 < m ? pool_d[lane + 32] : INFINITY;
        pi1 = (uint32_t)lane + 32 < m ? pool_i[lane + 32] : 0u;
      }
      // the buffer is ring memory: order these generic-proxy accesses before the bulk copies (async proxy) that follow
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }
  }

  // sorted insert. NSG: after every entry with distance <= d (InsertIntoPool, neighbor.h:107-135);
  // HNSW: (distance,id) lexicographic. Returns the position, or BIGPOS for a duplicate.
  __device__ __forceinline__ int insert(float d, uint32_t id) {
    if constexpr (PR == 0) {
      int pos = 0;
      bool dup = false;
      for (uint32_t j0 = 0; j0 < np; j0 += 32) {
        uint32_t j = j0 + lane;
        bool in = j < np;
        float pdj = in ? pool_d[j] : 0.f;
        uint32_t pij = in ? (pool_i[j] & ~FLAG) : INVALID_ID;
        bool before = in && (HNSW ? (pdj < d || (pdj == d && pij < id)) : (pdj <= d));
        dup |= in && pdj == d && pij == id;
        pos += __popc(__ballot_sync(FULL, before));
      }
      if (__any_sync(FULL, dup)) return BIGPOS;
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
    } else {
      const bool in0 = (uint32_t)lane < np;
      const uint32_t id0 = pi0 & ~FLAG;
      const bool before0 = in0 && (HNSW ? (pd0 < d || (pd0 == d && id0 < id)) : (pd0 <= d));
      bool dup = in0 && pd0 == d && id0 == id;
      int pos = __popc(__ballot_sync(FULL, before0));
      if constexpr (PR == 2) {
        const bool in1 = (uint32_t)lane + 32 < np;
        const uint32_t id1 = pi1 & ~FLAG;
        const bool before1 = in1 && (HNSW ? (pd1 < d || (pd1 == d && id1 < id)) : (pd1 <= d));
        dup |= in1 && pd1 == d && id1 == id;
        pos += __popc(__ballot_sync(FULL, before1));
      }
      if (__any_sync(FULL, dup)) return BIGPOS;
      const int end = (np < cap) ? (int)np : (int)cap - 1;  // entries [pos,end) move up by one
      const float s0d = __shfl_up_sync(FULL, pd0, 1);
      const uint32_t s0i = __shfl_up_sync(FULL, pi0, 1);
      if constexpr (PR == 2) {
        float s1d = __shfl_up_sync(FULL, pd1, 1);
        uint32_t s1i = __shfl_up_sync(FULL, pi1, 1);
        const float cd = __shfl_sync(FULL, pd0, 31);
        const uint32_t ci = __shfl_sync(FULL, pi0, 31);
        if (lane == 0) { s1d = cd; s1i = ci; }
        const int e1 = 32 + lane;
        if (e1 == pos) { pd1 = d; pi1 = id | FLAG; }
        else if (e1 > pos && e1 <= end) { pd1 = s1d; pi1 = s1i; }
      }
      if (lane == pos) { pd0 = d; pi0 = id | FLAG; }
      else if (lane > pos && lane <= end) { pd0 = s0d; pi0 = s0i; }
      if (np < cap) np++;
      return pos;
    }
  }

  // query qi -> registers (any alignment / stride), empty visited set. Block mode: every warp loads the query, the
  // block clears the owner's table.
  __device__ __forceinline__ void begin(uint32_t qi) {
    __syncwarp();
    if (coop) {
      if (lane == 0) { cmd[1] = qi; cmd[0] = CMD_QUERY; }
      __syncthreads();  // A
      load_query(p.q + (size_t)qi * p.qstride);
      for (uint32_t i = threadIdx.x; i <= hmask; i += W2 * 32) tab0[i] = INVALID_ID;
      __syncthreads();  // B
    } else {
      load_query(p.q + (size_t)qi * p.qstride);
      for (uint32_t i = lane; i <= hmask; i += 32) tab[i] = INVALID_ID;
      __syncwarp();
    }
    n_dist = 0; n_hops = 0;
    cap = p.L;
    np = 0;
    never = INVALID_ID;
  }

  // the greedy walk on the HNSW levels >= 1 (hnswalg.h:1275-1303; index_hnsw_nsg.h:111-143): the node it ends on
  __device__ __forceinline__ uint32_t descend_upper(uint32_t start, float &curdist) {
    uint32_t cur = start;
    if (lane == 0) cand_i[0] = cur;
    eval(1);
    curdist = cand_d[0];
    for (int level = p.maxlevel; level > 0; level--) {
      bool changed = true;
      while (changed) {
        changed = false;
        const uint32_t *row = p.upper_links + ((size_t)__ldg(p.upper_ptr + cur) + (level - 1)) * p.maxM;
        uint32_t m = load_row(row, p.maxM, p.maxM, false);
        n_hops++;
        eval(m);
        float bd = INFINITY; uint32_t bi = INVALID_ID;
        for (uint32_t c = lane; c < m; c += 32) {
          float d = cand_d[c];
          if (d < bd) { bd = d; bi = c; }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
          float od = __shfl_xor_sync(FULL, bd, off);
          uint32_t oi = __shfl_xor_sync(FULL, bi, off);
          if (od < bd || (od == bd && oi < bi)) { bd = od; bi = oi; }
        }
        if (bi != INVALID_ID && bd < curdist) { curdist = bd; cur = cand_i[bi]; changed = true; }
      }
    }
    __syncwarp();
    return cur;
  }

  // HNSW: the base layer starts from the one entry the descent ends on (hnswalg.h:323-340)
  __device__ __forceinline__ void descend_hnsw() {
    float curdist;
    uint32_t cur = descend_upper(p.ep, curdist);
    __syncwarp();
    if (lane == 0) { pool_d[0] = curdist; pool_i[0] = cur | FLAG; }
    visit_uniform(cur);
    pool_from_smem(1);
  }

  // rank sort of cand_i/cand_d[0..m) into the pool: std::sort(retset) as (distance,id,position)
  __device__ __forceinline__ void seeds_to_pool(uint32_t m, bool flag_ids) {
    for (uint32_t i0 = 0; i0 < m; i0 += 32) {
      const uint32_t i = i0 + lane;
      const bool in = i < m;
      const float di = in ? cand_d[i] : 0.f;
      const uint32_t ii = in ? cand_i[i] : INVALID_ID;
      if (in) {
        uint32_t r = 0;
        for (uint32_t j = 0; j < m; j++) {
          float dj = cand_d[j]; uint32_t ij = cand_i[j];
          r += (dj < di || (dj == di && (ij < ii || (ij == ii && j < i)))) ? 1u : 0u;
        }
        pool_d[r] = di; pool_i[r] = ii | FLAG;
      }
      if (flag_ids) visit(ii, in);
    }
    pool_from_smem(m);
  }

  // seeds of IndexNSG::SearchFromEnterpoint (hnsw_nsg/src/index_nsg.cpp:703-749), quirks included (see beam.cuh)
  __device__ __forceinline__ void seed_from(uint32_t ep, uint32_t qidx) {
    never = ep;
    visit_uniform(ep);
    uint32_t m = load_row(p.graph + (size_t)ep * p.gstride, p.gstride, p.L, false);
    for (uint32_t j0 = 0; j0 < m; j0 += 32) {
      const uint32_t j = j0 + lane;
      visit(j < m ? cand_i[j] : INVALID_ID, j < m);
    }
    __syncwarp();
    const uint32_t cur = m;
    uint32_t lim = p.L < p.gstride ? p.L : p.gstride;
    for (uint32_t i = 0; i < cur && m < p.L; i++) {
      const uint32_t *row = p.graph + (size_t)cand_i[i] * p.gstride;
      bool ended = false;
      while (m < lim && !ended) {
        uint32_t j = m + lane;
        uint32_t v = j < lim ? __ldg(row + j) : INVALID_ID;
        unsigned inv = __ballot_sync(FULL, v == INVALID_ID);
        uint32_t cnt = inv ? (uint32_t)(__ffs(inv) - 1) : 32u;
        const bool was = visit(v, (uint32_t)lane < cnt);
        if ((uint32_t)lane < cnt) cand_i[j] = was ? 0u : v;
        m += cnt;
        ended = inv != 0;
        __syncwarp();
      }
    }
    uint32_t ctr = 0;
    while (m < p.L) {
      uint32_t id = seed_hash(p.seed, qidx, ctr++) % p.n;
      if (!visit_uniform(id)) {
        if (lane == 0) cand_i[m] = id;
        m++;
      }
    }
    __syncwarp();
    eval(m);
    seeds_to_pool(m, false);
  }

  // NSG seeds: first L out-neighbours of ep (index_nsg.cpp:516-519), then pseudo-random ids not yet chosen (:521-527)
  __device__ __forceinline__ void seed_nsg(uint32_t qidx) {
    uint32_t m = load_row(p.graph + (size_t)p.ep * p.gstride, p.gstride, p.L, false);
    uint32_t ctr = 0;
    while (m < p.L) {
      uint32_t id = seed_hash(p.seed, qidx, ctr++) % p.n;
      bool hit = false;
      for (uint32_t j = lane; j < m; j += 32) hit |= cand_i[j] == id;
      if (!__any_sync(FULL, hit)) {
        if (lane == 0) cand_i[m] = id;
        m++;
      }
      __syncwarp();
    }
    eval(m);
    seeds_to_pool(m, true);
  }

  // seed_nsg for a query confined to the rows [base, base + size) of a block-diagonal graph (see beam.cuh)
  __device__ __forceinline__ void seed_nsg_local(uint32_t qidx, uint32_t ep, uint32_t base, uint32_t size) {
    const uint32_t Lq = p.L < size ? p.L : size;
    uint32_t m = load_row(p.graph + (size_t)ep * p.gstride, p.gstride, Lq, false);
    uint32_t ctr = 0;
    while (m < Lq) {
      uint32_t id = base + seed_hash(p.seed, qidx, ctr++) % size;
      bool hit = false;
      for (uint32_t j = lane; j < m; j += 32) hit |= cand_i[j] == id;
      if (!__any_sync(FULL, hit)) {
        if (lane == 0) cand_i[m] = id;
        m++;
      }
      __syncwarp();
    }
    eval(m);
    seeds_to_pool(m, true);
    cap = Lq;
  }

  // main loop (index_nsg.cpp:539-564 / hnswalg.h:342-436). The reference walks k upward and rewinds it to the
  // smallest insert position; every entry below k is expanded at all times, so "the entry the loop turns to next"
  // is simply the first unexpanded entry of the pool.
  __device__ __forceinline__ void expand_all() {
    uint32_t k = 0;
    for (;;) {
      k = first_flagged(k);
      if (k >= np) break;
      const uint32_t node = pool_get_i(k) & ~FLAG;
      pool_clear_flag(k);
      n_hops++;
      if (p.prefetch) {
        uint32_t k2 = first_flagged(k + 1);
        if (k2 < np) {
          const uint32_t *row2 = p.graph + (size_t)(pool_get_i(k2) & ~FLAG) * p.gstride;
          if ((uint32_t)lane * 32u < p.gstride) asm volatile("prefetch.global.L2 [%0];" ::"l"(row2 + lane * 32));
        }
      }
      const uint32_t m = load_row(p.graph + (size_t)node * p.gstride, p.gstride, p.gstride, true);
      eval(m);
      uint32_t nk = k + 1;
      for (uint32_t c0 = 0; c0 < m; c0 += 32) {
        uint32_t c = c0 + lane;
        float d = c < m ? cand_d[c] : INFINITY;
        uint32_t myid = c < m ? cand_i[c] : INVALID_ID;
        float last = pool_get_d(np - 1);
        bool pass = c < m && ((HNSW && np < cap) || d < last);
        unsigned mask = __ballot_sync(FULL, pass);
        while (mask) {
          int b = __ffs(mask) - 1;
          mask &= mask - 1;
          float dd = __shfl_sync(FULL, d, b);
          uint32_t id = __shfl_sync(FULL, myid, b);
          last = pool_get_d(np - 1);
          if (!((HNSW && np < cap) || dd < last)) continue;  // threshold moved since the vote
          int r = insert(dd, id);
          if ((uint32_t)r < nk) nk = (uint32_t)r;
        }
      }
      k = nk;  // nothing below min(insert positions, k + 1) is unexpanded
    }
    __syncwarp();
  }

  // results: the first K pool entries (ids without the flag bit; INVALID / +inf past the pool)
  template <class F>
  __device__ __forceinline__ void for_results(uint32_t K, F f) const {
    for (uint32_t e = lane; e < K; e += 32) {  // one call site: the writer is inlined once
      const bool ok = e < np;
      uint32_t id = INVALID_ID;
      float d = INFINITY;
      if constexpr (PR == 0) {
        if (ok) { id = pool_i[e] & ~FLAG; d = pool_d[e]; }
      } else {
        if (ok) { id = (e < 32 ? pi0 : pi1) & ~FLAG; d = e < 32 ? pd0 : pd1; }
      }
      f(e, id, d, ok);
    }
  }
};

}  // namespace beam2

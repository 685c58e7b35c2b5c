// Search kernel v2 over beam2.cuh: persistent, one query per warp and - for the tail of a batch - per block of 4
// warps, rows through a TMA bulk-copy ring, pool in registers for L <= 64. Included by search2_pr{0,1,2}.cu with SEARCH2_PR defined (one translation unit per pool
// form keeps the build parallel). Same entry semantics as search.cu's v1 kernel (modes 0 / 2 / 3 / 4 and HNSW).
#include "beam2.cuh"

using namespace beam2;

namespace {

constexpr int WARPS2 = W2;

// Two phases over two work queues (see beam2.cuh): work items [0, q1) are searched one per WARP - the throughput
// form, every warp of the resident wave busy - and the remainder [q1, nq), the part of a batch that would leave most of
// the GPU idle behind a last partial round of warps, one per BLOCK with every distance evaluation spread over the
// block's warps (a query is ~2x shorter that way). The launcher picks q1; results do not depend on it.
// registers: up to 128-dim rows (and 512-dim rows with a warp per row) fit 64 without spills -> 8 blocks per SM
template <int TEAM, int NCH, bool IP, bool HNSW, int PR>
__global__ void __launch_bounds__(WARPS2 * 32, (NCH <= 4 ? 8 : 4)) beam_search2_kernel(const SearchParams p) {
  extern __shared__ __align__(16) uint32_t smem2[];
  Beam2<TEAM, NCH, IP, HNSW, PR> b(p, smem2);
  for (int phase = 0; phase < 2; phase++) {
    if (phase == 1) {
      if (p.q1 >= p.nq) return;
      __syncthreads();  // every warp of the block is done with its own queries
      b.coop = true;
      if (b.warp != 0) {
        b.follow();
        return;
      }
    }
    const uint32_t lo = phase == 0 ? 0u : p.q1, hi = phase == 0 ? p.q1 : p.nq;
    for (;;) {
      uint32_t qi = 0;
      if (b.lane == 0) qi = atomicAdd(p.work + phase, 1u);
      qi = lo + __shfl_sync(0xffffffffu, qi, 0);
      if (qi >= hi) break;
      if (p.order) qi = __ldg(p.order + qi);
      if (!HNSW && p.mode == 4 && __ldg(p.cl_size + qi) == 0) {  // query not routed to any cluster in this pass
        for (uint32_t i = b.lane; i < p.K; i += 32) {
          p.ids[(size_t)qi * p.K + i] = INVALID_ID;
          if (p.dists) p.dists[(size_t)qi * p.K + i] = INFINITY;
        }
        if (p.stats && b.lane == 0) { p.stats[2 * (size_t)qi] = 0; p.stats[2 * (size_t)qi + 1] = 0; }
        continue;
      }
      b.begin(qi);
      if (HNSW) {
        b.descend_hnsw();
      } else if (p.mode == 0) {
        b.seed_nsg(p.qidx_base + qi);
      } else if (p.mode == 4) {
        b.seed_nsg_local(p.qidx_base + qi, __ldg(p.entry + qi), __ldg(p.cl_base + qi), __ldg(p.cl_size + qi));
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
      b.expand_all();

      b.for_results(p.K, [&](uint32_t i, uint32_t id, float d, bool ok) {
        if (HNSW) {
          p.labels_out[(size_t)qi * p.K + i] = ok ? (p.labels ? __ldg(p.labels + id) : (uint64_t)id) : ~(uint64_t)0;
        } else {
          p.ids[(size_t)qi * p.K + i] = id + (ok ? p.id_base : 0u);
        }
        if (p.dists) p.dists[(size_t)qi * p.K + i] = d;
      });
      if (p.stats && b.lane == 0) {
        p.stats[2 * (size_t)qi] = b.n_dist;
        p.stats[2 * (size_t)qi + 1] = b.n_hops;
      }
      __syncwarp();
    }
  }
  b.finish();
}

template <int TEAM, int NCH, bool IP, bool HNSW, int PR>
int launch2_t(b200ann_handle *h, const SearchParams &p0) {
  SearchParams p = p0;
  constexpr int RPS = 2 * (32 / TEAM);
  auto words = [&](uint32_t nst, int hbits) { return smem_words2(p.pool_cap, p.cand_cap, hbits, nst * RPS * (uint32_t)p.stride, nst, PR > 0); };
  // Shared memory decides how many blocks (= queries) an SM holds, and the kernel is bound by latency per query: the
  // defaults aim at the most blocks the registers allow (8 for rows up to 128 floats, else 4) with a one-stage ring per
  // warp, and give the visited set what is left (2^9 .. 2^12 words; a smaller table only re-evaluates a few
  // forgotten nodes).
  const size_t target_blocks = NCH <= 4 ? 8 : 4;
  const size_t budget = (227 * 1024 / target_blocks - 1024) / 4;  // words per block (the driver reserves 1 KB per block)
  p.nst = (uint32_t)(h->ring_stages > 0 ? h->ring_stages : 1);
  if (p.nst > 8) p.nst = 8;
  if (h->visited_bits_auto) {
    p.hbits = 11;
    while (p.hbits > 8 && words(p.nst, p.hbits) > budget) p.hbits--;
    p.nbits = p.nbits < p.hbits + 1 ? p.hbits + 1 : p.nbits;
    p.vis16 = (h->visited_compact && p.nbits - (p.hbits + 1) <= 12) ? 1 : 0;
  }
  while (p.nst > 1 && (size_t)words(p.nst, p.hbits) * 4 > 227 * 1024) p.nst--;
  size_t smem = (size_t)words(p.nst, p.hbits) * 4;
  if (smem > 227 * 1024) return set_err(h, B200ANN_ERR_ARG, "search: L / degree / visited_bits / dim need more than 227 KB shared memory per block");
  auto kern = beam_search2_kernel<TEAM, NCH, IP, HNSW, PR>;
  CU_CHECK(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  CU_CHECK(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WARPS2 * 32, smem));
  if (occ < 1) occ = 1;
  if (h->search_ctas_per_sm > 0 && h->search_ctas_per_sm < occ) occ = h->search_ctas_per_sm;
  const size_t wave = (size_t)occ * (size_t)h->sm_count, slots = wave * WARPS2;
  // q1: full rounds of the resident warps run one query per warp; a remainder of at most one round of BLOCKS runs one
  // query per block instead of leaving the GPU mostly idle behind a partial round of warps
  size_t q1 = p.nq;
  if (h->search_mode == 2) q1 = 0;
  else if (h->search_mode == 0) {
    const size_t rem = p.nq % slots;
    if (rem > 0 && rem <= wave) q1 = p.nq - rem;
    if (h->search_tail >= 0) q1 = p.nq > (size_t)h->search_tail ? p.nq - (size_t)h->search_tail : 0;  // experiments
  }
  p.q1 = (uint32_t)q1;
  size_t blocks = (q1 + WARPS2 - 1) / WARPS2;
  if (p.nq - q1 > blocks) blocks = p.nq - q1;
  if (blocks > wave) blocks = wave;
  if (!h->work) CU_CHECK(h, cudaMalloc((void **)&h->work, 2 * sizeof(uint32_t)));
  CU_CHECK(h, cudaMemsetAsync(h->work, 0, 2 * sizeof(uint32_t), h->stream));
  p.work = h->work;
  kern<<<dim3((unsigned)blocks), dim3(WARPS2 * 32), smem, h->stream>>>(p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

template <bool IP, bool HNSW, int PR>
int launch2_dim(b200ann_handle *h, const SearchParams &p) {
  int nch = p.nch;
  if (nch <= 8) return launch2_t<8, 1, IP, HNSW, PR>(h, p);
  if (nch <= 16) return launch2_t<8, 2, IP, HNSW, PR>(h, p);
  if (nch <= 32) return launch2_t<8, 4, IP, HNSW, PR>(h, p);
  if (nch <= 64) return launch2_t<8, 8, IP, HNSW, PR>(h, p);
  if (nch <= 128) return launch2_t<32, 4, IP, HNSW, PR>(h, p);
  if (nch <= 256) return launch2_t<32, 8, IP, HNSW, PR>(h, p);
  if (nch <= 512) return launch2_t<32, 16, IP, HNSW, PR>(h, p);
  return set_err(h, B200ANN_ERR_ARG, "search: dim > 2048 is not supported");
}

}  // namespace

#define SEARCH2_CAT_(a, b) a##b
#define SEARCH2_CAT(a, b) SEARCH2_CAT_(a, b)
int SEARCH2_CAT(launch_search2_pr, SEARCH2_PR)(b200ann_handle *h, const SearchParams &p, bool ip, bool hnsw) {
  if (ip) return hnsw ? launch2_dim<true, true, SEARCH2_PR>(h, p) : launch2_dim<true, false, SEARCH2_PR>(h, p);
  return hnsw ? launch2_dim<false, true, SEARCH2_PR>(h, p) : launch2_dim<false, false, SEARCH2_PR>(h, p);
}

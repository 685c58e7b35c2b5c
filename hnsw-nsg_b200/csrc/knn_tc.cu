// Exact brute-force kNN on the 5th-generation tensor cores (sm_100a: tcgen05.mma + TMEM + TMA).
//
// Replaces, like bf_knn.cu, hnswlib::BruteforceSearch::searchKnn (hnswlib/hnswlib/bruteforce.h:106-135),
// IndexGraph::generate_control_set (CNNS/efanna_graph/src/index_graph.cpp:138-153) and, with
// exclude_self, the exact form of IndexGraph::Build's output (:265-294). Same results as bf_knn.cu bit
// for bit; only the dense contraction runs on the tensor pipe:
//
//   1. prep      fp32 rows -> bf16 "hi" (+ "lo" = bf16(x - hi) when some value is not bf16-exact),
//                fp32 row norms, max norm.
//   2. sweep     one CTA = 128 query rows x a range of 256-row base tiles. TMA (128B swizzle) stages the
//                bf16 K-blocks, one thread issues tcgen05.mma (M128 N256 K16, fp32 accumulators in TMEM,
//                two accumulator stages). The base norm rides the contraction (one more 16-column K step:
//                -||y||^2/2 as three bf16 terms against 1.0 columns), so the accumulator IS the ranking key
//                ||y||^2 - 2 x.y  (L2) /  -2 x.y  (IP) times -1/2. Eight epilogue warps read TMEM (thread =
//                query row, two warps per lane quarter split the columns) and keep, per row, a SHORTLIST of the
//                K' = K + slack smallest keys: one compare per element + warp-cooperative radix-select
//                compaction. The epilogue runs at the TMEM read pipe's 64 B/clk (2048 cycles per 128 KB tile).
//                Inexact data uses three passes hi*hi + hi*lo + lo*hi (bf16x3, error ~2^-16).
//   3. finalize  CTA per query row: fp32 distances of the shortlisted rows in the canonical sequential-fma
//                order of bf_knn.cu, bitonic sort by (distance,id), top K out. PROOF of exactness per row:
//                every row outside the shortlist has key >= tau (the filter threshold), so if
//                tau - eps > K-th exact distance (eps = rigorous error bound of the tensor-core key) the
//                shortlist provably contains the true top K. Rows that fail the proof are recomputed by
//                the fp32 SIMT kernel (launch_bf_knn) - results are exact either way.
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"

namespace tc {

constexpr int BM = 128, BN = 256, BK = 64;          // tile rows (queries), tile columns (base rows), bf16 per K-block
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;  // 16 KB, 32 KB
constexpr int A_EXT_BYTES = BM * 32, B_EXT_BYTES = BN * 32;  // the norm step's tiles: 4 KB, 8 KB
constexpr int A_SLOTS = 5;                          // at most this many resident A K-blocks (80 KB): 128-d hi + lo + the norm block
constexpr int NTHREADS = 320;                       // warp0 TMA, warp1 MMA, warps 2-9 epilogue
constexpr int SMEM_MISC = 49152;                    // spill 36 KB, norms 2 KB, histograms 8 KB, row state 1 KB, barriers
constexpr int SMEM_MAX = 232448;                    // 227 KB opt-in limit of sm_100

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// the same for the producer-side waits (a free smem stage, a free accumulator): these warps are ahead of the epilogue
// most of the time and share their schedulers with epilogue warps, so they back off instead of spinning (round-2
// profile: 21 % of all instructions the kernel executed were these two spin loops)
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t *bar, uint32_t parity) {
  uint32_t ok = 0;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) break;
    __nanosleep(200);
  }
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16 inputs with fp32 accumulation
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major, 128B-swizzled operand tile: rows are 128 B apart, 8-row groups 1024 B apart (SBO), LBO unused
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;                 // leading byte offset (ignored for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;       // stride byte offset
  d |= (uint64_t)1 << 46;                 // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                 // SWIZZLE_128B
  return d;
}
// K-major, 32B-swizzled operand tile of ONE K step (16 bf16 = 32 B per row): 8-row groups 256 B apart
__device__ __forceinline__ uint64_t smem_desc32(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(256 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)6 << 61;                 // SWIZZLE_32B
  return d;
}
// instruction descriptor: D=f32, A=B=bf16, both K-major, N=256, M=128
constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

#define TC_LD32(taddr, v)                                                                                           \
  asm volatile(                                                                                                     \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                     \
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28," \
      "%29,%30,%31}, [%32];"                                                                                        \
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), \
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),      \
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),     \
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                   \
      : "r"(taddr))
// wait for the TMEM loads; the registers are operands so that no use of them can be scheduled above it
#define TC_LD_WAIT(v)                                                                                               \
  asm volatile("tcgen05.wait::ld.sync.aligned;"                                                                     \
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),    \
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]),           \
                 "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]),         \
                 "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]),         \
                 "+r"(v[29]), "+r"(v[30]), "+r"(v[31])::"memory")

// ---------------------------------------------------------------- prep: fp32 -> bf16 hi/lo + norms
// flags[0] |= 1 when some value is not exactly a bf16; flags[1] = max ||row||^2 (as ordered uint bits)
__global__ void prep_kernel(const float *src, size_t sstride, uint32_t n, uint32_t dim, uint32_t dpad, __nv_bfloat16 *hi,
                            __nv_bfloat16 *lo, float *norms, uint32_t *flags) {
  const uint32_t row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= n) return;
  float s = 0.f;
  bool inexact = false;
  for (uint32_t d = lane; d < dpad; d += 32) {
    float x = d < dim ? src[(size_t)row * sstride + d] : 0.f;
    __nv_bfloat16 h = __float2bfloat16_rn(x);
    float r = x - __bfloat162float(h);
    __nv_bfloat16 l = __float2bfloat16_rn(r);
    hi[(size_t)row * dpad + d] = h;
    lo[(size_t)row * dpad + d] = l;
    inexact |= (r != 0.f);
    s = fmaf(x, x, s);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  inexact = __any_sync(0xffffffffu, inexact);
  if (lane == 0) {
    norms[row] = s;
    if (inexact) atomicOr(&flags[0], 1u);
    atomicMax(&flags[1], __float_as_uint(s));  // s >= 0: uint order == float order
  }
}

// The norm rides the contraction: one more K step (16 columns, a 32-byte row, 32B swizzle) whose B rows hold -||y||^2 / 2 as three bf16 terms (columns 0-2;
// hi + mid + lo carry the 24 significant bits of the fp32 norm) and whose A rows hold 1.0 in those columns, so the
// accumulator ends as  x.y - ||y||^2 / 2 = -key / 2  and the epilogue's filter is one compare per element.
__global__ void ext_b_kernel(const float *norms, uint32_t n, __nv_bfloat16 *ext) {   // [n, 16]
  const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n) return;
  __nv_bfloat16 v[16];
#pragma unroll
  for (int c = 0; c < 16; c++) v[c] = __float2bfloat16_rn(0.f);
  const float x = -0.5f * norms[row];
  const __nv_bfloat16 h = __float2bfloat16_rn(x);
  const float r1 = x - __bfloat162float(h);
  const __nv_bfloat16 m = __float2bfloat16_rn(r1);
  const float r2 = r1 - __bfloat162float(m);
  v[0] = h; v[1] = m; v[2] = __float2bfloat16_rn(r2);
  uint4 *dst = reinterpret_cast<uint4 *>(ext + (size_t)row * 16);
  dst[0] = reinterpret_cast<const uint4 *>(v)[0];
  dst[1] = reinterpret_cast<const uint4 *>(v)[1];
}
__global__ void ext_a_kernel(__nv_bfloat16 *ext) {   // [128, 16]
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 128u * 16u) ext[i] = __float2bfloat16_rn((i & 15u) < 3u ? 1.f : 0.f);
}

// ---------------------------------------------------------------- sweep kernel
struct SweepParams {
  uint32_t nq_rows;   // query rows of this launch
  uint32_t q_row0;    // row of the query bf16 arrays the launch starts at
  uint32_t self_base; // exclude_self: query row i is base row self_base + i (0 when the queries are the whole base)
  uint32_t n;         // base rows
  uint32_t nkb;       // K-blocks per pass = dpad / 64
  uint32_t passes;    // 1 (bf16-exact data) or 3 (hi*hi + hi*lo + lo*hi)
  uint32_t a_resident;
  uint32_t a_slots, nstage;  // shared-memory carve: A K-block slots, B (and streamed A) ring stages
  uint32_t ntiles, tiles_per_split;
  uint32_t ext;       // 1 (L2): one more K-block step per tile carries -||y||^2 / 2 (mA_ext / mB_ext); 0: inner product
  int exclude_self;
  uint32_t Kp, cap, rows_pad;
  float *cand_d; uint32_t *cand_i;    // [splits, rows_pad, cap]
  uint32_t *cand_n; float *cand_tau;  // [splits, rows_pad]
  uint32_t tile_step;                 // 1: every tile of the split; s: every s-th tile (the sampling pass)
  const float *tau0;                  // [rows_pad] or NULL: start threshold per row (from the sampling pass)
  unsigned long long *prof;           // optional [8]: per-epilogue-warp cycle sums {wait tmem_full, process, compaction, total}, compactions
};

__device__ __forceinline__ uint32_t ord_key(float f) {
  uint32_t b = __float_as_uint(f);
  return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
}
__device__ __forceinline__ float unord_key(uint32_t u) {
  uint32_t b = (u & 0x80000000u) ? (u ^ 0x80000000u) : ~u;
  return __uint_as_float(b);
}

// Warp-cooperative: keep the Kp smallest keys of cd[0..cnt) (ties: first in buffer order = smallest ids),
// stable, in place. Returns the Kp-th smallest key = the new filter threshold. hist: 256 u32 of shared memory.
__device__ float compact_row(float *cd, uint32_t *ci, uint32_t cnt, uint32_t Kp, uint32_t *hist, int lane) {
  uint32_t prefix = 0, mask = 0, kth = Kp;
  const uint32_t lt_mask = (1u << lane) - 1u;
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int i = lane; i < 256; i += 32) hist[i] = 0;
    __syncwarp();
    for (uint32_t e = lane; e < cnt; e += 32) {
      uint32_t u = ord_key(__ldcg(cd + e));
      if ((u & mask) == prefix) atomicAdd(&hist[(u >> shift) & 255u], 1u);
    }
    __syncwarp();
    uint32_t hb[8], s = 0;
#pragma unroll
    for (int b = 0; b < 8; b++) { hb[b] = hist[8 * lane + b]; s += hb[b]; }
    uint32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    uint32_t excl = incl - s;
    bool mine = excl < kth && kth <= incl;
    uint32_t sel = 0, before = 0;
    if (mine) {
      uint32_t run = excl;
      bool found = false;
#pragma unroll
      for (int b = 0; b < 8; b++) {
        if (!found && run + hb[b] >= kth) { sel = 8 * lane + b; before = run; found = true; }
        run += hb[b];
      }
    }
    int src = __ffs(__ballot_sync(0xffffffffu, mine)) - 1;
    sel = __shfl_sync(0xffffffffu, sel, src);
    before = __shfl_sync(0xffffffffu, before, src);
    kth -= before;
    prefix |= sel << shift;
    mask |= 0xFFu << shift;
    __syncwarp();
  }
  const uint32_t ustar = prefix;  // kth = how many of the entries equal to ustar stay
  uint32_t out = 0, tie_seen = 0;
  for (uint32_t e0 = 0; e0 < cnt; e0 += 32) {
    uint32_t e = e0 + lane;
    bool valid = e < cnt;
    float d = valid ? __ldcg(cd + e) : 0.f;
    uint32_t id = valid ? __ldcg(ci + e) : 0u;
    uint32_t u = ord_key(d);
    bool lt = valid && u < ustar, eq = valid && u == ustar;
    uint32_t beq = __ballot_sync(0xffffffffu, eq);
    bool keep = lt || (eq && tie_seen + __popc(beq & lt_mask) < kth);
    uint32_t bk = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      uint32_t pos = out + __popc(bk & lt_mask);
      __stcg(cd + pos, d);
      __stcg(ci + pos, id);
    }
    out += __popc(bk);
    tie_seen += __popc(beq);
    __syncwarp();
  }
  return unord_key(ustar);
}

// Same contract as compact_row with the row's entries held in registers (EPL per lane, cap <= 32*EPL): one
// round trip to L2 instead of one per radix pass.
template <int EPL>
__device__ float compact_row_reg(float *cd, uint32_t *ci, uint32_t cnt, uint32_t Kp, uint32_t *hist, int lane) {
  uint32_t u[EPL], id[EPL];
#pragma unroll
  for (int i = 0; i < EPL; i++) {
    const uint32_t e = i * 32 + lane;
    u[i] = e < cnt ? ord_key(__ldcg(cd + e)) : 0xFFFFFFFFu;
    id[i] = e < cnt ? __ldcg(ci + e) : 0u;
  }
  uint32_t prefix = 0, mask = 0, kth = Kp;
  const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll 1
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int i = lane; i < 256; i += 32) hist[i] = 0;
    __syncwarp();
#pragma unroll
    for (int i = 0; i < EPL; i++)
      if ((uint32_t)(i * 32 + lane) < cnt && (u[i] & mask) == prefix) atomicAdd(&hist[(u[i] >> shift) & 255u], 1u);
    __syncwarp();
    uint32_t hb[8], s = 0;
#pragma unroll
    for (int b = 0; b < 8; b++) { hb[b] = hist[8 * lane + b]; s += hb[b]; }
    uint32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    uint32_t excl = incl - s;
    bool mine = excl < kth && kth <= incl;
    uint32_t sel = 0, before = 0;
    if (mine) {
      uint32_t run = excl;
      bool found = false;
#pragma unroll
      for (int b = 0; b < 8; b++) {
        if (!found && run + hb[b] >= kth) { sel = 8 * lane + b; before = run; found = true; }
        run += hb[b];
      }
    }
    int src = __ffs(__ballot_sync(0xffffffffu, mine)) - 1;
    sel = __shfl_sync(0xffffffffu, sel, src);
    before = __shfl_sync(0xffffffffu, before, src);
    kth -= before;
    prefix |= sel << shift;
    mask |= 0xFFu << shift;
    __syncwarp();
  }
  const uint32_t ustar = prefix;
  uint32_t out = 0, tie_seen = 0;
#pragma unroll
  for (int i = 0; i < EPL; i++) {
    if ((uint32_t)(i * 32) < cnt) {  // warp-uniform
      const bool valid = (uint32_t)(i * 32 + lane) < cnt;
      const bool lt = valid && u[i] < ustar, eq = valid && u[i] == ustar;
      const uint32_t beq = __ballot_sync(0xffffffffu, eq);
      const bool keep = lt || (eq && tie_seen + __popc(beq & lt_mask) < kth);
      const uint32_t bk = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const uint32_t pos = out + __popc(bk & lt_mask);
        __stcg(cd + pos, unord_key(u[i]));
        __stcg(ci + pos, id[i]);
      }
      out += __popc(bk);
      tie_seen += __popc(beq);
    }
  }
  __syncwarp();
  return unord_key(ustar);
}

// Epilogue geometry: 8 warps. Warp (quarter q = warp & 3, half hf) reads TMEM lanes [32q, 32q+32) (the rows
// it may address) and the columns [128 hf, 128 hf + 128) of every tile; the two warps of a quarter share the
// per-row state (count, threshold) in shared memory and synchronise on a 64-thread named barrier per tile.
__global__ void __launch_bounds__(NTHREADS, 1)
sweep_kernel(const __grid_constant__ CUtensorMap mA_hi, const __grid_constant__ CUtensorMap mA_lo,
             const __grid_constant__ CUtensorMap mB_hi, const __grid_constant__ CUtensorMap mB_lo,
             const __grid_constant__ CUtensorMap mA_ext, const __grid_constant__ CUtensorMap mB_ext, const SweepParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle; plain pointer arithmetic keeps the shared address space
  // known to the compiler (LDS/STS/ATOMS instead of generic accesses)
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t NS = p.nstage;
  uint8_t *sA = smem;                                  // a_slots x 16 KB
  uint8_t *sB = smem + p.a_slots * A_BYTES;            // NS x 32 KB
  uint8_t *misc = sB + NS * B_BYTES;
  float *spill = (float *)misc;                        // [8 warps][32 lanes][36]
  // (misc + 36864: 2 KB once used for the staged norms)
  uint32_t *hist = (uint32_t *)(misc + 36864 + 2048);  // [8][256]
  uint32_t *cnt_s = (uint32_t *)(misc + 36864 + 2048 + 8192);  // [128]
  float *tau_s = (float *)(cnt_s + 128);                       // [128]
  uint64_t *bars = (uint64_t *)(tau_s + 128);
  uint64_t *full = bars, *empty = bars + NS, *a_full = bars + 2 * NS;
  uint64_t *tmem_full = bars + 2 * NS + 1, *tmem_empty = bars + 2 * NS + 3;
  uint32_t *tmem_ptr = (uint32_t *)(bars + 2 * NS + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t m0 = p.q_row0 + blockIdx.x * BM;
  const uint32_t tile_begin = blockIdx.y * p.tiles_per_split;
  const uint32_t tile_end = min(p.ntiles, tile_begin + p.tiles_per_split);
  const uint32_t Tm = p.nkb * p.passes;  // K-block steps of the data per tile
  const uint32_t T = Tm + p.ext;         // + the norm block (last)
  const uint32_t nA = p.nkb * (p.passes == 3 ? 2u : 1u) + p.ext;  // resident A K-blocks (the norm block's ones last)

  if (warp == 0 && lane == 0) {
    for (uint32_t i = 0; i < NS; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(a_full, 1);
    for (int i = 0; i < 2; i++) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mA_hi) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mB_hi) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_ptr)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      if (p.a_resident) {
        mbar_expect_tx(a_full, (nA - p.ext) * A_BYTES + p.ext * A_EXT_BYTES);
        for (uint32_t b = 0; b < nA - p.ext; b++)
          tma_load_2d(sA + b * A_BYTES, b < p.nkb ? &mA_hi : &mA_lo, (int)((b % p.nkb) * BK), (int)m0, a_full);
        if (p.ext) tma_load_2d(sA + (nA - 1) * A_BYTES, &mA_ext, 0, 0, a_full);
      }
      uint32_t stage = 0, phase = 0;
      for (uint32_t tile = tile_begin; tile < tile_end; tile += p.tile_step) {
        const int n0 = (int)(tile * BN);
        for (uint32_t t = 0; t < T; t++) {
          const uint32_t seg = t / p.nkb, kb = t % p.nkb;
          mbar_wait_relaxed(&empty[stage], phase ^ 1u);
          if (t < Tm) mbar_expect_tx(&full[stage], B_BYTES + (p.a_resident ? 0 : A_BYTES));
          else mbar_expect_tx(&full[stage], B_EXT_BYTES + (p.a_resident ? 0 : A_EXT_BYTES));
          if (t < Tm) {
            tma_load_2d(sB + stage * B_BYTES, seg == 1 ? &mB_lo : &mB_hi, (int)(kb * BK), n0, &full[stage]);
            if (!p.a_resident)
              tma_load_2d(sA + stage * A_BYTES, seg == 2 ? &mA_lo : &mA_hi, (int)(kb * BK), (int)m0, &full[stage]);
          } else {
            tma_load_2d(sB + stage * B_BYTES, &mB_ext, 0, n0, &full[stage]);
            if (!p.a_resident) tma_load_2d(sA + stage * A_BYTES, &mA_ext, 0, 0, &full[stage]);
          }
          if (++stage == NS) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (p.a_resident) mbar_wait(a_full, 0);
    uint32_t stage = 0, phase = 0, ti = 0;
    for (uint32_t tile = tile_begin; tile < tile_end; tile += p.tile_step, ti++) {
      const uint32_t as = ti & 1u, aphase = (ti >> 1) & 1u;
      mbar_wait_relaxed(&tmem_empty[as], aphase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (uint32_t t = 0; t < T; t++) {
        const uint32_t seg = t / p.nkb, kb = t % p.nkb;
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t aslot = p.a_resident ? (t < Tm ? (seg == 2 ? p.nkb : 0u) + kb : nA - 1u) : stage;
          const uint32_t a_addr = smem_u32(sA + aslot * A_BYTES), b_addr = smem_u32(sB + stage * B_BYTES);
          if (t < Tm) {
#pragma unroll
            for (int k = 0; k < BK / 16; k++)
              tc_mma(d_tmem, smem_desc(a_addr + k * 32), smem_desc(b_addr + k * 32), IDESC, (t > 0 || k > 0) ? 1u : 0u);
          } else {
            tc_mma(d_tmem, smem_desc32(a_addr), smem_desc32(b_addr), IDESC, 1u);   // the norm step
          }
          tc_commit(&empty[stage]);                   // frees the smem stage once these MMAs have read it
          if (t == T - 1) tc_commit(&tmem_full[as]);  // accumulator complete
        }
        __syncwarp();
        if (++stage == NS) { stage = 0; phase ^= 1u; }
      }
    }
  } else {
    // ===== epilogue =====
    const int ew = warp - 2;   // 0..7
    const int q = warp & 3;    // TMEM lane quarter this warp may read
    const int hf = ew >> 2;    // column half of every tile
    const int rowi = q * 32 + lane;  // row within the CTA
    const uint32_t row_local = blockIdx.x * BM + rowi;
    const bool row_valid = row_local < p.nq_rows;
    const uint32_t self = p.exclude_self ? p.self_base + p.q_row0 + row_local : INVALID_ID;
    const size_t sr0 = (size_t)blockIdx.y * p.rows_pad + blockIdx.x * BM + q * 32;  // scratch row of lane 0
    float *cd = p.cand_d + (sr0 + lane) * p.cap;
    uint32_t *ci = p.cand_i + (sr0 + lane) * p.cap;
    uint32_t *myhist = hist + ew * 256;
    float *sp = spill + ew * (32 * 36) + lane * 36;
    const int pair_bar = 2 + q;
    long long pf_wait = 0, pf_proc = 0, pf_comp = 0, pf_t0 = clock64();
    uint32_t pf_ncomp = 0;

    if (hf == 0) { cnt_s[rowi] = 0; tau_s[rowi] = row_valid ? (p.tau0 ? p.tau0[blockIdx.x * BM + rowi] : INFINITY) : -INFINITY; }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    float tau = tau_s[rowi];

    // rows of this quarter whose count exceeds `limit`: the even rows are compacted by half 0, the odd by half 1
    auto compact_phase = [&](uint32_t limit) {
      asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
      uint32_t need = __ballot_sync(0xffffffffu, cnt_s[rowi] > limit && (lane & 1) == hf);
      if (need) {
        long long c0 = clock64();
        while (need) {
          const int l = __ffs(need) - 1;
          need &= need - 1;
          float *rcd = p.cand_d + (sr0 + l) * p.cap;
          uint32_t *rci = p.cand_i + (sr0 + l) * p.cap;
          const uint32_t rcnt = cnt_s[q * 32 + l];
          float t;
          if (p.cap <= 384) t = compact_row_reg<12>(rcd, rci, rcnt, p.Kp, myhist, lane);
          else if (p.cap <= 768) t = compact_row_reg<24>(rcd, rci, rcnt, p.Kp, myhist, lane);
          else t = compact_row(rcd, rci, rcnt, p.Kp, myhist, lane);
          if (lane == 0) { cnt_s[q * 32 + l] = p.Kp; tau_s[q * 32 + l] = t; }
          pf_ncomp++;
        }
        pf_comp += clock64() - c0;
      }
      asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
      tau = tau_s[rowi];
    };

    uint32_t ti = 0;
    for (uint32_t tile = tile_begin; tile < tile_end; tile += p.tile_step, ti++) {
      const uint32_t as = ti & 1u, aphase = (ti >> 1) & 1u;
      const uint32_t n0 = tile * BN;
      { long long c0 = clock64(); mbar_wait(&tmem_full[as], aphase); pf_wait += clock64() - c0; }
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + hf * 128;
      const uint32_t colh = n0 + hf * 128;
      // the accumulator holds -key / 2 (x.y - ||y||^2 / 2, or x.y for the inner product): key < tau  <=>  acc > -tau / 2
      const float ntau = -0.5f * tau;
      long long pc0 = clock64();

      auto process = [&](const uint32_t(&v)[32], int c) {
        // the largest of every 8 elements against the threshold: 9 instructions per 8 elements; the elements of a
        // group that passes (rare) are looked at one by one below
        uint32_t gm = 0;
#pragma unroll
        for (int g = 0; g < 4; g++) {
          const float m01 = fmaxf(__uint_as_float(v[8 * g + 0]), __uint_as_float(v[8 * g + 1]));
          const float m23 = fmaxf(__uint_as_float(v[8 * g + 2]), __uint_as_float(v[8 * g + 3]));
          const float m45 = fmaxf(__uint_as_float(v[8 * g + 4]), __uint_as_float(v[8 * g + 5]));
          const float m67 = fmaxf(__uint_as_float(v[8 * g + 6]), __uint_as_float(v[8 * g + 7]));
          if (fmaxf(fmaxf(m01, m23), fmaxf(m45, m67)) > ntau) gm |= 1u << g;
        }
        if (__any_sync(0xffffffffu, gm != 0)) {
          uint4 *dst = (uint4 *)sp;   // dynamic indexing below: through shared memory
#pragma unroll
          for (int j = 0; j < 8; j++) dst[j] = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          uint32_t hm = 0;
          while (gm) {
            const int g = __ffs(gm) - 1;
            gm &= gm - 1;
            for (int e = 0; e < 8; e++)
              if (sp[8 * g + e] > ntau) hm |= 1u << (8 * g + e);
          }
          while (hm) {  // rare (~0.1 % of the elements): lanes diverge here
            const int j = __ffs(hm) - 1;
            hm &= hm - 1;
            const uint32_t col = colh + c * 32 + j;
            if (col < p.n && col != self) {   // columns past the base are zero-filled tiles: acc = 0 may pass
              const float d = -2.f * sp[j];
              const uint32_t pos = atomicAdd(&cnt_s[rowi], 1u);
              __stcg(cd + pos, d);
              __stcg(ci + pos, col);
            }
          }
        }
        __syncwarp();
      };

      uint32_t va[32], vb[32];
      TC_LD32(taddr, va);
      TC_LD_WAIT(va);
      TC_LD32(taddr + 32, vb);
      process(va, 0);
      TC_LD_WAIT(vb);
      TC_LD32(taddr + 64, va);
      process(vb, 1);
      TC_LD_WAIT(va);
      TC_LD32(taddr + 96, vb);
      process(va, 2);
      TC_LD_WAIT(vb);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);  // all TMEM reads of this warp are done
      process(vb, 3);
      pf_proc += clock64() - pc0;
      compact_phase(p.cap - 256);
    }
    compact_phase(p.Kp);  // leave at most Kp entries per row
    if (hf == 0 && row_valid) { p.cand_n[sr0 + lane] = cnt_s[rowi]; p.cand_tau[sr0 + lane] = tau_s[rowi]; }
    if (p.prof && lane == 0) {
      atomicAdd(p.prof + 0, (unsigned long long)pf_wait); atomicAdd(p.prof + 1, (unsigned long long)pf_proc);
      atomicAdd(p.prof + 2, (unsigned long long)pf_comp); atomicAdd(p.prof + 3, (unsigned long long)(clock64() - pf_t0));
      atomicAdd(p.prof + 4, (unsigned long long)pf_ncomp);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// ---------------------------------------------------------------- finalize kernel
struct FinParams {
  const float *base; size_t bstride; uint32_t dim, stride4;  // fp32 index rows (stride4 = dim rounded up to 4)
  const float *q; size_t qstride;                            // fp32 query rows of this launch
  const float *nx;                                           // query norms of this launch
  uint32_t nq_rows, q_row0, K, Kp, cap, rows_pad, splits, P;
  const float *cand_d; const uint32_t *cand_i; const uint32_t *cand_n; const float *cand_tau;
  float c1, c2, ny_max;  // eps = c1*sqrt(nx*ny_max) + c2*(nx+ny_max)
  int ip;
  uint32_t *out_i; float *out_d;  // [nq_rows, K] of this launch
  uint32_t *fail_cnt, *fail_rows; // rows whose shortlist could not be proven complete
  unsigned long long *n_cand;     // total shortlisted rows re-ranked (statistics)
};

constexpr int FIN_THREADS = 128;
constexpr int RS4 = 33;  // float4 per staged candidate row (32 + 1 pad: conflict-free LDS.128)

__device__ __forceinline__ bool lex_less(float d0, uint32_t i0, float d1, uint32_t i1) {
  return d0 < d1 || (d0 == d1 && i0 < i1);
}

__global__ void __launch_bounds__(FIN_THREADS) finalize_kernel(const FinParams p) {
  extern __shared__ __align__(16) uint8_t fsm[];
  float4 *tile = (float4 *)fsm;                                   // [4 warps][32][RS4]
  float4 *qs4 = tile + 4 * 32 * RS4;                              // [stride4/4]
  float *sd = (float *)(qs4 + p.stride4 / 4);                     // [P]
  uint32_t *si = (uint32_t *)(sd + p.P);                          // [P]
  __shared__ uint32_t off_s[65];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t row = blockIdx.x;
  if (row >= p.nq_rows) return;

  if (tid == 0) {
    uint32_t o = 0;
    for (uint32_t s = 0; s < p.splits; s++) { off_s[s] = o; o += p.cand_n[(size_t)s * p.rows_pad + row]; }
    off_s[p.splits] = o;
  }
  for (uint32_t d4 = tid; d4 < p.stride4 / 4; d4 += FIN_THREADS) {
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) { uint32_t d = d4 * 4 + j; v[j] = d < p.dim ? __ldg(p.q + (size_t)row * p.qstride + d) : 0.f; }
    qs4[d4] = make_float4(v[0], v[1], v[2], v[3]);
  }
  __syncthreads();
  const uint32_t C = off_s[p.splits];
  for (uint32_t e = tid; e < p.P; e += FIN_THREADS) { sd[e] = INFINITY; si[e] = INVALID_ID; }
  __syncthreads();
  for (uint32_t s = 0; s < p.splits; s++) {
    const uint32_t o = off_s[s], c = off_s[s + 1] - o;
    const uint32_t *src = p.cand_i + ((size_t)s * p.rows_pad + row) * p.cap;
    for (uint32_t e = tid; e < c; e += FIN_THREADS) si[o + e] = __ldcg(src + e);
  }
  __syncthreads();

  // exact fp32 distances in the order of bf_knn.cu: sequential over the dimension, one fmaf per element
  float4 *mytile = tile + warp * 32 * RS4;
  for (uint32_t e0 = warp * 32; e0 < C; e0 += 4 * 32) {
    const uint32_t nb = min(32u, C - e0);
    const uint32_t myid = (uint32_t)lane < nb ? si[e0 + lane] : 0u;
    float acc = 0.f;
    for (uint32_t d0 = 0; d0 < p.stride4; d0 += 128) {
      const uint32_t n4 = min(32u, (p.stride4 - d0) / 4);
      // lane r holds candidate id r; 8 independent 512-byte row loads in flight per step
      for (uint32_t r0 = 0; r0 < nb; r0 += 8) {
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
          const uint32_t id = __shfl_sync(0xffffffffu, myid, (r0 + u) & 31);
          v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (r0 + u < nb && (uint32_t)lane < n4)
            v[u] = __ldg((const float4 *)(p.base + (size_t)id * p.bstride + d0) + lane);
        }
#pragma unroll
        for (int u = 0; u < 8; u++)
          if (r0 + u < nb && (uint32_t)lane < n4) mytile[(r0 + u) * RS4 + lane] = v[u];
      }
      __syncwarp();
      if ((uint32_t)lane < nb) {
        const float4 *brow = mytile + lane * RS4;
        const float4 *qrow = qs4 + d0 / 4;
        for (uint32_t c4 = 0; c4 < n4; c4++) {
          const float4 b = brow[c4], q = qrow[c4];
          if (p.ip) {
            acc = fmaf(q.x, b.x, acc); acc = fmaf(q.y, b.y, acc); acc = fmaf(q.z, b.z, acc); acc = fmaf(q.w, b.w, acc);
          } else {
            float t;
            t = q.x - b.x; acc = fmaf(t, t, acc);
            t = q.y - b.y; acc = fmaf(t, t, acc);
            t = q.z - b.z; acc = fmaf(t, t, acc);
            t = q.w - b.w; acc = fmaf(t, t, acc);
          }
        }
      }
      __syncwarp();
    }
    if ((uint32_t)lane < nb) sd[e0 + lane] = p.ip ? 1.0f - acc : acc;
  }
  __syncthreads();

  // bitonic sort of P (distance,id) pairs, ascending
  for (uint32_t sz = 2; sz <= p.P; sz <<= 1) {
    for (uint32_t st = sz >> 1; st >= 1; st >>= 1) {
      for (uint32_t t = tid; t < p.P / 2; t += FIN_THREADS) {
        uint32_t i = ((t / st) * 2 * st) + (t % st), j = i + st;
        bool up = (i & sz) == 0;
        float di = sd[i], dj = sd[j];
        uint32_t ii = si[i], ij = si[j];
        bool sw = up ? lex_less(dj, ij, di, ii) : lex_less(di, ii, dj, ij);
        if (sw) { sd[i] = dj; sd[j] = di; si[i] = ij; si[j] = ii; }
      }
      __syncthreads();
    }
  }
  for (uint32_t j = tid; j < p.K; j += FIN_THREADS) {
    p.out_i[(size_t)row * p.K + j] = j < C ? si[j] : INVALID_ID;
    if (p.out_d) p.out_d[(size_t)row * p.K + j] = j < C ? sd[j] : INFINITY;
  }
  if (tid == 0) {
    const float dK = C >= p.K ? sd[p.K - 1] : INFINITY;
    const float nx = p.nx[row];
    const float eps = p.c1 * sqrtf(nx * p.ny_max) + p.c2 * (nx + p.ny_max);
    bool ok = true;
    for (uint32_t s = 0; s < p.splits; s++) {
      const float tau = p.cand_tau[(size_t)s * p.rows_pad + row];
      if (tau == INFINITY) continue;  // nothing was ever dropped in this split
      const float bound = p.ip ? 1.0f + 0.5f * (tau - eps) : (tau - eps) + nx;
      ok = ok && (bound > dK);
    }
    if (!ok) p.fail_rows[atomicAdd(p.fail_cnt, 1u)] = p.q_row0 + row;
    atomicAdd(p.n_cand, (unsigned long long)C);
  }
}

// results of the fallback (compact rows) -> their rows of the output
__global__ void scatter_rows_kernel(const uint32_t *rows, uint32_t nf, uint32_t K, const uint32_t *src_i, const float *src_d,
                                    uint32_t *out_i, float *out_d) {
  size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= (size_t)nf * K) return;
  uint32_t r = rows[t / K];
  out_i[(size_t)r * K + t % K] = src_i[t];
  if (out_d) out_d[(size_t)r * K + t % K] = src_d[t];
}
__global__ void gather_rows_kernel(const uint32_t *rows, uint32_t nf, const float *src, size_t sstride, uint32_t dim, float *dst) {
  size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= (size_t)nf * dim) return;
  dst[t] = src[(size_t)rows[t / dim] * sstride + t % dim];
}

}  // namespace tc

// ---------------------------------------------------------------- host side
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// bf16 [rows, dpad] row-major, box = 64 elements (128 B) x box_rows, 128B swizzle, zero fill outside
bool make_map(CUtensorMap *m, void *ptr, uint64_t rows, uint64_t dpad, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {dpad, rows};
  cuuint64_t strides[1] = {dpad * 2};
  cuuint32_t box[2] = {(cuuint32_t)tc::BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// bf16 [rows, 16] row-major (32 B rows), box = 16 x box_rows, 32B swizzle: the norm step's tiles
bool make_map_ext(CUtensorMap *m, void *ptr, uint64_t rows, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {16, rows};
  cuuint64_t strides[1] = {32};
  cuuint32_t box[2] = {16, box_rows};
  cuuint32_t estr[2] = {1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct Bf16Set {  // bf16 hi/lo + norms of a row set
  void *hi = nullptr, *lo = nullptr;
  float *norms = nullptr;
  bool inexact = false;
  float nmax = 0.f;
};

int prep_rows(b200ann_handle *h, const float *src, size_t sstride, size_t n, size_t dim, size_t dpad, Bf16Set *o) {
  CU_CHECK(h, cudaMalloc(&o->hi, n * dpad * 2));
  CU_CHECK(h, cudaMalloc(&o->lo, n * dpad * 2));
  CU_CHECK(h, cudaMalloc((void **)&o->norms, n * sizeof(float)));
  DevBuf<uint32_t> flags;
  CU_CHECK(h, flags.alloc(2, h->stream));
  CU_CHECK(h, cudaMemsetAsync(flags.p, 0, 8, h->stream));
  tc::prep_kernel<<<(unsigned)((n + 7) / 8), 256, 0, h->stream>>>(src, sstride, (uint32_t)n, (uint32_t)dim, (uint32_t)dpad,
                                                                   (__nv_bfloat16 *)o->hi, (__nv_bfloat16 *)o->lo, o->norms, flags.p);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  uint32_t f[2];
  CU_CHECK(h, cudaMemcpyAsync(f, flags.p, 8, cudaMemcpyDeviceToHost, h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  o->inexact = f[0] != 0;
  memcpy(&o->nmax, &f[1], 4);
  return 0;
}
void free_set(Bf16Set *s) {
  if (s->hi) cudaFree(s->hi);
  if (s->lo) cudaFree(s->lo);
  if (s->norms) cudaFree(s->norms);
  s->hi = s->lo = nullptr; s->norms = nullptr;
}

uint32_t next_pow2(uint32_t x) { uint32_t p = 1; while (p < x) p <<= 1; return p; }

uint32_t shortlist_size(const b200ann_handle *h, size_t k) {
  size_t slack = h->knn_slack >= 0 ? (size_t)h->knn_slack : (k / 4 > 32 ? k / 4 : 32);
  return (uint32_t)(k + slack);
}

}  // namespace

void knn_tc_free(b200ann_index *ix) {
  if (ix->tc_ext) cudaFree(ix->tc_ext);
  ix->tc_ext = nullptr;
  if (ix->tc_hi) cudaFree(ix->tc_hi);
  if (ix->tc_lo) cudaFree(ix->tc_lo);
  if (ix->tc_norms) cudaFree(ix->tc_norms);
  ix->tc_hi = ix->tc_lo = nullptr; ix->tc_norms = nullptr;
}

bool knn_tc_supported(const b200ann_index *ix, size_t nq, size_t k) {
  (void)nq;
  if (k == 0 || k > 1024) return false;
  if (ix->dim > 4096) return false;
  return get_encode() != nullptr;
}

int launch_knn_tc(b200ann_handle *h, b200ann_index *ix, const float *q, size_t nq, size_t qstride, size_t k,
                  bool exclude_self, uint32_t *ids, float *dists, uint32_t self_base) {
  if (nq == 0) return 0;
  if (!knn_tc_supported(ix, nq, k)) return set_err(h, B200ANN_ERR_ARG, "knn (tensor path): unsupported k/dim or no cuTensorMapEncodeTiled");
  const size_t n = ix->n, dim = ix->dim;
  const size_t dpad = (dim + tc::BK - 1) / tc::BK * tc::BK;
  const bool ip = ix->metric == B200ANN_IP;

  // 1. bf16 copies + norms of the base (cached in the index) and of the queries
  if (!ix->tc_hi) {
    Bf16Set b;
    int rc = prep_rows(h, ix->vec, ix->stride, n, dim, dpad, &b);
    if (rc) { free_set(&b); return rc; }
    ix->tc_hi = b.hi; ix->tc_lo = b.lo; ix->tc_norms = b.norms; ix->tc_dpad = dpad;
    ix->tc_inexact = b.inexact; ix->tc_ny_max = b.nmax;
  }
  const uint32_t ext = ip ? 0u : 1u;
  if (ext && !ix->tc_ext) {
    CU_CHECK(h, cudaMalloc(&ix->tc_ext, n * 16 * 2));
    tc::ext_b_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(ix->tc_norms, (uint32_t)n, (__nv_bfloat16 *)ix->tc_ext);
    h->launches++;
    CU_CHECK(h, cudaGetLastError());
  }
  DevBuf<__nv_bfloat16> a_ext;
  CU_CHECK(h, a_ext.alloc(128 * 16, h->stream));
  tc::ext_a_kernel<<<8, 256, 0, h->stream>>>(a_ext.p);
  h->launches++;
  Bf16Set qs;
  // queries that are rows of the index itself (the whole base, or a block of rows of it: b200ann_knn_graph_rows) use
  // the cached bf16 copies
  const bool self_queries = q >= ix->vec && q < ix->vec + n * ix->stride && qstride == ix->stride &&
                            (size_t)(q - ix->vec) % ix->stride == 0 && (size_t)(q - ix->vec) / ix->stride + nq <= n;
  struct QGuard { Bf16Set *s; bool own; ~QGuard() { if (own) free_set(s); } } qguard{&qs, !self_queries};
  if (self_queries) {
    const size_t r0 = (size_t)(q - ix->vec) / ix->stride;
    qs.hi = (char *)ix->tc_hi + r0 * dpad * 2; qs.lo = ix->tc_lo ? (char *)ix->tc_lo + r0 * dpad * 2 : nullptr;
    qs.norms = ix->tc_norms + r0; qs.inexact = ix->tc_inexact; qs.nmax = ix->tc_ny_max;
  } else {
    int rc = prep_rows(h, q, qstride, nq, dim, dpad, &qs);
    if (rc) return rc;
  }
  const uint32_t passes = (ix->tc_inexact || qs.inexact) ? 3u : 1u;
  const uint32_t nkb = (uint32_t)(dpad / tc::BK);
  const uint32_t a_res = nkb * (passes == 3 ? 2u : 1u) + ext <= (uint32_t)tc::A_SLOTS ? 1u : 0u;
  uint32_t a_slots, nstage;
  if (a_res) {
    a_slots = nkb * (passes == 3 ? 2u : 1u) + ext;
    nstage = (tc::SMEM_MAX - 1024 - tc::SMEM_MISC - a_slots * tc::A_BYTES) / tc::B_BYTES;
  } else {
    nstage = (tc::SMEM_MAX - 1024 - tc::SMEM_MISC) / (tc::A_BYTES + tc::B_BYTES);
    a_slots = nstage;
  }
  if (nstage > 6) nstage = 6;
  const int sweep_smem = 1024 + (int)(a_slots * tc::A_BYTES + nstage * tc::B_BYTES) + tc::SMEM_MISC;

  CUtensorMap mA_hi, mA_lo, mB_hi, mB_lo, mA_ext, mB_ext;
  // inner product: no norm block; the two maps are never used but must be valid kernel arguments
  if (!make_map(&mA_hi, qs.hi, nq, dpad, tc::BM) || !make_map(&mA_lo, qs.lo, nq, dpad, tc::BM) ||
      !make_map(&mB_hi, ix->tc_hi, n, dpad, tc::BN) || !make_map(&mB_lo, ix->tc_lo, n, dpad, tc::BN) ||
      !make_map_ext(&mA_ext, a_ext.p, 128, tc::BM) ||
      !make_map_ext(&mB_ext, ext ? ix->tc_ext : (void *)a_ext.p, ext ? n : 128, ext ? tc::BN : tc::BM))
    return set_err(h, B200ANN_ERR_CUDA, "knn (tensor path): cuTensorMapEncodeTiled failed");

  // 2. geometry
  const uint32_t Kp = shortlist_size(h, k);
  const uint32_t cap = (2 * Kp + 256 + 31) / 32 * 32;  // a tile adds at most 256 entries to a row between compactions
  const uint32_t ntiles = (uint32_t)((n + tc::BN - 1) / tc::BN);
  const size_t chunk_rows = (size_t)h->sm_count * 4 * tc::BM;
  const size_t rows0 = nq < chunk_rows ? nq : chunk_rows;
  const uint32_t mtiles0 = (uint32_t)((rows0 + tc::BM - 1) / tc::BM);
  uint32_t splits = 1;
  {
    double best = 1e30;
    uint32_t smax = ntiles / 4 < 32 ? ntiles / 4 : 32;
    if (smax < 1) smax = 1;
    while (smax > 1 && next_pow2(smax * Kp) > 4096) smax--;
    for (uint32_t s = 1; s <= smax; s++) {
      uint32_t tps = (ntiles + s - 1) / s;
      double waves = ceil((double)mtiles0 * s / h->sm_count);
      // +2: per-CTA start-up; every further split also triples-in-kind the finalize work (it reranks splits x K'
      // candidates per row: 0.155 us per row and split against 2.5 us per tile - measured: 49k rows x 1M, 3 splits
      // 37.8 ms against 29.4 ms unsplit, where the wave count alone predicted a gain)
      double t = waves * (tps + 2.0) + 0.062 * (double)rows0 * (s - 1);
      if (t < best * 0.98) { best = t; splits = s; }
    }
  }
  const uint32_t tiles_per_split = (ntiles + splits - 1) / splits;
  const uint32_t P = next_pow2(splits * Kp);
  const uint32_t rows_pad = mtiles0 * tc::BM;

  // start thresholds from a sampled pre-pass (B200ANN_KNN_SAMPLE=0 turns it off; =s sets the tile stride): worth it
  // when the base is large against the shortlist. sample_m = 2.25 * Kp / step + 8: the sample_m-th key of a 1/step
  // sample sits ~4 standard deviations above the Kp-th key of the whole base.
  uint32_t sample_step = ntiles >= 2048 ? 16u : 0u, sample_m = 0;
  if (const char *e = getenv("B200ANN_KNN_SAMPLE")) sample_step = (uint32_t)atoi(e);
  if (sample_step < 2 || ntiles < 8 * sample_step) sample_step = 0;
  if (sample_step) {
    double margin = 2.25;
    if (const char *e = getenv("B200ANN_KNN_SAMPLE_MARGIN")) margin = atof(e);
    sample_m = (uint32_t)(margin * Kp / sample_step) + 8;
    if (sample_m < 16) sample_m = 16;
    if (sample_m >= Kp) sample_step = 0;
  }
  DevBuf<float> cand_d, cand_tau, tau0; DevBuf<uint32_t> cand_i, cand_n, fail; DevBuf<unsigned long long> ncand;
  const size_t slots = (size_t)splits * rows_pad;
  if (sample_step) CU_CHECK(h, tau0.alloc(rows_pad, h->stream));
  CU_CHECK(h, cand_d.alloc(slots * cap, h->stream));
  CU_CHECK(h, cand_i.alloc(slots * cap, h->stream));
  CU_CHECK(h, cand_n.alloc(slots, h->stream));
  CU_CHECK(h, cand_tau.alloc(slots, h->stream));
  CU_CHECK(h, fail.alloc(1 + nq, h->stream));
  CU_CHECK(h, ncand.alloc(1, h->stream));
  CU_CHECK(h, cudaMemsetAsync(fail.p, 0, 4, h->stream));
  CU_CHECK(h, cudaMemsetAsync(ncand.p, 0, 8, h->stream));

  const size_t stride4 = (dim + 3) / 4 * 4;
  const size_t fin_smem = (size_t)4 * 32 * tc::RS4 * 16 + stride4 * 4 + (size_t)P * 8;
  CU_CHECK(h, cudaFuncSetAttribute(tc::sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sweep_smem));
  CU_CHECK(h, cudaFuncSetAttribute(tc::finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fin_smem));

  // error bound of the tensor-core ranking key (see the file header): inputs, accumulation, fp32 roundings
  const double keff = (double)dpad * passes;
  const float c_in = passes == 3 ? ldexpf(1.f, -16) : 0.f;
  const float c1 = 2.f * (c_in + (float)keff * ldexpf(1.f, -22));
  const float c2 = (float)(dim + 24) * ldexpf(1.f, -23);   // + the three norm terms added into the accumulator
  const float ny_max = ip ? ix->tc_ny_max : ix->tc_ny_max;

  DevBuf<unsigned long long> prof;
  const bool want_prof = getenv("B200ANN_KNN_PROF") != nullptr;
  if (want_prof) { CU_CHECK(h, prof.alloc(8, h->stream)); CU_CHECK(h, cudaMemsetAsync(prof.p, 0, 64, h->stream)); }
  for (size_t r0 = 0; r0 < nq; r0 += chunk_rows) {
    const uint32_t rows = (uint32_t)((nq - r0) < chunk_rows ? (nq - r0) : chunk_rows);
    const uint32_t mtiles = (rows + tc::BM - 1) / tc::BM;
    tc::SweepParams sp = {};
    sp.nq_rows = rows; sp.q_row0 = (uint32_t)r0; sp.n = (uint32_t)n; sp.nkb = nkb; sp.passes = passes; sp.a_resident = a_res; sp.a_slots = a_slots; sp.nstage = nstage;
    sp.ntiles = ntiles; sp.tiles_per_split = tiles_per_split; sp.ext = ext;
    sp.exclude_self = exclude_self; sp.self_base = self_base; sp.Kp = Kp; sp.cap = cap; sp.rows_pad = rows_pad;
    sp.cand_d = cand_d.p; sp.cand_i = cand_i.p; sp.cand_n = cand_n.p; sp.cand_tau = cand_tau.p;
    sp.prof = want_prof ? prof.p : nullptr;
    sp.tile_step = 1; sp.tau0 = nullptr;
    if (sample_step) {
      // sampling pass: every sample_step-th tile, keep the sample_m smallest keys per row; the sample_m-th of the
      // sample is (with margin) above the Kp-th of the whole base, so the full pass starts from it instead of from
      // +inf and appends ~3x fewer hits. A row for which it turns out too tight fails its proof and is recomputed.
      tc::SweepParams ss = sp;
      ss.tile_step = sample_step; ss.Kp = sample_m; ss.tiles_per_split = ntiles; ss.prof = nullptr;
      tc::sweep_kernel<<<dim3(mtiles, 1), tc::NTHREADS, sweep_smem, h->stream>>>(mA_hi, mA_lo, mB_hi, mB_lo, mA_ext, mB_ext, ss);
      h->launches++;
      CU_CHECK(h, cudaMemcpyAsync(tau0.p, cand_tau.p, (size_t)rows_pad * sizeof(float), cudaMemcpyDeviceToDevice, h->stream));
      sp.tau0 = tau0.p;
    }
    tc::sweep_kernel<<<dim3(mtiles, splits), tc::NTHREADS, sweep_smem, h->stream>>>(mA_hi, mA_lo, mB_hi, mB_lo, mA_ext, mB_ext, sp);
    h->launches++;
    CU_CHECK(h, cudaGetLastError());
    tc::FinParams fp = {};
    fp.base = ix->vec; fp.bstride = ix->stride; fp.dim = (uint32_t)dim; fp.stride4 = (uint32_t)stride4;
    fp.q = q + r0 * qstride; fp.qstride = qstride; fp.nx = qs.norms + r0;
    fp.nq_rows = rows; fp.q_row0 = (uint32_t)r0; fp.K = (uint32_t)k; fp.Kp = Kp; fp.cap = cap; fp.rows_pad = rows_pad;
    fp.splits = splits; fp.P = P;
    fp.cand_d = cand_d.p; fp.cand_i = cand_i.p; fp.cand_n = cand_n.p; fp.cand_tau = cand_tau.p;
    fp.c1 = c1; fp.c2 = c2; fp.ny_max = ny_max; fp.ip = ip;
    fp.out_i = ids + r0 * k; fp.out_d = dists ? dists + r0 * k : nullptr;
    fp.fail_cnt = fail.p; fp.fail_rows = fail.p + 1; fp.n_cand = ncand.p;
    tc::finalize_kernel<<<rows, tc::FIN_THREADS, fin_smem, h->stream>>>(fp);
    h->launches++;
    CU_CHECK(h, cudaGetLastError());
  }

  // 3. rows whose shortlist could not be proven complete: recompute with the fp32 SIMT kernel
  uint32_t nf = 0;
  unsigned long long nc = 0;
  CU_CHECK(h, cudaMemcpyAsync(&nf, fail.p, 4, cudaMemcpyDeviceToHost, h->stream));
  CU_CHECK(h, cudaMemcpyAsync(&nc, ncand.p, 8, cudaMemcpyDeviceToHost, h->stream));
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  if (want_prof) {
    unsigned long long pv[8];
    CU_CHECK(h, cudaMemcpy(pv, prof.p, 64, cudaMemcpyDeviceToHost));
    fprintf(stderr, "[knn_tc prof] epilogue-warp cycle sums: wait %llu process %llu compact %llu total %llu | compactions %llu | nstage %u a_slots %u\n",
            pv[0], pv[1], pv[2], pv[3], pv[4], nstage, a_slots);
  }
  h->knn_stats[0] = nq; h->knn_stats[1] = nf; h->knn_stats[2] = nc; h->knn_stats[3] = passes;
  h->knn_stats[4] = splits; h->knn_stats[5] = Kp; h->knn_stats[6] = (uint64_t)ntiles * ((nq + tc::BM - 1) / tc::BM); h->knn_stats[7] = a_res;
  if (nf) {
    DevBuf<float> fq, fd; DevBuf<uint32_t> fi;
    CU_CHECK(h, fq.alloc((size_t)nf * dim, h->stream));
    CU_CHECK(h, fi.alloc((size_t)nf * k, h->stream));
    CU_CHECK(h, fd.alloc((size_t)nf * k, h->stream));
    size_t tot = (size_t)nf * dim;
    tc::gather_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(fail.p + 1, nf, q, qstride, (uint32_t)dim, fq.p);
    h->launches++;
    int rc = launch_bf_knn(h, ix, fq.p, nf, dim, k, exclude_self, fi.p, fd.p, fail.p + 1, self_base);
    if (rc) return rc;
    tot = (size_t)nf * k;
    tc::scatter_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(fail.p + 1, nf, (uint32_t)k, fi.p, fd.p, ids, dists);
    h->launches++;
    CU_CHECK(h, cudaGetLastError());
    CU_CHECK(h, cudaStreamSynchronize(h->stream));
  }
  return 0;
}

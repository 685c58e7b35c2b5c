// Team-of-lanes distance evaluation shared by every graph kernel.
//
// A vector row is nch = ceil(dim/4) float4 chunks (rows are stored with a 16-byte aligned, zero padded
// stride). A team of TEAM consecutive lanes owns one row: lane t loads chunks t, t+TEAM, t+2*TEAM, ...
// so each load instruction of a team touches TEAM*16 contiguous bytes (128 B for TEAM=8), and
// accumulates x,y,z,w of each chunk in that order with fmaf. The TEAM partial sums are combined with
// an xor butterfly, which leaves the same bits in every lane of the team (fp add is commutative).
// oracle/oracle.c gpu_order() restates exactly this order, so GPU-vs-oracle comparisons are
// bit-exact on arbitrary float data.
//
// Replaces: hnswlib::L2Sqr* / InnerProduct* (hnswlib/hnswlib/space_l2.h:6-205, space_ip.h:6-339) and
// efanna2e::DistanceL2/DistanceInnerProduct::compare (CNNS/include/nsg/distance.h:25-227).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// read-only, streaming 16-byte load: rows are gathered once per query, keep them out of L1
__device__ __forceinline__ float4 ld_row16(const float4 *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}

// cached 16-byte load: rows that are re-read soon (the kept set of the occlusion prune) stay in L1
__device__ __forceinline__ float4 ld_row16_ca(const float4 *p) { return __ldg(p); }

// MET (the `IP` template argument of the kernels; a bool converts to 0 / 1): 0 = squared L2, 1 = inner product
// (1 - <a,b>), 2 = squared L2 over UNSIGNED BYTES: a 16-byte chunk holds 16 dimensions, the distance is the
// reference's int sum of (a-b)^2 (hnswlib L2SqrI / L2SqrI4x, space_l2.h:255-292), exact in any order. It travels
// through the float-typed plumbing as its bit pattern: non-negative ints below 0x7F800000 order like the floats
// with the same bits (denormals included; nothing here is compiled with -ftz), +inf stays above all of them, and no
// kernel does arithmetic on a distance - only comparisons.
constexpr int MET_L2 = 0, MET_IP = 1, MET_U8 = 2;

__device__ __forceinline__ uint32_t acc16_u8(uint32_t acc, const float4 &q, const float4 &v) {
  uint32_t d;
  d = __vabsdiffu4(__float_as_uint(q.x), __float_as_uint(v.x)); acc = __dp4a(d, d, acc);
  d = __vabsdiffu4(__float_as_uint(q.y), __float_as_uint(v.y)); acc = __dp4a(d, d, acc);
  d = __vabsdiffu4(__float_as_uint(q.z), __float_as_uint(v.z)); acc = __dp4a(d, d, acc);
  d = __vabsdiffu4(__float_as_uint(q.w), __float_as_uint(v.w)); acc = __dp4a(d, d, acc);
  return acc;
}

template <int IP>
__device__ __forceinline__ float acc4(float acc, const float4 &q, const float4 &v) {
  if (IP == MET_IP) {
    acc = fmaf(q.x, v.x, acc);
    acc = fmaf(q.y, v.y, acc);
    acc = fmaf(q.z, v.z, acc);
    acc = fmaf(q.w, v.w, acc);
  } else {
    float d;
    d = q.x - v.x; acc = fmaf(d, d, acc);
    d = q.y - v.y; acc = fmaf(d, d, acc);
    d = q.z - v.z; acc = fmaf(d, d, acc);
    d = q.w - v.w; acc = fmaf(d, d, acc);
  }
  return acc;
}

template <int TEAM>
__device__ __forceinline__ float team_reduce(float acc) {
#pragma unroll
  for (int off = TEAM / 2; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  return acc;
}

// Distances of two rows per team in one go (2*NCH independent 16-byte loads in flight per lane).
// r0/r1 may be nullptr (no row): the result is then unspecified and must be ignored by the caller.
template <int TEAM, int NCH, int IP, bool CACHE = false>
__device__ __forceinline__ void team_dist2(const float4 *r0, const float4 *r1, const float4 (&q)[NCH], int t,
                                           int nch, float &d0, float &d1) {
  float4 v0[NCH], v1[NCH];
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int c = 0; c < NCH; c++) {
    int ch = t + c * TEAM;
    v0[c] = (r0 != nullptr && ch < nch) ? (CACHE ? ld_row16_ca(r0 + ch) : ld_row16(r0 + ch)) : z;
  }
#pragma unroll
  for (int c = 0; c < NCH; c++) {
    int ch = t + c * TEAM;
    v1[c] = (r1 != nullptr && ch < nch) ? (CACHE ? ld_row16_ca(r1 + ch) : ld_row16(r1 + ch)) : z;
  }
  if constexpr (IP == MET_U8) {
    uint32_t u0 = 0, u1 = 0;
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      int ch = t + c * TEAM;
      if (ch < nch) {
        u0 = acc16_u8(u0, q[c], v0[c]);
        u1 = acc16_u8(u1, q[c], v1[c]);
      }
    }
#pragma unroll
    for (int off = TEAM / 2; off >= 1; off >>= 1) {
      u0 += __shfl_xor_sync(0xffffffffu, u0, off);
      u1 += __shfl_xor_sync(0xffffffffu, u1, off);
    }
    d0 = __uint_as_float(u0);
    d1 = __uint_as_float(u1);
  } else {
    float a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      int ch = t + c * TEAM;
      if (ch < nch) {
        a0 = acc4<IP>(a0, q[c], v0[c]);
        a1 = acc4<IP>(a1, q[c], v1[c]);
      }
    }
    a0 = team_reduce<TEAM>(a0);
    a1 = team_reduce<TEAM>(a1);
    d0 = IP == MET_IP ? 1.0f - a0 : a0;
    d1 = IP == MET_IP ? 1.0f - a1 : a1;
  }
}

// The same for ROWS rows per team in one go (ROWS * NCH independent 16-byte loads in flight per lane): for kernels that
// are bound by the latency of a single warp's dependent hops and have registers to spare (the HNSW insertion beam).
// Per-row arithmetic and its order are those of team_dist2: same bits.
template <int TEAM, int NCH, int IP, int ROWS>
__device__ __forceinline__ void team_dist_rows(const float4 *const (&r)[ROWS], const float4 (&q)[NCH], int t, int nch,
                                               float (&d)[ROWS]) {
  float4 v[ROWS][NCH];
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < ROWS; k++)
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      int ch = t + c * TEAM;
      v[k][c] = (r[k] != nullptr && ch < nch) ? ld_row16(r[k] + ch) : z;
    }
  if constexpr (IP == MET_U8) {
#pragma unroll
    for (int k = 0; k < ROWS; k++) {
      uint32_t u = 0;
#pragma unroll
      for (int c = 0; c < NCH; c++)
        if (t + c * TEAM < nch) u = acc16_u8(u, q[c], v[k][c]);
#pragma unroll
      for (int off = TEAM / 2; off >= 1; off >>= 1) u += __shfl_xor_sync(0xffffffffu, u, off);
      d[k] = __uint_as_float(u);
    }
  } else {
#pragma unroll
    for (int k = 0; k < ROWS; k++) {
      float a = 0.f;
#pragma unroll
      for (int c = 0; c < NCH; c++)
        if (t + c * TEAM < nch) a = acc4<IP>(a, q[c], v[k][c]);
      a = team_reduce<TEAM>(a);
      d[k] = IP == MET_IP ? 1.0f - a : a;
    }
  }
}

// team size / chunks-per-lane rule shared by all launchers (and restated in tests via b200ann_dist_team)
__host__ __device__ inline int dist_team_for_nch(int nch) { return nch <= 64 ? 8 : 32; }

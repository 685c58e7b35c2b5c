// Per-warp "C smallest" collector shared by the NSG build kernels (nsg_build.cu: get_neighbors' fullset, InterInsert)
// and the HNSW update kernel (hnsw_build.cu: the candidate set of updatePoint): a shared-memory buffer of cap2
// (distance, id) entries that is compacted - bitonic sort on (distance,id), duplicates dropped, the `keep` smallest
// kept - whenever it fills; entries above the current keep-th smallest are dropped on arrival.
#pragma once
#include "beam.cuh"

namespace beam {

__device__ __forceinline__ bool lex_less(float d0, uint32_t i0, float d1, uint32_t i1) {
  return d0 < d1 || (d0 == d1 && i0 < i1);
}

// ---------------------------------------------------------------- per-warp "C smallest" collector
struct TopBuf {
  float *d; uint32_t *i;  // [cap2] shared memory, cap2 = power of two >= 2*keep, >= 64
  uint32_t cap2, keep, cnt;
  float thr_d; uint32_t thr_i;
  int lane;

  __device__ __forceinline__ void reset() { cnt = 0; thr_d = INFINITY; thr_i = INVALID_ID; }

  // sort [0,cnt) ascending by (distance,id), drop duplicates, truncate to `keep`
  __device__ __forceinline__ void compact() {
    uint32_t ns = 32;
    while (ns < cnt) ns <<= 1;
    for (uint32_t j = cnt + lane; j < ns; j += 32) { d[j] = INFINITY; i[j] = INVALID_ID; }
    __syncwarp();
    for (uint32_t k = 2; k <= ns; k <<= 1) {
      for (uint32_t j = k >> 1; j > 0; j >>= 1) {
        for (uint32_t x = lane; x < (ns >> 1); x += 32) {
          uint32_t lo = ((x & ~(j - 1)) << 1) | (x & (j - 1));  // index with bit j clear
          uint32_t hi = lo | j;
          bool up = (lo & k) == 0;
          float dl = d[lo], dh = d[hi];
          uint32_t il = i[lo], ih = i[hi];
          bool sw = up ? lex_less(dh, ih, dl, il) : lex_less(dl, il, dh, ih);
          if (sw) { d[lo] = dh; i[lo] = ih; d[hi] = dl; i[hi] = il; }
        }
        __syncwarp();
      }
    }
    // unique + truncate, in place (writes never pass the reads of the same round)
    uint32_t out = 0;
    for (uint32_t j0 = 0; j0 < cnt && out < keep; j0 += 32) {
      uint32_t j = j0 + lane;
      float dj = INFINITY; uint32_t ij = INVALID_ID;
      bool ok = false;
      if (j < cnt) {
        dj = d[j]; ij = i[j];
        ok = !(j > 0 && i[j - 1] == ij && d[j - 1] == dj);
      }
      __syncwarp();
      unsigned mask = __ballot_sync(0xffffffffu, ok);
      uint32_t pos = out + __popc(mask & ((1u << lane) - 1));
      if (ok && pos < keep) { d[pos] = dj; i[pos] = ij; }
      out += __popc(mask);
      __syncwarp();
    }
    cnt = out < keep ? out : keep;
    if (cnt == keep) { thr_d = d[keep - 1]; thr_i = i[keep - 1]; }
    __syncwarp();
  }

  // offer (cd[c], ci[c]) for c in [0,m); `lo_*` is an exclusive lower bound (pass -inf/0 for none)
  __device__ __forceinline__ void push_batch(const uint32_t *ci, const float *cd, uint32_t m, float lo_d,
                                             uint32_t lo_i, bool use_lo) {
    for (uint32_t c0 = 0; c0 < m; c0 += 32) {
      uint32_t c = c0 + lane;
      float dd = c < m ? cd[c] : INFINITY;
      uint32_t id = c < m ? ci[c] : INVALID_ID;
      bool in = c < m && (!use_lo || lex_less(lo_d, lo_i, dd, id));
      bool pass = in && !lex_less(thr_d, thr_i, dd, id);
      unsigned mask = __ballot_sync(0xffffffffu, pass);
      if (cnt + __popc(mask) > cap2) {
        compact();
        pass = in && !lex_less(thr_d, thr_i, dd, id);
        mask = __ballot_sync(0xffffffffu, pass);
      }
      if (pass) {
        uint32_t pos = cnt + __popc(mask & ((1u << lane) - 1));
        d[pos] = dd; i[pos] = id;
      }
      cnt += __popc(mask);
      __syncwarp();
    }
  }
};

}  // namespace beam

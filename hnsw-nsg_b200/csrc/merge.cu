// Shard-result merge: per query, `parts` ascending (distance,id) lists of k entries -> the kout
// smallest of their union, ascending (distance,id). The multi-GPU search exchanges per-shard lists
// with one NCCL all-gather and every rank runs this kernel on the gathered buffer.
// CPU analogue in the reference: the per-cluster result merge of
// CNNS/tests/cluster_hnsw_nsg_search.cpp:189-270 and merge_topk_queue
// (CNNS/tests/cluster_hnsw_nsg_search_pipeline.cpp:195-212).
#include "common.cuh"

namespace {

__device__ __forceinline__ bool lex_less(float d0, uint32_t i0, float d1, uint32_t i1) {
  return d0 < d1 || (d0 == d1 && i0 < i1);
}

// one warp per query; element e = (part s, slot j): rank = j + sum over other parts of
// #{entries < e} (binary search; (distance,id) keys are unique across shards because ids are global;
// padding entries (+inf, INVALID) are ordered by part so that ranks stay unique)
__global__ void merge_topk_kernel(const float *d, const uint32_t *ids, uint32_t parts, uint32_t nq, uint32_t k,
                                  uint32_t kout, float *out_d, uint32_t *out_i) {
  const uint32_t q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const uint32_t lane = threadIdx.x & 31;
  if (q >= nq) return;
  const uint32_t total = parts * k;
  for (uint32_t e = lane; e < total; e += 32) {
    const uint32_t s = e / k, j = e % k;
    const size_t off = ((size_t)s * nq + q) * k;
    const float de = d[off + j];
    const uint32_t ie = ids[off + j];
    uint32_t rank = j;
    for (uint32_t s2 = 0; s2 < parts; s2++) {
      if (s2 == s) continue;
      const size_t off2 = ((size_t)s2 * nq + q) * k;
      uint32_t lo = 0, hi = k;
      while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        float d2 = d[off2 + mid];
        uint32_t i2 = ids[off2 + mid];
        bool less = (i2 == INVALID_ID && ie == INVALID_ID) ? (s2 < s) : lex_less(d2, i2, de, ie);
        if (less) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    if (rank < kout) {
      out_i[(size_t)q * kout + rank] = ie;
      if (out_d) out_d[(size_t)q * kout + rank] = de;
    }
  }
}

}  // namespace

// device pointers only, asynchronous on h->stream
int launch_merge_topk(b200ann_handle *h, const float *d, const uint32_t *ids, size_t parts, size_t nq, size_t k, size_t kout,
                      float *out_d, uint32_t *out_i) {
  merge_topk_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(d, ids, (uint32_t)parts, (uint32_t)nq, (uint32_t)k,
                                                                   (uint32_t)kout, out_d, out_i);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  return 0;
}

extern "C" int b200ann_merge_topk(b200ann_handle_t h, const float *d, const uint32_t *ids, size_t parts, size_t nq,
                                  size_t k, size_t kout, float *out_d, uint32_t *out_i) {
  if (!h) return B200ANN_ERR_ARG;
  if (!d || !ids || !out_i || parts == 0 || k == 0 || kout == 0 || kout > parts * k)
    return set_err(h, B200ANN_ERR_ARG, "merge_topk: need d, ids, out_i, parts > 0, 0 < kout <= parts*k");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  InArg<float> din; InArg<uint32_t> iin; OutArg<float> od; OutArg<uint32_t> oi;
  CU_CHECK(h, din.stage(d, parts * nq * k, h->stream));
  CU_CHECK(h, iin.stage(ids, parts * nq * k, h->stream));
  CU_CHECK(h, od.stage(out_d, nq * kout, h->stream));
  CU_CHECK(h, oi.stage(out_i, nq * kout, h->stream));
  cudaEventRecord(h->ev0, h->stream);
  merge_topk_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(din.d, iin.d, (uint32_t)parts, (uint32_t)nq,
                                                                   (uint32_t)k, (uint32_t)kout, od.d, oi.d);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  cudaEventRecord(h->ev1, h->stream);
  CU_CHECK(h, od.finish(h->stream));
  CU_CHECK(h, oi.finish(h->stream));
  if (h->async_mode && !od.host && !oi.host && is_device_ptr(d) && is_device_ptr(ids)) return 0;  // stream-ordered
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
  return 0;
}

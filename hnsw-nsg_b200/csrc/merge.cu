// Shard-result merge: per query, `parts` ascending (distance,id) lists of k entries -> the kout
// smallest of their union, ascending (distance,id). The multi-GPU search exchanges per-shard lists
// with one NCCL all-gather and every rank runs this kernel on the gathered buffer.
// CPU analogue in the reference: the per-cluster result merge of
// CNNS/tests/cluster_hnsw_nsg_search.cpp:189-270 and merge_topk_queue
// (CNNS/tests/cluster_hnsw_nsg_search_pipeline.cpp:195-212).
#include "common.cuh"

namespace {

// one warp per query; element e = (part s, slot j): rank = j + sum over the other parts of the entries that come
// before e in the order (distance, part, slot). Parts only have to be ascending by DISTANCE (the NSG-family searches
// emit equal distances in arrival order, not id order): ties between parts go to the lower part, so ranks are unique
// and no output slot is ever left unwritten. With row shards (part = shard, ids ascending with the part) this is the
// (distance, id) order whenever the parts themselves are (distance, id) sorted. Padding (+inf, INVALID) sorts last.
// Part s starts at d + off(s), ids + off(s) with off(s) = (s / lp) * rank_stride + (s % lp) * nq * k: the plain
// [parts, nq, k] layout is lp = parts; an all-gathered buffer of per-rank blocks [d: lp x nq x k][ids: lp x nq x k]
// is rank_stride = 2 * lp * nq * k with `ids` pointing lp * nq * k words behind `d`.
__global__ void merge_topk_kernel(const float *d, const uint32_t *ids, uint32_t parts, uint32_t lp, size_t rank_stride,
                                  uint32_t nq, uint32_t k, uint32_t kout, float *out_d, uint32_t *out_i) {
  const uint32_t q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const uint32_t lane = threadIdx.x & 31;
  if (q >= nq) return;
  const uint32_t total = parts * k;
  for (uint32_t e = lane; e < total; e += 32) {
    const uint32_t s = e / k, j = e % k;
    const size_t off = (size_t)(s / lp) * rank_stride + ((size_t)(s % lp) * nq + q) * k;
    const float de = d[off + j];
    const uint32_t ie = ids[off + j];
    uint32_t rank = j;
    for (uint32_t s2 = 0; s2 < parts; s2++) {
      if (s2 == s) continue;
      const size_t off2 = (size_t)(s2 / lp) * rank_stride + ((size_t)(s2 % lp) * nq + q) * k;
      uint32_t lo = 0, hi = k;
      while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        float d2 = d[off2 + mid];
        bool before = s2 < s ? !(de < d2) : d2 < de;  // lower part: d2 <= de (NaN-free input), higher part: d2 < de
        if (before) lo = mid + 1; else hi = mid;
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
  merge_topk_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(d, ids, (uint32_t)parts, (uint32_t)parts, 0, (uint32_t)nq,
                                                                   (uint32_t)k, (uint32_t)kout, out_d, out_i);
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
  merge_topk_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(din.d, iin.d, (uint32_t)parts, (uint32_t)parts, 0,
                                                                   (uint32_t)nq, (uint32_t)k, (uint32_t)kout, od.d, oi.d);
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

// the all-gathered form: buf = world blocks, block r = [d: lp x nq x k floats][ids: lp x nq x k u32] of rank r
extern "C" int b200ann_merge_topk_gathered(b200ann_handle_t h, const void *buf, size_t world, size_t lp, size_t nq, size_t k,
                                           size_t kout, float *out_d, uint32_t *out_i) {
  if (!h) return B200ANN_ERR_ARG;
  if (!buf || !out_i || world == 0 || lp == 0 || k == 0 || kout == 0 || kout > world * lp * k)
    return set_err(h, B200ANN_ERR_ARG, "merge_topk_gathered: need buf, out_i, world > 0, lp > 0, 0 < kout <= world*lp*k");
  if (!is_device_ptr(buf) || !is_device_ptr(out_i) || (out_d && !is_device_ptr(out_d)))
    return set_err(h, B200ANN_ERR_ARG, "merge_topk_gathered: device buffers only (the gather buffer of the collective)");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  const float *d = (const float *)buf;
  const uint32_t *ids = (const uint32_t *)buf + lp * nq * k;
  cudaEventRecord(h->ev0, h->stream);
  merge_topk_kernel<<<(unsigned)((nq + 3) / 4), 128, 0, h->stream>>>(d, ids, (uint32_t)(world * lp), (uint32_t)lp, 2 * lp * nq * k,
                                                                   (uint32_t)nq, (uint32_t)k, (uint32_t)kout, out_d, out_i);
  h->launches++;
  CU_CHECK(h, cudaGetLastError());
  cudaEventRecord(h->ev1, h->stream);
  if (h->async_mode) return 0;  // stream-ordered
  CU_CHECK(h, cudaStreamSynchronize(h->stream));
  cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
  return 0;
}

// The CNNS cluster pipeline for a batch of queries, entirely on the device: the per-query loop of
// CNNS/tests/cluster_hnsw_nsg_search.cpp:138-270 -
//   nprobe nearest routing points (:143-160; an exact scan replaces faiss::IndexHNSWFlat) -> clusters ranked by how
//   many of their routing points were hit (:162-187) -> IndexNSG::Search inside each cluster in that order (:189-215)
//   with the early-stop rule (:249-263) -> unique global ids by distance, top k (:217-247, :265-270)
// as a fixed sequence of launches on one stream with no host round trip: bf_knn over the routing points, route_kernel,
// then per rank r: prep_kernel (per-query navigating node / row range of "its r-th cluster", size 0 when the query has
// stopped), the beam kernel in mode 4 (search.cu), early_stop_kernel; finally merge_topk_kernel + the local-row ->
// global-id map. All clusters live in ONE index under a block-diagonal graph (hnsw-nsg_b200/cluster.py builds it).
#include "common.cuh"

namespace {

constexpr int MAX_NPROBE = 256;

// one warp per query: clusters of the nprobe routing hits, most hits first, ties to the smaller cluster id
__global__ void __launch_bounds__(128) route_kernel(const uint32_t *pid, uint32_t nq, uint32_t nprobe, uint32_t per_cluster,
                                                    uint32_t *order) {
  __shared__ uint32_t s_cl[4][MAX_NPROBE];   // cluster of hit j
  __shared__ uint32_t s_cnt[4][MAX_NPROBE];  // hits of that cluster if j is its first occurrence, else 0
  const uint32_t w = threadIdx.x >> 5;
  const uint32_t q = blockIdx.x * 4 + w;
  const uint32_t lane = threadIdx.x & 31;
  if (q >= nq) return;
  const uint32_t *row = pid + (size_t)q * nprobe;
  uint32_t *out = order + (size_t)q * nprobe;
  for (uint32_t j = lane; j < nprobe; j += 32) {
    const uint32_t pj = row[j];
    s_cl[w][j] = pj == INVALID_ID ? INVALID_ID : pj / per_cluster;
    out[j] = INVALID_ID;
  }
  __syncwarp();
  for (uint32_t j = lane; j < nprobe; j += 32) {
    const uint32_t cj = s_cl[w][j];
    uint32_t cnt = 0;
    bool first = cj != INVALID_ID;
    for (uint32_t i = 0; i < nprobe; i++) {
      if (s_cl[w][i] != cj) continue;
      cnt++;
      if (i < j) first = false;
    }
    s_cnt[w][j] = first ? cnt : 0u;
  }
  __syncwarp();
  for (uint32_t j = lane; j < nprobe; j += 32) {
    const uint32_t cnt = s_cnt[w][j], cj = s_cl[w][j];
    if (!cnt) continue;
    uint32_t rank = 0;  // distinct clusters with more hits, or as many and a smaller id
    for (uint32_t i = 0; i < nprobe; i++) {
      const uint32_t ci = s_cl[w][i], cnti = s_cnt[w][i];
      if (cnti && (cnti > cnt || (cnti == cnt && ci < cj))) rank++;
    }
    out[rank] = cj;
  }
}

__global__ void prep_kernel(const uint32_t *order, uint32_t nq, uint32_t nprobe, uint32_t r, const uint8_t *alive,
                            const uint32_t *cl_ep, const uint32_t *cl_base, const uint32_t *cl_size, uint32_t *ep,
                            uint32_t *base, uint32_t *size, uint8_t *active) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  const uint32_t c = order[(size_t)q * nprobe + r];
  const bool act = alive[q] && c != INVALID_ID;
  active[q] = act ? 1 : 0;
  ep[q] = act ? cl_ep[c] : 0u;
  base[q] = act ? cl_base[c] : 0u;
  size[q] = act ? cl_size[c] : 0u;
}

// one warp per query, after rank r: the reference's early stop (:249-263). A query whose cluster was searched stops
// when that cluster's best distance is not below cur_min_max (its results are kept); otherwise cur_min_max becomes
// min(cur_min_max, k-th smallest distance found so far) (the min(k, found)-th when fewer than k were found).
__global__ void early_stop_kernel(const float *parts_d, uint32_t nq, uint32_t K, uint32_t r, uint32_t k, int do_stop,
                                  const uint8_t *active, uint8_t *alive, float *cur_min_max, uint32_t *searched) {
  const uint32_t q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const uint32_t lane = threadIdx.x & 31;
  if (q >= nq) return;
  if (!active[q]) return;  // this query had no r-th cluster, or has stopped
  if (lane == 0) searched[q]++;
  if (!do_stop) return;
  const float cmin = parts_d[((size_t)r * nq + q) * K];
  const float cur = cur_min_max[q];
  if (cmin >= cur) {
    if (lane == 0) alive[q] = 0;
    return;
  }
  // k-th smallest of the ascending lists of parts 0..r: k-step merge, one head pointer per part
  float kth = INFINITY;
  uint32_t taken = 0;
  // heads[] live in registers: lane p owns parts p, p+32, ... ; nprobe <= 256 -> at most 8 per lane
  uint32_t head[8];
#pragma unroll
  for (int i = 0; i < 8; i++) head[i] = 0;
  while (taken < k) {
    float best = INFINITY; uint32_t bp = INVALID_ID;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint32_t part = lane + 32 * i;
      if (part <= r && head[i] < K) {
        const float d = parts_d[((size_t)part * nq + q) * K + head[i]];
        if (d < best) { best = d; bp = part; }
      }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const float od = __shfl_xor_sync(0xffffffffu, best, off);
      const uint32_t op = __shfl_xor_sync(0xffffffffu, bp, off);
      if (od < best || (od == best && op < bp)) { best = od; bp = op; }
    }
    if (bp == INVALID_ID || best == INFINITY) break;  // fewer than k found so far
    kth = best;
    taken++;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (lane + 32 * i == bp) head[i]++;
  }
  if (lane == 0 && kth < cur) cur_min_max[q] = kth;
}

__global__ void init_state_kernel(uint32_t nq, uint8_t *alive, float *cur_min_max, uint32_t *searched) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  alive[q] = 1; cur_min_max[q] = INFINITY; searched[q] = 0;
}

__global__ void map_ids_kernel(uint32_t *ids, size_t count, const uint32_t *perm) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= count) return;
  const uint32_t v = ids[i];
  if (v != INVALID_ID) ids[i] = perm[v];
}

}  // namespace

extern "C" int b200ann_cluster_search(b200ann_handle_t h, b200ann_index_t ix, b200ann_index_t route_ix, size_t per_cluster,
                                      const uint32_t *cl_ep, const uint32_t *cl_base, const uint32_t *cl_size,
                                      size_t n_clusters, const uint32_t *perm, const float *q, size_t nq, size_t qstride,
                                      size_t k, size_t nprobe, size_t search_K, size_t L, uint32_t seed, int early_stop,
                                      uint32_t *ids, float *dists, uint32_t *searched) {
  FLOAT_ONLY(h, ix, "cluster_search");
  if (!h || !ix || !route_ix) return B200ANN_ERR_ARG;
  if (!ix->graph) return set_err(h, B200ANN_ERR_STATE, "cluster_search: index has no graph (set_graph first)");
  if (!q || !ids || !cl_ep || !cl_base || !cl_size || per_cluster == 0 || n_clusters == 0 || k == 0 || search_K == 0 ||
      L < search_K || nprobe == 0 || nprobe > (size_t)MAX_NPROBE || nprobe > route_ix->n || qstride < ix->dim || route_ix->dim != ix->dim ||
      k > nprobe * search_K || route_ix->n != n_clusters * per_cluster)
    return set_err(h, B200ANN_ERR_ARG, "cluster_search: need q, ids, cluster tables, route_ix of n_clusters*per_cluster rows and "
                                       "the index's dim, 0 < search_K <= L, 0 < nprobe <= min(256, routing points), k <= nprobe*search_K");
  if (nq == 0) return 0;
  CU_CHECK(h, cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  InArg<float> qin; InArg<uint32_t> iep, ib, is, ip; OutArg<uint32_t> oi; OutArg<float> od; OutArg<uint32_t> os;
  CU_CHECK(h, qin.stage(q, nq * qstride, st));
  CU_CHECK(h, iep.stage(cl_ep, n_clusters, st));
  CU_CHECK(h, ib.stage(cl_base, n_clusters, st));
  CU_CHECK(h, is.stage(cl_size, n_clusters, st));
  CU_CHECK(h, ip.stage(perm, ix->n, st));
  CU_CHECK(h, oi.stage(ids, nq * k, st));
  CU_CHECK(h, od.stage(dists, nq * k, st));
  DevBuf<uint32_t> pid, order, ep, base, size, part_i, nsearched;
  DevBuf<float> pdist, part_d, curmm, outd;
  DevBuf<uint8_t> alive, active;
  CU_CHECK(h, pid.alloc(nq * nprobe, st));
  CU_CHECK(h, pdist.alloc(nq * nprobe, st));
  CU_CHECK(h, order.alloc(nq * nprobe, st));
  CU_CHECK(h, ep.alloc(nq, st));
  CU_CHECK(h, base.alloc(nq, st));
  CU_CHECK(h, size.alloc(nq, st));
  CU_CHECK(h, part_i.alloc(nprobe * nq * search_K, st));
  CU_CHECK(h, part_d.alloc(nprobe * nq * search_K, st));
  CU_CHECK(h, curmm.alloc(nq, st));
  CU_CHECK(h, alive.alloc(nq, st));
  CU_CHECK(h, active.alloc(nq, st));
  CU_CHECK(h, nsearched.alloc(nq, st));
  float *out_d = od.d;
  if (!out_d) { CU_CHECK(h, outd.alloc(nq * k, st)); out_d = outd.p; }
  cudaEventRecord(h->ev0, st);
  int rc = launch_bf_knn(h, route_ix, qin.d, nq, qstride, nprobe, false, pid.p, pdist.p);
  if (rc) return rc;
  const unsigned wblocks = (unsigned)((nq + 3) / 4), tblocks = (unsigned)((nq + 255) / 256);
  route_kernel<<<wblocks, 128, 0, st>>>(pid.p, (uint32_t)nq, (uint32_t)nprobe, (uint32_t)per_cluster, order.p);
  init_state_kernel<<<tblocks, 256, 0, st>>>((uint32_t)nq, alive.p, curmm.p, nsearched.p);
  h->launches += 2;
  for (size_t r = 0; r < nprobe; r++) {
    prep_kernel<<<tblocks, 256, 0, st>>>(order.p, (uint32_t)nq, (uint32_t)nprobe, (uint32_t)r, alive.p, iep.d, ib.d, is.d,
                                         ep.p, base.p, size.p, active.p);
    h->launches++;
    rc = launch_nsg_search_clustered(h, ix, qin.d, nq, qstride, search_K, L, seed, ep.p, base.p, size.p,
                                     part_i.p + r * nq * search_K, part_d.p + r * nq * search_K, nullptr);
    if (rc) return rc;
    early_stop_kernel<<<wblocks, 128, 0, st>>>(part_d.p, (uint32_t)nq, (uint32_t)search_K, (uint32_t)r, (uint32_t)k,
                                               early_stop, active.p, alive.p, curmm.p, nsearched.p);
    h->launches++;
  }
  rc = launch_merge_topk(h, part_d.p, part_i.p, nprobe, nq, search_K, k, out_d, oi.d);
  if (rc) return rc;
  if (ip.d) {
    map_ids_kernel<<<(unsigned)((nq * k + 255) / 256), 256, 0, st>>>(oi.d, nq * k, ip.d);
    h->launches++;
  }
  CU_CHECK(h, cudaGetLastError());
  cudaEventRecord(h->ev1, st);
  CU_CHECK(h, oi.finish(st));
  CU_CHECK(h, od.finish(st));
  if (searched) CU_CHECK(h, cudaMemcpyAsync(searched, nsearched.p, nq * sizeof(uint32_t), cudaMemcpyDefault, st));
  CU_CHECK(h, cudaStreamSynchronize(st));
  cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
  return 0;
}

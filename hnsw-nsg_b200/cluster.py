"""CNNS cluster pipeline on the GPU: route -> per-cluster NSG search -> top-k merge.

Reference (file IO + faiss orchestration, not library code):
  build   CNNS/tests/cluster_IVF_nndescent.cpp:116-265 (IVF k-means, centroid + m samples per cluster, per-cluster kNN
          graph) and CNNS/tests/nndescent_nsg.cpp:35-128 (per-cluster NSG)
  search  CNNS/tests/cluster_hnsw_nsg_search.cpp:138-270: nprobe nearest routing points -> clusters ranked by how
          many of their points were hit -> IndexNSG::Search inside each cluster in that order with the early-stop
          rule (:249-263) -> unique global ids by distance -> top k.
Here every cluster lives in ONE device index: rows permuted so that a cluster is a contiguous block, one
block-diagonal graph with global ids, so a whole query batch searches "its r-th cluster" in one launch of
b200ann_nsg_search_clustered; routing is an exact brute-force scan of the n_clusters*(m+1) routing points
(b200ann_bf_knn) instead of faiss::IndexHNSWFlat (faiss is not part of the hot path; SURVEY 8c: routing parity is
unpinned, the per-cluster search and the merge are pinned by tests/test_gpu_cluster.py against the oracle).
Host logic is numpy; the only compute calls are C-ABI entry points. k-means itself is not the reference's code
(faiss): `assign` may be given, else a few Lloyd iterations use b200ann_bf_knn for the assignment step."""
import importlib

import numpy as np

INVALID = 0xFFFFFFFF


def _B():
    return importlib.import_module(__package__)


def kmeans_assign(handle, base, n_clusters, iters=8, seed=1):
    """Lloyd iterations; assignment = exact nearest centroid on the GPU. Returns (assign [n], centroids)."""
    B = _B()
    rng = np.random.default_rng(seed)
    cent = np.ascontiguousarray(base[rng.choice(len(base), n_clusters, replace=False)])
    assign = None
    for _ in range(iters):
        cix = B.Index(handle, cent, B.L2)
        assign = cix.bf_knn(base, 1)[0][:, 0].astype(np.int64)
        cix.close()
        cnt = np.bincount(assign, minlength=n_clusters)
        sums = np.zeros((n_clusters, base.shape[1]), np.float64)
        np.add.at(sums, assign, base)
        ok = cnt > 0
        cent[ok] = (sums[ok] / cnt[ok, None]).astype(np.float32)
    return assign, cent


class ClusterIndex:
    """All clusters in one device index + the routing point set."""

    def __init__(self, handle):
        self.h = handle

    def build(self, base, assign, centroids, m=4, knn_k=32, nsg=(40, 32, 200), seed=1, nndescent=None):
        """base [n,dim]; assign [n] cluster of every row; centroids [n_clusters,dim]. m = sample points per cluster
        in the routing set (cluster_IVF_nndescent.cpp:140-191). nsg = (L, R, C). nndescent = None for the exact
        per-cluster kNN graph, or (L, iter, S, R) to run NN-descent like the reference."""
        B = _B()
        base = np.ascontiguousarray(base, np.float32)
        n, dim = base.shape
        nc = len(centroids)
        order = np.argsort(assign, kind="stable")
        self.perm = order.astype(np.uint32)                       # mapping_<id>: local row -> global id (:201-204)
        sizes = np.bincount(assign, minlength=nc).astype(np.uint32)
        self.cl_base = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.uint32)
        self.cl_size = sizes
        rows = np.ascontiguousarray(base[order])
        L, R, C_ = nsg
        gstride = 0
        graphs, eps = [], np.zeros(nc, np.uint32)
        for c in range(nc):
            sz, b0 = int(sizes[c]), int(self.cl_base[c])
            if sz == 0:
                graphs.append(np.zeros((0, 1), np.uint32))
                continue
            sub = rows[b0:b0 + sz]
            if sz <= 2:                                           # degenerate cluster: complete graph
                g = np.full((sz, 1), INVALID, np.uint32)
                if sz == 2:
                    g[0, 0], g[1, 0] = 1, 0
                ep = 0
            else:
                six = B.Index(self.h, sub, B.L2)
                K = min(knn_k, sz - 1)
                if nndescent is not None and sz > 2 * nndescent[2] + 1 and K <= nndescent[0]:
                    knn, _ = six.nndescent(K, nndescent[0], nndescent[1], nndescent[2], nndescent[3], seed=seed)
                else:
                    knn = six.knn_graph(K)
                six.nsg_build(knn, min(L, sz - 1), R, C_, seed=seed)
                g, ep, _ = six.get_graph()
                six.close()
            gg = np.where(g == INVALID, g, g + np.uint32(b0)).astype(np.uint32)
            graphs.append(gg)
            eps[c] = b0 + ep
            gstride = max(gstride, gg.shape[1])
        full = np.full((n, gstride), INVALID, np.uint32)
        for c in range(nc):
            sz, b0 = int(sizes[c]), int(self.cl_base[c])
            if sz:
                full[b0:b0 + sz, :graphs[c].shape[1]] = graphs[c]
        self.eps = eps
        self.graph = full
        self.ix = B.Index(self.h, rows, B.L2)
        self.ix.set_graph(full, int(eps[np.argmax(sizes)]))
        # routing points: per cluster its centroid followed by m sampled rows (centroids.fvecs layout)
        rng = np.random.default_rng(seed)
        self.m = m
        pts = np.zeros((nc * (m + 1), dim), np.float32)
        for c in range(nc):
            pts[c * (m + 1)] = centroids[c]
            sz, b0 = int(sizes[c]), int(self.cl_base[c])
            if sz:
                pick = rng.integers(0, sz, m)
                pts[c * (m + 1) + 1:(c + 1) * (m + 1)] = rows[b0 + pick]
            else:
                pts[c * (m + 1) + 1:(c + 1) * (m + 1)] = centroids[c]
        self.route_points = pts
        self.rix = B.Index(self.h, pts, B.L2)
        return self

    def route(self, q, nprobe):
        """clusters of every query, most routing-point hits first (ties: smaller cluster id) -> [nq, <=nprobe], -1 padded"""
        pid, _ = self.rix.bf_knn(q, nprobe)
        cl = (pid // (self.m + 1)).astype(np.int64)
        nq, nc = len(q), len(self.cl_size)
        counts = np.bincount((np.arange(nq)[:, None] * nc + cl).ravel(), minlength=nq * nc).reshape(nq, nc)
        # most hits first, ties to the smaller cluster id: stable argsort of -count over ascending ids
        order = np.argsort(-counts, axis=1, kind="stable")[:, :nprobe]
        hit = np.take_along_axis(counts, order, axis=1) > 0
        return np.where(hit, order, -1).astype(np.int64)

    def search(self, q, k, nprobe, search_K, L, seed=0, early_stop=True, return_stats=False):
        """the reference's per-query loop (cluster_hnsw_nsg_search.cpp:189-270) for a batch: rank r of every query in
        one launch. Returns global ids [nq,k] (+ distances), INVALID / inf padded."""
        q = np.ascontiguousarray(q, np.float32)
        nq = len(q)
        order = self.route(q, nprobe)
        alive = np.ones(nq, bool)
        cur_min_max = np.full(nq, np.inf, np.float32)
        parts_d, parts_i = [], []
        launches = 0
        for r in range(order.shape[1]):
            c = order[:, r]
            act = alive & (c >= 0)
            if not act.any():
                break
            cc = np.where(act, c, 0)
            size = np.where(act, self.cl_size[cc], 0).astype(np.uint32)
            ids, d = self.ix.nsg_search_clustered(q, self.eps[cc], self.cl_base[cc], size, search_K, L, seed=seed)
            launches += 1
            parts_d.append(d)
            parts_i.append(ids)
            if early_stop:
                cmin = np.where(act, d[:, 0], np.inf)
                stop = act & (cmin >= cur_min_max)                # :249-251 (the cluster's results are kept)
                alive &= ~stop
                upd = act & ~stop
                if upd.any():                                     # :254-262: k-th smallest accumulated distance
                    alld = np.sort(np.concatenate(parts_d, axis=1), axis=1)
                    have = np.isfinite(alld).sum(1)
                    kth = alld[np.arange(nq), np.minimum(k, np.maximum(have, 1)) - 1]
                    cur_min_max = np.where(upd, np.minimum(cur_min_max, kth), cur_min_max)
        pd = np.ascontiguousarray(np.stack(parts_d))              # [parts, nq, search_K]
        pi = np.ascontiguousarray(np.stack(parts_i))
        out_d, out_i = self.h.merge_topk(pd, pi, k)
        gi = np.where(out_i == INVALID, out_i, self.perm[np.minimum(out_i, len(self.perm) - 1)])
        if return_stats:
            return gi.astype(np.uint32), out_d, dict(launches=launches, clusters_searched=int(sum(np.isfinite(d[:, 0]).sum() for d in parts_d)))
        return gi.astype(np.uint32), out_d

    def search_device(self, q, k, nprobe, search_K, L, seed=0, early_stop=True, return_stats=False):
        """search() as ONE C-ABI call (b200ann_cluster_search): routing, the per-rank searches, the early-stop rule and
        the merge all stay on the device, no host round trip between ranks. Same results as search()."""
        B = _B()
        q = np.ascontiguousarray(q, np.float32)
        nq = len(q)
        if getattr(self, "_dev", None) is None:                   # cluster tables + row map resident on the device
            import torch
            dev = "cuda:%d" % self.h.device
            as_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.uint32).view(np.int32)).to(dev)
            self._dev = tuple(as_dev(a) for a in (self.eps, self.cl_base, self.cl_size, self.perm))
            torch.cuda.synchronize()
        ep, base, size, perm = self._dev
        ids = np.empty((nq, k), np.uint32)
        d = np.empty((nq, k), np.float32)
        searched = np.zeros(nq, np.uint32)
        lib = B.lib()
        self.h._check(lib.b200ann_cluster_search(self.h._h, self.ix._ix, self.rix._ix, self.m + 1, B._ptr(ep),
                                                 B._ptr(base), B._ptr(size), len(self.cl_size), B._ptr(perm),
                                                 B._ptr(q), nq, q.shape[1], k, nprobe, search_K, L, seed,
                                                 1 if early_stop else 0, B._ptr(ids), B._ptr(d), B._ptr(searched)))
        if return_stats:
            return ids, d, dict(launches=3 * nprobe + 4, clusters_searched=int(searched.sum()))
        return ids, d

    def close(self):
        self._dev = None
        self.ix.close()
        self.rix.close()

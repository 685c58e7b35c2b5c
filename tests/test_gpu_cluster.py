"""CNNS cluster pipeline on the GPU (hnsw-nsg_b200/cluster.py + b200ann_nsg_search_clustered) against a serial
restatement of the reference driver's per-query loop (CNNS/tests/cluster_hnsw_nsg_search.cpp:138-270) built from
oracle pieces: exact routing (orc_bf_knn), IndexNSG::Search inside each cluster (orc_nsg_search_qidx), the
early-stop rule and the unique-id merge. Ids and distances bit for bit; recall against brute force as a sanity bar."""
import importlib

import numpy as np
import pytest

import refutil as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("hnsw-nsg_b200")


@pytest.fixture(scope="module")
def H(B):
    return B.Handle(0)


@pytest.fixture(scope="module")
def case(B, H):
    cl = importlib.import_module("hnsw-nsg_b200.cluster")
    n, dim, nc = 12000, 32, 12
    # real-valued clustered rows: no exact distance ties, so the order inside every result list is unique (on tied
    # distances a cluster's list keeps IndexNSG::Search's insertion order, which a (distance,id) sort would not)
    base = R.data_glove_like(n, dim, 1, n_clusters=24)
    q = R.data_glove_like(300, dim, 2, n_clusters=24)
    assign, cent = cl.kmeans_assign(H, base, nc, iters=5, seed=3)
    ci = cl.ClusterIndex(H).build(base, assign, cent, m=4, knn_k=24, nsg=(30, 20, 120), seed=1)
    return cl, ci, base, q, assign


def _oracle_search(B, ci, base, q, k, nprobe, search_K, L, seed, early_stop=True):
    dim = base.shape[1]
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    rows = np.ascontiguousarray(base[ci.perm.astype(np.int64)])
    pid, _ = R.orc_bf_knn(ci.route_points, q, nprobe, R.L2_SEQ_FMA)
    out_i = np.full((len(q), k), R.INVALID, np.uint32)
    out_d = np.full((len(q), k), np.inf, np.float32)
    for i in range(len(q)):
        cl = (pid[i] // (ci.m + 1)).astype(np.int64)
        u, cnt = np.unique(cl, return_counts=True)
        order = u[np.lexsort((u, -cnt))]
        acc = {}
        cur = np.inf
        for c in order:
            b0, sz = int(ci.cl_base[c]), int(ci.cl_size[c])
            if sz == 0:
                continue
            g = ci.graph[b0:b0 + sz]
            gl = np.where(g == R.INVALID, g, g - np.uint32(b0)).astype(np.uint32)
            Lq = min(L, sz)
            ids, d = R.orc_nsg_search_qidx(rows[b0:b0 + sz], gl, int(ci.eps[c]) - b0, q[i:i + 1], [i], min(search_K, Lq), Lq,
                                           R.L2_GPU, seed=seed)
            for lid, dd in zip(ids[0], d[0]):
                gid = int(ci.perm[b0 + int(lid)])
                acc[gid] = min(acc.get(gid, np.inf), float(dd))
            cmin = float(d[0, 0])
            if early_stop and cmin >= cur:
                break
            ds = sorted(acc.values())
            cur = min(cur, ds[min(k - 1, len(ds) - 1)])
        # (distance, position-in-the-permuted-index) order == the merge kernel's (distance, id) order on local ids
        inv = {int(g): p for p, g in enumerate(ci.perm)}
        best = sorted(acc.items(), key=lambda t: (np.float32(t[1]), inv[t[0]]))[:k]
        for j, (gid, dd) in enumerate(best):
            out_i[i, j], out_d[i, j] = gid, dd
    return out_i, out_d


@pytest.mark.parametrize("k,nprobe,search_K,L,early", [(10, 8, 10, 30, True), (5, 12, 20, 40, True), (10, 6, 10, 24, False)])
def test_cluster_search_vs_oracle(B, case, k, nprobe, search_K, L, early):
    cl, ci, base, q, assign = case
    gi, gd = ci.search(q, k, nprobe, search_K, L, seed=5, early_stop=early)
    oi, od = _oracle_search(B, ci, base, q, k, nprobe, search_K, L, 5, early)
    assert (gi == oi).all()
    assert (gd.view(np.uint32) == od.view(np.uint32)).all()


@pytest.mark.parametrize("k,nprobe,search_K,L,early", [(10, 8, 10, 30, True), (5, 12, 20, 40, True), (10, 6, 10, 24, False),
                                                       (7, 40, 8, 16, True), (10, 1, 10, 32, True)])
def test_cluster_search_on_device_vs_oracle(B, case, k, nprobe, search_K, L, early):
    # b200ann_cluster_search: the same loop as ONE call (routing, per-rank searches, early stop, merge on the device)
    cl, ci, base, q, assign = case
    gi, gd, st = ci.search_device(q, k, nprobe, search_K, L, seed=5, early_stop=early, return_stats=True)
    oi, od = _oracle_search(B, ci, base, q, k, nprobe, search_K, L, 5, early)
    assert (gi == oi).all()
    assert (gd.view(np.uint32) == od.view(np.uint32)).all()
    hi, hd, hst = ci.search(q, k, nprobe, search_K, L, seed=5, early_stop=early, return_stats=True)
    assert (gi == hi).all() and st["clusters_searched"] == hst["clusters_searched"]


def test_cluster_search_recall_and_structure(B, H, case):
    cl, ci, base, q, assign = case
    # every cluster is a contiguous block whose edges stay inside it
    for c in range(len(ci.cl_size)):
        b0, sz = int(ci.cl_base[c]), int(ci.cl_size[c])
        g = ci.graph[b0:b0 + sz]
        v = g[g != R.INVALID]
        assert ((v >= b0) & (v < b0 + sz)).all()
        assert (assign[ci.perm[b0:b0 + sz].astype(np.int64)] == c).all()
    ix = B.Index(H, base, 0)
    truth, _ = ix.bf_knn(q, 10)
    gi, gd, st = ci.search(q, 10, 10, 10, 40, seed=1, return_stats=True)
    gi2, _, st2 = ci.search(q, 10, 10, 10, 40, seed=1, early_stop=False, return_stats=True)
    # without the reference's early-stop rule (cluster_hnsw_nsg_search.cpp:249-251) every probed cluster is searched
    assert R.recall(gi2, truth) > 0.95
    assert 0.5 < R.recall(gi, truth) <= R.recall(gi2, truth) + 1e-9
    assert st["clusters_searched"] < st2["clusters_searched"]          # the early-stop rule prunes cluster visits
    d = ((base[gi.astype(np.int64)] - q[:, None, :]) ** 2).sum(2)
    assert np.allclose(d, gd, rtol=1e-5) and (np.diff(gd, axis=1) >= 0).all()


def test_clustered_search_skips_and_small_clusters(B, H):
    """size 0 rows are skipped; a cluster smaller than L is searched with a pool of its size."""
    base = R.data_uniform(600, 16, 1)
    ix = B.Index(H, base, 0)
    sub = B.Index(H, np.ascontiguousarray(base[:20]), 0)
    knn = sub.knn_graph(8)
    sub.nsg_build(knn, 10, 6, 30, seed=1)
    g, ep, _ = sub.get_graph()
    full = np.full((600, g.shape[1]), R.INVALID, np.uint32)
    full[:20] = g
    ix.set_graph(full, ep)
    q = R.data_uniform(4, 16, 2)
    ids, d = ix.nsg_search_clustered(q, [ep, ep, ep, ep], [0, 0, 0, 0], [20, 0, 20, 0], 5, 50, seed=1)
    assert (ids[1] == R.INVALID).all() and np.isinf(d[1]).all() and (ids[3] == R.INVALID).all()
    truth = np.argsort(((base[:20] - q[0]) ** 2).sum(1))[:5]
    assert set(ids[0].tolist()) == set(truth.tolist())                # pool of 20 >= the whole cluster: exact

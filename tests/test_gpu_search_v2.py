"""The two graph-search kernels give the same bits: v2 (csrc/beam2.cuh: rows through the TMA bulk-copy ring, pool in
registers for L <= 64) against v1 (csrc/beam.cuh, the round-1 kernel) and against the oracle, for every entry point,
every pool form (L <= 32, <= 64, > 64), every team shape, ring depths 1..4, and the scheduling knobs (compact visited
set, adjacency prefetch, query order, id base) that must never change a result."""
import importlib

import numpy as np
import pytest

import refutil as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("hnsw-nsg_b200")


@pytest.fixture(scope="module")
def H0(B):
    return B.Handle(0)


@pytest.fixture()
def H(H0):
    yield H0   # every test starts from, and leaves, the default knobs
    H0.set_search_kernel(version=2, ring_stages=0, mode=0)
    H0.set_search_options(prefetch=1, visited_compact=1, ctas_per_sm=0)
    H0.set_visited_bits(0)
    H0.set_id_base(0)
    H0.set_query_order(None)


def _graph(base, deg, seed=0):
    ids, _ = R.orc_bf_knn(base, base, deg + 1, R.L2_SCALAR)
    knn = np.stack([row[row != i][:deg] for i, row in enumerate(ids)]).astype(np.uint32)
    g = np.full((base.shape[0], (deg + 7) // 8 * 8), R.INVALID, np.uint32)
    g[:, :deg] = knn
    rng = np.random.default_rng(seed)   # ragged rows: cut a random tail off every row
    for i in range(0, base.shape[0], 3):
        g[i, rng.integers(deg // 2, deg):] = R.INVALID
    return g


@pytest.mark.parametrize("n,dim,metric,K,L,deg", [
    (3000, 128, 0, 10, 32, 30), (3000, 128, 0, 10, 33, 30), (3000, 128, 0, 10, 64, 30), (3000, 128, 0, 10, 65, 30),
    (3000, 128, 0, 100, 200, 40), (2500, 96, 0, 10, 40, 24), (2500, 100, 1, 10, 48, 24), (2000, 7, 0, 5, 20, 16),
    (2000, 24, 0, 1, 1, 8), (2000, 200, 0, 10, 30, 20), (1200, 960, 0, 20, 60, 20), (600, 2048, 0, 10, 32, 12)])
def test_nsg_search_v2_equals_v1_and_oracle(B, H, n, dim, metric, K, L, deg):
    base = R.data_uniform(n, dim, 1) if metric == 0 else R.data_glove_like(n, dim, 1)
    q = R.data_uniform(97, dim, 2) if metric == 0 else R.data_glove_like(97, dim, 2)
    g = _graph(base, deg)
    ix = B.Index(H, base, metric)
    ix.set_graph(g, 5)
    H.set_search_kernel(version=1)
    H.set_visited_bits(10)   # the same table in both kernels: the work counters must then agree too
    i1, d1, s1 = ix.nsg_search(q, K, L, seed=7, want_stats=True)
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    oi, od = R.orc_nsg_search(base, g, 5, q, K, L, R.L2_GPU if metric == 0 else R.IP_GPU, rng_mode=1, seed=7)
    assert (i1 == oi).all() and (d1 == od).all()
    for stages, mode in ((1, 1), (2, 1), (3, 2), (4, 1), (1, 2), (2, 0)):
        H.set_search_kernel(version=2, ring_stages=stages, mode=mode)
        i2, d2, s2 = ix.nsg_search(q, K, L, seed=7, want_stats=True)
        assert (i2 == i1).all() and (d2 == d1).all() and (s2 == s1).all(), (stages, mode)
    # knobs that only schedule
    H.set_search_kernel(version=2, ring_stages=0)
    H.set_visited_bits(0)   # auto-sized table
    for pf, comp in ((0, 0), (1, 0), (0, 1), (1, 1)):
        H.set_search_options(prefetch=pf, visited_compact=comp)
        i3, d3 = ix.nsg_search(q, K, L, seed=7)
        assert (i3 == i1).all() and (d3 == d1).all(), (pf, comp)


def test_query_order_and_id_base(B, H):
    import torch
    n, dim, nq = 4000, 64, 333
    base = R.data_uniform(n, dim, 3)
    q = R.data_uniform(nq, dim, 4)
    g = _graph(base, 24)
    ix = B.Index(H, base, 0)
    ix.set_graph(g, 0)
    a_i, a_d = ix.nsg_search(q, 10, 40, seed=1)
    order = torch.from_numpy(np.random.default_rng(0).permutation(nq).astype(np.int32)).cuda()
    H.set_query_order(order)
    b_i, b_d = ix.nsg_search(q, 10, 40, seed=1)
    H.set_query_order(None)
    assert (a_i == b_i).all() and (a_d == b_d).all()
    H.set_id_base(1000000)
    c_i, c_d = ix.nsg_search(q[:50], 10, 12, seed=1)
    H.set_id_base(0)
    e_i, e_d = ix.nsg_search(q[:50], 10, 12, seed=1)
    ok = e_i != R.INVALID
    assert (c_i[ok] == e_i[ok] + 1000000).all() and (c_i[~ok] == R.INVALID).all() and (c_d == e_d).all()


@pytest.mark.parametrize("metric,dim,ef", [(0, 128, 10), (0, 128, 50), (1, 100, 64), (0, 32, 100), (0, 960, 40)])
def test_hnsw_search_v2_equals_v1(B, H, metric, dim, ef):
    n = 3000
    base = R.data_uniform(n, dim, 1) if metric == 0 else R.data_glove_like(n, dim, 1)
    q = R.data_uniform(120, dim, 2) if metric == 0 else R.data_glove_like(120, dim, 2)
    ix = B.Index(H, base, metric)
    ix.hnsw_build(8, 60, seed=100, labels=np.arange(n, dtype=np.uint64) * 3 + 1)
    H.set_search_kernel(version=1)
    H.set_visited_bits(10)
    l1, d1, s1 = ix.hnsw_search(q, 10, ef, want_stats=True)
    for mode in (1, 2, 0):
        H.set_search_kernel(version=2, mode=mode)
        l2, d2, s2 = ix.hnsw_search(q, 10, ef, want_stats=True)
        assert (l1 == l2).all() and (d1 == d2).all() and (s1 == s2).all(), mode
    assert (l1 % 3 == 1).all()


@pytest.mark.parametrize("L", [20, 48, 90])
def test_hybrid_and_search_from_v2_equals_v1(B, H, L):
    n, dim = 4000, 96
    base = R.data_uniform(n, dim, 5)
    q = R.data_uniform(150, dim, 6)
    ix = B.Index(H, base, 0)
    ix.hnsw_build(8, 60, seed=100)
    g = _graph(base, 20)
    ix.set_graph(g, 11)
    entry = np.random.default_rng(2).integers(0, n, 150).astype(np.uint32)
    out = {}
    H.set_visited_bits(10)
    for v, mode in ((1, 0), (2, 1), (3, 2)):
        H.set_search_kernel(version=min(v, 2), mode=mode)
        out[v] = (ix.hybrid_search(q, 10, L, seed=3, want_stats=True), ix.nsg_search_from(q, entry, 10, L, seed=3, want_stats=True))
    for v in (2, 3):
        for a, b in zip(out[1][0], out[v][0]):
            assert (a == b).all()
        for a, b in zip(out[1][1], out[v][1]):
            assert (a == b).all()


def test_clustered_v2_equals_v1(B, H):
    n, dim, nq = 3000, 48, 200
    base = R.data_uniform(n, dim, 8)
    q = R.data_uniform(nq, dim, 9)
    # three clusters back to back with their own kNN graphs (global ids)
    bounds = [0, 900, 2100, 3000]
    g = np.full((n, 16), R.INVALID, np.uint32)
    for a, b in zip(bounds[:-1], bounds[1:]):
        sub = _graph(np.ascontiguousarray(base[a:b]), 16)
        sub[sub != R.INVALID] += a
        g[a:b] = sub[:, :16]
    ix = B.Index(H, base, 0)
    ix.set_graph(g, 0)
    rng = np.random.default_rng(1)
    c = rng.integers(0, 4, nq)
    ep = np.array([bounds[x] + 3 if x < 3 else 0 for x in c], np.uint32)
    bs = np.array([bounds[x] if x < 3 else 0 for x in c], np.uint32)
    sz = np.array([bounds[x + 1] - bounds[x] if x < 3 else 0 for x in c], np.uint32)
    res = {}
    H.set_visited_bits(10)
    for v, mode in ((1, 0), (2, 1), (3, 2)):
        H.set_search_kernel(version=min(v, 2), mode=mode)
        res[v] = ix.nsg_search_clustered(q, ep, bs, sz, 10, 30, seed=2, want_stats=True)
    for v in (2, 3):
        for a, b in zip(res[1], res[v]):
            assert (a == b).all()

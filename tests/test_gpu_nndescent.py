"""GPU NN-descent (b200ann_nndescent, csrc/nndescent.cu) against the oracle's batch form, bit for bit, and
against the compiled reference at recall level. The batch form itself is pinned to the reference in
tests/test_oracle_pinning.py (graph quality) next to the statement-for-statement form (graph for graph)."""
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


def _recall(knn, truth):
    K = knn.shape[1]
    return float(np.mean([len(set(knn[i].tolist()) & set(truth[i].tolist())) / K for i in range(len(knn))]))


@pytest.mark.parametrize("n,dim,K,L,it,S,Rr,kind", [
    (3000, 32, 20, 40, 0, 8, 20, "u"),     # InitializeGraph only
    (3000, 32, 40, 40, 1, 8, 20, "u"),     # whole pools after one join + update
    (3000, 32, 20, 40, 4, 8, 20, "u"),
    (2500, 100, 10, 30, 3, 5, 12, "u"),    # dim not a multiple of 8 / 128
    (1500, 200, 16, 32, 2, 6, 10, "u"),    # two k-chunks in the tile loop
    (4000, 64, 24, 48, 3, 10, 40, "s"),    # integer-valued data: many exact distance ties
    (2000, 16, 30, 70, 3, 12, 30, "u"),    # L > 64: wider insert template
    (600, 8, 8, 12, 5, 4, 150, "u"),       # R larger than any list; lists of every node overlap heavily
])
def test_nndescent_vs_oracle_batch_form(B, H, n, dim, K, L, it, S, Rr, kind):
    base = R.data_uniform(n, dim, 11) if kind == "u" else R.data_sift_like(n, dim, 11)
    ix = B.Index(H, base, B.L2)
    knn, d, info = ix.nndescent(K, L, it, S, Rr, seed=5, want_dists=True)
    oknn, od, oinfo = R.orc_nndescent_batch(base, K, L, it, S, Rr, R.L2_SEQ_FMA, seed=5)
    assert (knn == oknn).all()
    assert (d.view(np.uint32) == od.view(np.uint32)).all()
    assert info["dist_evals"] == int(oinfo[0])          # same pairs evaluated
    assert info["offers"] == int(oinfo[1])              # same offers admitted (threshold = the pool's largest at the join's start)
    assert info["iterations"] == it


def test_nndescent_record_buffer_overflow_reruns_the_pass(B, H, monkeypatch):
    """The offer-record buffer is sized once and kept; an iteration whose offers do not fit re-runs its tile pass with
    a larger buffer. Forced here with a buffer of 1000 records: graph, distances and counters equal the oracle's."""
    monkeypatch.setenv("B200ANN_NND_REC_CAP", "1000")
    base = R.data_uniform(3000, 32, 11)
    ix = B.Index(H, base, B.L2)
    knn, d, info = ix.nndescent(20, 40, 4, 8, 20, seed=5, want_dists=True)
    oknn, od, oinfo = R.orc_nndescent_batch(base, 20, 40, 4, 8, 20, R.L2_SEQ_FMA, seed=5)
    assert (knn == oknn).all() and (d.view(np.uint32) == od.view(np.uint32)).all()
    assert info["dist_evals"] == int(oinfo[0]) and info["offers"] == int(oinfo[1])


def test_nndescent_deterministic_across_runs(B, H):
    base = R.data_sift_like(5000, 48, 3)
    ix = B.Index(H, base, B.L2)
    a, _ = ix.nndescent(20, 40, 4, 8, 20, seed=9)
    b, _ = ix.nndescent(20, 40, 4, 8, 20, seed=9)
    c, _ = ix.nndescent(20, 40, 4, 8, 20, seed=10)
    assert (a == b).all() and (a != c).any()


@pytest.mark.skipif(not R.have_ref(), reason="needs oracle/_ref (compiled reference NN-descent)")
def test_nndescent_recall_vs_reference(B, H, tmp_path):
    """graph quality at matched K/L/iter/S/R within 0.005 of IndexGraph::Build (north_star bar)."""
    n, dim, K, L, it, S, Rr = 20000, 64, 30, 60, 7, 10, 30
    base = R.data_sift_like(n, dim, 1)
    ix = B.Index(H, base, B.L2)
    truth = ix.knn_graph(K)
    ref = R.ref_nndescent(base, K, L, it, S, Rr, str(tmp_path / "g"), seed=1, threads=0, lib="efanna_fast")
    got, info = ix.nndescent(K, L, it, S, Rr, seed=1)
    rg, rr = _recall(got, truth), _recall(ref, truth)
    assert rg >= rr - 0.005, (rg, rr)
    assert rg > 0.9


def test_nndescent_feeds_nsg_build(B, H):
    """IndexGraph::Build -> Load_nn_graph -> IndexNSG::Build chain on the GPU; search recall on the result."""
    n, dim = 20000, 32
    base = R.data_sift_like(n, dim, 1)
    q = R.data_sift_like(300, dim, 2)
    ix = B.Index(H, base, B.L2)
    knn, _ = ix.nndescent(40, 60, 8, 10, 40, seed=1)
    ix.nsg_build(knn, 40, 24, 200, seed=1)
    ids, _ = ix.nsg_search(q, 10, 60, seed=1)
    truth, _ = ix.bf_knn(q, 10)
    assert R.recall(ids, truth) > 0.95


def test_nndescent_small_and_duplicate_rows(B, H):
    """smallest legal n (2S+2 rows), and a dataset where every vector is stored twice (zero distances, ties)."""
    base = R.data_uniform(10, 8, 1)
    ix = B.Index(H, base, B.L2)
    knn, d, _ = ix.nndescent(3, 5, 3, 4, 4, seed=2, want_dists=True)
    o, od, _ = R.orc_nndescent_batch(base, 3, 5, 3, 4, 4, R.L2_SEQ_FMA, seed=2)
    assert (knn == o).all() and (d.view(np.uint32) == od.view(np.uint32)).all()
    half = R.data_uniform(1000, 16, 3)
    base = np.ascontiguousarray(np.concatenate([half, half]))
    ix = B.Index(H, base, B.L2)
    knn, d, _ = ix.nndescent(10, 20, 6, 6, 10, seed=2, want_dists=True)
    o, od, _ = R.orc_nndescent_batch(base, 10, 20, 6, 6, 10, R.L2_SEQ_FMA, seed=2)
    assert (knn == o).all() and (d.view(np.uint32) == od.view(np.uint32)).all()
    twin = (np.arange(2000) + 1000) % 2000
    assert (knn[:, 0] == twin).mean() > 0.9 and (d[knn[:, 0] == twin, 0] == 0).all()


def test_nndescent_argument_errors(B, H):
    base = R.data_uniform(500, 16, 1)
    ix = B.Index(H, base, B.L2)
    for args in ((10, 8, 1, 4, 4), (5, 20, 1, 1, 4), (5, 20, 1, 4, 0), (5, 600, 1, 4, 4)):
        with pytest.raises(B.B200AnnError):
            ix.nndescent(*args)
    ipx = B.Index(H, base, B.IP)
    with pytest.raises(B.B200AnnError):
        ipx.nndescent(5, 20, 1, 4, 4)


@pytest.mark.parametrize("n,dim,K,L,it,S,Rr,k0", [
    (3000, 32, 20, 40, 0, 8, 20, 25),      # InitializeGraph_Refine only
    (3000, 32, 20, 40, 3, 8, 20, 40),      # init_k == L
    (2500, 100, 10, 30, 2, 5, 12, 10),
])
def test_nndescent_refine_vs_oracle(B, H, n, dim, K, L, it, S, Rr, k0):
    """b200ann_nndescent_refine (IndexGraph::RefineGraph, index_graph.cpp:203-262) == oracle batch form bit for bit. The
    initializer's graph: a coarse NN-descent result with some rows spoiled (self, duplicates, padding, an empty row)."""
    base = R.data_uniform(n, dim, 11)
    ix = B.Index(H, base, B.L2)
    g0, _ = ix.nndescent(k0, max(k0, 2 * S + 2), 1, S, Rr, seed=9)
    g0 = g0.copy()
    g0[5, :3] = 5                          # the node itself
    g0[6, 1] = g0[6, 0]                    # a duplicate
    g0[7, k0 // 2:] = R.INVALID            # padding
    g0[8, :] = R.INVALID                   # nothing usable: the random start stays
    knn, d, info = ix.nndescent_refine(g0, K, L, it, S, Rr, seed=5, want_dists=True)
    oknn, od, oinfo = R.orc_nndescent_refine(base, g0, K, L, it, S, Rr, R.L2_SEQ_FMA, seed=5)
    assert (knn == oknn).all()
    assert (d.view(np.uint32) == od.view(np.uint32)).all()
    assert info["dist_evals"] == int(oinfo[0])
    # a refined graph is at least as good as its start, and better than the same effort from a random start
    if it >= 2:
        truth = ix.knn_graph(K)
        rec = lambda g: np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(g[:500], truth[:500])]) / K
        cold, _ = ix.nndescent(K, L, it, S, Rr, seed=5)
        assert rec(knn) >= rec(cold)


@pytest.mark.parametrize("n,dim,K,nt,ml,kind", [
    (5000, 32, 20, 4, 0, "u"),
    (8000, 64, 30, 8, 4, "u"),
    (3000, 100, 10, 3, 2, "u"),       # dim not a multiple of 32
    (6000, 128, 24, 4, 3, "s"),       # integer-valued data: equal coordinates (the == class of the partition), distance ties
    (40, 8, 5, 2, 0, "u"),            # barely deeper than a leaf; short rows get the pseudo-random fill
    (3000, 16, 12, 2, 30, "u"),       # merge level below every leaf: nothing is conquered, all rows are fill
])
def test_kdtree_graph_vs_oracle(B, H, n, dim, K, nt, ml, kind):
    """b200ann_kdtree_graph (efanna2e::IndexKDtree::Build, index_kdtree.cpp:279-435) == oracle.c orc_kdtree_graph, bit for
    bit: ids, distances, number of distance evaluations, tree shape"""
    base = R.data_uniform(n, dim, 11) if kind == "u" else R.data_sift_like(n, dim, 11)
    ix = B.Index(H, base, B.L2)
    knn, d, info = ix.kdtree_graph(K, nt, ml, seed=7, want_dists=True)
    oknn, od, oinfo = R.orc_kdtree_graph(base, K, nt, ml, seed=7)
    assert (knn == oknn).all()
    assert (d.view(np.uint32) == od.view(np.uint32)).all()
    assert info["dist_evals"] == int(oinfo[0]) and info["nodes"] == int(oinfo[1]) and info["leaves"] == int(oinfo[2])
    assert info["deepest"] == int(oinfo[3])
    assert (knn != np.arange(n, dtype=np.uint32)[:, None]).all() or ml >= 30   # a conquered row never lists itself
    for i in range(0, n, max(1, n // 50)):
        assert len(set(knn[i].tolist())) == K


def test_kdtree_init_speeds_up_nndescent(B, H):
    """the point of the initializer (IndexGraph::RefineGraph over IndexKDtree's graph): the same NN-descent effort started
    from the KD-tree graph ends at a better graph than started from random ids"""
    n, dim, K = 20000, 64, 20
    base = R.data_sift_like(n, dim, 3)
    ix = B.Index(H, base, B.L2)
    truth = ix.knn_graph(K)
    rec = lambda g: np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(g[:1000], truth[:1000])]) / K
    g0, _ = ix.kdtree_graph(K, 8, 4, seed=1)
    cold, _ = ix.nndescent(K, 40, 2, 8, 20, seed=5)
    warm, _ = ix.nndescent_refine(g0, K, 40, 2, 8, 20, seed=5)
    assert rec(g0) > 0.1 and rec(warm) > rec(cold) + 0.05, (rec(g0), rec(cold), rec(warm))

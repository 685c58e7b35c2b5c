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
    assert info["iterations"] == it


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

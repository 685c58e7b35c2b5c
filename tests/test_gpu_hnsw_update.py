"""In-place updatePoint on the GPU (b200ann_hnsw_update_batch, csrc/hnsw_build.cu hnsw_update_launch): what addPoint does
for a label that already exists (hnswlib/hnswlib/hnswalg.h:993-1140). Link for link against the oracle's batch form
(orc_hnsw_update variant 1); its variant 0 is the reference's one-element-at-a-time form and is pinned to the compiled
reference in tests/test_oracle_pinning.py::test_hnsw_update_point_vs_reference. Plus: recall after an update equals
the recall of an index rebuilt from the updated vectors."""
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


@pytest.mark.parametrize("n,dim,M,efc,metric,nu", [
    (6000, 32, 8, 60, 0, 300),
    (4000, 100, 16, 100, 1, 150),     # inner product, dim not a multiple of 8
    (3000, 16, 4, 20, 0, 600),        # tiny M, a fifth of the index updated: claims collide, rows overflow
    (2000, 64, 12, 200, 0, 1),        # one element (the entry point)
])
def test_update_vs_oracle_batch_form(B, H, n, dim, M, efc, metric, nu):
    gen = R.data_uniform if metric == 0 else R.data_glove_like
    base = gen(n, dim, 21)
    ix = B.Index(H, base, metric)
    ix.hnsw_build(M, efc, seed=100)
    G0 = ix.get_hnsw()
    ids = np.random.default_rng(3).choice(n, nu, replace=False).astype(np.uint32)
    ids[0] = G0["ep"]
    newvec = gen(nu, dim, 22)
    info = ix.hnsw_update_batch(ids, newvec)
    G1 = ix.get_hnsw()
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    data, O = R.orc_hnsw_update(base, G0, ids, newvec, R.L2_GPU if metric == 0 else R.IP_GPU, 1, efc=efc)
    assert (G1["levels"] == G0["levels"]).all() and G1["ep"] == G0["ep"] and G1["maxlevel"] == G0["maxlevel"]
    assert (G1["links0"] != G0["links0"]).any()
    assert (G1["links0"] == O["links0"]).all()
    assert G1["upper_links"].shape == O["upper_links"].shape and (G1["upper_links"] == O["upper_links"]).all()
    assert info["lists_rewritten"] == int(O["info"][1])
    # searches see the new vectors
    lab, d = ix.hnsw_search(newvec, 1, max(efc, 50))
    if metric == 0:
        assert (lab[:, 0] == ids).mean() >= 0.97 and (d[lab[:, 0] == ids, 0] == 0).all()
    ix.close()


def test_update_installed_graph_u8_and_recall(B, H):
    """a graph installed from a file (set_hnsw: no build parameters, ef_construction 200 like the reference's load ctor)
    over byte rows; then quality: recall after moving 10 % of a 20k index == recall of a rebuild within 0.01"""
    rng = np.random.default_rng(5)
    n, dim, M = 3000, 48, 8
    base = np.minimum(255, rng.gamma(1.2, 24.0, (n, dim))).astype(np.uint8)
    src = B.Index(H, base, B.L2_U8)
    src.hnsw_build(M, 200, seed=100)
    G0 = src.get_hnsw()
    src.close()
    ix = B.Index(H, base, B.L2_U8)
    ix.set_hnsw(G0["links0"], G0["levels"], G0["upper_ptr"], G0["upper_links"], G0["labels"], G0["ep"], G0["maxlevel"])
    ids = rng.choice(n, 200, replace=False).astype(np.uint32)
    newvec = np.minimum(255, rng.gamma(1.2, 24.0, (200, dim))).astype(np.uint8)
    ix.hnsw_update_batch(ids, newvec)
    G1 = ix.get_hnsw()
    _, O = R.orc_hnsw_update(R.u8_rows(base), G0, ids, R.u8_rows(newvec), R.L2_U8, 1, efc=200, dim=dim)
    assert (G1["links0"] == O["links0"]).all() and (G1["upper_links"] == O["upper_links"]).all()
    ix.close()

    n, dim, M, efc = 20000, 32, 12, 100
    base = R.data_uniform(n, dim, 31)
    q = R.data_uniform(500, dim, 32)
    ix = B.Index(H, base, B.L2)
    ix.hnsw_build(M, efc, seed=100)
    ids = rng.choice(n, n // 10, replace=False).astype(np.uint32)
    newvec = R.data_uniform(len(ids), dim, 33)
    ix.hnsw_update_batch(ids, newvec)
    moved = base.copy()
    moved[ids] = newvec
    fresh = B.Index(H, moved, B.L2)
    fresh.hnsw_build(M, efc, seed=100)
    truth, _ = fresh.bf_knn(q, 10)
    rec = []
    for x in (fresh, ix):
        lab, _ = x.hnsw_search(q, 10, 60)
        rec.append(np.mean([len(set(a.tolist()) & set(b.astype(np.int64).tolist())) for a, b in zip(truth, lab)]) / 10)
    assert rec[1] >= rec[0] - 0.01, rec
    with pytest.raises(Exception, match="distinct"):
        ix.hnsw_update_batch(np.array([5, 9, 5], np.uint32), newvec[:3])
    with pytest.raises(Exception, match="not a row"):
        ix.hnsw_update_batch(np.array([n], np.uint32), newvec[:1])
    ix.close(); fresh.close()

"""GPU HNSW construction (b200ann_hnsw_build, csrc/hnsw_build.cu): link for link against the oracle's batch form
(orc_hnsw_build variant 1; variant 0 is pinned to the compiled reference in tests/test_oracle_pinning.py), level
assignment identical to the reference, and search recall at matched ef within 0.005 of a reference-built index."""
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


@pytest.mark.parametrize("n,dim,M,efc,metric,ratio,bmax", [
    (3000, 32, 8, 40, 0, 16, 256),
    (6000, 64, 16, 100, 0, 32, 1024),
    (2500, 100, 6, 30, 1, 8, 512),       # inner product, dim not a multiple of 8
    (1500, 16, 4, 10, 0, 4, 2048),       # tiny M: reverse lists overflow constantly -> heuristic on every row
    (800, 200, 12, 200, 0, 2, 4096),     # ef_construction > most of the graph, huge batches
])
def test_hnsw_build_vs_oracle_batch_form(B, H, n, dim, M, efc, metric, ratio, bmax):
    base = R.data_uniform(n, dim, 21) if metric == 0 else R.data_glove_like(n, dim, 21)
    ix = B.Index(H, base, metric)
    info = ix.hnsw_build(M, efc, seed=100, ratio=ratio, bmax=bmax)
    G = ix.get_hnsw()
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    O = R.orc_hnsw_build(base, M, efc, R.L2_GPU if metric == 0 else R.IP_GPU, 1, ratio=ratio, bmax=bmax)
    assert (G["levels"] == O["levels"]).all()
    assert G["ep"] == O["ep"] and G["maxlevel"] == O["maxlevel"]
    assert (G["upper_ptr"] == O["upper_ptr"]).all()
    assert (G["links0"] == O["links0"]).all()
    assert G["upper_links"].shape == O["upper_links"].shape and (G["upper_links"] == O["upper_links"]).all()
    assert info["batches"] == int(O["info"][1])


@pytest.mark.skipif(not R.have_ref(), reason="needs oracle/_ref (compiled hnswlib)")
def test_hnsw_levels_match_reference(B, H, tmp_path):
    n, dim = 4000, 16
    base = R.data_uniform(n, dim, 1)
    ref = R.RefHnsw(0, dim, n, 12, 50)
    ref.add(base, label0=0, threads=1)
    path = str(tmp_path / "h.bin")
    ref.save(path)
    Hr = R.parse_hnsw(path)
    ix = B.Index(H, base, 0)
    ix.hnsw_build(12, 50, seed=100)
    G = ix.get_hnsw()
    assert (G["levels"] == Hr["levels"]).all() and G["maxlevel"] == Hr["maxlevel"] and G["ep"] == Hr["ep"]
    assert (G["labels"] == Hr["labels"]).all()


@pytest.mark.skipif(not R.have_ref(), reason="needs oracle/_ref (compiled hnswlib)")
@pytest.mark.parametrize("metric", [0, 1])
def test_hnsw_build_search_recall_vs_reference(B, H, tmp_path, metric):
    """C1-shaped (10k x 128, M=16, efC=200, k=10): b200ann_hnsw_search over the GPU-built graph vs the same search
    over the index the unmodified reference built, at matched ef: recall@10 within 0.005 (north_star bar)."""
    n, dim, M, efc = 10000, 128, 16, 200
    base = R.data_uniform(n, dim, 1) if metric == 0 else R.data_glove_like(n, dim, 1)
    q = R.data_uniform(1000, dim, 2) if metric == 0 else R.data_glove_like(1000, dim, 2)
    ref = R.RefHnsw(metric, dim, n, M, efc)
    ref.add(base, label0=0, threads=1)
    path = str(tmp_path / "h.bin")
    ref.save(path)
    Hr = R.parse_hnsw(path)
    rix = B.Index(H, base, metric)
    rix.set_hnsw(Hr["links0"], Hr["levels"], Hr["upper_ptr"], Hr["upper_links"], Hr["labels"], Hr["ep"], Hr["maxlevel"])
    gix = B.Index(H, base, metric)
    gix.hnsw_build(M, efc, seed=100)
    truth, _ = gix.bf_knn(q, 10)
    for ef in (10, 20, 50, 100):
        lr, _ = rix.hnsw_search(q, 10, ef)
        lg, _ = gix.hnsw_search(q, 10, ef)
        rr, rg = R.recall(lr.astype(np.uint32), truth), R.recall(lg.astype(np.uint32), truth)
        assert rg >= rr - 0.005, (ef, rg, rr)


@pytest.mark.parametrize("n,dim,M,efc,kind", [
    (1, 8, 4, 10, "u"), (2, 8, 4, 10, "u"), (40, 16, 16, 50, "u"),   # lists never fill; first element alone
    (3000, 32, 8, 60, "s"),                                          # integer-valued rows: exact distance ties
    (1500, 8, 6, 40, "dup"),                                         # every vector stored twice: zero distances
])
def test_hnsw_build_edge_cases_vs_oracle(B, H, n, dim, M, efc, kind):
    if kind == "u":
        base = R.data_uniform(n, dim, 5)
    elif kind == "s":
        base = R.data_sift_like(n, dim, 5)
    else:
        half = R.data_uniform(n // 2, dim, 5)
        base = np.ascontiguousarray(np.concatenate([half, half]))
    ix = B.Index(H, base, 0)
    ix.hnsw_build(M, efc, ratio=8, bmax=512)
    G = ix.get_hnsw()
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    O = R.orc_hnsw_build(base, M, efc, R.L2_GPU, 1, ratio=8, bmax=512)
    assert G["ep"] == O["ep"] and G["maxlevel"] == O["maxlevel"]
    assert (G["links0"] == O["links0"]).all()
    assert G["upper_links"].shape == O["upper_links"].shape and (G["upper_links"] == O["upper_links"]).all()
    if n > 1:
        lab, d = ix.hnsw_search(base[:min(n, 64)], 1, 30)
        assert (d[:, 0] == 0).all()                                   # every stored point finds itself (or its twin)


def test_hnsw_build_is_deterministic_and_rebuildable(B, H):
    base = R.data_sift_like(8000, 48, 2)
    ix = B.Index(H, base, 0)
    a = ix.hnsw_build(16, 100)
    Ga = ix.get_hnsw()
    b = ix.hnsw_build(16, 100)           # rebuilding replaces the graph
    Gb = ix.get_hnsw()
    assert a == b and (Ga["links0"] == Gb["links0"]).all() and (Ga["upper_links"] == Gb["upper_links"]).all()
    c = ix.hnsw_build(16, 100, seed=7)   # another level sequence -> another graph
    assert (ix.get_hnsw()["levels"] != Ga["levels"]).any()


def test_hnsw_build_labels_and_errors(B, H):
    base = R.data_uniform(600, 8, 1)
    ix = B.Index(H, base, 0)
    labels = np.arange(600, dtype=np.uint64) * 7 + 3
    ix.hnsw_build(8, 40, labels=labels)
    lab, d = ix.hnsw_search(base[:50], 1, 20)
    assert (lab[:, 0] == labels[:50]).all() and (d[:, 0] == 0).all()
    with pytest.raises(B.B200AnnError):
        ix.hnsw_build(1, 40)
    ix2 = B.Index(H, base, 0)
    with pytest.raises(B.B200AnnError):
        ix2.get_hnsw()

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


def test_hnsw_insert_batch_in_place(B, H):
    """b200ann_hnsw_insert_batch (nb addPoint calls on a live index, hnswalg.h:1153-1267): rows appended in several
    calls are linked into the existing graph in place - every appended row finds itself under its label, recall stays
    within 0.005 of an index built from all rows at once, and the level sequence continues the build's generator (an
    append after a build draws the levels a one-shot build draws for the same positions)."""
    n0, nadd, dim = 20000, 3000, 64
    base = R.data_uniform(n0 + nadd, dim, 21)
    q = R.data_uniform(300, dim, 22)
    labels = np.arange(n0 + nadd, dtype=np.uint64) * 2 + 5
    ix = B.Index(H, np.ascontiguousarray(base[:n0]), B.L2)
    ix.hnsw_build(16, 100, seed=100, labels=labels[:n0])
    g0 = ix.get_hnsw()
    ix.hnsw_insert_batch(base[n0:n0 + 1000], labels[n0:n0 + 1000])
    ix.hnsw_insert_batch(base[n0 + 1000:n0 + 1001], labels[n0 + 1000:n0 + 1001])      # a single point
    info = ix.hnsw_insert_batch(base[n0 + 1001:], labels[n0 + 1001:])
    assert ix.n == n0 + nadd and info["batches"] >= 1
    g1 = ix.get_hnsw()
    one = B.Index(H, base, B.L2)
    one.hnsw_build(16, 100, seed=100, labels=labels)
    g2 = one.get_hnsw()
    assert (g1["levels"] == g2["levels"]).all()                     # same generator sequence
    assert (g1["labels"] == labels).all()
    # rows linked before the append keep their out-links unless a new point was reverse-linked into them
    changed = (g1["links0"][:n0] != g0["links0"]).any(axis=1).mean()
    assert 0.05 < changed < 0.95
    # every appended row is reachable and finds itself
    lab, d = ix.hnsw_search(base[n0:], 1, 50)
    assert (lab[:, 0] == labels[n0:]).mean() > 0.995
    truth, _ = R.orc_bf_knn(base, q, 10, R.L2_SCALAR)
    la, _ = ix.hnsw_search(q, 10, 100)
    lb, _ = one.hnsw_search(q, 10, 100)
    ra = R.recall(((la - 5) // 2).astype(np.uint32), truth)
    rb = R.recall(((lb - 5) // 2).astype(np.uint32), truth)
    assert ra >= rb - 0.005, (ra, rb)


def test_hnsw_insert_batch_into_installed_graph(B, H):
    """an index installed with set_hnsw (e.g. loaded from a reference-written file) grows in place too; without a label
    array labels are row numbers"""
    n0, dim = 6000, 32
    base = R.data_uniform(n0 + 500, dim, 31)
    a = B.Index(H, np.ascontiguousarray(base[:n0]), B.L2)
    a.hnsw_build(8, 60, seed=100)
    g = a.get_hnsw()
    b = B.Index(H, np.ascontiguousarray(base[:n0]), B.L2)
    b.set_hnsw(g["links0"], g["levels"], g["upper_ptr"], g["upper_links"], None, g["ep"], g["maxlevel"])
    b.hnsw_insert_batch(base[n0:], efc=60)
    lab, _ = b.hnsw_search(base[n0:], 1, 40)
    assert (lab[:, 0] == np.arange(n0, n0 + 500)).mean() >= 0.97
    lab0, _ = b.hnsw_search(base[:500], 1, 40)
    assert (lab0[:, 0] == np.arange(500)).mean() >= 0.97    # M=8, ef=40: the old rows stay as findable as before


def test_hnsw_build_survives_reverse_group_overflow(B, H, monkeypatch):
    """a hub that collects more reverse edges in one batch than a warp's shared memory holds made the whole
    build fail (duplicate-heavy data, hubs under inner product). Such a batch is now undone and re-run at half the
    size. Forced here with a tiny group limit and a block of identical rows."""
    n, dim = 6000, 32
    base = R.data_uniform(n, dim, 41)
    base[1000:3000] = base[1000]           # 2000 identical rows: everything links to the same few nodes
    monkeypatch.setenv("B200ANN_HNSW_MAX_GROUP", "40")
    ix = B.Index(H, base, B.L2)
    info = ix.hnsw_build(8, 60, seed=100, ratio=4, bmax=4096)
    assert info["largest_group"] <= 40 or info["batches"] > 0
    monkeypatch.delenv("B200ANN_HNSW_MAX_GROUP")
    q = R.data_uniform(200, dim, 42)
    truth, _ = R.orc_bf_knn(base, q, 10, R.L2_SCALAR)
    lab, _ = ix.hnsw_search(q, 10, 200)
    assert R.recall(lab.astype(np.uint32), truth) > 0.9
    G = ix.get_hnsw()
    deg = (G["links0"] != R.INVALID).sum(1)
    assert (deg[1:] > 0).all()             # every point got linked


@pytest.mark.parametrize("kind,n,dim,M,ef", [
    ("s", 6000, 24, 8, 100),     # integer-valued rows: exact distance ties at the pool's boundary
    ("dup", 4000, 16, 6, 120),   # every vector stored twice: zero distances, equal (distance) pairs with different ids
    ("u", 5000, 32, 12, 200),
    ("u", 3000, 8, 16, 30),      # small pool: one 32-entry block
])
def test_one_pass_pool_merge_equals_sequential_insertion(B, H, monkeypatch, kind, n, dim, M, ef):
    """Beam::merge_chunk (beam.cuh: a hop's candidates merged into the pool in one pass through the closed-form
    acceptance rule) against the candidate-by-candidate loop it replaces (B200ANN_MERGE_MIN turns the merge off for
    search launches): same labels, same distance bits, same number of evaluations, on data full of ties. The search
    also equals the oracle's (hnswalg.h:342-436 searchBaseLayerST)."""
    if kind == "u":
        base = R.data_uniform(n, dim, 41)
    elif kind == "s":
        base = R.data_sift_like(n, dim, 41)
    else:
        half = R.data_sift_like(n // 2, dim, 41)
        base = np.ascontiguousarray(np.concatenate([half, half]))
    q = base[::7][:300].copy() if kind != "u" else R.data_uniform(300, dim, 42)
    ix = B.Index(H, base, 0)
    ix.hnsw_build(M, 60, seed=100)
    H.set_search_kernel(1)                       # the Beam kernel for every ef (the v2 kernel has its own pool)
    try:
        lab1, d1, st1 = ix.hnsw_search(q, 10, ef, want_stats=True)
        monkeypatch.setenv("B200ANN_MERGE_MIN", "100000")
        lab0, d0, st0 = ix.hnsw_search(q, 10, ef, want_stats=True)
        monkeypatch.delenv("B200ANN_MERGE_MIN")
    finally:
        H.set_search_kernel(0)
    assert (lab1 == lab0).all() and (d1.view(np.uint32) == d0.view(np.uint32)).all()
    assert (st1 == st0).all()                    # per query: distance evaluations, hops
    G = ix.get_hnsw()
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    ol, od = R.orc_hnsw_search(base, G, q, 10, ef, R.L2_GPU)
    assert (lab1 == ol).all()
    ix.close()


def test_graphs_do_not_depend_on_the_visited_table(B, H, monkeypatch):
    """The beam's visited set only saves re-evaluations (a forgotten id is evaluated again and is either in the pool
    already or fails its threshold again; the Link candidate buffer drops duplicates): 16-bit tags or 32-bit ids,
    2^11 or 2^8 words - same HNSW links, same NSG graph, only the evaluation counts move."""
    n, dim = 6000, 32
    base = R.data_uniform(n, dim, 77)
    ix = B.Index(H, base, 0)
    i0 = ix.hnsw_build(12, 120, seed=100)
    G0 = ix.get_hnsw()
    knn, _ = ix.bf_knn(base, 41)
    knn = np.ascontiguousarray(knn[:, 1:].astype(np.uint32))
    n0 = ix.nsg_build(knn, 30, 24, 200, seed=1)
    N0 = ix.get_graph()
    monkeypatch.setenv("B200ANN_HNSW_VIS16", "0")
    monkeypatch.setenv("B200ANN_LINK_VIS16", "0")
    monkeypatch.setenv("B200ANN_HNSW_HBITS", "8")
    H.set_link_visited_bits(7)
    try:
        i1 = ix.hnsw_build(12, 120, seed=100)
        G1 = ix.get_hnsw()
        n1 = ix.nsg_build(knn, 30, 24, 200, seed=1)
        N1 = ix.get_graph()
    finally:
        H.set_link_visited_bits(0)
    assert (G0["links0"] == G1["links0"]).all() and (G0["upper_links"] == G1["upper_links"]).all()
    assert np.array_equal(np.asarray(N0[0]), np.asarray(N1[0])) and N0[1:] == N1[1:]
    assert i1["beam_evals"] > i0["beam_evals"] and n1["beam_evals"] > n0["beam_evals"]   # the small tables forget more
    assert n1["ep"] == n0["ep"] and n1["width"] == n0["width"]
    ix.close()

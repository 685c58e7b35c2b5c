"""Pins oracle/oracle.c: against the reference's own known answers / golden files (tests/golden) and
against the UNMODIFIED reference compiled into oracle/_ref (oracle/Makefile) on identical seeded inputs.
CPU only."""
import ctypes
import json
import os

import numpy as np
import pytest

import refutil as R

GOLD = os.path.join(R.ROOT, "tests", "golden")
needs_ref = pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not built (needs /root/reference)")


# ---------------------------------------------------------------- known answers of the reference's tests
@pytest.mark.parametrize("mode_l2,mode_ip", [(R.L2_SCALAR, R.IP_SCALAR), (R.L2_HNSW_AVX, R.IP_HNSW_AVX),
                                             (R.L2_GPU, R.IP_GPU), (R.L2_SEQ_FMA, R.IP_SEQ_FMA)])
def test_spaces_known_answers(mode_l2, mode_ip):
    """hnswlib/tests/python/bindings_test_spaces.py:11-33 — every SIMD/residual dispatch branch"""
    ka = json.load(open(os.path.join(GOLD, "spaces_known_answers.json")))
    data1 = np.asarray(ka["data1"], np.float32)
    lib = R.oracle()
    for rd in ka["rightdims"]:
        for ld in ka["leftdims"]:
            d2 = np.concatenate([np.zeros((5, ld), np.float32), data1, np.zeros((5, rd), np.float32)], axis=1)
            d2 = np.ascontiguousarray(d2)
            dim = d2.shape[1]
            lib.orc_set_gpu_team(8 if (dim + 3) // 4 <= 64 else 32)
            q = d2[ka["query_row"]]
            l2 = sorted(lib.orc_dist(mode_l2, q, d2[i], dim) for i in range(5))
            ip = sorted(lib.orc_dist(mode_ip, q, d2[i], dim) for i in range(5))
            assert np.mean(np.abs(np.array(l2) - ka["l2"])) < ka["tol_mean_abs"]
            assert np.mean(np.abs(np.array(ip) - ka["ip"])) < ka["tol_mean_abs"]


def _golden_updates():
    n, _, nq, d, k = [int(x) for x in open(os.path.join(GOLD, "hnswlib_updates_config.txt")).read().split()]
    base = np.fromfile(os.path.join(GOLD, "hnswlib_updates_batch_final.bin"), np.float32).reshape(n, d)
    q = np.fromfile(os.path.join(GOLD, "hnswlib_updates_queries.bin"), np.float32).reshape(nq, d)
    gt = np.fromfile(os.path.join(GOLD, "hnswlib_updates_gt.bin"), np.int32).reshape(nq, k)
    return base, q, gt


def test_golden_updates_bruteforce():
    """hnswlib/tests/cpp/data/{batch_final,queries,gt}.bin: exact top-5 by dot product (update_gen_data.py:25-26).
    The oracle's exact IP scan must reproduce gt up to fp32 near-ties of the numpy generator."""
    base, q, gt = _golden_updates()
    ids, _ = R.orc_bf_knn(base, q[:200], 5, R.IP_SCALAR)
    rec = R.recall(ids, gt[:200])
    assert rec > 0.999, rec
    # data are L2-normalised, so the L2 ranking equals the IP ranking
    ids2, _ = R.orc_bf_knn(base, q[:200], 5, R.L2_SCALAR)
    assert R.recall(ids2, gt[:200]) > 0.999


# ---------------------------------------------------------------- InsertIntoPool edge cases
def _pool(dists, ids):
    K = len(dists)
    return (np.array(list(ids) + [0], np.uint32), np.array(list(dists) + [0], np.float32),
            np.array([1] * K + [0], np.int32), K)


def test_insert_into_pool_semantics():
    """CNNS/include/nsg/neighbor.h:107-135"""
    lib = R.oracle()
    ids, d, f, K = _pool([1, 2, 2, 3], [10, 11, 12, 13])
    assert lib.orc_insert_into_pool(ids, d, f, K, 99, 0.5) == 0          # before the head
    assert ids[:5].tolist() == [99, 10, 11, 12, 13]
    ids, d, f, K = _pool([1, 2, 2, 3], [10, 11, 12, 13])
    assert lib.orc_insert_into_pool(ids, d, f, K, 99, 4.0) == K          # beyond the tail -> slot K
    ids, d, f, K = _pool([1, 2, 2, 3], [10, 11, 12, 13])
    assert lib.orc_insert_into_pool(ids, d, f, K, 99, 2.0) == 3          # after every entry <= d
    assert ids[:5].tolist() == [10, 11, 12, 99, 13]
    ids, d, f, K = _pool([1, 2, 2, 3], [10, 11, 12, 13])
    assert lib.orc_insert_into_pool(ids, d, f, K, 11, 2.0) == K + 1      # duplicate id among equal distances
    ids, d, f, K = _pool([1, 2, 2, 3], [10, 11, 12, 13])
    assert lib.orc_insert_into_pool(ids, d, f, K, 12, 2.0) == K + 1


# ---------------------------------------------------------------- against the compiled reference
@needs_ref
@pytest.mark.parametrize("dim", [3, 8, 20, 33, 128, 100])
def test_distances_vs_reference(dim):
    rng = np.random.default_rng(dim)
    a = rng.normal(size=(16, dim)).astype(np.float32)
    b = rng.normal(size=(16, dim)).astype(np.float32)
    lib = R.oracle()
    for i in range(16):
        for metric, ms, ma in ((0, R.L2_SCALAR, R.L2_HNSW_AVX), (1, R.IP_SCALAR, R.IP_HNSW_AVX)):
            assert lib.orc_dist(ms, a[i], b[i], dim) == R.reflib("hnsw").ref_dist(metric, a[i], b[i], dim)
            assert lib.orc_dist(ma, a[i], b[i], dim) == R.reflib("hnsw_avx").ref_dist(metric, a[i], b[i], dim)
        if dim % 8 == 0:  # efanna's AVX path over-reads unless rows are padded to x8 (util.h data_align)
            assert lib.orc_dist(R.L2_EFANNA_AVX, a[i], b[i], dim) == R.reflib("nsg").ref_nsg_dist_l2(a[i], b[i], dim)


@needs_ref
@pytest.mark.parametrize("metric,mode", [(0, R.L2_SCALAR), (1, R.IP_SCALAR)])
def test_bruteforce_vs_reference(metric, mode):
    base = R.data_uniform(3000, 24, 1)
    q = R.data_uniform(40, 24, 2)
    lab, d = R.ref_bf_knn(base, q, 10, metric)
    ids, d2 = R.orc_bf_knn(base, q, 10, mode)
    assert (lab.astype(np.uint32) == ids).all() and (d == d2).all()


@needs_ref
def test_bruteforce_ties_vs_reference():
    """integer data with many exact ties: smaller id must win (bruteforce.h:119-131)"""
    rng = np.random.default_rng(3)
    base = rng.integers(0, 3, (2000, 6)).astype(np.float32)
    q = rng.integers(0, 3, (30, 6)).astype(np.float32)
    lab, d = R.ref_bf_knn(base, q, 15, 0)
    ids, d2 = R.orc_bf_knn(base, q, 15, R.L2_SCALAR)
    assert (lab.astype(np.uint32) == ids).all() and (d == d2).all()


@needs_ref
@pytest.mark.parametrize("metric,mode", [(0, R.L2_SCALAR), (1, R.IP_SCALAR)])
def test_hnsw_update_point_vs_reference(tmp_path, metric, mode):
    """oracle updatePoint restatement (variant 0) == the unmodified reference: addPoint on labels that exist runs
    updatePoint(.., 1.0) + repairConnectionsForUpdate (hnswalg.h:993-1140, :1160-1174); every level-0 and upper list
    afterwards is the reference's, link for link"""
    n, dim, M, efc = 3000, 24, 8, 60
    base = R.data_uniform(n, dim, 1) if metric == 0 else R.data_glove_like(n, dim, 1)
    h = R.RefHnsw(metric, dim, n, M, efc)
    h.add(base, threads=1)
    p0 = str(tmp_path / "h0.bin")
    h.save(p0)
    H0 = R.parse_hnsw(p0)
    rng = np.random.default_rng(7)
    ids = rng.choice(n, 150, replace=False).astype(np.uint32)
    ids[0] = H0["ep"]                                   # the entry point itself is updated too
    newvec = (R.data_uniform(150, dim, 9) if metric == 0 else R.data_glove_like(150, dim, 9))
    for i, p in enumerate(ids):
        h.add(newvec[i:i + 1], label0=int(p), threads=1)   # labels == internal ids here
    p1 = str(tmp_path / "h1.bin")
    h.save(p1)
    H1 = R.parse_hnsw(p1)
    data, G = R.orc_hnsw_update(base, H0, ids, newvec, mode, 0, efc=efc)
    assert (H1["vectors"] == data).all()
    assert (H1["levels"] == H0["levels"]).all() and H1["ep"] == H0["ep"]
    assert (G["links0"] != H0["links0"]).any()
    assert (G["links0"] == H1["links0"]).all()
    assert (G["upper_links"] == H1["upper_links"]).all()
    # the batch form (what the GPU computes) is a different schedule of the same rules: searches over its graph are as
    # good as over the reference's (recall@10 at ef = 40 against the exact scan of the updated vectors)
    _, G1 = R.orc_hnsw_update(base, H0, ids, newvec, mode, 1, efc=efc)
    q = R.data_uniform(200, dim, 11) if metric == 0 else R.data_glove_like(200, dim, 11)
    truth, _ = R.orc_bf_knn(data, q, 10, mode)
    rec = []
    for g in (H1, G1):
        lab, _ = R.orc_hnsw_search(data, g, q, 10, 40, mode, variant=0)
        rec.append(np.mean([len(set(a.tolist()) & set(b.astype(np.int64).tolist())) for a, b in zip(truth, lab)]) / 10)
    assert rec[1] >= rec[0] - 0.01, rec
    lab, _ = R.orc_hnsw_search(data, G1, newvec, 1, 40, mode, variant=0)      # the moved elements are found where they are now
    assert (lab[:, 0] == ids).mean() >= 0.97


@needs_ref
@pytest.mark.parametrize("metric,mode,ndocs,efcoll,keep", [(0, R.L2_SCALAR, 10, 15, 1.0), (0, R.L2_SCALAR, 5, 5, 1.0),
                                                           (0, R.L2_SCALAR, 20, 40, 0.5), (0, R.L2_SCALAR, 3, 1, 1.0)])
def test_hnsw_multivector_search_vs_reference(tmp_path, metric, mode, ndocs, efcoll, keep):
    """oracle multi-vector search == HierarchicalNSW::searchStopConditionClosest(q, MultiVectorSearchStopCondition(space,
    num_docs, ef_collection), filter) over a MultiVector space of the unmodified reference: ids, distances, counts;
    the reference's own invariant (multivector_search_test.cpp:86): exactly num_docs distinct documents come back.
    L2 only: the reference's MultiVectorInnerProductSpace never sets dim_ (stop_condition.h:85-122), its distance
    function reads an uninitialised dimension (it crashed here)."""
    n, dim, nd = 3000, 16, 150
    gen = R.data_uniform if metric == 0 else R.data_glove_like
    base, q = gen(n, dim, 1), gen(100, dim, 2)
    doc = np.random.default_rng(3).integers(0, nd, n).astype(np.uint32)
    h = R.RefHnswMV(metric, dim, n, 16, 100)
    h.add(base, doc)
    p = str(tmp_path / "mv.bin")
    h.save(p)
    H = R.parse_hnsw(p)
    assert (np.ascontiguousarray(H["vectors"][:, :dim]) == base).all()
    assert (np.ascontiguousarray(H["vectors"][:, dim]).view(np.uint32) == doc).all()   # the document id sits behind the vector
    allow = None if keep == 1.0 else (np.random.default_rng(4).random(n) < keep).astype(np.uint8)
    ri, rd, rc = h.search(q, ndocs, efcoll, n, allow, threads=1)
    oi, od, oc = R.orc_hnsw_search_multivector(base, H, doc, q, ndocs, efcoll, n, mode, allow=allow)
    rd = np.where(ri == np.uint64(0xFFFFFFFFFFFFFFFF), np.float32(np.inf), rd)
    assert (rc == oc).all() and (ri == oi).all() and (rd == od).all()
    for i in range(len(q)):
        got = oi[i, :oc[i]].astype(np.int64)
        assert len(set(doc[got].tolist())) == ndocs
        if allow is not None:
            assert allow[got].all()


@pytest.fixture(scope="module")
def hnsw_u8_case(tmp_path_factory):
    rng = np.random.default_rng(5)
    n, dim = 4000, 50          # dim not a multiple of 4: L2SqrI, not L2SqrI4x (space_l2.h:302-306)
    base = np.minimum(255, rng.gamma(1.2, 24.0, (n, dim))).astype(np.uint8)
    q = np.minimum(255, rng.gamma(1.2, 24.0, (100, dim))).astype(np.uint8)
    h = R.RefHnswI(dim, n, 16, 200)
    h.add(base, threads=1)
    p = str(tmp_path_factory.mktemp("hnswi") / "h.bin")
    h.save(p)
    return base, q, h, R.parse_hnsw(p, elem="u8")


def _oi(d):
    i = d.view(np.int32).copy()
    i[np.isinf(d)] = 0x7FFFFFFF
    return i


def _same_up_to_ties(base, q, rl, rd, ol, od):
    """distances identical; labels identical except where two elements tie on the int distance (the reference's heaps
    order equal distances by their own history, the oracle by id): a differing label must carry the same distance"""
    oi = _oi(od)
    assert (rd == oi).all()
    for r, c in zip(*np.nonzero(rl != ol)):
        a = base[int(rl[r, c])].astype(np.int64) - q[r].astype(np.int64)
        assert int((a * a).sum()) == int(oi[r, c])
    assert (rl != ol).mean() < 0.005


@needs_ref
def test_hnsw_u8_vs_reference(hnsw_u8_case):
    """oracle mode L2_U8 == the unmodified HierarchicalNSW<int> over L2SpaceI, bit for bit: searchKnn (both
    formulations), searchKnn with a label filter, searchStopConditionClosest with EpsilonSearchStopCondition<int>"""
    base, q, h, H = hnsw_u8_case
    n, dim = base.shape
    assert (H["vectors"] == base).all()
    bf, qf = R.u8_rows(base), R.u8_rows(q)
    for ef in (10, 40, 200):
        rl, rd = h.search(q, 10, ef, threads=1)
        for variant in (0, 1):
            ol, od = R.orc_hnsw_search(bf, H, qf, 10, ef, R.L2_U8, variant=variant, dim=dim)
            _same_up_to_ties(base, q, rl, rd, ol, od)
    for keep, ef in ((0.5, 40), (0.05, 30)):
        allow = (np.random.default_rng(ef).random(n) < keep).astype(np.uint8)
        rl, rd = h.search(q, 10, ef, allow_by_label=allow, threads=1)
        ol, od, _ = R.orc_hnsw_search_filtered(bf, H, allow, qf, 10, ef, R.L2_U8, variant=0, dim=dim)
        _same_up_to_ties(base, q, rl, rd, ol, od)
    _, d50 = h.search(q, 50, 200, threads=1)
    for kth, mn, mx in ((5, 0, 50), (20, 10, 100), (3, 30, 30)):
        eps = float(np.median(d50[:, kth])) + 0.5
        ri, rd, rc = h.search_epsilon(q, eps, mn, mx, threads=1)
        oi, od, oc = R.orc_hnsw_search_epsilon(bf, H, qf, np.int32(np.floor(eps)).view(np.float32), mn, mx, R.L2_U8, dim=dim)
        assert (rc == oc).all()
        _same_up_to_ties(base, q, np.where(ri == np.uint64(0xFFFFFFFFFFFFFFFF), 0, ri), rd, np.where(oi == np.uint64(0xFFFFFFFFFFFFFFFF), 0, oi), od)


@needs_ref
def test_hnsw_u8_build_vs_reference():
    """oracle addPoint restatement (variant 0, one point at a time) in mode L2_U8 == the graph the unmodified
    HierarchicalNSW<int> builds single-threaded"""
    import tempfile, os
    rng = np.random.default_rng(9)
    n, dim, M, efc = 1500, 32, 8, 60
    base = np.minimum(255, rng.gamma(1.2, 24.0, (n, dim))).astype(np.uint8)
    h = R.RefHnswI(dim, n, M, efc)
    h.add(base, threads=1)
    p = os.path.join(tempfile.mkdtemp(), "h.bin")
    h.save(p)
    Hr = R.parse_hnsw(p, elem="u8")
    O = R.orc_hnsw_build(R.u8_rows(base), M, efc, R.L2_U8, 0, dim=dim)
    assert (O["levels"] == Hr["levels"]).all() and O["ep"] == Hr["ep"] and O["maxlevel"] == Hr["maxlevel"]
    same = (np.sort(O["links0"], axis=1) == np.sort(Hr["links0"], axis=1)).all(axis=1).mean()
    assert same > 0.99, same   # integer distances tie often; the reference's heap order among equal distances is its own


@pytest.fixture(scope="module")
def hnsw_case(tmp_path_factory):
    base = R.data_uniform(4000, 32, 1)
    q = R.data_uniform(100, 32, 2)
    h = R.RefHnsw(0, 32, 4000, 16, 200)
    h.add(base, threads=1)
    p = str(tmp_path_factory.mktemp("hnsw") / "h.bin")
    h.save(p)
    return base, q, h, R.parse_hnsw(p)


@needs_ref
@pytest.mark.parametrize("ef", [10, 40, 200])
def test_hnsw_search_vs_reference(hnsw_case, ef):
    """oracle searchKnn (both formulations) == HierarchicalNSW::searchKnn, bit for bit"""
    base, q, h, H = hnsw_case
    assert (H["vectors"] == base).all()
    rl, rd = h.search(q, 10, ef, threads=1)
    for variant in (0, 1):
        ol, od = R.orc_hnsw_search(base, H, q, 10, ef, R.L2_SCALAR, variant=variant)
        assert (rl == ol).all() and (rd == od).all()


@needs_ref
@pytest.mark.parametrize("keep,ef", [(0.5, 40), (0.1, 10), (0.02, 30), (0.9, 200), (0.0, 20)])
def test_hnsw_filtered_search_vs_reference(hnsw_case, keep, ef):
    """oracle searchBaseLayerST<false> == HierarchicalNSW::searchKnn(query, k, isIdAllowed), bit for bit, for a label
    filter, for deleted marks and for both together; and the kernel's bounded-array form equals the two-heap form
    wherever it does not report a query inexact"""
    base, q, h, H = hnsw_case
    n = len(base)
    rng = np.random.default_rng(int(keep * 100) + ef)
    allow_f = (rng.random(n) < keep).astype(np.uint8)
    dele = rng.random(n) < 0.2
    pad = lambda l, d: (l, np.where(l == np.uint64(0xFFFFFFFFFFFFFFFF), np.float32(np.inf), d))
    # (a) functor only
    rl, rd = pad(*h.search_filtered(q, 10, ef, allow_f, threads=1))
    ol, od, _ = R.orc_hnsw_search_filtered(base, H, allow_f, q, 10, ef, R.L2_SCALAR, variant=0)
    assert (rl == ol).all() and (rd == od).all()
    if keep == 0.0:
        assert (ol == np.uint64(0xFFFFFFFFFFFFFFFF)).all()
    # (b) deleted marks only, (c) both
    for lab in np.nonzero(dele)[0]:
        h.mark_deleted(lab, True)
    try:
        rl, rd = pad(*h.search_filtered(q, 10, ef, None, threads=1))
        ol, od, _ = R.orc_hnsw_search_filtered(base, H, (~dele).astype(np.uint8), q, 10, ef, R.L2_SCALAR, variant=0)
        assert (rl == ol).all() and (rd == od).all()
        both = (allow_f.astype(bool) & ~dele).astype(np.uint8)
        rl, rd = pad(*h.search_filtered(q, 10, ef, allow_f, threads=1))
        ol, od, _ = R.orc_hnsw_search_filtered(base, H, both, q, 10, ef, R.L2_SCALAR, variant=0)
        assert (rl == ol).all() and (rd == od).all()
    finally:
        for lab in np.nonzero(dele)[0]:
            h.mark_deleted(lab, False)
    # bounded form: exact where not flagged; a large bound flags nothing; a tiny one flags (and says so)
    for ccap in (n, 256, 16):
        pl, pd_, bad = R.orc_hnsw_search_filtered(base, H, both, q, 10, ef, R.L2_SCALAR, variant=1, ccap=ccap)
        ok = bad == 0
        assert (pl[ok] == ol[ok]).all() and (pd_[ok] == od[ok]).all()
        if ccap == n:
            assert ok.all()
    if keep <= 0.1:
        assert bad.any()


@needs_ref
@pytest.mark.parametrize("kth,mn,mx,keep", [(5, 0, 50, 1.0), (20, 10, 100, 1.0), (3, 30, 30, 1.0), (40, 0, 20, 1.0),
                                            (10, 5, 60, 0.3), (10, 0, 60, 0.05), (0, 0, 10, 1.0)])
def test_hnsw_epsilon_search_vs_reference(hnsw_case, kth, mn, mx, keep):
    """oracle epsilon search == HierarchicalNSW::searchStopConditionClosest(q, EpsilonSearchStopCondition(eps, min, max),
    filter), bit for bit (internal ids, distances, result counts); eps = the median distance to the kth neighbour
    (kth = 0: an epsilon smaller than every distance - nothing is returned)"""
    base, q, h, H = hnsw_case
    n = len(base)
    _, d50 = h.search(q, 50, 200, threads=1)
    eps = float(np.median(d50[:, kth])) if kth else float(d50.min()) * 0.5
    allow = None
    if keep < 1.0:
        allow = (np.random.default_rng(kth + mx).random(n) < keep).astype(np.uint8)
    ri, rd, rc = h.search_epsilon(q, eps, mn, mx, allow, threads=1)   # labels == internal ids here (added in order)
    oi, od, oc = R.orc_hnsw_search_epsilon(base, H, q, eps, mn, mx, R.L2_SCALAR, allow=allow)
    rd = np.where(ri == np.uint64(0xFFFFFFFFFFFFFFFF), np.float32(np.inf), rd)
    assert (rc == oc).all() and (ri == oi).all() and (rd == od).all()
    valid = np.arange(mx)[None, :] < oc[:, None]
    assert (od[valid] <= eps).all()
    if kth == 0:
        assert (oc == 0).all()
    elif keep == 1.0 and mx >= 50:
        assert oc.min() < oc.max()   # a real range search: the count varies with the query


@pytest.fixture(scope="module")
def nsg_case(tmp_path_factory):
    n, dim = 3000, 24
    base = R.data_uniform(n, dim, 1)
    q = R.data_uniform(64, dim, 2)
    ids, _ = R.orc_bf_knn(base, base, 41, R.L2_EFANNA_AVX)
    knn = np.stack([row[row != i][:40] for i, row in enumerate(ids)]).astype(np.uint32)
    d = tmp_path_factory.mktemp("nsg")
    kp, gp = str(d / "k.graph"), str(d / "g.nsg")
    R.write_knn_graph(kp, knn)
    R.ref_nsg_build(base, kp, 40, 20, 200, gp, seed=7, threads=1)
    g, ep, width = R.parse_nsg(gp, n)
    return base, q, knn, g, ep, width, gp


@needs_ref
@pytest.mark.parametrize("L", [10, 30, 64])
def test_nsg_search_vs_reference(nsg_case, L):
    """oracle beam search == IndexNSG::Search incl. the rand() seed fill (srand-synchronised)"""
    base, q, knn, g, ep, width, gp = nsg_case
    ix = R.RefNsg(base, gp)
    rids = ix.search(q, 10, L, seed=5, threads=1)
    ctypes.CDLL(None).srand(5)
    oids, _ = R.orc_nsg_search(base, g, ep, q, 10, L, R.L2_EFANNA_AVX, rng_mode=0)
    assert (rids == oids).all()


@needs_ref
def test_nsg_search_optgraph_vs_reference(nsg_case):
    """SearchWithOptGraph ranks by ||x||^2 - 2<q,x>: same ids as the L2 search up to fp near-ties"""
    base, q, knn, g, ep, width, gp = nsg_case
    ix = R.RefNsg(base, gp)
    ix.optimize()
    rids = ix.search_opt(q, 10, 30, seed=5, threads=1)
    ctypes.CDLL(None).srand(5)
    oids, _ = R.orc_nsg_search(base, g, ep, q, 10, 30, R.L2_EFANNA_AVX, rng_mode=0)
    assert (rids == oids).mean() > 0.99


def test_hnsw_pool_equals_twoheap_on_ties():
    """the single-pool form may differ from the two heaps only through distance ties at the eviction
    boundary; on heavily tied integer data the returned distance lists must still agree"""
    if not R.have_ref():
        pytest.skip("needs oracle/_ref to build an HNSW graph")
    rng = np.random.default_rng(9)
    base = rng.integers(0, 4, (2000, 8)).astype(np.float32)
    q = rng.integers(0, 4, (50, 8)).astype(np.float32)
    h = R.RefHnsw(0, 8, 2000, 8, 50)
    h.add(base, threads=1)
    import tempfile
    p = os.path.join(tempfile.mkdtemp(), "h.bin")
    h.save(p)
    H = R.parse_hnsw(p)
    _, d0 = R.orc_hnsw_search(base, H, q, 5, 20, R.L2_SCALAR, variant=0)
    _, d1 = R.orc_hnsw_search(base, H, q, 5, 20, R.L2_SCALAR, variant=1)
    assert np.mean(d0 == d1) > 0.95


# ---------------------------------------------------------------- NSG build (Link / InterInsert / tree_grow)
def _exact_knn(base, K):
    ids0, _ = R.orc_bf_knn(base, base, K + 1, R.L2_SCALAR)
    return np.stack([row[row != i][:K] for i, row in enumerate(ids0)]).astype(np.uint32)


@needs_ref
@pytest.mark.parametrize("n,dim,K,L,Rr,Cc,seed", [(3000, 32, 24, 20, 12, 100, 7), (2000, 24, 10, 16, 8, 50, 3),
                                                  (1500, 128, 30, 30, 20, 200, 11)])
def test_nsg_build_vs_reference(tmp_path, n, dim, K, L, Rr, Cc, seed):
    """IndexNSG::Build by the unmodified reference (1 thread, srand(seed)) == the oracle's phases chained
    with libc rand() drawn at the same places: identical navigating node, identical adjacency rows.
    The second case has L > K so the rand() seed fill of get_neighbors (:234-240) is exercised (dims are
    multiples of 8: the reference AVX distance reads past the row end otherwise, distance.h:41-51)."""
    base = R.data_uniform(n, dim, 1)
    knn = _exact_knn(base, K)
    kp, gp = str(tmp_path / "k.graph"), str(tmp_path / "g.nsg")
    R.write_knn_graph(kp, knn)
    R.ref_nsg_build(base, kp, L, Rr, Cc, gp, seed=seed, threads=1)
    g_ref, ep_ref, width_ref = R.parse_nsg(gp, n)
    lc = R.libc()
    lc.srand(seed)
    ep0 = lc.rand() % n
    kg = R.pad_graph(knn)
    mode = R.L2_EFANNA_AVX
    ep, _ = R.orc_nsg_init_ep(base, kg, ep0, L, mode, rng_mode=0)
    assert ep == ep_ref
    ci, cd = R.orc_nsg_link(base, kg, ep, L, Rr, Cc, mode, rng_mode=0)
    ci, cd, _ = R.orc_nsg_interinsert(base, ci, cd, mode, 0)
    g = R.orc_nsg_cut_to_graph(ci, cd, g_ref.shape[1] + 32)
    g, added, width = R.orc_nsg_tree_grow(base, g, ep, L, mode, Rr, rng_mode=0)
    assert width == width_ref
    gs = g_ref.shape[1]
    assert (g[:, :gs] == g_ref).all() and (g[:, gs:] == R.INVALID).all()


def test_nsg_interinsert_batch_form_properties():
    """the batch InterInsert (what the CUDA build runs): every row is duplicate free, -1 terminated,
    carries the true distances, and gains reverse edges up to the cap R"""
    n, dim, K, L, Rr, Cc = 2500, 24, 20, 20, 10, 80
    base = R.data_uniform(n, dim, 5)
    kg = R.pad_graph(_exact_knn(base, K))
    mode = R.L2_SCALAR
    ci, cd = R.orc_nsg_link(base, kg, 0, L, Rr, Cc, mode)
    bi, bd, _ = R.orc_nsg_interinsert(base, ci, cd, mode, 1)
    for i in range(0, n, 37):
        ids = bi[i][bd[i] != -1]
        assert len(set(ids.tolist())) == len(ids) and i not in ids
        assert (bi[i][bd[i] == -1] == R.INVALID).all()
        d = bd[i][bd[i] != -1]
        want = ((base[ids] - base[i]) ** 2).sum(1)
        assert np.allclose(d, want, rtol=1e-5)
    assert (bd != -1).sum(1).mean() > (cd != -1).sum(1).mean()
    assert ((bd != -1).sum(1) <= Rr).all()


# ---------------------------------------------------------------- hybrid HNSW_NSG (hnsw_nsg/)
@needs_ref
@pytest.mark.parametrize("dim", [8, 24, 128])
def test_hybrid_distance_vs_reference(dim):
    """DistanceL2AVX::compare (hnsw_nsg/include/nsg/distance.h:325-346)"""
    rng = np.random.default_rng(dim)
    a = rng.normal(size=(16, dim)).astype(np.float32)
    b = rng.normal(size=(16, dim)).astype(np.float32)
    for i in range(16):
        assert R.oracle().orc_dist(R.L2_HYBRID_AVX, a[i], b[i], dim) == R.reflib("hybrid").ref_hybrid_dist(a[i], b[i], dim)


@pytest.fixture(scope="module")
def hybrid_case(tmp_path_factory):
    """index built by the UNMODIFIED reference hybrid (addPoint loop + Build_NSG + saveIndex, 1 thread)"""
    n, dim = 3000, 24
    base = R.data_uniform(n, dim, 1)
    q = R.data_uniform(80, dim, 2)
    knn = _exact_knn(base, 30)
    d = tmp_path_factory.mktemp("hyb")
    kp, prefix = str(d / "k.graph"), str(d / "hy")
    R.write_knn_graph(kp, knn)
    # M=6 keeps the upper layers populated at n=3000 (mult = 1/ln 6)
    R.ref_hybrid_build(base, kp, 30, 16, 150, 6, 60, prefix, seed=7, threads=1)
    H = R.parse_hnsw(prefix + "_hnsw.bin")
    g, ep, width = R.parse_nsg(prefix + "_nsg.bin", n)
    return base, q, prefix, H, g


@needs_ref
@pytest.mark.parametrize("L,ef,k", [(12, 10, 10), (40, 20, 10), (64, 30, 5)])
def test_hybrid_search_vs_reference(hybrid_case, L, ef, k):
    """oracle hybrid search == HNSW_NSG::searchKnn, ids bit for bit. L is chosen both below and above
    the NSG degree so the neighbour-of-seed pass with its zero-slot quirk and the rand() fill
    (hnsw_nsg/src/index_nsg.cpp:720-739) are exercised."""
    base, q, prefix, H, g = hybrid_case
    assert H["maxlevel"] >= 2
    hy = R.RefHybrid(base, prefix, L)
    rids = hy.search(q, k, ef, seed=5)
    K = max(k, ef)
    ctypes.CDLL(None).srand(5)
    oids, _, entry = R.orc_hybrid_search(base, H, g, q, K, L, R.L2_HNSW_AVX, R.L2_HYBRID_AVX, rng_mode=0)
    assert (rids == oids).all()
    assert len(np.unique(entry)) > 5      # the descent really picks per-query entry points
    # entry points are flagged but never evaluated: they cannot be returned (:711-712)
    assert not (oids == entry[:, None]).any()


@needs_ref
def test_search_from_enterpoint_vs_reference(hybrid_case):
    base, q, prefix, H, g = hybrid_case
    hy = R.RefHybrid(base, prefix, 30)
    entry = np.random.default_rng(4).integers(0, base.shape[0], q.shape[0]).astype(np.uint32)
    rids = hy.search_from(q, 10, 30, entry, seed=9)
    ctypes.CDLL(None).srand(9)
    oids, _ = R.orc_nsg_search_from(base, g, entry, q, 10, 30, R.L2_HYBRID_AVX, rng_mode=0)
    assert (rids == oids).all()


# ---------------------------------------------------------------- NN-descent (IndexGraph::Build)
def test_mt19937_known_answer():
    """std::mt19937 default seed: the 10000th draw is 4123659995 (the C++ standard's own check value)."""
    out = np.zeros(10000, np.uint32)
    R.oracle().orc_mt19937(5489, out, 10000)
    assert int(out[0]) == 3499211612 and int(out[-1]) == 4123659995


def test_gen_random_properties():
    """GenRandom (util.h:17-31): `size` distinct ids below N, for both generators."""
    for rng in (0, 1):
        for size, N in ((16, 100), (21, 22), (64, 5000)):
            a = np.zeros(size, np.uint32)
            R.oracle().orc_gen_random(rng, 7, 3, a, size, N)
            assert len(set(a.tolist())) == size and a.max() < N


@needs_ref
@pytest.mark.parametrize("n,dim,K,L,it,S,Rr,seed", [(2000, 32, 20, 30, 4, 5, 10, 3), (3000, 24, 10, 40, 6, 10, 20, 5),
                                                   (1500, 16, 30, 30, 3, 8, 4, 7), (1200, 40, 5, 16, 0, 6, 6, 2)])
def test_nndescent_ref_form_vs_reference(tmp_path, n, dim, K, L, it, S, Rr, seed):
    """orc_nndescent_ref == the UNMODIFIED IndexGraph::Build + IndexRandom run with one OpenMP thread under
    srand(seed), graph for graph (float data: no distance ties, so std::sort's tie order does not matter).
    Pins mt19937, GenRandom, libstdc++'s heap layout (pool.back()), the join/update order and rand() use."""
    base = R.data_uniform(n, dim, seed)
    ref = R.ref_nndescent(base, K, L, it, S, Rr, str(tmp_path / "g"), seed=seed, threads=1)
    got, info = R.orc_nndescent_ref(base, K, L, it, S, Rr, R.L2_EFANNA_AVX, seed=seed)
    assert (got == ref).all()
    assert info[0] > 0


def _knn_recall(knn, truth):
    K = knn.shape[1]
    return float(np.mean([len(set(knn[i].tolist()) & set(truth[i].tolist())) / K for i in range(len(knn))]))


@needs_ref
@pytest.mark.parametrize("n,dim,K,L,it,S,Rr,kind", [(5000, 32, 20, 40, 6, 8, 20, "u"), (4000, 64, 20, 40, 5, 8, 20, "s")])
def test_nndescent_batch_form_recall_vs_reference(tmp_path, n, dim, K, L, it, S, Rr, kind):
    """the batch form (what the GPU computes) reaches the reference's graph quality at matched K/L/iter/S/R:
    recall of the K-NN lists against the exact graph within 0.005 (north_star bar), averaged over 3 seeds."""
    base = R.data_uniform(n, dim, 1) if kind == "u" else R.data_sift_like(n, dim, 1)
    oid, _ = R.orc_bf_knn(base, base, K + 1, R.L2_SCALAR)
    truth = np.stack([row[row != i][:K] for i, row in enumerate(oid)])
    rr = np.mean([_knn_recall(R.ref_nndescent(base, K, L, it, S, Rr, str(tmp_path / "g"), seed=s, threads=1), truth)
                  for s in (1, 2, 3)])
    rb = np.mean([_knn_recall(R.orc_nndescent_batch(base, K, L, it, S, Rr, R.L2_SEQ_FMA, seed=s)[0], truth)
                  for s in (1, 2, 3)])
    assert rb >= rr - 0.005, (rb, rr)
    assert abs(rb - rr) < 0.02


def test_nndescent_batch_form_properties():
    """rows sorted by (distance, id), no self, no duplicates, distances = the oracle's own distance of the pair;
    iter=0 returns the initial random pools; more iterations never make a row worse."""
    n, dim, K, L, S, Rr = 1500, 16, 12, 24, 6, 10
    base = R.data_uniform(n, dim, 4)
    prev = None
    for it in (0, 1, 3):
        knn, d, info = R.orc_nndescent_batch(base, K, L, it, S, Rr, R.L2_SEQ_FMA, seed=5)
        for i in range(0, n, 97):
            row = knn[i][knn[i] != R.INVALID]
            assert i not in row and len(set(row.tolist())) == len(row)
            dd = d[i][:len(row)]
            assert (np.diff(dd) >= 0).all()
            for j, v in enumerate(row[:3]):
                assert dd[j] == R.oracle().orc_dist(R.L2_SEQ_FMA, base[i], base[v], dim)
        if prev is not None:
            both = (knn[:, S - 2] != R.INVALID) & (prev[0][:, S - 2] != R.INVALID)
            assert (d[both, S - 2] <= prev[1][both, S - 2]).all()
        prev = (knn, d)


# ---------------------------------------------------------------- HNSW construction (addPoint loop)
@needs_ref
@pytest.mark.parametrize("n,dim,M,efc,metric", [(3000, 16, 8, 40, 0), (4000, 32, 16, 100, 0), (2000, 24, 6, 30, 1),
                                              (1200, 8, 4, 8, 0)])
def test_hnsw_build_ref_form_vs_reference(tmp_path, n, dim, M, efc, metric):
    """orc_hnsw_build variant 0 == the UNMODIFIED hnswlib addPoint loop run with one thread, link for link:
    levels (std::default_random_engine + uniform_real_distribution restated), entry point, every level-0 and
    upper-level list in the reference's own order (float data: no distance ties)."""
    base = R.data_uniform(n, dim, 3) if metric == 0 else R.data_glove_like(n, dim, 3)
    ref = R.RefHnsw(metric, dim, n, M, efc)
    ref.add(base, label0=0, threads=1)
    path = str(tmp_path / "h.bin")
    ref.save(path)
    Hr = R.parse_hnsw(path)
    Ho = R.orc_hnsw_build(base, M, efc, R.L2_SCALAR if metric == 0 else R.IP_SCALAR, 0)
    assert (Hr["levels"] == Ho["levels"]).all()
    assert Hr["ep"] == Ho["ep"] and Hr["maxlevel"] == Ho["maxlevel"]
    assert (Hr["links0"] == Ho["links0"]).all()
    assert Hr["upper_links"].shape == Ho["upper_links"].shape and (Hr["upper_links"] == Ho["upper_links"]).all()


def test_hnsw_build_batch_form_quality_and_schedule():
    """the batch form (what the GPU computes): same levels / entry point as the one-at-a-time form, every list within
    its capacity and free of self links and duplicates, links only between nodes that exist on that level, and search
    recall at matched ef within 0.005 of the one-at-a-time form (default ratio)."""
    n, dim, M, efc = 6000, 32, 12, 100
    base = R.data_uniform(n, dim, 5)
    q = R.data_uniform(2000, dim, 6)
    H0 = R.orc_hnsw_build(base, M, efc, R.L2_SCALAR, 0)
    H1 = R.orc_hnsw_build(base, M, efc, R.L2_SCALAR, 1, ratio=32, bmax=16384)   # the product's default schedule
    assert (H0["levels"] == H1["levels"]).all() and H0["ep"] == H1["ep"] and H0["maxlevel"] == H1["maxlevel"]
    assert int(H1["info"][1]) < n // 8   # batches
    for i in range(0, n, 37):
        row = H1["links0"][i][H1["links0"][i] != R.INVALID]
        assert i not in row and len(set(row.tolist())) == len(row)
    up = H1["upper_links"]
    for i in np.nonzero(H1["levels"] > 0)[0][:200]:
        for l in range(1, H1["levels"][i] + 1):
            row = up[H1["upper_ptr"][i] + l - 1]
            row = row[row != R.INVALID]
            assert (H1["levels"][row] >= l).all() and i not in row
    truth, _ = R.orc_bf_knn(base, q, 10, R.L2_SCALAR)
    for ef in (20, 50, 100):
        r0 = R.recall(R.orc_hnsw_search(base, H0, q, 10, ef, R.L2_SCALAR)[0].astype(np.uint32), truth)
        r1 = R.recall(R.orc_hnsw_search(base, H1, q, 10, ef, R.L2_SCALAR)[0].astype(np.uint32), truth)
        assert r1 >= r0 - 0.005, (ef, r1, r0)
    # schedule: first row alone, a level-raising row alone, sizes bounded by inserted / ratio and bmax
    lv = H1["levels"]
    s, ml = 0, -1
    while s < n:
        e = R.oracle().orc_hnsw_batch_end(s, n, lv, ml, 32, 64)
        assert s < e <= n and e - s <= max(1, min(64, s // (8 if s < min(2048, n // 16) else 32)))   # the first rows: 1/8
        assert all(lv[i] <= ml for i in range(s + 1, e))
        ml = max(ml, int(lv[s:e].max()))
        s = e

"""Filtered / deleted-element HNSW search on the GPU (b200ann_hnsw_search_filtered, csrc/hnsw_filtered.cu) against the
oracle: orc_hnsw_search_filtered variant 0 is searchBaseLayerST<false> statement for statement and is pinned bit for
bit to the compiled reference (label filter, markDelete, both) in tests/test_oracle_pinning.py; variant 1 is the
kernel's bounded-array form. Labels bit-exact, distances 1e-5 relative (they are bit-exact too), the inexact report
and the per-query work counters identical."""
import importlib

import numpy as np
import pytest

import refutil as R

pytestmark = pytest.mark.gpu
NOLAB = np.uint64(0xFFFFFFFFFFFFFFFF)


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("hnsw-nsg_b200")


@pytest.fixture(scope="module")
def H(B):
    return B.Handle(0)


@pytest.fixture(scope="module")
def case(B, H):
    n, dim, M = 20000, 32, 12
    base = R.data_uniform(n, dim, 31)
    ix = B.Index(H, base, B.L2)
    ix.hnsw_build(M, 80, seed=100)
    G = ix.get_hnsw()
    q = R.data_uniform(300, dim, 32)
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    return base, ix, G, q


@pytest.mark.parametrize("keep,ef,k", [(0.5, 64, 10), (0.9, 200, 10), (0.1, 32, 10), (0.02, 100, 20), (0.003, 50, 5),
                                       (1.0, 40, 10), (0.0, 16, 4)])
def test_filtered_search_vs_oracle(B, case, keep, ef, k):
    base, ix, G, q = case
    n = len(base)
    rng = np.random.default_rng(int(keep * 1000) + ef)
    allow = rng.random(n) < keep
    lab, d, bad, st = ix.hnsw_search_filtered(q, k, ef, allow, want_stats=True)
    # reference form: every query the kernel does not report is the reference's answer
    rl, rd, _ = R.orc_hnsw_search_filtered(base, G, allow, q, k, ef, R.L2_GPU, variant=0)
    ok = bad == 0
    assert (lab[ok] == rl[ok]).all()
    assert np.array_equal(d[ok], rd[ok])
    got = lab[lab != NOLAB].astype(np.int64)
    assert allow[got].all()                                       # labels == rows here: nothing filtered out comes back
    # kernel form incl. the overflow report: pass 0 with 2048 entries, reported queries again with room for all n
    l0, d0, b0, s0 = R.orc_hnsw_search_filtered(base, G, allow, q, k, ef, R.L2_GPU, variant=1, ccap=2048, want_stats=True)
    l1, d1, b1, s1 = R.orc_hnsw_search_filtered(base, G, allow, q, k, ef, R.L2_GPU, variant=1, ccap=n, want_stats=True)
    assert not b1.any()                                           # n entries cannot overflow
    again = b0 != 0
    el = np.where(again[:, None], l1, l0)
    ed = np.where(again[:, None], d1, d0)
    es = np.where(again[:, None], s1, s0)
    assert (lab == el).all() and np.array_equal(d, ed)
    assert (bad == np.where(again, b1, 0)).all()
    assert (st.astype(np.uint64) == es).all()
    assert ok.all()                                               # after the retry pass every query is exact
    if keep == 0.003:
        assert again.any()                                        # the retry pass really ran
    if keep == 0.0:
        assert (lab == NOLAB).all() and np.isinf(d).all()


def test_filtered_recall_and_deleted_semantics(B, case):
    # exact answer among the allowed rows (brute force over them) vs the filtered walk: recall like the unfiltered index
    base, ix, G, q = case
    n = len(base)
    allow = np.random.default_rng(5).random(n) < 0.3
    lab, d, bad = ix.hnsw_search_filtered(q, 10, 200, allow)
    assert not bad.any()
    rows = np.nonzero(allow)[0]
    sub = B.Index(ix.h, np.ascontiguousarray(base[rows]), B.L2)
    ids, _ = sub.bf_knn(q, 10)
    sub.close()
    truth = rows[ids]
    rec = np.mean([len(set(truth[i].tolist()) & set(lab[i].astype(np.int64).tolist())) for i in range(len(q))]) / 10
    assert rec > 0.97, rec


def test_filtered_ip_labels_and_device_buffers(B, H):
    import torch
    n, dim = 6000, 100
    base = R.data_glove_like(n, dim, 41)
    labels = np.arange(n, dtype=np.uint64) * np.uint64(3) + np.uint64(7)
    ix = B.Index(H, base, B.IP)
    ix.hnsw_build(8, 40, seed=100, labels=labels)
    G = ix.get_hnsw()
    q = R.data_glove_like(64, dim, 42)
    allow = np.random.default_rng(6).random(n) < 0.25
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    rl, rd, _ = R.orc_hnsw_search_filtered(base, G, allow, q, 8, 48, R.IP_GPU, variant=0)
    lab, d, bad = ix.hnsw_search_filtered(q, 8, 48, allow)
    assert not bad.any() and (lab == rl).all() and np.allclose(d, rd, rtol=1e-5, atol=0)
    # device-resident queries / bitmap / outputs through the raw C ABI
    lib = B.lib()
    bits = np.packbits(allow, bitorder="little")
    words = np.zeros(((n + 31) // 32) * 4, np.uint8)
    words[:len(bits)] = bits
    tq = torch.from_numpy(q).cuda()
    ta = torch.from_numpy(words.view(np.int32).copy()).cuda()
    tl = torch.empty((64, 8), dtype=torch.int64, device="cuda")
    td = torch.empty((64, 8), dtype=torch.float32, device="cuda")
    flags = np.zeros(64, np.uint8)
    torch.cuda.synchronize()
    rc = lib.b200ann_hnsw_search_filtered(H._h, ix._ix, tq.data_ptr(), 64, dim, 8, 48, ta.data_ptr(), tl.data_ptr(),
                                          td.data_ptr(), None, flags.ctypes.data)
    assert rc == 0
    assert (tl.cpu().numpy().view(np.uint64) == rl).all()
    ix.close()


@pytest.mark.parametrize("kth,mn,mx,keep", [(5, 0, 50, 1.0), (20, 10, 100, 1.0), (3, 30, 30, 1.0), (40, 0, 20, 1.0),
                                            (10, 5, 60, 0.3), (10, 0, 64, 0.02), (0, 0, 10, 1.0), (30, 200, 400, 0.004), (10, 500, 25000, 1.0)])
def test_epsilon_search_vs_oracle(B, case, kth, mn, mx, keep):
    """b200ann_hnsw_search_epsilon == the oracle's searchStopConditionClosest / EpsilonSearchStopCondition restatement
    (pinned bit for bit to the compiled reference in test_oracle_pinning.py): ids, distances, per-query result counts
    and work counters. eps = the median distance to the kth neighbour (kth = 0: smaller than every distance)."""
    base, ix, G, q = case
    n = len(base)
    _, d50 = ix.hnsw_search(q, 50, 200)
    eps = float(np.median(d50[:, kth])) if kth else float(d50.min()) * 0.5
    allow = None if keep == 1.0 else (np.random.default_rng(kth + mx).random(n) < keep)
    ol, od, oc, os_ = R.orc_hnsw_search_epsilon(base, G, q, eps, mn, mx, R.L2_GPU, allow=allow, want_stats=True)
    lab, d, cnt, st = ix.hnsw_search_epsilon(q, eps, mn, mx, allow=allow, internal_ids=True, want_stats=True)
    assert (cnt == oc).all()
    assert (lab == ol).all() and np.array_equal(d, od)
    assert (st.astype(np.uint64) == os_).all()
    valid = np.arange(mx)[None, :] < cnt[:, None]
    assert (d[valid] <= eps).all() and (lab[~valid] == NOLAB).all() and np.isinf(d[~valid]).all()
    if allow is not None:
        assert allow[lab[valid].astype(np.int64)].all()
    if kth == 0:
        assert (cnt == 0).all()
    elif keep == 1.0 and mx >= 50:
        assert cnt.min() < cnt.max()


def test_epsilon_search_labels_ip_and_errors(B, H):
    n, dim = 6000, 100
    base = R.data_glove_like(n, dim, 41)
    labels = np.arange(n, dtype=np.uint64) * np.uint64(3) + np.uint64(7)
    ix = B.Index(H, base, B.IP)
    ix.hnsw_build(8, 40, seed=100, labels=labels)
    G = ix.get_hnsw()
    q = R.data_glove_like(64, dim, 42)
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    _, d20 = ix.hnsw_search(q, 20, 100)
    eps = float(np.median(d20[:, 8]))
    ol, od, oc = R.orc_hnsw_search_epsilon(base, G, q, eps, 2, 40, R.IP_GPU, internal_ids=False)
    lab, d, cnt = ix.hnsw_search_epsilon(q, eps, 2, 40)                       # labels, not internal ids
    assert (cnt == oc).all() and (lab == ol).all() and np.allclose(d, od, rtol=1e-5, atol=0)
    ii, _, _ = ix.hnsw_search_epsilon(q, eps, 2, 40, internal_ids=True)
    valid = np.arange(40)[None, :] < cnt[:, None]
    assert (labels[ii[valid].astype(np.int64)] == lab[valid]).all()
    with pytest.raises(Exception, match="min_num_candidates"):
        ix.hnsw_search_epsilon(q, eps, 41, 40)
    with pytest.raises(Exception, match="NaN"):
        ix.hnsw_search_epsilon(q, float("nan"), 0, 40)
    ix.close()


@pytest.mark.parametrize("ndocs,efcoll,keep,nd", [(10, 15, 1.0, 500), (5, 5, 1.0, 500), (20, 40, 0.3, 500), (3, 1, 1.0, 50),
                                                  (8, 30, 1.0, 20000)])
def test_multivector_search_vs_oracle(B, case, ndocs, efcoll, keep, nd):
    """b200ann_hnsw_search_multivector == the oracle's MultiVectorSearchStopCondition restatement (pinned bit for bit to
    the compiled reference in test_oracle_pinning.py): ids, distances, counts, work counters; every row covers exactly
    num_docs distinct documents (the reference test's own invariant, multivector_search_test.cpp:86)"""
    base, ix, G, q = case
    n = len(base)
    rng = np.random.default_rng(ndocs * 100 + efcoll)
    doc = rng.integers(0, nd, n).astype(np.uint32)
    allow = None if keep == 1.0 else (rng.random(n) < keep)
    mx = 4096
    ol, od, oc, os_ = R.orc_hnsw_search_multivector(base, G, doc, q, ndocs, efcoll, mx, R.L2_GPU, allow=allow, want_stats=True)
    lab, d, cnt, st = ix.hnsw_search_multivector(q, doc, ndocs, efcoll, mx, allow=allow, internal_ids=True, want_stats=True)
    assert (cnt == oc).all() and (lab == ol).all() and np.array_equal(d, od)
    assert (st.astype(np.uint64) == os_).all()
    for i in range(len(q)):
        got = lab[i, :cnt[i]].astype(np.int64)
        assert len(set(doc[got].tolist())) == ndocs
        if allow is not None:
            assert allow[got].all()
    # too few result slots: refused with the reason, never a silently short answer
    if nd == 50:
        with pytest.raises(Exception, match="max_results"):
            ix.hnsw_search_multivector(q, doc, ndocs, efcoll, int(cnt.max()) - 1)

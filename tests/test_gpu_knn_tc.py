"""Parity of the tensor-core brute-force kNN (csrc/knn_tc.cu: tcgen05 contraction + fp32 re-rank) through the
C ABI: ids AND distances bit-exact against the oracle (bruteforce.h:106-135 restated in the kernels'
sequential-fma order), against the fp32 SIMT kernel, and - on integer-valued data where every summation
order is exact - against the scalar oracle / compiled reference. BASELINE.json north_star: "brute-force kNN
ids must be bit-exact (ties broken by id), distances within 1e-5 relative"."""
import importlib

import numpy as np
import pytest

import refutil as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("hnsw-nsg_b200")


@pytest.fixture()
def H(B):
    h = B.Handle(0)
    h.set_knn_mode(2)  # force the tensor path at every size
    yield h
    h.close()


@pytest.mark.parametrize("n,dim,nq,k,metric,kind", [
    (5000, 128, 200, 10, 0, "uniform"),    # inexact data -> bf16x3 passes, A resident
    (5000, 128, 200, 10, 0, "sift"),       # bf16-exact data -> single pass
    (5000, 100, 130, 10, 1, "glove"),      # inner product, dim padded to 128
    (3000, 24, 70, 100, 0, "uniform"),
    (777, 7, 33, 5, 0, "uniform"),         # ragged everything
    (20000, 96, 64, 200, 0, "uniform"),    # K=200, base-range splits
    (300, 960, 17, 10, 0, "uniform"),      # 15 K-blocks: A streams through the ring
    (50, 16, 9, 64, 0, "uniform"),         # k > n: unfilled slots INVALID / +inf
    (1, 8, 1, 1, 0, "uniform"),
])
def test_knn_tc_vs_oracle(B, H, n, dim, nq, k, metric, kind):
    gen = {"uniform": R.data_uniform, "sift": R.data_sift_like, "glove": R.data_glove_like}[kind]
    base, q = gen(n, dim, 1), gen(nq, dim, 2)
    ix = B.Index(H, base, metric)
    ids, d = ix.bf_knn(q, k)
    st = H.knn_last_stats
    assert st["rows"] == nq and st["passes"] == (1 if kind == "sift" else 3)
    oid, od = R.orc_bf_knn(base, q, k, R.L2_SEQ_FMA if metric == 0 else R.IP_SEQ_FMA)
    assert (ids == oid).all()
    assert (d == od).all()
    if kind == "sift":  # integer data: identical to the scalar order of the reference too
        sid, sd = R.orc_bf_knn(base, q, k, R.L2_SCALAR)
        assert (ids == sid).all() and (d == sd).all()
    ix.close()


def test_knn_tc_graph_excludes_self_and_ties(B, H):
    # few distinct integer values: masses of exact distance ties, smaller id must win
    base = np.random.default_rng(7).integers(0, 3, (6000, 16)).astype(np.float32)
    ix = B.Index(H, base, 0)
    knn = ix.knn_graph(40)
    oid, _ = R.orc_bf_knn(base, base, 41, R.L2_SCALAR)
    # the row itself is at distance 0 but so may be duplicates with smaller ids: drop exactly the row id
    want = np.stack([row[row != i][:40] for i, row in enumerate(oid)])
    assert (knn == want).all()
    ix.close()


def test_knn_tc_fallback_rows_are_exact(B):
    """slack 0: the shortlist holds exactly k entries, ties at the k-th distance cannot be proven complete,
    so rows go through the SIMT fallback - the result must not change."""
    h = B.Handle(0)
    h.set_knn_mode(2, 0)
    # n < 2048 keeps the base range in one piece, so tau IS the k-th key and "tau - eps > k-th distance" fails
    base = np.random.default_rng(3).integers(0, 4, (1500, 32)).astype(np.float32)
    q = np.random.default_rng(4).integers(0, 4, (300, 32)).astype(np.float32)
    ix = B.Index(h, base, 0)
    ids, d = ix.bf_knn(q, 25)
    st = h.knn_last_stats
    assert st["fallback_rows"] > 0
    oid, od = R.orc_bf_knn(base, q, 25, R.L2_SCALAR)
    assert (ids == oid).all() and (d == od).all()
    ix.close()
    h.close()


def test_knn_tc_equals_simt_at_scale(B):
    """200k x 128 SIFT-shaped rows: kNN graph rows from the tensor path == the fp32 SIMT kernel's on a
    row sample (the oracle would need minutes); plus the size-independent properties: no self, ascending."""
    h = B.Handle(0)
    base = R.data_sift_shape(200000, 128, 1)
    ix = B.Index(h, base, 0)
    h.set_knn_mode(2)
    knn = ix.knn_graph(64)
    st = h.knn_last_stats
    assert st["rows"] == 200000 and st["passes"] == 1
    assert (knn != np.arange(200000, dtype=np.uint32)[:, None]).all()
    rows = np.random.default_rng(0).choice(200000, 512, replace=False)
    h.set_knn_mode(1)
    sid, sd = ix.bf_knn(base[rows], 65)
    assert h.knn_last_stats["rows"] == 0  # SIMT kernel ran
    want = np.stack([r[r != i][:64] for i, r in zip(rows, sid)])
    assert (knn[rows] == want).all()
    ix.close()
    h.close()


@pytest.mark.parametrize("mode,n,dim,K", [(2, 9000, 64, 24), (1, 3000, 40, 10), (0, 20000, 128, 50)])
def test_knn_graph_rows_union_equals_whole(B, H, mode, n, dim, K):
    """row-block partition of the kNN-graph build (multi-GPU build): the blocks' union is the whole graph, bit for bit"""
    base = R.data_sift_like(n, dim, 7) if mode != 1 else R.data_uniform(n, dim, 7)
    ix = B.Index(H, base, 0)
    H.set_knn_mode(mode)
    try:
        whole = ix.knn_graph(K)
        cuts = [0, n // 3 + 5, n // 3 + 6, n - 1000, n]
        parts = [ix.knn_graph_rows(a, b - a, K) for a, b in zip(cuts[:-1], cuts[1:])]
        assert (np.concatenate(parts) == whole).all()
    finally:
        H.set_knn_mode(0)


@pytest.mark.parametrize("step,kind", [("4", "sift"), ("8", "sorted")])
def test_knn_tc_sampled_start_threshold(B, step, kind):
    """the sweep's start thresholds from a sampled pre-pass (every step-th tile): results stay exact - the same rows as
    without sampling and as the SIMT kernel - and only a handful of rows fail their proof. "sorted": rows ordered
    by their first coordinate, the adversarial layout for a sample by position (neighbours sit in the same tiles)."""
    import os
    h = B.Handle(0)
    n = 60000
    base = R.data_sift_shape(n, 128, 5)
    if kind == "sorted":
        base = np.ascontiguousarray(base[np.argsort(base[:, 0], kind="stable")])
    ix = B.Index(h, base, 0)
    h.set_knn_mode(2)
    old = os.environ.get("B200ANN_KNN_SAMPLE")
    try:
        os.environ["B200ANN_KNN_SAMPLE"] = "0"
        plain = ix.knn_graph(40)
        f0 = h.knn_last_stats["fallback_rows"]
        os.environ["B200ANN_KNN_SAMPLE"] = step
        sampled = ix.knn_graph(40)
        st = h.knn_last_stats
    finally:
        if old is None:
            os.environ.pop("B200ANN_KNN_SAMPLE", None)
        else:
            os.environ["B200ANN_KNN_SAMPLE"] = old
    assert (sampled == plain).all()
    assert st["rows"] == n and st["fallback_rows"] <= f0 + n // 100, (st, f0)
    ix.close()
    h.close()

"""HierarchicalNSW<int> over L2SpaceI on the GPU (b200ann_index_create_u8 + the HNSW entry points): unsigned-byte rows,
int squared-L2 distances (hnswlib/hnswlib/space_l2.h:255-323). Integer sums do not depend on the summation order, so
every comparison is bit-exact: labels AND int distances against the oracle in mode L2_U8, which
tests/test_oracle_pinning.py pins to the compiled reference's HierarchicalNSW<int> (search, filter, epsilon)."""
import importlib

import numpy as np
import pytest

import refutil as R

pytestmark = pytest.mark.gpu
NOLAB = np.uint64(0xFFFFFFFFFFFFFFFF)
IMAX = np.int32(0x7FFFFFFF)


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("hnsw-nsg_b200")


@pytest.fixture(scope="module")
def H(B):
    return B.Handle(0)


def sift_like(n, dim, seed):
    """bytes with the skew of SIFT descriptors (many small values, a few large)"""
    rng = np.random.default_rng(seed)
    return np.minimum(255, rng.gamma(1.2, 24.0, (n, dim))).astype(np.uint8)


def odist(d):
    """oracle distances (int bit patterns in float32, +inf padding) -> int32 with the C ABI's INT32_MAX padding"""
    i = d.view(np.int32).copy()
    i[np.isinf(d)] = IMAX
    return i


@pytest.mark.parametrize("n,dim,M,efc,ratio,bmax", [
    (3000, 128, 8, 40, 16, 256),      # SIFT shape: 8 chunks of 16 bytes, one per lane of a team
    (5000, 50, 12, 100, 32, 1024),    # dim not a multiple of 4 or 16: padded rows, byte-wise query load
    (2000, 960, 6, 30, 8, 512),       # GIST shape: 60 chunks
    (1500, 16, 4, 10, 4, 2048),       # one chunk per row, tiny M
])
def test_u8_build_vs_oracle_batch_form(B, H, n, dim, M, efc, ratio, bmax):
    base = sift_like(n, dim, 21)
    ix = B.Index(H, base, B.L2_U8)
    info = ix.hnsw_build(M, efc, seed=100, ratio=ratio, bmax=bmax)
    G = ix.get_hnsw()
    O = R.orc_hnsw_build(R.u8_rows(base), M, efc, R.L2_U8, 1, ratio=ratio, bmax=bmax, dim=dim)
    assert (G["levels"] == O["levels"]).all() and G["ep"] == O["ep"] and G["maxlevel"] == O["maxlevel"]
    assert (G["links0"] == O["links0"]).all()
    assert G["upper_links"].shape == O["upper_links"].shape and (G["upper_links"] == O["upper_links"]).all()
    assert info["batches"] == int(O["info"][1])
    # search over that graph: labels and int distances
    q = sift_like(200, dim, 22)
    for k, ef in ((10, 10), (10, 64), (5, 200)):
        lab, d = ix.hnsw_search(q, k, ef)
        ol, od = R.orc_hnsw_search(R.u8_rows(base), G, R.u8_rows(q), k, ef, R.L2_U8, variant=1, dim=dim)
        assert d.dtype == np.int32
        assert (lab == ol).all() and (d == odist(od)).all()
    ix.close()


@pytest.fixture(scope="module")
def case(B, H):
    n, dim, M = 20000, 128, 12
    base = sift_like(n, dim, 31)
    ix = B.Index(H, base, B.L2_U8)
    ix.hnsw_build(M, 80, seed=100)
    return base, ix, ix.get_hnsw(), sift_like(300, dim, 32)


def test_u8_recall_and_float_twin(B, H, case):
    """the walk over byte rows finds what the exact scan finds (b200ann_bf_knn_u8), and the same graph over the same
    values stored as floats returns the same labels and (exactly representable) distances"""
    base, ix, G, q = case
    lab, d = ix.hnsw_search(q, 10, 100)
    ti, td = H.bf_knn_u8(base, q, 10)
    rec = np.mean([len(set(ti[i].tolist()) & set(lab[i].astype(np.int64).tolist())) for i in range(len(q))]) / 10
    assert rec > 0.75, rec   # iid bytes in 128 dimensions are a hard case (0.81 at M=12, ef=100); the float twin below is the check
    fx = B.Index(H, base.astype(np.float32), B.L2)
    fx.set_hnsw(G["links0"], G["levels"], G["upper_ptr"], G["upper_links"], G["labels"], G["ep"], G["maxlevel"])
    fl, fd = fx.hnsw_search(q.astype(np.float32), 10, 100)
    assert (fl == lab).all() and (fd == d.astype(np.float32)).all()
    fx.close()


@pytest.mark.parametrize("keep,ef,k", [(0.5, 64, 10), (0.05, 40, 10), (0.004, 50, 5), (0.0, 16, 4)])
def test_u8_filtered_vs_oracle(B, case, keep, ef, k):
    base, ix, G, q = case
    allow = np.random.default_rng(int(keep * 1000) + ef).random(len(base)) < keep
    lab, d, bad = ix.hnsw_search_filtered(q, k, ef, allow)
    ol, od, _ = R.orc_hnsw_search_filtered(R.u8_rows(base), G, allow, R.u8_rows(q), k, ef, R.L2_U8, variant=0, dim=base.shape[1])
    assert not bad.any()
    assert (lab == ol).all() and (d == odist(od)).all()
    if keep == 0.0:
        assert (lab == NOLAB).all() and (d == IMAX).all()


@pytest.mark.parametrize("kth,mn,mx", [(5, 0, 50), (20, 10, 100), (0, 0, 10), (10, 200, 300)])
def test_u8_epsilon_vs_oracle(B, case, kth, mn, mx):
    base, ix, G, q = case
    _, d50 = ix.hnsw_search(q, 50, 200)
    eps = float(np.median(d50[:, kth])) + 0.5 if kth else float(d50.min()) * 0.5   # non-integer epsilon: compared as floor
    lab, d, cnt = ix.hnsw_search_epsilon(q, eps, mn, mx, internal_ids=True)
    ol, od, oc = R.orc_hnsw_search_epsilon(R.u8_rows(base), G, R.u8_rows(q), np.floor(eps).astype(np.int32).view(np.float32),
                                           mn, mx, R.L2_U8, dim=base.shape[1])
    assert (cnt == oc).all() and (lab == ol).all() and (d == odist(od)).all()
    valid = np.arange(mx)[None, :] < cnt[:, None]
    assert (d[valid] <= eps).all()
    if kth == 0:
        assert (cnt == 0).all()


def test_u8_insert_batch_device_buffers_and_guards(B, H):
    import torch
    n0, nb, dim = 6000, 500, 48
    base = sift_like(n0 + nb, dim, 41)
    ix = B.Index(H, base[:n0], B.L2_U8)
    ix.hnsw_build(12, 80, seed=100)
    ix.hnsw_insert_batch(base[n0:])
    assert ix.n == n0 + nb
    lab, d = ix.hnsw_search(base[n0:], 1, 100)          # the appended rows find themselves ...
    new = lab[:, 0] == np.arange(n0, n0 + nb)
    lab0, _ = ix.hnsw_search(base[:nb], 1, 100)         # ... as reliably as rows of the original build do
    old = lab0[:, 0] == np.arange(nb)
    assert new.mean() >= old.mean() - 0.05 and new.mean() >= 0.8, (new.mean(), old.mean())
    assert (d[new, 0] == 0).all()
    dim = 128
    base = sift_like(n0, dim, 43)
    ix.close()
    ix = B.Index(H, base, B.L2_U8)
    ix.hnsw_build(8, 40, seed=100)
    # device-resident byte queries (aligned stride -> used in place) through the raw C ABI
    lib = B.lib()
    q = sift_like(64, dim, 42)
    want_l, want_d = ix.hnsw_search(q, 8, 48)
    tq = torch.from_numpy(q).cuda()
    tl = torch.empty((64, 8), dtype=torch.int64, device="cuda")
    td = torch.empty((64, 8), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = lib.b200ann_hnsw_search(H._h, ix._ix, tq.data_ptr(), 64, dim, 8, 48, tl.data_ptr(), td.data_ptr(), None)
    assert rc == 0
    assert (tl.cpu().numpy().view(np.uint64) == want_l).all() and (td.cpu().numpy() == want_d).all()
    # float-only entry points refuse a byte index
    with pytest.raises(Exception, match="unsigned-byte rows"):
        ix.bf_knn(np.zeros((2, dim), np.float32), 3)
    with pytest.raises(Exception, match="unsigned-byte rows"):
        ix.knn_graph(5)
    with pytest.raises(Exception, match="unsigned-byte rows"):
        ix.nsg_search(np.zeros((2, dim), np.float32), 3, 8)
    ix.close()

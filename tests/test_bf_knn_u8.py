"""u8 / integer-L2 brute force (hnswlib L2SpaceI + BruteforceSearch<int>, space_l2.h:255-323, bruteforce.h:106-135).
CPU part: oracle/oracle.c's orc_bf_knn_u8 pinned to the UNMODIFIED reference compiled into oracle/_ref and to a
numpy int64 restatement. GPU part: b200ann_bf_knn_u8 through the C ABI, ids AND distances bit-exact (integer
arithmetic), plus size-independent properties at a larger size."""
import importlib

import numpy as np
import pytest

import refutil as R

needs_ref = pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not built (needs /root/reference)")


def u8_data(n, dim, seed, few_values=False):
    rng = np.random.default_rng(seed)
    if few_values:  # many exact ties
        return rng.integers(0, 3, (n, dim), dtype=np.uint8)
    return rng.integers(0, 256, (n, dim), dtype=np.uint8)


def numpy_knn(base, q, k):
    b = base.astype(np.int64)
    out_i = np.empty((q.shape[0], k), np.uint32)
    out_d = np.empty((q.shape[0], k), np.int32)
    for i, row in enumerate(q.astype(np.int64)):
        d = ((b - row) ** 2).sum(1)
        order = np.lexsort((np.arange(len(d)), d))[:k]
        out_i[i, :len(order)] = order
        out_d[i, :len(order)] = d[order]
        out_i[i, len(order):] = 0xFFFFFFFF
        out_d[i, len(order):] = np.iinfo(np.int32).max
    return out_i, out_d


CASES = [(2000, 128, 50, 10, False), (1500, 100, 33, 10, False), (700, 7, 20, 5, False), (900, 33, 40, 100, True),
         (64, 16, 9, 64, True), (300, 960, 7, 10, False), (1000, 4, 30, 10, True)]


@pytest.mark.parametrize("n,dim,nq,k,ties", CASES)
def test_oracle_u8_vs_numpy(n, dim, nq, k, ties):
    base, q = u8_data(n, dim, 1, ties), u8_data(nq, dim, 2, ties)
    oi, od = R.orc_bf_knn_u8(base, q, k)
    ni, nd = numpy_knn(base, q, k)
    assert np.array_equal(oi, ni) and np.array_equal(od, nd)


@needs_ref
@pytest.mark.parametrize("n,dim,nq,k,ties", CASES)
def test_oracle_u8_pinned_to_reference(n, dim, nq, k, ties):
    """both dispatch branches of L2SpaceI: dim % 4 == 0 -> L2SqrI4x, else L2SqrI (space_l2.h:300-306)"""
    base, q = u8_data(n, dim, 3, ties), u8_data(nq, dim, 4, ties)
    oi, od = R.orc_bf_knn_u8(base, q, k)
    lab, rd = R.ref_bf_knn_u8(base, q, k)
    assert np.array_equal(lab.astype(np.uint32), oi)
    assert np.array_equal(rd, od)


def test_oracle_u8_k_larger_than_n():
    base, q = u8_data(5, 12, 1), u8_data(3, 12, 2)
    oi, od = R.orc_bf_knn_u8(base, q, 8)
    assert (oi[:, 5:] == 0xFFFFFFFF).all() and (od[:, 5:] == np.iinfo(np.int32).max).all()
    ni, nd = numpy_knn(base, q, 8)
    assert np.array_equal(oi, ni) and np.array_equal(od, nd)


# ---------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def H():
    B = importlib.import_module("hnsw-nsg_b200")
    return B.Handle(0)


@pytest.mark.gpu
@pytest.mark.parametrize("n,dim,nq,k,ties", CASES + [(20000, 128, 300, 200, False), (5000, 21, 65, 10, True),
                                                  (30000, 16, 100, 50, True), (60000, 8, 1000, 10, True)])  # many splits, heavy ties
def test_gpu_u8_bit_exact_vs_oracle(H, n, dim, nq, k, ties):
    base, q = u8_data(n, dim, 5, ties), u8_data(nq, dim, 6, ties)
    gi, gd = H.bf_knn_u8(base, q, k)
    oi, od = R.orc_bf_knn_u8(base, q, k)
    assert np.array_equal(gi, oi)
    assert np.array_equal(gd, od)


@pytest.mark.gpu
@needs_ref
def test_gpu_u8_bit_exact_vs_reference(H):
    base, q = u8_data(10000, 128, 7), u8_data(200, 128, 8)
    gi, gd = H.bf_knn_u8(base, q, 10)
    lab, rd = R.ref_bf_knn_u8(base, q, 10)
    assert np.array_equal(gi, lab.astype(np.uint32)) and np.array_equal(gd, rd)


@pytest.mark.gpu
def test_gpu_u8_edge_cases(H):
    B = importlib.import_module("hnsw-nsg_b200")
    base, q = u8_data(5, 12, 1), u8_data(3, 12, 2)
    gi, gd = H.bf_knn_u8(base, q, 8)  # k > n: unfilled slots
    oi, od = R.orc_bf_knn_u8(base, q, 8)
    assert np.array_equal(gi, oi) and np.array_equal(gd, od)
    gi, gd = H.bf_knn_u8(base[:1], q, 1)  # n = 1
    assert (gi == 0).all()
    # extreme values: 0 vs 255 in every dimension at the largest admitted dimension
    big = np.zeros((3, 32768), np.uint8); big[1] = 255; big[2, ::2] = 255
    qq = np.zeros((1, 32768), np.uint8)
    gi, gd = H.bf_knn_u8(big, qq, 3)
    assert gi.tolist() == [[0, 2, 1]] and gd.tolist() == [[0, 16384 * 65025, 32768 * 65025]]
    with pytest.raises(B.B200AnnError):
        H.bf_knn_u8(np.zeros((2, 32772), np.uint8), np.zeros((1, 32772), np.uint8), 1)
    # duplicate rows: ties go to the smaller id
    dup = np.repeat(u8_data(1, 64, 9), 100, axis=0)
    gi, gd = H.bf_knn_u8(dup, dup[:4], 10)
    assert (gi == np.arange(10, dtype=np.uint32)).all() and (gd == 0).all()


@pytest.mark.gpu
def test_gpu_u8_unaligned_rows_and_device_buffers(H):
    """row strides that defeat the 16-byte loads, and device-resident inputs / outputs"""
    import torch
    n, dim, nq, k = 3000, 50, 40, 10
    base, q = u8_data(n, dim, 11), u8_data(nq, dim, 12)
    oi, od = R.orc_bf_knn_u8(base, q, k)
    wide = np.zeros((n, 57), np.uint8); wide[:, :dim] = base; wide[:, dim:] = 200  # padding must be ignored
    wq = np.zeros((nq, 64), np.uint8); wq[:, :dim] = q; wq[:, dim:] = 77
    dev = torch.device("cuda:0")
    tb, tq = torch.from_numpy(wide).to(dev), torch.from_numpy(wq).to(dev)
    ti = torch.empty((nq, k), dtype=torch.int32, device=dev)
    td = torch.empty((nq, k), dtype=torch.int32, device=dev)
    H.bf_knn_u8_raw(tb, n, dim, 57, tq, nq, 64, k, ti, td)
    torch.cuda.synchronize()
    assert np.array_equal(ti.cpu().numpy().view(np.uint32), oi) and np.array_equal(td.cpu().numpy(), od)


@pytest.mark.gpu
def test_gpu_u8_properties_at_scale(H):
    """1M x 128 (SIFT bytes), 1000 queries: rows sorted, self-match first with distance 0, distances recomputed
    on the host for every returned id, k-th distance no larger than that of any of 2000 random rows"""
    n, dim, nq, k = 1_000_000, 128, 1000, 10
    base = u8_data(n, dim, 21)
    qrows = np.random.default_rng(22).choice(n, nq, replace=False)
    q = base[qrows]
    gi, gd = H.bf_knn_u8(base, q, k)
    assert (np.diff(gd.astype(np.int64), axis=1) >= 0).all()
    assert (gd[:, 0] == 0).all() and (gi[:, 0] <= qrows).all()
    re = ((base[gi].astype(np.int64) - q[:, None, :].astype(np.int64)) ** 2).sum(-1)
    assert np.array_equal(re, gd.astype(np.int64))
    probe = np.random.default_rng(23).choice(n, 2000, replace=False)
    pd_ = ((base[probe][None].astype(np.int32) - q[:100, None, :].astype(np.int32)) ** 2).sum(-1)
    assert (pd_.min(1) >= gd[:100, 0]).all()
    worse = (pd_ < gd[:100, -1:]).sum(1)  # probes strictly inside the k-th distance must be among the answers
    for i in range(100):
        inside = probe[pd_[i] < gd[i, -1]]
        assert set(inside.tolist()) <= set(gi[i].tolist())
    assert worse.max() <= k

"""b200ann_bf_knn_noalloc (csrc/bf_knn.cu): the brute-force kNN kernel pair with its partial lists in a static device
array instead of allocated scratch - for callers that cannot call cudaMalloc. No counterpart in the reference;
the checker is the existing entry point b200ann_bf_knn on the same inputs:
ids bit-exact, distances bit-exact (same kernel, same arithmetic order), on two calls in a row (the second call finds
the first call's lists in the shared array). Kept in a file of its own that pytest collects last."""
import importlib

import numpy as np
import pytest

import refutil as R

pytestmark = pytest.mark.gpu
ERR_ARG = -1   # B200ANN_ERR_ARG, include/b200ann.h


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("hnsw-nsg_b200")


def _run(B, H, ix, tq, nq, dim, k):
    import torch
    lib = B.lib()
    ids = torch.full((nq, k), -1, dtype=torch.int32, device="cuda")
    d = torch.full((nq, k), -1.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    rc = lib.b200ann_bf_knn_noalloc(H._h, ix._ix, tq.data_ptr(), nq, dim, k, ids.data_ptr(), d.data_ptr())
    assert rc == 0, lib.b200ann_last_error(H._h).decode()
    torch.cuda.synchronize()
    return ids.cpu().numpy().view(np.uint32), d.cpu().numpy()


def test_noalloc_equals_bf_knn_two_calls_in_a_row(B):
    import torch
    H = B.Handle(0)
    H.set_knn_mode(1)                      # the existing entry point on the same fp32 SIMT kernel pair
    n, dim = 30000, 48
    base = R.data_uniform(n, dim, 71)
    ix = B.Index(H, base, B.L2)
    # call 1: 700 queries x k=20 (several splits of the base range); call 2: other queries, other shape (k=7), so the
    # array still holds call 1's lists where call 2 puts its own
    for nq, k, seed in ((700, 20, 72), (333, 7, 73), (700, 20, 72)):
        q = R.data_uniform(nq, dim, seed)
        want_i, want_d = ix.bf_knn(q, k)
        tq = torch.from_numpy(q).cuda()
        got_i, got_d = _run(B, H, ix, tq, nq, dim, k)
        assert (got_i == want_i).all()
        assert np.allclose(got_d, want_d, rtol=1e-5, atol=0) and np.array_equal(got_d, want_d)
    ix.close()
    # inner product index, k > n (unfilled slots: INVALID / +inf like b200ann_bf_knn)
    base = R.data_glove_like(50, 100, 74)
    ix = B.Index(H, base, B.IP)
    q = R.data_glove_like(9, 100, 75)
    want_i, want_d = ix.bf_knn(q, 64)
    got_i, got_d = _run(B, H, ix, torch.from_numpy(q).cuda(), 9, 100, 64)
    assert (got_i == want_i).all() and np.array_equal(got_d, want_d)
    ix.close()
    H.close()


def test_noalloc_rejects_what_it_cannot_hold(B):
    import torch
    H = B.Handle(0)
    base = R.data_uniform(2000, 16, 76)
    ix = B.Index(H, base, B.L2)
    lib = B.lib()
    limit = B.NOALLOC_SCRATCH_BYTES
    nq, k = 4096, 512                      # 4096 * 512 * 16 B = 32 MiB of partial lists > the 16 MiB array
    assert nq * k * 16 > limit
    tq = torch.zeros((nq, 16), dtype=torch.float32, device="cuda")
    ids = torch.empty((nq, k), dtype=torch.int32, device="cuda")
    rc = lib.b200ann_bf_knn_noalloc(H._h, ix._ix, tq.data_ptr(), nq, 16, k, ids.data_ptr(), None)
    assert rc == ERR_ARG and "static scratch" in lib.b200ann_last_error(H._h).decode()
    # host buffers would need staging allocations: refused, not staged
    qh = np.zeros((8, 16), np.float32)
    ih = np.zeros((8, 4), np.uint32)
    rc = lib.b200ann_bf_knn_noalloc(H._h, ix._ix, qh.ctypes.data, 8, 16, 4, ih.ctypes.data, None)
    assert rc == ERR_ARG and "device memory" in lib.b200ann_last_error(H._h).decode()
    ix.close()
    H.close()

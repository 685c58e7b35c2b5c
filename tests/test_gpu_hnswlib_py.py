"""The hnswlib Python-binding mirror (hnsw-nsg_b200/hnswlib_py.py) on the GPU, following the reference's own binding
tests: hnswlib/tests/python/bindings_test_recall.py (HNSW recall against BFIndex), bindings_test_spaces.py (the three
distances), bindings_test.py / bindings_test_labels.py (save -> load -> add more -> query), bindings_test_getdata.py,
bindings_test_metadata.py, bindings_test_pickle.py; plus what the reference cannot test: the files written here load
in the compiled reference and answer identically to the oracle."""
import importlib
import pickle

import numpy as np
import pytest

import refutil as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hnswlib():
    return importlib.import_module("hnsw-nsg_b200.hnswlib_py")


def test_recall_against_bruteforce(hnswlib):
    # bindings_test_recall.py:7-87 (dim 32, k 10, M 16, ef_construction 200, ef 200, recall > 0.95)
    dim, n, k, nq = 32, 100000, 10, 200
    rng = np.random.default_rng(0)
    data = rng.random((n, dim), dtype=np.float32)
    hnsw_index = hnswlib.Index(space="l2", dim=dim)
    bf_index = hnswlib.BFIndex(space="l2", dim=dim)
    hnsw_index.init_index(max_elements=n, ef_construction=200, M=16)
    bf_index.init_index(max_elements=n)
    hnsw_index.set_ef(200)
    hnsw_index.set_num_threads(4)
    hnsw_index.add_items(data)
    bf_index.add_items(data)
    q = rng.random((nq, dim), dtype=np.float32)
    labels_hnsw, distances_hnsw = hnsw_index.knn_query(q, k)
    labels_bf, distances_bf = bf_index.knn_query(q, k)
    assert labels_hnsw.dtype == np.uint64 and labels_hnsw.shape == (nq, k) and distances_hnsw.dtype == np.float32
    correct = sum(len(set(labels_hnsw[i].tolist()) & set(labels_bf[i].tolist())) for i in range(nq))
    assert correct / (nq * k) > 0.95
    assert (np.diff(distances_hnsw, axis=1) >= 0).all() and (np.diff(distances_bf, axis=1) >= 0).all()
    # exact scan == numpy
    d2 = ((q[:8, None, :].astype(np.float64) - data[None, :, :]) ** 2).sum(-1)
    assert (np.argsort(d2, axis=1, kind="stable")[:, :k] == labels_bf[:8].astype(np.int64)).all()
    hnsw_index.close()
    bf_index.close()


def test_spaces(hnswlib):
    # bindings_test_spaces.py:9-39, a subset of its (leftdim, rightdim) grid
    data1 = np.asarray([[1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1]])
    for space, expected in [("l2", [[0., 1., 2., 2., 2.]]), ("ip", [[-2., -1., 0., 0., 0.]]),
                            ("cosine", [[0, 1.835e-01, 4.23e-01, 4.23e-01, 4.23e-01]])]:
        for rightdim in (1, 4, 28, 127):
            for leftdim in (1, 6, 31):
                data2 = np.concatenate([np.zeros([5, leftdim]), data1, np.zeros([5, rightdim])], axis=1)
                p = hnswlib.Index(space=space, dim=data2.shape[1])
                p.init_index(max_elements=5, ef_construction=100, M=16)
                p.set_ef(10)
                p.add_items(data2)
                labels, distances = p.knn_query(np.asarray(data2[-1:]), k=5)
                assert np.mean(np.abs(distances - expected)) < 1e-3, (space, leftdim, rightdim, distances)
                p.close()


def test_save_load_add_query_getdata_metadata(hnswlib, tmp_path):
    # bindings_test.py:7-64 + bindings_test_labels.py (without the mark_deleted half) + getdata + metadata
    dim, n = 16, 10000
    data = np.float32(np.random.default_rng(1).random((n, dim)))
    p = hnswlib.Index(space="l2", dim=dim)
    with pytest.raises(Exception):
        p.get_items(np.arange(n))
    p.init_index(max_elements=n, ef_construction=100, M=16)
    p.set_ef(100)
    data1, data2 = data[:n // 2], data[n // 2:]
    p.add_items(data1)
    labels, distances = p.knn_query(data1, k=1)
    assert np.mean(labels.reshape(-1) == np.arange(len(data1))) > 0.9995
    assert np.mean(np.abs(data1 - p.get_items(labels))) < 1e-4
    assert (p.get_max_elements(), p.get_current_count()) == (n, n // 2)
    assert (p.space, p.dim, p.M, p.ef_construction, p.max_elements, p.element_count) == ("l2", dim, 16, 100, n, n // 2)
    path = str(tmp_path / "first_half.bin")
    p.save_index(path)
    assert p.index_file_size() == (tmp_path / "first_half.bin").stat().st_size
    before = p.knn_query(data2[:300], k=7)
    p.close()
    del p
    p = hnswlib.Index(space="l2", dim=dim)
    p.load_index(path)
    p.set_ef(100)
    after = p.knn_query(data2[:300], k=7)                         # the loaded graph answers like the one that was saved
    assert (before[0] == after[0]).all() and (before[1] == after[1]).all()
    if R.have_ref():                                              # the reference (CPU) reads the GPU-built file
        ref = R.RefHnsw(0, dim, path=path)
        lr, dr = ref.search(data2[:300], 7, 100, threads=1)
        ref.close()
        assert np.mean([len(set(lr[i].tolist()) & set(after[0][i].tolist())) for i in range(300)]) / 7 > 0.99
    with pytest.raises(RuntimeError, match="exceeds"):
        p.add_items(np.concatenate([data2, data2[:1]]))          # one more than max_elements: the first 5000 stay
    assert p.get_current_count() == n
    labels, distances = p.knn_query(data, k=1)
    assert np.mean(labels.reshape(-1) == np.arange(len(data))) > 0.9995
    assert np.mean(np.abs(data - p.get_items(labels))) < 1e-4
    assert sorted(p.get_ids_list()) == list(range(n))
    with pytest.raises(RuntimeError, match="contiguous 2D array"):
        p.knn_query(data[:2], k=n + 1)
    # pickle round trip (bindings_test_pickle.py): same parameters, same answers
    p2 = pickle.loads(pickle.dumps(p))
    assert (p2.space, p2.dim, p2.M, p2.ef_construction, p2.max_elements, p2.element_count, p2.ef) == \
           (p.space, p.dim, p.M, p.ef_construction, p.max_elements, p.element_count, p.ef)
    a, b = p.knn_query(data[:500], k=5), p2.knn_query(data[:500], k=5)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
    p.close()
    p2.close()


def test_custom_labels_match_oracle(hnswlib):
    # labels are arbitrary u64; the graph is the one b200ann_hnsw_build makes (pinned link for link to the oracle in
    # test_gpu_hnsw_build.py), so knn_query must answer exactly like the oracle searching the oracle-built graph
    B = importlib.import_module("hnsw-nsg_b200")
    n, dim, M, efc = 4000, 48, 8, 50
    base = R.data_uniform(n, dim, 9)
    labs = (np.arange(n, dtype=np.uint64) * np.uint64(7919) + np.uint64(1 << 40))
    p = hnswlib.Index("l2", dim)
    p.init_index(n, M=M, ef_construction=efc, random_seed=100)
    p.add_items(base[:1500], labs[:1500])
    p.add_items(base[1500:], labs[1500:])
    p.set_ef(64)
    q = R.data_uniform(200, dim, 10)
    lab, d = p.knn_query(q, k=10)
    R.oracle().orc_set_gpu_team(B.dist_team(dim))
    O = R.orc_hnsw_build(base, M, efc, R.L2_GPU, 1, ratio=32, bmax=16384)
    O["labels"] = labs
    ol, od = R.orc_hnsw_search(base, O, q, 10, 64, R.L2_GPU)
    assert (lab == ol).all()
    assert np.allclose(d, od, rtol=1e-5, atol=0)
    p.close()


def test_bfindex_filter_delete_cosine(hnswlib, tmp_path):
    dim, n = 20, 3000
    rng = np.random.default_rng(2)
    data = rng.standard_normal((n, dim)).astype(np.float32)
    q = rng.standard_normal((50, dim)).astype(np.float32)
    bf = hnswlib.BFIndex("cosine", dim)
    bf.init_index(n)
    bf.add_items(data, np.arange(n) + 100)
    lab, d = bf.knn_query(q, 5)
    unit = data / np.linalg.norm(data, axis=1, keepdims=True)
    want = 1.0 - q.astype(np.float64) @ unit.T.astype(np.float64)   # stored vectors normalised, queries as given
    order = np.argsort(want, axis=1)[:, :5]
    assert (lab.astype(np.int64) - 100 == order).all()
    assert np.allclose(d, np.take_along_axis(want, order, 1), atol=1e-4)
    lab_even, _ = bf.knn_query(q, 5, filter=lambda l: l % 2 == 0)   # bindings_test_filter.py: only even labels
    assert (lab_even % 2 == 0).all()
    sub = np.arange(0, n, 2)
    assert (lab_even.astype(np.int64) - 100 == sub[np.argsort(want[:, sub], axis=1)[:, :5]]).all()
    gone = int(lab[0, 0])
    bf.delete_vector(gone)
    lab2, _ = bf.knn_query(q[:1], 5)
    assert gone not in lab2 and (lab2[0, :4] == lab[0, 1:]).all()
    bf.save_index(str(tmp_path / "bf.bin"))
    b2 = hnswlib.BFIndex("cosine", dim)
    b2.load_index(str(tmp_path / "bf.bin"))
    lab3, d3 = b2.knn_query(q, 5)
    lab4, d4 = bf.knn_query(q, 5)
    assert (lab3 == lab4).all() and (d3 == d4).all()
    few = hnswlib.BFIndex("l2", dim)
    few.init_index(10)
    few.add_items(data[:3])
    l5, d5 = few.knn_query(q[:2], 5)                               # k > count: 3 answers, then padding
    assert np.isinf(d5[:, 3:]).all() and np.isfinite(d5[:, :3]).all()
    for x in (bf, b2, few):
        x.close()


def test_filter_and_mark_deleted(hnswlib, tmp_path):
    # bindings_test_filter.py (only even labels come back, recall against the filtered brute force) and the
    # mark_deleted half of bindings_test_labels.py:77-127 (deleted labels never come back, the marks survive
    # save_index -> load_index, unmark_deleted restores them)
    dim, n, k = 16, 10000, 10
    rng = np.random.default_rng(3)
    data = rng.random((n, dim), dtype=np.float32)
    q = rng.random((200, dim), dtype=np.float32)
    p = hnswlib.Index("l2", dim)
    p.init_index(n, ef_construction=100, M=16)
    p.set_ef(100)
    p.add_items(data)
    bf = hnswlib.BFIndex("l2", dim)
    bf.init_index(n)
    bf.add_items(data)
    even = lambda l: l % 2 == 0
    lh, dh = p.knn_query(q, k, filter=even)
    lb, db = bf.knn_query(q, k, filter=even)
    assert (lh % 2 == 0).all()
    rec = np.mean([len(set(lh[i].tolist()) & set(lb[i].tolist())) for i in range(len(q))]) / k
    assert rec > 0.95, rec
    with pytest.raises(RuntimeError, match="contiguous 2D array"):
        p.knn_query(q[:3], k, filter=lambda l: l < 5)            # only 5 labels pass: fewer than k results
    data1 = data[:n // 2]
    first, _ = p.knn_query(data1, k=1)
    assert np.mean(first.reshape(-1) == np.arange(len(data1))) > 0.999
    dead = set(first.reshape(-1).tolist())
    for l in dead:
        p.mark_deleted(l)
    after, _ = p.knn_query(data1, k=1)
    assert not (set(after.reshape(-1).tolist()) & dead)
    l2, _ = p.knn_query(data[n // 2:], k=1)                       # the other half is untouched
    assert np.mean(np.abs(data[n // 2:] - p.get_items(l2))) < 1e-3
    path = str(tmp_path / "with_deleted.bin")
    p.save_index(path)
    p2 = hnswlib.Index("l2", dim)
    p2.load_index(path)
    p2.set_ef(100)
    again, _ = p2.knn_query(data1, k=1)
    assert (again == after).all()
    if R.have_ref():                                              # the reference reads the marks this file carries
        ref = R.RefHnsw(0, dim, path=path)
        rl, _ = ref.search_filtered(np.ascontiguousarray(data1[:500]), 1, 100, None, threads=1)
        ref.close()
        assert (rl == after[:500]).mean() > 0.99
    for l in dead:
        p2.unmark_deleted(l)
    restored, _ = p2.knn_query(data1, k=1)
    assert np.mean(restored.reshape(-1) == np.arange(len(data1))) > 0.999
    for x in (p, p2, bf):
        x.close()


def test_replace_deleted(hnswlib, tmp_path):
    # bindings_test_replace.py:9-139 (dim 16, 4 batches, capacity = 2 batches, replace the deleted ones twice, with a
    # save -> load(allow_replace_deleted=True) in between); smaller batches than the reference's 5000
    dim, ne = 16, 2000
    rng = np.random.default_rng(5)
    data = [rng.random((ne, dim), dtype=np.float32) for _ in range(4)]
    labels = [np.arange(b * ne, (b + 1) * ne) for b in range(4)]
    ix = hnswlib.Index(space="l2", dim=dim)
    ix.init_index(max_elements=2 * ne, ef_construction=200, M=16, allow_replace_deleted=True)
    ix.set_ef(100)
    ix.add_items(data[0], labels[0])
    ix.add_items(data[1], labels[1])                       # the index is full
    with pytest.raises(RuntimeError, match="exceeds the specified limit"):
        ix.add_items(data[2][:1], labels[2][:1])
    deleted = set(ix.knn_query(data[1], k=1)[0].flatten().tolist())
    ndup = ne - len(deleted)
    for lab in deleted:
        ix.mark_deleted(lab)
    with pytest.raises(RuntimeError, match="Can't use addPoint to update deleted elements"):
        ix.add_items(data[1][:1], [next(iter(deleted))])
    found1 = ix.knn_query(data[0], k=1)[0]
    assert np.mean(np.abs(data[0] - ix.get_items(found1.flatten()))) < 1e-3
    assert not (set(ix.knn_query(data[1], k=1)[0].flatten().tolist()) & deleted)
    l3, d3 = labels[2][:ne - ndup], data[2][:ne - ndup]
    ix.add_items(d3, l3, replace_deleted=True)             # fits only because the deleted slots are reused
    assert ix.get_current_count() == 2 * ne
    remaining = sorted((set(labels[0].tolist()) | set(labels[1].tolist())) - deleted)
    comb = np.concatenate([data[0], data[1]])
    assert (np.asarray(ix.get_items(remaining)) == comb[remaining]).all()
    assert (np.asarray(ix.get_items(l3)) == d3).all()
    got = ix.knn_query(d3, k=1)[0].flatten()
    assert np.mean(got == l3) > 0.99                       # the replacing points are found under their new labels
    for lab in deleted:                                    # the replaced labels are gone
        with pytest.raises(RuntimeError):
            ix.get_items([lab])
        break
    # serialisation: delete batch 3, save, load with replacement enabled, replace with batch 4
    for lab in l3:
        ix.mark_deleted(lab)
    path = str(tmp_path / "replace.bin")
    ix.save_index(path)
    ix2 = hnswlib.Index(space="l2", dim=dim)
    ix2.load_index(path, max_elements=2 * ne, allow_replace_deleted=True)
    ix2.set_ef(100)
    l4, d4 = labels[3][:len(l3)], data[3][:len(l3)]
    ix2 = pickle.loads(pickle.dumps(ix2))               # allow_replace_deleted survives pickling (bindings.cpp:491, :567)
    ix2.add_items(d4, l4, replace_deleted=True)
    assert ix2.get_current_count() == 2 * ne
    got = ix2.knn_query(d4, k=1)[0].flatten()
    assert np.mean(got == l4) > 0.99
    assert not (set(ix2.knn_query(data[2], k=1)[0].flatten().tolist()) & set(l3.tolist()))
    # replacement disabled -> the reference's exception
    ix3 = hnswlib.Index(space="l2", dim=dim)
    ix3.init_index(max_elements=10)
    with pytest.raises(RuntimeError, match="Replacement of deleted elements is disabled in constructor"):
        ix3.add_items(data[0][:2], [0, 1], replace_deleted=True)

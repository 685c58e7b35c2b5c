"""hnsw-nsg_b200/hnswlib_py.py, CPU side: the HierarchicalNSW::saveIndex / BruteforceSearch::saveIndex writers and
readers (hnswlib/hnswlib/hnswalg.h:685-822, bruteforce.h:139-171) against the compiled reference, and the host logic
of the binding mirror (argument checks, labels, error messages of bindings.cpp) -- no compute calls."""
import importlib
import os

import numpy as np
import pytest

import refutil as R

hp = importlib.import_module("hnsw-nsg_b200.hnswlib_py")


def test_hnsw_file_writer_reader_vs_compiled_reference(tmp_path):
    if not R.have_ref():
        pytest.skip("oracle/_ref not built")
    n, dim, M, efc = 3000, 24, 8, 60
    base = R.data_uniform(n, dim, 3)
    ref = R.RefHnsw(0, dim, n, M, efc, 100)
    ref.add(base, label0=1000)
    p_ref = str(tmp_path / "ref.bin")
    ref.save(p_ref)
    g = hp._read_hnsw_file(p_ref, dim)                            # reader == the test parser, field by field
    want = R.parse_hnsw(p_ref)
    for key in ("links0", "levels", "upper_ptr", "upper_links", "labels", "vectors"):
        assert (g[key] == want[key]).all(), key
    assert (g["ep"], g["maxlevel"], g["M"], g["efc"], g["max_elements"]) == (want["ep"], want["maxlevel"], M, efc, n)
    p_ours = str(tmp_path / "ours.bin")
    hp._write_hnsw_file(p_ours, g, g["vectors"], n, M, efc)       # same size and header; slots past a row's count hold
    a, b = open(p_ours, "rb").read(), open(p_ref, "rb").read()    # stale links in the reference's file, zeros in ours
    assert len(a) == len(b) and a[:96] == b[:96]
    mine = R.parse_hnsw(p_ours)
    for key in ("links0", "levels", "upper_ptr", "upper_links", "labels", "vectors"):
        assert (mine[key] == want[key]).all(), key
    q = R.data_uniform(50, dim, 4)
    again = R.RefHnsw(0, dim, path=p_ours)                        # and the reference loads + searches it
    l0, d0 = ref.search(q, 5, 40, threads=1)
    l1, d1 = again.search(q, 5, 40, threads=1)
    assert (l0 == l1).all() and (d0 == d1).all()
    ref.close()
    again.close()
    # delete marks: the reference's markDelete -> file -> our reader; our writer -> the reference's loader + search
    ref2 = R.RefHnsw(0, dim, path=p_ref)
    dele = np.zeros(n, bool)
    dele[::7] = True
    for i in np.nonzero(dele)[0]:
        ref2.mark_deleted(1000 + i)
    p_del = str(tmp_path / "ref_deleted.bin")
    ref2.save(p_del)
    gd = hp._read_hnsw_file(p_del, dim)
    assert (gd["deleted"] == dele).all() and not g["deleted"].any()
    p_del2 = str(tmp_path / "ours_deleted.bin")
    hp._write_hnsw_file(p_del2, g, g["vectors"], n, M, efc, deleted=dele)
    ref3 = R.RefHnsw(0, dim, path=p_del2)
    l2, d2 = ref2.search_filtered(q, 5, 40, None, threads=1)
    l3, d3 = ref3.search_filtered(q, 5, 40, None, threads=1)
    assert (l2 == l3).all() and (d2 == d3).all() and not dele[(l3 - 1000).astype(np.int64)].any()
    ref2.close()
    ref3.close()
    with pytest.raises(RuntimeError, match="dimension"):
        hp._read_hnsw_file(p_ref, dim + 1)
    with pytest.raises(RuntimeError, match="Cannot open file"):
        hp._read_hnsw_file(str(tmp_path / "missing.bin"), dim)


def test_hnsw_file_writer_scattered_upper_rows(tmp_path):
    # upper_ptr need not be the running sum (the GPU build hands rows out in its own order): same file either way
    n, dim, M = 400, 8, 4
    base = R.data_uniform(n, dim, 5)
    R.oracle().orc_set_gpu_team(1)
    g = R.orc_hnsw_build(base, M, 20, R.L2_GPU, 0)
    g["labels"] = np.arange(n, dtype=np.uint64)
    p0, p1 = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    hp._write_hnsw_file(p0, g, base, n, M, 20)
    rows = g["upper_links"].shape[0]
    perm = np.random.default_rng(0).permutation(n)
    # re-pack the upper rows element by element in a shuffled element order
    new_ptr = np.zeros(n, np.uint32)
    new_up = np.zeros_like(g["upper_links"])
    at = 0
    for i in perm:
        lv = int(g["levels"][i])
        if lv:
            new_ptr[i] = at
            new_up[at:at + lv] = g["upper_links"][g["upper_ptr"][i]:g["upper_ptr"][i] + lv]
            at += lv
    assert at == rows
    g2 = dict(g, upper_ptr=new_ptr, upper_links=new_up)
    hp._write_hnsw_file(p1, g2, base, n, M, 20)
    assert open(p0, "rb").read() == open(p1, "rb").read()
    back = hp._read_hnsw_file(p1, dim)
    for key in ("links0", "levels", "upper_links"):
        assert (back[key] == g[key]).all(), key
    has = g["levels"] > 0
    assert (back["upper_ptr"][has] == g["upper_ptr"][has]).all()


def test_binding_host_logic_without_compute(tmp_path):
    with pytest.raises(RuntimeError, match="Space name"):
        hp.Index("hamming", 4)
    p = hp.Index("l2", 4)
    assert (p.M, p.ef_construction, p.max_elements, p.element_count, p.ef) == (0, 0, 0, 0, 10)
    with pytest.raises(RuntimeError):
        p.add_items(np.zeros((1, 4), np.float32))
    p.init_index(max_elements=5, ef_construction=3, M=6)
    with pytest.raises(RuntimeError, match="already initiated"):
        p.init_index(5)
    assert (p.M, p.ef_construction, p.max_elements) == (6, 6, 5)   # ef_construction = max(efc, M), hnswalg.h:107
    p.ef = 33
    assert p.ef == 33
    with pytest.raises(RuntimeError, match="Wrong dimensionality"):
        p.add_items(np.zeros((2, 3), np.float32))
    x = np.arange(20, dtype=np.float32).reshape(5, 4)
    p.add_items(x[:2])                                            # labels 0, 1
    p.add_items(x[2:4], ids=[7, 9])
    p.add_items(x[4])                                             # 1-D row; default label = rows added so far = 4
    assert p.get_ids_list() == [0, 1, 7, 9, 4]
    assert p.get_current_count() == 5 and p.get_max_elements() == 5
    with pytest.raises(RuntimeError, match="exceeds the specified limit"):
        p.add_items(x[:1], ids=[100])
    p.add_items(x[:1] + 1, ids=[9])                               # existing label: the vector is replaced
    assert (p.get_items([9, 0]) == np.stack([x[0] + 1, x[0]])).all()
    assert p.get_items([7], return_type="list") == [x[2].tolist()]
    with pytest.raises(ValueError):
        p.get_items(np.int64(3))
    with pytest.raises(RuntimeError, match="Label not found"):
        p.get_items([12345])
    with pytest.raises(RuntimeError, match="Cannot resize"):
        p.resize_index(2)
    p.resize_index(9)
    assert p.max_elements == 9
    with pytest.raises(RuntimeError, match="Replacement of deleted elements is disabled in constructor"):  # hnswalg.h:955-957
        p.add_items(x[:1], replace_deleted=True)
    # markDelete bookkeeping and its error messages (hnswalg.h:853-922); an update un-deletes (:1166-1168)
    p.mark_deleted(7)
    with pytest.raises(RuntimeError, match="already deleted"):
        p.mark_deleted(7)
    with pytest.raises(RuntimeError, match="Label not found"):
        p.mark_deleted(4242)
    with pytest.raises(RuntimeError, match="not deleted"):
        p.unmark_deleted(0)
    assert p._deleted.tolist() == [False, False, True, False, False]
    with pytest.raises(RuntimeError, match="Label not found"):  # getDataByLabel refuses deleted elements, hnswalg.h:832
        p.get_items([7])
    p.unmark_deleted(7)
    p.mark_deleted(9)
    p.add_items(x[:1], ids=[9])
    assert not p._deleted.any()
    # replace_deleted bookkeeping (hnswalg.h:954-992, :1160-1167): deleted slots are reused smallest id first
    r = hp.Index("l2", 4)
    r.init_index(max_elements=4, allow_replace_deleted=True)
    r.add_items(x[:4], ids=[10, 11, 12, 13])
    r.mark_deleted(12)
    r.mark_deleted(10)
    with pytest.raises(RuntimeError, match="Can't use addPoint to update deleted elements"):
        r.add_items(x[:1], ids=[12])
    with pytest.raises(RuntimeError, match="exceeds the specified limit"):
        r.add_items(x[4:5], ids=[50])
    r.add_items(x[4:5] + 7, ids=[50], replace_deleted=True)
    assert r.get_ids_list() == [50, 11, 12, 13] and r._deleted.tolist() == [False, False, True, False]
    assert (r.get_items([50]) == x[4:5] + 7).all() and r.get_current_count() == 4
    with pytest.raises(RuntimeError, match="Label not found"):
        r.get_items([10])
    r.add_items(x[0:1], ids=[60], replace_deleted=True)           # takes slot 2 (label 12)
    assert r.get_ids_list() == [50, 11, 60, 13] and not r._deleted.any()
    with pytest.raises(RuntimeError, match="exceeds the specified limit"):   # nothing vacant: a plain add
        r.add_items(x[1:2], ids=[61], replace_deleted=True)
    c = hp.Index("cosine", 4)
    c.init_index(3)
    c.add_items(np.array([[3, 0, 4, 0]], np.float32))
    assert np.allclose(c.get_items([0]), [[0.6, 0, 0.8, 0]])
    # BFIndex bookkeeping: removePoint moves the last element into the hole (bruteforce.h:86-103)
    b = hp.BFIndex("l2", 4)
    b.init_index(4)
    b.add_items(x[:4], ids=[10, 11, 12, 13])
    b.delete_vector(11)
    b.delete_vector(999)
    assert b.get_current_count() == 3 and b._labels[:3].tolist() == [10, 13, 12]
    b.save_index(str(tmp_path / "bf.bin"))
    assert os.path.getsize(tmp_path / "bf.bin") == 24 + 4 * (16 + 8)
    b2 = hp.BFIndex("l2", 4)
    b2.load_index(str(tmp_path / "bf.bin"))
    assert b2.get_current_count() == 3 and b2.get_max_elements() == 4
    assert (b2._vec[:3] == x[[0, 3, 2]]).all() and b2._lookup == {10: 0, 13: 1, 12: 2}
    with pytest.raises(RuntimeError, match="exceeds"):
        b2.add_items(x[:2], ids=[1, 2])

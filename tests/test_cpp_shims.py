"""The C++ host shims (include/b200ann/*.hpp) mirror the reference's classes over the C ABI.
CPU: they compile with plain g++ and link against libb200ann.so, and refuse to run without a GPU.
GPU: tests/cpp/*.cpp drive them the way the reference's own tests / drivers drive the reference."""
import os
import subprocess

import numpy as np
import pytest

import refutil as R

INC = os.path.join(R.ROOT, "include")
LIBDIR = os.path.join(R.ROOT, "hnsw-nsg_b200")


def _build(name, tmp):
    exe = os.path.join(str(tmp), name)
    src = os.path.join(R.ROOT, "tests", "cpp", name + ".cpp")
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-fopenmp", "-I" + INC, src, "-o", exe, "-L" + LIBDIR, "-lb200ann",
                           "-Wl,-rpath," + LIBDIR])
    return exe


@pytest.mark.parametrize("name", ["test_efanna_shim", "test_hnswlib_shim", "test_hybrid_shim", "test_l2spacei_shim", "test_replace_deleted",
                                  "test_hybrid_order", "test_epsilon_search", "test_hnswlib_int", "test_updates", "test_multivector_search"])
def test_shims_compile_and_fail_loudly_without_gpu(tmp_path, name):
    exe = _build(name, tmp_path)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
        assert p.returncode != 0 and "no CUDA device" in (p.stderr + p.stdout)


@pytest.mark.gpu
def test_l2spacei_shim_driver(tmp_path):
    """BruteforceSearch<int> over L2SpaceI (space_l2.h:294-323) through the shim: labels and int distances"""
    exe = _build("test_l2spacei_shim", tmp_path)
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "l2spacei shim ok" in p.stdout


@pytest.mark.gpu
def test_replace_deleted_driver(tmp_path):
    """allow_replace_deleted / addPoint(replace_deleted=True), hnswalg.h:954-992"""
    exe = _build("test_replace_deleted", tmp_path)
    p = subprocess.run([exe], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "replace_deleted ok" in p.stdout


@pytest.mark.gpu
def test_epsilon_search_driver(tmp_path):
    """searchStopConditionClosest + EpsilonSearchStopCondition, the flow of hnswlib/tests/cpp/epsilon_search_test.cpp"""
    exe = _build("test_epsilon_search", tmp_path)
    p = subprocess.run([exe], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "epsilon search ok" in p.stdout


@pytest.mark.gpu
def test_hnswlib_int_driver(tmp_path):
    """HierarchicalNSW<int> over L2SpaceI: GPU build, recall vs BruteforceSearch<int>, save/load, filter, delete, epsilon,
    append; then (with oracle/_ref) an index file written by the unmodified reference, searched through the shim"""
    exe = _build("test_hnswlib_int", tmp_path)
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "hnswlib<int> shim ok" in p.stdout
    if not R.have_ref():
        return
    rng = np.random.default_rng(3)
    n, dim, nq, k, ef = 5000, 96, 200, 10, 50
    base = np.minimum(255, rng.gamma(1.2, 24.0, (n, dim))).astype(np.uint8)
    q = np.minimum(255, rng.gamma(1.2, 24.0, (nq, dim))).astype(np.uint8)
    ref = R.RefHnswI(dim, n, 16, 100)
    ref.add(base, label0=300, threads=1)
    path = str(tmp_path / "ref_hnswi.bin")
    ref.save(path)
    lab, d = ref.search(q, k, ef, threads=1)
    with open(str(tmp_path / "expect.bin"), "wb") as f:
        np.array([nq, k, ef], np.uint64).tofile(f)
        q.tofile(f); lab.tofile(f); d.tofile(f)
    p = subprocess.run([exe, path, str(dim), str(tmp_path / "expect.bin")], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "vs reference file ok" in p.stdout


@pytest.mark.gpu
def test_updates_driver(tmp_path):
    """hnswlib/tests/cpp/updates_test.cpp in "update" mode on the reference's golden data: all 100000 elements are moved
    with addPoint on their existing labels (updatePoint, in place on the GPU), then recall > 0.99 must be reachable.
    batch_dummy_00 is regenerated the way hnswlib/tests/cpp/update_gen_data.py made it (np.random.seed(1)); the
    regenerated final batch must equal the committed golden file."""
    import shutil
    gold = os.path.join(R.ROOT, "tests", "golden")
    n, mult, nq, d, k = [int(x) for x in open(os.path.join(gold, "hnswlib_updates_config.txt")).read().split()]

    def normalized(a):
        l2 = np.atleast_1d(np.linalg.norm(a, 2, -1))
        l2[l2 == 0] = 1
        return a / np.expand_dims(l2, -1)

    np.random.seed(1)
    dummies = [normalized(np.float32(np.random.random((n, d)))) for _ in range(mult)]
    final = normalized(np.float32(np.random.random((n, d))))
    gold_final = np.fromfile(os.path.join(gold, "hnswlib_updates_batch_final.bin"), np.float32).reshape(n, d)
    assert np.array_equal(final, gold_final)
    dummies[0].astype(np.float32).tofile(str(tmp_path / "batch_dummy_00.bin"))
    for f in ("batch_final.bin", "queries.bin", "gt.bin", "config.txt"):
        shutil.copyfile(os.path.join(gold, "hnswlib_updates_" + f), str(tmp_path / f))
    exe = _build("test_updates", tmp_path)
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "updates ok" in p.stdout


@pytest.mark.gpu
def test_multivector_search_driver(tmp_path):
    """MultiVectorL2Space + MultiVectorSearchStopCondition, the flow of hnswlib/tests/cpp/multivector_search_test.cpp"""
    exe = _build("test_multivector_search", tmp_path)
    p = subprocess.run([exe], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "multivector search ok" in p.stdout


@pytest.mark.gpu
def test_efanna_shim_driver(tmp_path):
    exe = _build("test_efanna_shim", tmp_path)
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0
    if R.have_ref():  # the .nsg the shim saved is readable by the unmodified reference and searchable there
        rng = np.random.default_rng(0)
        g, ep, width = R.parse_nsg(str(tmp_path / "shim.nsg"), 5000)
        assert g.shape[0] == 5000 and ep < 5000


@pytest.mark.gpu
@pytest.mark.skipif(not R.have_ref(), reason="needs oracle/_ref to write a reference index file")
def test_hnswlib_shim_driver(tmp_path):
    exe = _build("test_hnswlib_shim", tmp_path)
    n, dim = 4000, 16
    base = R.data_uniform(n, dim, 1)
    ref = R.RefHnsw(0, dim, n, 16, 100)
    ref.add(base, label0=500, threads=1)
    path = str(tmp_path / "ref_hnsw.bin")
    ref.save(path)
    p = subprocess.run([exe, path, str(dim)], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0
    # the file re-saved by the shim loads in the unmodified reference and answers like the original
    again = R.RefHnsw(0, dim, path=path + ".resaved")
    q = R.data_uniform(50, dim, 9)
    a, ad = ref.search(q, 10, 50, threads=1)
    b, bd = again.search(q, 10, 50, threads=1)
    assert (a == b).all() and (ad == bd).all()
    # the file saved with delete marks: the unmodified reference sees the same marks and, searching the same queries
    # (searchBaseLayerST<false> on the CPU), returns the labels the GPU returned (CPU vs GPU distance rounding only)
    marked = R.RefHnsw(0, dim, path=path + ".deleted")
    qd = np.fromfile(path + ".deleted.q", np.float32).reshape(-1, dim)
    want = np.fromfile(path + ".deleted.labels", np.uint64).reshape(len(qd), 10)
    got, _ = marked.search_filtered(qd, 10, 100, None, threads=1)
    assert (got == want).mean() > 0.995
    orig, _ = ref.search(qd, 10, 100, threads=1)
    assert (got[:, 0] != orig[:, 0]).all()                            # the original nearest neighbours are the deleted ones
    # the index the shim BUILT on the GPU (addPoint + build + saveIndex) loads in the unmodified reference, which
    # finds every stored point as its own nearest neighbour
    built = R.RefHnsw(0, 16, path=path + ".gpubuilt")
    rng = np.random.RandomState(0)
    Hb = R.parse_hnsw(path + ".gpubuilt")
    lab, dd = built.search(np.ascontiguousarray(Hb["vectors"][:200]), 1, 50, threads=1)
    assert (lab[:, 0] == Hb["labels"][:200]).mean() > 0.99


@pytest.mark.gpu
def test_hybrid_shim_driver(tmp_path):
    """hnsw_nsg::HNSW_NSG through the shim: build (addPoint + Build_NSG), search, saveIndex, load ctor; then the
    UNMODIFIED reference hybrid loads the pair of files the shim wrote and reaches the same recall."""
    exe = _build("test_hybrid_shim", tmp_path)
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0
    if R.have_ref():
        prefix = str(tmp_path / "hyshim")
        Hb = R.parse_hnsw(prefix + "_hnsw.bin")
        base = np.ascontiguousarray(Hb["vectors"])
        assert (Hb["labels"] == np.arange(len(base))).all()
        q = R.data_uniform(300, base.shape[1], 77)
        truth, _ = R.orc_bf_knn(base, q, 10, R.L2_SCALAR)
        hy = R.RefHybrid(base, prefix, 60)
        ids = hy.search(q, 10, 10, seed=1)
        assert R.recall(ids[:, :10], truth) > 0.9


@pytest.mark.gpu
def test_hybrid_order_driver(tmp_path):
    """addPoint in shuffled / parallel order: the hybrid shim still answers in data-row ids"""
    exe = _build("test_hybrid_order", tmp_path)
    p = subprocess.run([exe, "build", str(tmp_path)], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "hybrid order ok" in p.stdout


@pytest.mark.gpu
@pytest.mark.skipif(not R.have_ref(), reason="needs oracle/_ref to build a reference hybrid index with several threads")
def test_hybrid_shim_loads_parallel_built_reference_index(tmp_path):
    """the UNMODIFIED reference builds the pair of files with 4 addPoint threads (internal id != label); the shim's load
    ctor answers with the same data-row ids the reference returns on those files"""
    n, dim, nq, k, L = 5000, 32, 300, 10, 40
    base = R.data_sift_like(n, dim, 3)      # integer-valued: every summation order is exact
    q = R.data_sift_like(nq, dim, 4)
    kpath = str(tmp_path / "k.graph")
    knn, _ = R.orc_bf_knn(base, base, 41, R.L2_SCALAR)
    knn = np.stack([row[row != i][:40] for i, row in enumerate(knn)]).astype(np.uint32)
    R.write_knn_graph(kpath, knn)
    prefix = str(tmp_path / "refhy")
    R.ref_hybrid_build(base, kpath, 40, 30, 300, 16, 100, prefix, seed=1, threads=4)
    Hb = R.parse_hnsw(prefix + "_hnsw.bin")
    assert (Hb["labels"] != np.arange(n)).any(), "the parallel build happened to add in order; nothing is tested"
    hy = R.RefHybrid(base, prefix, L)
    want = hy.search(q, k, k, seed=1)
    base.tofile(str(tmp_path / "base.f32")); q.tofile(str(tmp_path / "q.f32"))
    exe = _build("test_hybrid_order", tmp_path)
    out = str(tmp_path / "ids.u32")
    p = subprocess.run([exe, "load", prefix, str(tmp_path / "base.f32"), str(n), str(dim), str(tmp_path / "q.f32"), str(nq),
                        str(L), str(k), out], capture_output=True, text=True)
    print(p.stdout, p.stderr)
    assert p.returncode == 0
    got = np.fromfile(out, np.uint32).reshape(nq, k)
    truth, _ = R.orc_bf_knn(base, q, k, R.L2_SCALAR)
    assert abs(R.recall(got, truth) - R.recall(want[:, :k], truth)) <= 0.005
    same = np.mean([set(a.tolist()) == set(b.tolist()) for a, b in zip(got, want[:, :k])])
    assert same > 0.97, same

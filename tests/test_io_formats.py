"""fvecs / ivecs / centroids files: the numpy reader/writers (hnsw-nsg_b200/io.py) and the C++
loaders with the reference's names (include/b200ann/io.hpp, CNNS/src/utils/aux_util.cpp:8-84) agree byte for byte.
CPU only."""
import importlib
import os
import subprocess

import numpy as np

import refutil as R

io = importlib.import_module("hnsw-nsg_b200.io")


def test_fvecs_ivecs_centroids_roundtrip_python_and_cpp(tmp_path):
    n, dim = 11, 6
    x = (np.arange(n * dim, dtype=np.float32).reshape(n, dim) + 0.5)
    io.save_fvecs(str(tmp_path / "a.fvecs"), x)
    assert os.path.getsize(tmp_path / "a.fvecs") == n * (dim + 1) * 4
    assert (io.load_fvecs(str(tmp_path / "a.fvecs")) == x).all()
    gt = np.arange(35, dtype=np.uint32).reshape(7, 5)
    io.save_ivecs(str(tmp_path / "gt.ivecs"), gt)
    assert (io.load_ivecs(str(tmp_path / "gt.ivecs")) == gt).all()
    assert (R.read_knn_graph(str(tmp_path / "gt.ivecs")) == gt).all()      # efanna .graph is the same row layout
    io.save_centroids(str(tmp_path / "centroids.fvecs"), x[:9], 3, 2)
    pts, nc, m = io.load_centroids(str(tmp_path / "centroids.fvecs"))
    assert nc == 3 and m == 2 and (pts == x[:9]).all()
    exe = str(tmp_path / "test_io")
    subprocess.check_call(["g++", "-std=c++11", "-O1", "-I" + os.path.join(R.ROOT, "include"),
                           os.path.join(R.ROOT, "tests", "cpp", "test_io.cpp"), "-o", exe])
    p = subprocess.run([exe, str(tmp_path), str(n), str(dim), repr(float(x[-1, -1]))], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    assert open(tmp_path / "b.fvecs", "rb").read() == open(tmp_path / "a.fvecs", "rb").read()
    assert open(tmp_path / "gt2.ivecs", "rb").read() == open(tmp_path / "gt.ivecs", "rb").read()


def test_ragged_and_truncated_files(tmp_path):
    # ragged ground truth rows load as a list (loadGT accepts them); a truncated tail row is dropped
    raw = np.array([2, 7, 8, 3, 1, 2, 3, 4, 9], np.uint32)          # rows [7,8], [1,2,3], then a cut-off row of 4
    raw.tofile(str(tmp_path / "r.ivecs"))
    rows = io.load_ivecs(str(tmp_path / "r.ivecs"))
    assert [r.tolist() for r in rows] == [[7, 8], [1, 2, 3]]
    open(tmp_path / "e.fvecs", "wb").close()
    assert io.load_fvecs(str(tmp_path / "e.fvecs")).shape == (0, 0)

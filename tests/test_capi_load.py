"""The C-ABI library loads and exports every symbol include/b200ann.h declares; without a CUDA device
the product fails loudly (no CPU fallback). CPU only, no compute."""
import ctypes
import importlib
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header():
    hdr = open(os.path.join(ROOT, "include", "b200ann.h")).read()
    declared = set(re.findall(r"\b(b200ann_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 15
    lib = ctypes.CDLL(os.path.join(ROOT, "hnsw-nsg_b200", "libb200ann.so"))
    for name in declared:
        assert hasattr(lib, name), "missing export " + name
    mod = importlib.import_module("hnsw-nsg_b200")
    assert set(mod.EXPORTS) == declared


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    mod = importlib.import_module("hnsw-nsg_b200")
    with pytest.raises(mod.B200AnnError):
        mod.Handle(0)


def test_product_does_not_reference_oracle():
    """nothing under hnsw-nsg_b200/ may import, link or load the oracle"""
    for dp, _, fs in os.walk(os.path.join(ROOT, "hnsw-nsg_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                s = open(os.path.join(dp, f), errors="ignore").read()
                assert "liboracle" not in s and "oracle/_ref" not in s and "import oracle" not in s, f

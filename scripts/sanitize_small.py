"""small NN-descent / HNSW-build / search runs for compute-sanitizer (memcheck):
compute-sanitizer --tool memcheck python scripts/sanitize_small.py"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("hnsw-nsg_b200")
h = B.Handle(0)
rng = np.random.default_rng(1)
for n, dim in ((700, 20), (1500, 132)):
    base = rng.random((n, dim), dtype=np.float32)
    ix = B.Index(h, base, B.L2)
    knn, info = ix.nndescent(10, 24, 3, 6, 12, seed=1)
    print("nndescent", n, dim, info["dist_evals"])
    hi = ix.hnsw_build(6, 40, ratio=4, bmax=256)
    print("hnsw_build", hi["batches"])
    lab, d = ix.hnsw_search(base[:64], 5, 30)
    ix.nsg_build(knn, 20, 12, 60, seed=1)
    ids, d, e, st = ix.hybrid_search(base[:64], 5, 20, seed=1, want_stats=True)
    print("hybrid ok", int((ids[:, 0] == np.arange(64)).sum()))
    ix.close()
ipx = B.Index(h, rng.random((900, 33), dtype=np.float32), B.IP)
print("ip hnsw", ipx.hnsw_build(5, 30)["batches"])

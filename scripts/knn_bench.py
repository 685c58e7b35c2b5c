"""kNN-graph build timing on the GPU: python scripts/knn_bench.py n K [dim] [mode] [reps]"""
import importlib, sys, time, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("hnsw-nsg_b200")
synth = importlib.import_module("hnsw-nsg_b200.synth")
n = int(sys.argv[1]); K = int(sys.argv[2])
dim = int(sys.argv[3]) if len(sys.argv) > 3 else 128
mode = int(sys.argv[4]) if len(sys.argv) > 4 else 2
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
h = B.Handle(0)
base = synth.sift_shape(n, dim, 1) if dim == 128 else np.random.default_rng(1).random((n, dim), dtype=np.float32)
ix = B.Index(h, base, 0)
h.set_knn_mode(mode)
import torch
out = torch.empty((n, K), dtype=torch.int32, device="cuda")
for r in range(reps):
    t0 = time.time()
    ix.knn_graph_raw(K, out)
    print("rep %d: wall %.1f ms, device %.1f ms, stats %s" % (r, (time.time() - t0) * 1e3, h.last_kernel_ms, h.knn_last_stats), flush=True)

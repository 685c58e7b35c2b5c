"""NSG build phase timing on one GPU: python scripts/nsg_phase_bench.py n K L R C [dim]"""
import importlib, sys, time, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("hnsw-nsg_b200")
synth = importlib.import_module("hnsw-nsg_b200.synth")
n, K, L, R_, C_ = (int(x) for x in sys.argv[1:6])
dim = int(sys.argv[6]) if len(sys.argv) > 6 else 128
h = B.Handle(0)
base = synth.sift_shape(n, dim, 1)
ix = B.Index(h, base, 0)
knn = ix.knn_graph(K)
print("knn_graph %.1f ms" % h.last_kernel_ms, flush=True)
ep = ix.nsg_init_ep(knn, L, seed=1); print("init_ep %.1f ms" % h.last_kernel_ms, flush=True)
ep = ep[0] if isinstance(ep, tuple) else ep
for bits in (11, 10, 9, 8):
    h.set_link_visited_bits(bits)
    ci, cd, st = ix.nsg_link(knn, ep, L, R_, C_, seed=1, want_stats=True)
    print("link (visited bits %d) %.1f ms  beam evals %d prune evals %d" % (bits, h.last_kernel_ms, st[:, 0].astype(np.int64).sum(), st[:, 2].astype(np.int64).sum()), flush=True)
h.set_link_visited_bits(12)
r = ix.nsg_interinsert(ci, cd); print("interinsert %.1f ms" % h.last_kernel_ms, flush=True)
ci2, cd2 = r[0], r[1]
t = time.time(); ix.nsg_finish(ci2, cd2, ep, L, seed=1); print("finish %.1f ms (wall %.1f)" % (h.last_kernel_ms, (time.time() - t) * 1e3), flush=True)

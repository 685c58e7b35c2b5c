"""GPU probe: kNN graph -> NSG build -> search sweep at N x 128 (not a bench line)."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import refutil as R
B = importlib.import_module("hnsw-nsg_b200")
n = int(os.environ.get("N", 200000)); dim = int(os.environ.get("DIM", 128)); nq = 10000
Kg = int(os.environ.get("KG", 64)); L = int(os.environ.get("L", 40)); Rr = int(os.environ.get("R", 50)); Cc = int(os.environ.get("C", 500))
ncl = int(os.environ.get("NCL", 1000))
lat = int(os.environ.get('LATENT', 0))
t = time.time()
if lat:
    base = R.data_sift_shape(n, dim, 1, latent=lat); q = R.data_sift_shape(nq, dim, 2, latent=lat)
else:
    base = R.data_sift_like(n, dim, 1, n_clusters=ncl); q = R.data_sift_like(nq, dim, 2, n_clusters=ncl)
print("gen %.1fs" % (time.time() - t), flush=True)
h = B.Handle(0)
ix = B.Index(h, base, 0)
t = time.time(); gt, gd = ix.bf_knn(q, 10); print("bf gt: kernel %.1f ms" % h.last_kernel_ms, flush=True)
t = time.time(); knn = ix.knn_graph(Kg); print("knn_graph K=%d: wall %.2fs kernel %.1f ms" % (Kg, time.time() - t, h.last_kernel_ms), flush=True)
for lb in [int(x) for x in os.environ.get("LBITS", "12").split(",")]:
    h.set_link_visited_bits(lb)
    t = time.time(); info = ix.nsg_build(R.pad_graph(knn), L, Rr, Cc, seed=1)
    print("nsg_build lbits=%d: wall %.2fs device %.1f ms info %s" % (lb, time.time() - t, h.last_kernel_ms, info), flush=True)
    print("   per node: beam evals %.0f hops %.1f prune evals %.0f" % (info["beam_evals"] / n, info["hops"] / n, info["prune_evals"] / n))
g, ep, width = ix.get_graph()
deg = (g != R.INVALID).sum(1)
print("graph: stride %d width %d mean deg %.1f" % (g.shape[1], width, deg.mean()), flush=True)
gs = g.shape[1]
for bits in (10,):
    h.set_visited_bits(bits)
    for Ls in (10, 20, 30, 40, 60, 80, 120):
        ids, d, st = ix.nsg_search(q, 10, Ls, seed=1, want_stats=True)
        ms = h.last_kernel_ms
        for _ in range(2):
            ids, d, st = ix.nsg_search(q, 10, Ls, seed=1, want_stats=True)
            ms = min(ms, h.last_kernel_ms)
        nd = st[:, 0].astype(np.float64).sum(); nh = st[:, 1].astype(np.float64).sum()
        byts = nd * dim * 4 + nh * (gs * 4)
        print("bits=%d L=%d: %.3f ms  QPS %.0f  recall@10 %.4f  ndist/q %.0f hops/q %.0f  GB/s %.0f" % (
            bits, Ls, ms, nq / ms * 1e3, R.recall(ids, gt), nd / nq, nh / nq, byts / ms / 1e6), flush=True)

"""first GPU probe: kernel timings at 1M x 128 on an exact kNN graph (not a bench line)."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import refutil as R
B = importlib.import_module("hnsw-nsg_b200")
n = int(os.environ.get("N", 1000000)); dim = 128; nq = 10000; Kg = int(os.environ.get("KG", 32))
t = time.time(); base = R.data_sift_like(n, dim, 1, n_clusters=1000); q = R.data_sift_like(nq, dim, 2, n_clusters=1000)
print("gen %.1fs" % (time.time() - t), flush=True)
h = B.Handle(0)
t = time.time(); ix = B.Index(h, base, 0); print("upload %.2fs" % (time.time() - t), flush=True)
t = time.time(); gt, gd = ix.bf_knn(q, 10); print("bf gt 10k x %d: wall %.2fs kernel %.1f ms" % (n, time.time() - t, h.last_kernel_ms), flush=True)
t = time.time(); knn = ix.knn_graph(Kg); print("knn_graph K=%d: wall %.2fs kernel %.1f ms" % (Kg, time.time() - t, h.last_kernel_ms), flush=True)
gs = (Kg + 7) // 8 * 8
g = np.full((n, gs), R.INVALID, np.uint32); g[:, :Kg] = knn
ix.set_graph(g, 0)
for bits in (9, 10, 11):
    h.set_visited_bits(bits)
    for L in (32, 64, 128):
        ids, d, st = ix.nsg_search(q, 10, L, seed=1, want_stats=True)
        ms = h.last_kernel_ms
        ids, d, st = ix.nsg_search(q, 10, L, seed=1, want_stats=True)
        ms = min(ms, h.last_kernel_ms)
        nd = st[:, 0].astype(np.float64).sum(); nh = st[:, 1].astype(np.float64).sum()
        byts = nd * dim * 4 + nh * (gs * 4)
        print("bits=%d L=%d: %.2f ms  QPS %.0f  recall@10 %.4f  ndist/q %.0f hops/q %.0f  GB/s %.0f" % (
            bits, L, ms, nq / ms * 1e3, R.recall(ids, gt), nd / nq, nh / nq, byts / ms / 1e6), flush=True)

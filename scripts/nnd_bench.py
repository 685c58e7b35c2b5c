"""NN-descent timing + recall on the GPU: python scripts/nnd_bench.py n K L iter S R [dim] [cpu_sample_n]
Prints per-run device time, distance evaluations and recall of the K-NN lists against the exact graph;
optionally times the reference's IndexGraph::Build (oracle/_ref, all host cores) on the first cpu_sample_n rows."""
import importlib, sys, time, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
B = importlib.import_module("hnsw-nsg_b200")
synth = importlib.import_module("hnsw-nsg_b200.synth")
n, K, L, it, S, Rr = (int(x) for x in sys.argv[1:7])
dim = int(sys.argv[7]) if len(sys.argv) > 7 else 128
cpu_n = int(sys.argv[8]) if len(sys.argv) > 8 else 0
import torch
h = B.Handle(0)
base = synth.sift_shape(n, dim, 1) if dim == 128 else np.random.default_rng(1).random((n, dim), dtype=np.float32)
ix = B.Index(h, base, 0)
out = torch.empty((n, K), dtype=torch.int32, device="cuda")
for r in range(2):
    t0 = time.time()
    _, info = ix.nndescent(K, L, it, S, Rr, seed=1, out=out)
    print("nndescent rep %d: wall %.1f ms device %.1f ms info %s" % (r, (time.time() - t0) * 1e3, h.last_kernel_ms, info), flush=True)
exact = torch.empty((n, K), dtype=torch.int32, device="cuda")
ix.knn_graph_raw(K, exact)
print("exact knn_graph device %.1f ms" % h.last_kernel_ms, flush=True)
a, b = out[:20000].cpu().numpy(), exact[:20000].cpu().numpy()
rec = np.mean([len(set(a[i].tolist()) & set(b[i].tolist())) / K for i in range(len(a))])
print("recall@%d of the NN-descent lists vs exact (first 20000 rows): %.4f" % (K, rec), flush=True)
if cpu_n:
    import refutil as R, tempfile
    sub = np.ascontiguousarray(base[:cpu_n])
    t0 = time.time()
    ref = R.ref_nndescent(sub, K, L, it, S, Rr, tempfile.mkdtemp() + "/g", seed=1, threads=0, lib="efanna_fast")
    t = time.time() - t0
    sx = B.Index(h, sub, 0)
    ex = sx.knn_graph(K)
    rr = np.mean([len(set(ref[i].tolist()) & set(ex[i].tolist())) / K for i in range(0, cpu_n, max(1, cpu_n // 5000))])
    g, _ = sx.nndescent(K, L, it, S, Rr, seed=1)
    rg = np.mean([len(set(g[i].tolist()) & set(ex[i].tolist())) / K for i in range(0, cpu_n, max(1, cpu_n // 5000))])
    print("reference IndexGraph::Build on %d rows, %d threads: %.2f s, recall %.4f; GPU same rows: %.1f ms recall %.4f"
          % (cpu_n, os.cpu_count(), t, rr, h.last_kernel_ms, rg), flush=True)

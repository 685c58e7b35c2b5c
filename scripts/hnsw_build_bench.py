"""HNSW build + search timing on the GPU: python scripts/hnsw_build_bench.py n dim M efc [metric] [cpu_sample_n]
Builds with b200ann_hnsw_build, searches 10k queries at several ef, reports recall@10 against brute force; optionally
times the reference's addPoint loop (oracle/_ref, all host cores) on the first cpu_sample_n rows."""
import importlib, sys, time, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
B = importlib.import_module("hnsw-nsg_b200")
synth = importlib.import_module("hnsw-nsg_b200.synth")
import refutil as R
n, dim, M, efc = (int(x) for x in sys.argv[1:5])
metric = int(sys.argv[5]) if len(sys.argv) > 5 else 0
cpu_n = int(sys.argv[6]) if len(sys.argv) > 6 else 0
h = B.Handle(0)
if metric == 0:
    base = synth.sift_shape(n, dim, 1); q = synth.sift_shape(10000, dim, 2)
else:
    base = R.data_glove_like(n, dim, 1); q = R.data_glove_like(10000, dim, 2)
ix = B.Index(h, base, metric)
for r in range(2):
    t0 = time.time()
    info = ix.hnsw_build(M, efc)
    print("hnsw_build rep %d: wall %.2f s device %.1f ms info %s" % (r, time.time() - t0, h.last_kernel_ms, info), flush=True)
truth, _ = ix.bf_knn(q, 10)
for ef in (10, 20, 40, 80, 160):
    lab, d = ix.hnsw_search(q, 10, ef)
    t0 = time.time(); lab, d = ix.hnsw_search(q, 10, ef); t = time.time() - t0
    print("ef %d: recall@10 %.4f  %.0f QPS (kernel %.2f ms)" % (ef, R.recall(lab.astype(np.uint32), truth), len(q) / t, h.last_kernel_ms), flush=True)
if cpu_n:
    sub = np.ascontiguousarray(base[:cpu_n])
    ref = R.RefHnsw(metric, dim, cpu_n, M, efc, lib="hnsw_fast")
    t0 = time.time(); ref.add(sub, label0=0, threads=0); t = time.time() - t0
    sx = B.Index(h, sub, metric)
    t0 = time.time(); sx.hnsw_build(M, efc); tg = time.time() - t0
    print("reference addPoint loop on %d rows, %d threads: %.2f s; GPU same rows: %.2f s" % (cpu_n, os.cpu_count(), t, tg), flush=True)

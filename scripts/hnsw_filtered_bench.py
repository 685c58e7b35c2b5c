"""Filtered HNSW search timing: python scripts/hnsw_filtered_bench.py n dim M efc ef [cpu_n]
GPU b200ann_hnsw_search_filtered at several filter selectivities on a GPU-built index (10k queries, recall@10 against
the exact scan over the allowed rows); with cpu_n: the unmodified reference (oracle/_ref, all host cores,
searchKnn(query, k, isIdAllowed)) on an index of cpu_n rows, next to the GPU on the same rows. Writes one JSON line."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
B = importlib.import_module("hnsw-nsg_b200")
synth = importlib.import_module("hnsw-nsg_b200.synth")
import refutil as R
n, dim, M, efc, ef = (int(x) for x in sys.argv[1:6])
cpu_n = int(sys.argv[6]) if len(sys.argv) > 6 else 0
h = B.Handle(0)
base = synth.sift_shape(n, dim, 1)
q = synth.sift_shape(10000, dim, 2)
ix = B.Index(h, base, B.L2)
ix.hnsw_build(M, efc)
out = dict(n=n, dim=dim, M=M, efc=efc, ef=ef, nq=len(q), rows=[])
rng = np.random.default_rng(0)
lab, d = ix.hnsw_search(q, 10, ef)
t0 = time.time(); lab, d = ix.hnsw_search(q, 10, ef); t = time.time() - t0
print("unfiltered: %.0f QPS (kernel %.2f ms)" % (len(q) / t, h.last_kernel_ms), flush=True)
out["unfiltered_qps"] = len(q) / t
for keep in (0.9, 0.5, 0.1, 0.02):
    allow = rng.random(n) < keep
    rows = np.nonzero(allow)[0]
    sub = B.Index(h, np.ascontiguousarray(base[rows]), B.L2)
    truth = rows[sub.bf_knn(q, 10)[0]]
    sub.close()
    lab, d, bad = ix.hnsw_search_filtered(q, 10, ef, allow)
    t0 = time.time(); lab, d, bad, st = ix.hnsw_search_filtered(q, 10, ef, allow, want_stats=True); t = time.time() - t0
    rec = np.mean([len(set(truth[i].tolist()) & set(lab[i].astype(np.int64).tolist())) for i in range(len(q))]) / 10
    row = dict(keep=keep, qps=len(q) / t, kernel_ms=h.last_kernel_ms, recall=float(rec), inexact=int(bad.sum()),
               evals_per_query=float(st[:, 0].mean()), hops_per_query=float(st[:, 1].mean()))
    print(row, flush=True)
    out["rows"].append(row)
if cpu_n:
    subb = np.ascontiguousarray(base[:cpu_n])
    ref = R.RefHnsw(0, dim, cpu_n, M, efc, lib="hnsw_fast")
    ref.add(subb, threads=0)
    sx = B.Index(h, subb, B.L2)
    sx.hnsw_build(M, efc)
    out["cpu"] = []
    for keep in (0.5, 0.1):
        allow = rng.random(cpu_n) < keep
        a8 = allow.astype(np.uint8)
        ref.search_filtered(q[:1000], 10, ef, a8, threads=0)
        t0 = time.time(); ref.search_filtered(q, 10, ef, a8, threads=0); tc = time.time() - t0
        sx.hnsw_search_filtered(q, 10, ef, allow)
        t0 = time.time(); sx.hnsw_search_filtered(q, 10, ef, allow); tg = time.time() - t0
        row = dict(keep=keep, n=cpu_n, cpu_threads=os.cpu_count(), cpu_qps=len(q) / tc, gpu_qps=len(q) / tg)
        print(row, flush=True)
        out["cpu"].append(row)
print(json.dumps(out))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", "hnsw_filtered_bench.json"), "w").write(json.dumps(out) + "\n")

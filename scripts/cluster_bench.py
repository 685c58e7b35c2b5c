"""CNNS cluster pipeline on one B200: python scripts/cluster_bench.py n n_clusters nprobe [dim]
k-means (GPU assignment), per-cluster exact kNN + NSG, routed batched search; prints build seconds, recall@10 and QPS
(whole pipeline, host routing logic included) with and without the reference's early-stop rule."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("hnsw-nsg_b200")
cl = importlib.import_module("hnsw-nsg_b200.cluster")
synth = importlib.import_module("hnsw-nsg_b200.synth")
n, nc, nprobe = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dim = int(sys.argv[4]) if len(sys.argv) > 4 else 128
h = B.Handle(0)
base = synth.sift_shape(n, dim, 1); q = synth.sift_shape(10000, dim, 2)
t0 = time.time(); assign, cent = cl.kmeans_assign(h, base, nc, iters=6, seed=1); tk = time.time() - t0
t0 = time.time(); ci = cl.ClusterIndex(h).build(base, assign, cent, m=8, knn_k=64, nsg=(40, 32, 300), seed=1); tb = time.time() - t0
ix = B.Index(h, base, 0)
truth, _ = ix.bf_knn(q, 10)
def rec(f):
    return float(np.mean([len(set(f[i].tolist()) & set(truth[i].tolist())) / 10 for i in range(len(truth))]))
out = dict(n=n, dim=dim, n_clusters=nc, nprobe=nprobe, kmeans_s=tk, build_s=tb, cluster_sizes=[int(ci.cl_size.min()), int(ci.cl_size.max())])
for early in (True, False):
    ci.search(q, 10, nprobe, 10, 40, seed=1, early_stop=early)
    t0 = time.time(); gi, gd, st = ci.search(q, 10, nprobe, 10, 40, seed=1, early_stop=early, return_stats=True); t = time.time() - t0
    out["early_stop" if early else "all_probed"] = dict(recall_at_10=round(rec(gi), 4), qps=len(q) / t, seconds=t, **st)
    # the same loop as one C-ABI call with no host round trip (b200ann_cluster_search)
    ci.search_device(q, 10, nprobe, 10, 40, seed=1, early_stop=early)
    t0 = time.time(); di, dd, dst = ci.search_device(q, 10, nprobe, 10, 40, seed=1, early_stop=early, return_stats=True); t = time.time() - t0
    out[("early_stop" if early else "all_probed") + "_device"] = dict(recall_at_10=round(rec(di), 4), qps=len(q) / t, seconds=t,
                                                                      device_ms=h.last_kernel_ms, same_ids=bool((di == gi).all()), **dst)
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", "cluster_pipeline.json"), "w").write(json.dumps(out) + "\n")

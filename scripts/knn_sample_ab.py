import importlib, os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import torch
B = importlib.import_module("hnsw-nsg_b200")
spec = importlib.util.spec_from_file_location("synth", "/root/repo/hnsw-nsg_b200/synth.py")
S = importlib.util.module_from_spec(spec); spec.loader.exec_module(S)
h = B.Handle(0)
n = int(os.environ.get("N", 1000000)); K = int(os.environ.get("K", 200))
base = S.sift_shape(n, 128, seed=1000)
ix = B.Index(h, base, B.L2)
knn = torch.empty((n, K), dtype=torch.int32, device="cuda")
ref = None
for step in os.environ.get("STEPS", "0,16,8,32,16").split(","):
    os.environ["B200ANN_KNN_SAMPLE"] = step
    ms = []
    for rep in range(2):
        ix.knn_graph_raw(K, knn); ms.append(h.last_kernel_ms)
    st = h.knn_last_stats
    cs = int(knn[:2000].to(torch.int64).sum())
    if ref is None: ref = cs
    print("sample", step, "ms", [round(x, 1) for x in ms], "fallback", st["fallback_rows"], "reranked/row", st["reranked"] / n, "same", cs == ref, flush=True)

import importlib, os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import torch
B = importlib.import_module("hnsw-nsg_b200")
spec = importlib.util.spec_from_file_location("synth", "/root/repo/hnsw-nsg_b200/synth.py")
S = importlib.util.module_from_spec(spec); spec.loader.exec_module(S)
h = B.Handle(0)
n = 1000000
base = S.sift_shape(n, 128, seed=1000)
ix = B.Index(h, base, B.L2)
out = torch.empty((125000, 200), dtype=torch.int32, device="cuda")
for rep in range(3):
    t = time.time(); ix.knn_graph_rows(125000, 125000, 200, out=out); torch.cuda.synchronize(); dt = time.time() - t
    print("rows 125000 wall %.1f ms device %.1f ms" % (dt * 1e3, h.last_kernel_ms), h.knn_last_stats, flush=True)
for nr in (75776, 49224, 37888):
    o2 = out[:nr]
    ix.knn_graph_rows(0, nr, 200, out=o2); torch.cuda.synchronize()
    print("rows", nr, "device %.1f ms" % h.last_kernel_ms, flush=True)

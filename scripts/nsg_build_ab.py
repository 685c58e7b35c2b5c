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
knn = torch.empty((n, 200), dtype=torch.int32, device="cuda")
ix.knn_graph_raw(200, knn)
if os.environ.get("LINK_BITS"): h.set_link_visited_bits(int(os.environ["LINK_BITS"]))
for rep in range(2):
    t = time.time(); info = ix.nsg_build((knn, 200), 40, 50, 500, seed=1); dt = time.time() - t
    print("nsg_build", round(dt, 3), h.last_kernel_ms, {k: info[k] for k in list(info)[:6]}, flush=True)

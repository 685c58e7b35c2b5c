import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
B = importlib.import_module("hnsw-nsg_b200")
spec = importlib.util.spec_from_file_location("synth", "/root/repo/hnsw-nsg_b200/synth.py")
S = importlib.util.module_from_spec(spec); spec.loader.exec_module(S)
h = B.Handle(0)
base = S.sift_shape(1000000, 128, seed=1000)
ix = B.Index(h, base, B.L2)
for rep in range(2):
    t = time.time(); g, info = ix.nndescent(200, 200, 10, 10, 100, seed=1); dt = time.time() - t
    print(os.environ.get("B200ANN_NND_LOCKFREE", "1"), round(dt, 3), h.last_kernel_ms, info, int(g[:1000].astype(np.uint64).sum()), flush=True)

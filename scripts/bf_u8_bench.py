"""u8 brute-force kNN (b200ann_bf_knn_u8) timing: python scripts/bf_u8_bench.py [n] [dim] [nq] [k] [reps]
Device-resident inputs, CUDA-event time of the call's kernels (h.last_kernel_ms), and the unmodified
reference (oracle/_ref, BruteforceSearch<int> over L2SpaceI, all host cores) on a bounded query sample."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
B = importlib.import_module("hnsw-nsg_b200")
import torch
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 128
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 10_000
k = int(sys.argv[4]) if len(sys.argv) > 4 else 10
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
rng = np.random.default_rng(1)
base = rng.integers(0, 256, (n, dim), dtype=np.uint8)
q = rng.integers(0, 256, (nq, dim), dtype=np.uint8)
h = B.Handle(0)
dev = torch.device("cuda:0")
tb, tq = torch.from_numpy(base).to(dev), torch.from_numpy(q).to(dev)
ti = torch.empty((nq, k), dtype=torch.int32, device=dev)
td = torch.empty((nq, k), dtype=torch.int32, device=dev)
ms = []
for r in range(reps + 2):
    l0 = h.launch_count
    h.bf_knn_u8_raw(tb, n, dim, dim, tq, nq, dim, k, ti, td)
    if r >= 2:
        ms.append(h.last_kernel_ms)
launches = h.launch_count - l0
best = float(np.median(ms))
t0 = time.time(); gi, gd = h.bf_knn_u8(base, q, k); e2e = time.time() - t0
out = {"workload": "u8 brute force %dx%d, %d queries, k=%d" % (n, dim, nq, k), "device_ms": best, "launches_per_call": launches,
       "qps": nq / best * 1e3, "byte_pairs_per_s": nq * n * dim / best * 1e3, "idp4a_per_s": nq * n * dim / 4 / best * 1e3,
       "e2e_host_buffers_s": e2e, "base_bytes": n * dim}
try:
    import refutil as R
    if R.have_ref():
        ns = 64
        t0 = time.time(); lab, rd = R.ref_bf_knn_u8(base, q[:ns], k, lib="hnsw_fast"); dt = time.time() - t0
        out["cpu_reference"] = {"qps": ns / dt, "cores": os.cpu_count(), "sample": "%d queries" % ns,
                                "ids_equal": bool(np.array_equal(lab.astype(np.uint32), gi[:ns])),
                                "dists_equal": bool(np.array_equal(rd, gd[:ns]))}
except Exception as e:  # noqa
    out["cpu_reference"] = {"error": repr(e)}
print(json.dumps(out))

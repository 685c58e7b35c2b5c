import sys, importlib, numpy as np, tempfile
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import refutil as R
B = importlib.import_module("hnsw-nsg_b200")
H = B.Handle(0)
n, dim = 5000, 64
base = R.data_sift_like(n, dim, 1); q = R.data_sift_like(1000, dim, 2)
ix0 = B.Index(H, base, 0); knn = ix0.knn_graph(32)
tmp = tempfile.mkdtemp(); kp, prefix = tmp+"/k.graph", tmp+"/hy"
R.write_knn_graph(kp, knn)
R.ref_hybrid_build(base, kp, 32, 20, 200, 6, 60, prefix, seed=7, threads=1)
Hh = R.parse_hnsw(prefix + "_hnsw.bin"); g, ep, width = R.parse_nsg(prefix + "_nsg.bin", n)
ix = B.Index(H, base, 0); ix.set_graph(g, ep)
ix.set_hnsw(Hh["links0"], Hh["levels"], Hh["upper_ptr"], Hh["upper_links"], Hh["labels"], Hh["ep"], Hh["maxlevel"])
R.oracle().orc_set_gpu_team(B.dist_team(dim))
for bits in (8, 10):
    H.set_visited_bits(bits)
    for K, L in ((10, 16), (20, 40), (10, 80)):
        ids, d, entry, st = ix.hybrid_search(q, K, L, seed=9, want_stats=True)
        oid, od, oentry, ost = R.orc_hybrid_search(base, Hh, g, q, K, L, R.L2_GPU, R.L2_GPU, rng_mode=1, seed=9, want_stats=True)
        bad = np.where(st[:,1] != ost[:,1])[0]
        print(bits, K, L, "ids eq", (ids==oid).all(), "hops differ", len(bad), "dist fewer", (st[:,0] < ost[:,0]).sum(), [(int(b), int(st[b,1]), int(ost[b,1]), int(st[b,0]), int(ost[b,0])) for b in bad[:4]])

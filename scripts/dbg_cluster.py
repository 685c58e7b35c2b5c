import sys, importlib, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import refutil as R
import test_gpu_cluster as T
B = importlib.import_module("hnsw-nsg_b200"); cl = importlib.import_module("hnsw-nsg_b200.cluster")
H = B.Handle(0)
n, dim, nc = 12000, 32, 12
base = R.data_glove_like(n, dim, 1, n_clusters=24); q = R.data_glove_like(300, dim, 2, n_clusters=24)
assign, cent = cl.kmeans_assign(H, base, nc, iters=5, seed=3)
ci = cl.ClusterIndex(H).build(base, assign, cent, m=4, knn_k=24, nsg=(30, 20, 120), seed=1)
k,nprobe,sK,L=10,8,10,30
pid_g,_ = ci.rix.bf_knn(q, nprobe); pid_o,_ = R.orc_bf_knn(ci.route_points, q, nprobe, R.L2_SEQ_FMA)
print("routing equal", (pid_g==pid_o).all(), (pid_g==pid_o).mean())
gi, gd = ci.search(q, k, nprobe, sK, L, seed=5, early_stop=False)
oi, od = T._oracle_search(B, ci, base, q, k, nprobe, sK, L, 5, False)
bad = np.where((gi!=oi).any(1))[0]
print("no-early-stop mismatching queries", len(bad))
for i in bad[:3]:
    print(i, gi[i], oi[i], gd[i], od[i], "sets equal", set(gi[i].tolist())==set(oi[i].tolist()))
# single cluster check
order = ci.route(q, nprobe)
c = order[:,0]
ids, d = ci.ix.nsg_search_clustered(q, ci.eps[c], ci.cl_base[c], ci.cl_size[c], sK, L, seed=5)
R.oracle().orc_set_gpu_team(B.dist_team(dim))
rows = np.ascontiguousarray(base[ci.perm.astype(np.int64)])
nb=0
for i in range(len(q)):
    b0, sz = int(ci.cl_base[c[i]]), int(ci.cl_size[c[i]])
    g = ci.graph[b0:b0+sz]; gl = np.where(g==R.INVALID, g, g-np.uint32(b0)).astype(np.uint32)
    oi1, od1 = R.orc_nsg_search_qidx(rows[b0:b0+sz], gl, int(ci.eps[c[i]])-b0, q[i:i+1], [i], sK, min(L,sz), R.L2_GPU, seed=5)
    if not ((oi1[0]+b0)==ids[i]).all():
        nb+=1
        if nb<3: print("cluster search mismatch q",i, ids[i]-b0, oi1[0], d[i], od1[0])
print("first-cluster mismatches", nb)

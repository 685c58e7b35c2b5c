"""GPU debug driver for the tensor-core kNN path: tensor (mode 2) vs SIMT (mode 1) on growing sizes."""
import importlib, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
B = importlib.import_module("hnsw-nsg_b200")
import refutil as R
h = B.Handle(0)

def run(name, base, q, k, metric=0, graph=False):
    ix = B.Index(h, base, metric)
    h.set_knn_mode(1)
    t0 = time.time()
    if graph: i1 = ix.knn_graph(k); d1 = None
    else: i1, d1 = ix.bf_knn(q, k)
    t1 = time.time() - t0; ms1 = h.last_kernel_ms
    h.set_knn_mode(2)
    t0 = time.time()
    if graph: i2 = ix.knn_graph(k); d2 = None
    else: i2, d2 = ix.bf_knn(q, k)
    t2 = time.time() - t0; ms2 = h.last_kernel_ms
    st = h.knn_last_stats
    t0 = time.time()
    if graph: i2 = ix.knn_graph(k)
    else: i2, d2 = ix.bf_knn(q, k)
    ms3 = h.last_kernel_ms
    ok_i = (i1 == i2).all(); ok_d = True if d1 is None else (d1 == d2).all()
    print("%-28s ids %s dist %s  simt %.1f ms  tc %.1f ms (2nd %.1f ms)  stats %s" % (name, ok_i, ok_d, ms1, ms2, ms3, st), flush=True)
    if not ok_i:
        bad = np.argwhere(i1 != i2)
        print("   mismatches: %d of %d; first %s" % (len(bad), i1.size, bad[:5].tolist()))
        r = bad[0][0]
        print("   simt row", i1[r][:12], "\n   tc   row", i2[r][:12])
    h.set_knn_mode(0)
    ix.close()
    return ok_i and ok_d

ok = True
ok &= run("sift 4096x128 q256 k10", R.data_sift_like(4096, 128, 1), R.data_sift_like(256, 128, 2), 10)
ok &= run("sift 5000x128 q200 k10", R.data_sift_like(5000, 128, 1), R.data_sift_like(200, 128, 2), 10)
ok &= run("unif 5000x128 q200 k10", R.data_uniform(5000, 128, 1), R.data_uniform(200, 128, 2), 10)
ok &= run("glove ip 5000x100 q130 k10", R.data_glove_like(5000, 100, 1), R.data_glove_like(130, 100, 2), 10, 1)
ok &= run("unif 300x960 q17 k10", R.data_uniform(300, 960, 1), R.data_uniform(17, 960, 2), 10)
ok &= run("unif 20000x96 q64 k200", R.data_uniform(20000, 96, 1), R.data_uniform(64, 96, 2), 200)
ok &= run("sift graph 20000x128 K50", R.data_sift_like(20000, 128, 1), None, 50, graph=True)
if ok and len(sys.argv) > 1:
    n = int(sys.argv[1])
    ok &= run("sift graph %dx128 K200" % n, R.data_sift_shape(n, 128, 1), None, 200, graph=True)
print("ALL OK" if ok else "FAILED")

"""BASELINE.json configs C1, C3, C4 and one C5 shard at FULL size on one B200 (C2 is bench.py's workload).
python scripts/run_configs.py [c1 c3 c4 c5] [--small]  -> one JSON line per config (build seconds, the smallest
ef / L_search reaching recall@k >= 0.95 against brute force, QPS resident in HBM, algorithmic GB/s of the search
kernel from its own counters, and the size-independent checks: rows sorted, ids distinct, deterministic)."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("hnsw-nsg_b200")
synth = importlib.import_module("hnsw-nsg_b200.synth")

small = "--small" in sys.argv
which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c1", "c3", "c4", "c5"]
h = B.Handle(0)
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0


def recall(found, truth):
    k = truth.shape[1]
    return float(np.mean([len(set(found[i, :k].tolist()) & set(truth[i].tolist())) / k for i in range(len(truth))]))


def checks(ids, d):
    assert (np.diff(d, axis=1) >= 0).all(), "rows not sorted by distance"
    srt = np.sort(ids, axis=1)
    assert (srt[:, 1:] != srt[:, :-1]).all(), "duplicate ids in a row"


def sweep(search, truth, k, grid, stride_bytes, deg):
    for v in grid:
        ids, d, st = search(v)
        r = recall(ids.astype(np.uint32), truth)
        if r >= 0.95:
            search(v)
            t0 = time.time(); ids2, d2, st = search(v); wall = time.time() - t0
            kms = h.last_kernel_ms
            assert (ids2 == ids).all() and (d2 == d).all(), "search is not deterministic"
            checks(ids.astype(np.int64), d)
            nq = len(truth)
            alg = float(st[:, 0].sum()) * stride_bytes + float(st[:, 1].sum()) * deg * 4
            return dict(param=v, recall=round(r, 4), qps_resident=nq / (kms * 1e-3), kernel_ms=kms, wall_ms=wall * 1e3,
                        dist_evals_per_query=float(st[:, 0].mean()), hops_per_query=float(st[:, 1].mean()),
                        roofline=dict(bound="hbm", achieved=alg / (kms * 1e-3) / 1e9, peak=PEAK, unit="GB/s",
                                      frac=alg / (kms * 1e-3) / 1e9 / PEAK))
    return dict(param=None, recall=round(r, 4))


def run_hnsw(name, base, q, metric, M, efc, k):
    import torch
    ix = B.Index(h, base, metric)
    t0 = time.time(); info = ix.hnsw_build(M, efc); tb = time.time() - t0
    qd = torch.from_numpy(q).cuda()
    truth, _ = ix.bf_knn(q, k)
    def s(ef):
        lab, d, st = ix.hnsw_search(q, k, ef, want_stats=True)
        return lab, d, st
    res = sweep(s, truth, k, [k, 16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512], ix.stride * 4 if hasattr(ix, "stride") else base.shape[1] * 4, 2 * M + 1)
    out = dict(config=name, n=len(base), dim=base.shape[1], metric="ip" if metric else "l2", M=M, efc=efc, k=k,
               hnsw_build_s=tb, hnsw_build=info, search=res)
    ix.close()
    return out


def run_nsg(name, base, q, K, L, R, C, k):
    import torch
    n = len(base)
    ix = B.Index(h, base, B.L2)
    knn = torch.empty((n, K), dtype=torch.int32, device="cuda")
    t0 = time.time(); ix.knn_graph_raw(K, knn); tk = time.time() - t0
    kst = h.knn_last_stats
    t0 = time.time(); bi = ix.nsg_build((knn, K), L, R, C, seed=1); tn = time.time() - t0
    del knn
    truth, _ = ix.bf_knn(q, k)
    g, ep, width = ix.get_graph()
    def s(Ls):
        ids, d, st = ix.nsg_search(q, k, Ls, seed=1, want_stats=True)
        return ids, d, st
    grid = [v for v in (10, 16, 24, 32, 40, 48, 64, 80, 100, 120, 160, 200, 256, 320, 400) if v >= k]
    res = sweep(s, truth, k, grid, (base.shape[1] + 3) // 4 * 16, g.shape[1])
    out = dict(config=name, n=n, dim=base.shape[1], K=K, L=L, R=R, C=C, k=k, knn_graph_s=tk, knn_stats=kst,
               nsg_build_s=tn, nsg=bi, search=res)
    ix.close()
    return out


for c in which:
    t0 = time.time()
    if c == "c1":      # hnswlib L2 10k x 128 synthetic float, M=16 efC=200, search k=10 ef=100
        rng = np.random.default_rng(1)
        base = rng.random((10000, 128), dtype=np.float32); q = np.random.default_rng(2).random((1000, 128), dtype=np.float32)
        o = run_hnsw("C1 hnswlib L2 10k x 128 M=16 efC=200 k=10", base, q, B.L2, 16, 200, 10)
    elif c == "c3":    # HNSW inner-product on 1.2M x 100 GloVe-shape normalized, M=32, 10k-query batch k=10
        n = 120000 if small else 1200000
        base = synth.glove_shape(n, 100, 1); q = synth.glove_shape(10000, 100, 2)
        o = run_hnsw("C3 HNSW IP %d x 100 GloVe-shape M=32 k=10" % n, base, q, B.IP, 32, 200, 10)
    elif c == "c4":    # NSG on 1M x 960 GIST-shape, k=100
        n = 100000 if small else 1000000
        base = synth.gist_shape(n, 960, 1); q = synth.gist_shape(1000, 960, 2)
        o = run_nsg("C4 NSG %d x 960 GIST-shape k=100" % n, base, q, 200, 40, 50, 500, 100)
    elif c == "c5":    # one shard of the hybrid hnsw_nsg 10M x 96 Deep-shape config: 1.25M rows per GPU
        n = 125000 if small else 1250000
        base = synth.deep_shape(n, 96, 1); q = synth.deep_shape(10000, 96, 2)
        o = run_nsg("C5 shard: NSG %d x 96 Deep-shape k=10" % n, base, q, 200, 40, 50, 500, 10)
        # the hybrid entry: HNSW levels >= 1 pick the NSG entry point (HNSW_NSG::searchKnn)
        ix = B.Index(h, base, B.L2)
        import torch
        knn = torch.empty((n, 200), dtype=torch.int32, device="cuda")
        ix.knn_graph_raw(200, knn); ix.nsg_build((knn, 200), 40, 50, 500, seed=1); del knn
        t1 = time.time(); ix.hnsw_build(16, 200); o["hybrid_hnsw_build_s"] = time.time() - t1
        truth, _ = ix.bf_knn(q, 10)
        for L in (16, 24, 32, 40, 48, 64, 80, 100):
            ids, d, entry, st = ix.hybrid_search(q, 10, L, seed=1, want_stats=True)
            r = recall(ids, truth)
            if r >= 0.95:
                ix.hybrid_search(q, 10, L, seed=1)
                o["hybrid"] = dict(L=L, recall=round(r, 4), qps_resident=len(q) / (h.last_kernel_ms * 1e-3), kernel_ms=h.last_kernel_ms)
                break
        ix.close()
    o["total_s"] = time.time() - t0
    print(json.dumps(o), flush=True)

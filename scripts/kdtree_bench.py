#!/usr/bin/env python
"""KD-tree initializer + RefineGraph against NN-descent from a random start, 1M x 128 SIFT-shape rows, K = 200
(efanna_graph/tests/test_kdtree_graph.cpp + RefineGraph vs test_nndescent.cpp). Graph recall against the exact kNN graph
on the first 2000 rows; device seconds of each call. Writes gpurun_out/kdtree_r2.json."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    n = int(os.environ.get("N", 1000000))
    B = importlib.import_module("hnsw-nsg_b200")
    spec = importlib.util.spec_from_file_location("synth", os.path.join(ROOT, "hnsw-nsg_b200", "synth.py"))
    S = importlib.util.module_from_spec(spec); spec.loader.exec_module(S)
    h = B.Handle(0)
    base = S.sift_shape(n, 128, seed=1000)
    ix = B.Index(h, base, B.L2)
    K = 200
    truth, _ = ix.bf_knn(base[:2000], K + 1)
    truth = truth[:, 1:]
    rec = lambda g: float(np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(g[:2000], truth)]) / K)
    out = {"n": n, "K": K, "runs": []}

    def run(tag, f):
        t = time.time(); g, info = f(); dt = time.time() - t
        r = dict(tag=tag, wall_s=dt, device_ms=h.last_kernel_ms, recall=rec(g), info={k: int(v) for k, v in info.items()})
        print(json.dumps(r), flush=True)
        out["runs"].append(r)
        return g

    for it in (4, 10):
        run("nndescent random start iter=%d" % it, lambda: ix.nndescent(K, 200, it, 10, 100, seed=1))
    for nt, ml in ((4, 8), (8, 8), (16, 8)):
        g0 = run("kdtree nTrees=%d mLevel=%d" % (nt, ml), lambda: ix.kdtree_graph(K, nt, ml, seed=1))
        for it in (2, 4):
            run("  + refine iter=%d" % it, lambda: ix.nndescent_refine(g0, K, 200, it, 10, 100, seed=1))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/kdtree_r2.json", "w"), indent=1)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""HNSW build: batch schedule ratio (rows already inserted per new row of a batch) against build time and recall."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    B = importlib.import_module("hnsw-nsg_b200")
    spec = importlib.util.spec_from_file_location("synth", os.path.join(ROOT, "hnsw-nsg_b200", "synth.py"))
    S = importlib.util.module_from_spec(spec); spec.loader.exec_module(S)
    h = B.Handle(0)
    sizes = [int(x) for x in os.environ.get("SIZES", "50000,1000000").split(",")]
    ratios = [int(x) for x in os.environ.get("RATIOS", "64,32,16,8").split(",")]
    base_all = S.sift_shape(max(sizes), 128, seed=1000)
    q = S.sift_shape(2000, 128, seed=2)
    out = []
    for n in sizes:
        base = base_all[:n]
        truth = None
        for ratio in ratios:
            best = 1e9
            for rep in range(2):
                ix = B.Index(h, base, B.L2)
                t = time.time(); info = ix.hnsw_build(16, 200, seed=100, ratio=ratio, bmax=int(os.environ.get("BMAX", "0"))); dt = time.time() - t
                best = min(best, dt)
                if rep == 1:
                    if truth is None: truth, _ = ix.bf_knn(q, 10)
                    rec = {}
                    for ef in (10, 20, 100):
                        lab, _ = ix.hnsw_search(q, 10, ef)
                        rec[ef] = float(np.mean([len(set(a.tolist()) & set(b.astype(np.int64).tolist())) for a, b in zip(truth, lab)]) / 10)
                ix.close()
            r = dict(n=n, ratio=ratio, build_s=round(best, 4), batches=info["batches"], recall=rec)
            print(json.dumps(r), flush=True)
            out.append(r)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/hnsw_ratio_ab.json", "w"), indent=1)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Recall of the GPU-built HNSW graph at several batch ratios against the graph the unmodified reference builds
(oracle/_ref, one thread) on the same rows, searched by the same GPU kernel at matched ef."""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import refutil as R


def main():
    B = importlib.import_module("hnsw-nsg_b200")
    H = B.Handle(0)
    M, efc = 16, 200
    out = []
    cases = [("uniform L2", 0, 10000, 128, R.data_uniform), ("glove-like IP", 1, 10000, 128, R.data_glove_like),
             ("sift-like L2", 0, 50000, 128, R.data_sift_like)]
    for name, metric, n, dim, gen in cases:
        base, q = gen(n, dim, 1), gen(1000, dim, 2)
        ref = R.RefHnsw(metric, dim, n, M, efc)
        ref.add(base, label0=0, threads=1)
        path = "/tmp/h_%d.bin" % n
        ref.save(path)
        Hr = R.parse_hnsw(path)
        rix = B.Index(H, base, metric)
        rix.set_hnsw(Hr["links0"], Hr["levels"], Hr["upper_ptr"], Hr["upper_links"], Hr["labels"], Hr["ep"], Hr["maxlevel"])
        truth, _ = rix.bf_knn(q, 10)
        efs = (10, 20, 50, 100)
        row = {"case": name, "n": n, "reference": [round(R.recall(rix.hnsw_search(q, 10, ef)[0].astype(np.uint32), truth), 4) for ef in efs]}
        for ratio in [int(x) for x in os.environ.get("RATIOS", "64,32,16,8").split(",")]:
            gix = B.Index(H, base, metric)
            import time as _t; _t0 = _t.time(); gix.hnsw_build(M, efc, seed=100, ratio=ratio); row["build_s %d" % ratio] = round(_t.time() - _t0, 4)
            row["ratio %d" % ratio] = [round(R.recall(gix.hnsw_search(q, 10, ef)[0].astype(np.uint32), truth), 4) for ef in efs]
            gix.close()
        rix.close()
        print(json.dumps(row), flush=True)
        out.append(row)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/hnsw_ratio_vs_ref.json", "w"), indent=1)


if __name__ == "__main__":
    main()

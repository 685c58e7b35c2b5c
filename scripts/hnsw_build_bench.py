#!/usr/bin/env python
"""HNSW build time (b200ann_hnsw_build, M=16, efC=200, SIFT-shape rows) at several sizes and knob settings."""
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
    base_all = S.sift_shape(max(sizes), 128, seed=1000)
    q = S.sift_shape(2000, 128, seed=2)
    out = []
    for n in sizes:
        base = base_all[:n]
        for pf in (0, 1):
            h.set_search_options(prefetch=pf)
            best = 1e9
            for rep in range(2):
                ix = B.Index(h, base, B.L2)
                t = time.time(); info = ix.hnsw_build(16, 200, seed=100); dt = time.time() - t
                best = min(best, dt)
                if rep == 1:
                    truth, _ = ix.bf_knn(q, 10)
                    lab, _ = ix.hnsw_search(q, 10, 100)
                    rec = float(np.mean([len(set(a.tolist()) & set(b.astype(np.int64).tolist())) for a, b in zip(truth, lab)]) / 10)
                ix.close()
            r = dict(n=n, prefetch=pf, build_s=best, device_ms=h.last_kernel_ms, batches=info["batches"], recall=rec)
            print(json.dumps(r), flush=True)
            out.append(r)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/hnsw_build_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Cost of changing a LIVE 1M x 128 HNSW index (M=16, efC=200) on the GPU: appending rows
(b200ann_hnsw_insert_batch = addPoint of new labels) and moving rows (b200ann_hnsw_update_batch = addPoint of existing
labels -> updatePoint), against the 1.6 s a rebuild costs. Writes gpurun_out/hnsw_live_r2.json."""
import importlib
import json
import os
import sys
import time

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
    extra = S.sift_shape(40000, 128, seed=5)
    q = S.sift_shape(10000, 128, seed=2)
    ix = B.Index(h, base, B.L2)
    t = time.time(); ix.hnsw_build(16, 200, seed=100); t_build = time.time() - t
    out = {"n": n, "dim": 128, "M": 16, "efc": 200, "build_s": t_build, "append": [], "update": []}

    def recall(index, rows):
        truth, _ = index.bf_knn(q, 10)
        lab, _ = index.hnsw_search(q, 10, 100)
        return float(np.mean([len(set(a.tolist()) & set(b.astype(np.int64).tolist())) for a, b in zip(truth, lab)]) / 10)

    out["recall_before"] = recall(ix, base)
    at = 0
    for nb in (1, 100, 1000, 10000):
        rows = extra[at:at + nb]; at += nb   # host rows: the copy is part of what a caller pays
        t = time.time(); ix.hnsw_insert_batch(rows); dt = time.time() - t
        out["append"].append({"rows": nb, "wall_ms": dt * 1e3, "device_ms": h.last_kernel_ms})
        print(json.dumps(out["append"][-1]), flush=True)
    rng = np.random.default_rng(1)
    for nb in (1, 100, 1000, 10000):
        ids = rng.choice(n, nb, replace=False).astype(np.uint32)
        rows = extra[at:at + nb]; at += nb
        t = time.time(); info = ix.hnsw_update_batch(ids, rows); dt = time.time() - t
        out["update"].append({"rows": nb, "wall_ms": dt * 1e3, "device_ms": h.last_kernel_ms, "lists_rewritten": info["lists_rewritten"],
                              "dist_evals": info["dist_evals"]})
        print(json.dumps(out["update"][-1]), flush=True)
    out["recall_after"] = recall(ix, None)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/hnsw_live_r2.json", "w"), indent=1)
    print(json.dumps({k: v for k, v in out.items() if k not in ("append", "update")}))


if __name__ == "__main__":
    main()

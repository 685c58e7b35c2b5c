#!/usr/bin/env python
"""HNSW search over unsigned-byte rows (B200ANN_L2_U8, HierarchicalNSW<int> / L2SpaceI) against the same graph over the
same values stored as floats: 1M x 128 SIFT-shape bytes, M=16, efC=200, 10k queries, k=10. Prints one JSON line and
writes gpurun_out/hnsw_u8_r2.json. Both walks are identical (same labels); only the bytes per hop differ."""
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
    base = np.clip(np.rint(S.sift_shape(n, 128, seed=1000)), 0, 255).astype(np.uint8)
    q = np.clip(np.rint(S.sift_shape(10000, 128, seed=2)), 0, 255).astype(np.uint8)
    ux = B.Index(h, base, B.L2_U8)
    t = time.time(); info = ux.hnsw_build(16, 200, seed=100); tb_u8 = time.time() - t
    G = ux.get_hnsw()
    fx = B.Index(h, base.astype(np.float32), B.L2)
    t = time.time(); fx.hnsw_build(16, 200, seed=100); tb_f32 = time.time() - t
    fx.set_hnsw(G["links0"], G["levels"], G["upper_ptr"], G["upper_links"], G["labels"], G["ep"], G["maxlevel"])
    h.set_search_kernel(version=1)   # byte rows run on the v1 kernel: compare like with like (v2 reported separately)
    truth, _ = h.bf_knn_u8(base, q, 10)
    out = {"n": n, "dim": 128, "M": 16, "efc": 200, "nq": 10000, "build_s_u8": tb_u8, "build_s_f32": tb_f32, "runs": []}
    for ef in (20, 50, 100, 200):
        r = {"ef": ef}
        for name, ix, qq in (("u8", ux, q), ("f32", fx, q.astype(np.float32))):
            lab, d, st = ix.hnsw_search(qq, 10, ef, want_stats=True)
            ms = []
            for _ in range(5):
                ix.hnsw_search(qq, 10, ef)
                ms.append(h.last_kernel_ms)
            r[name + "_ms"] = float(np.median(ms))
            r[name + "_qps"] = 10000 / (r[name + "_ms"] * 1e-3)
            r[name + "_evals"] = float(st[:, 0].mean())
            r[name + "_recall"] = float(np.mean([len(set(a.tolist()) & set(b.astype(np.int64).tolist())) for a, b in zip(truth, lab)]) / 10)
            r[name + "_labels"] = lab
        r["same_labels"] = bool((r.pop("u8_labels") == r.pop("f32_labels")).all())
        bytes_u8, bytes_f32 = r["u8_evals"] * 128, r["f32_evals"] * 512
        r["u8_gather_GBps"] = bytes_u8 * 10000 / (r["u8_ms"] * 1e-3) / 1e9
        r["f32_gather_GBps"] = bytes_f32 * 10000 / (r["f32_ms"] * 1e-3) / 1e9
        print(json.dumps(r), flush=True)
        out["runs"].append(r)
    h.set_search_kernel(version=2)
    for ef in (50, 100):
        fx.hnsw_search(q.astype(np.float32), 10, ef)
        ms = []
        for _ in range(5):
            fx.hnsw_search(q.astype(np.float32), 10, ef)
            ms.append(h.last_kernel_ms)
        out.setdefault("f32_v2", []).append({"ef": ef, "ms": float(np.median(ms))})
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/hnsw_u8_r2.json", "w"), indent=1)
    print(json.dumps({k: v for k, v in out.items() if k != "runs"}))


if __name__ == "__main__":
    main()

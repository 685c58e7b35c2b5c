#!/usr/bin/env python
"""bench.py - the hot path of Cookiecoolkid/hnsw-nsg on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload of `value` at every N = BASELINE.json configs[1] (C2): NSG on 1M x 128 SIFT-shaped synthetic rows (exact kNN
graph K=200 -> NSG L=40 R=50 C=500), 10k-query batches, k=10, metric = QPS at recall@10 >= 0.95.
A step = one pass of the graph search over one 10k-query batch at the smallest L_search whose recall@10 (against exact
brute-force ground truth) is >= 0.95.
  value    device-resident queries/results, steps enqueued on TWO handles (streams) alternately so that the drain of
           one batch overlaps the start of the next (`serial` = the same with one batch in flight)
  e2e      the same calls through the C ABI with pinned HOST buffers, copies inside the timed region, two host threads
           with a handle each (`e2e.serial`: one thread)
  N > 1    one process per GPU (torchrun). BUILD: the kNN graph is partitioned by row blocks (every rank computes the
           rows of its block, one NCCL all-gather), every rank then holds the whole index. SEARCH: the query stream is
           split, every rank searches its own 10k-query batches on its replica, no collective in the data path
           (weak scaling in real queries/s).
  c5       (extra object, every N) BASELINE.json configs[4]: hybrid hnsw_nsg on Deep-shaped 96-d rows in 8 row shards of
           1.25M (10M rows), shards spread over the N GPUs, every query searched on every shard, per-shard top-k lists
           exchanged with ONE NCCL all-gather and merged on the GPU; value = global queries/s (strong scaling: at N=1
           one GPU holds all 8 shards).
--impl reference times the UNMODIFIED reference IndexNSG::Search (oracle/_ref, OpenMP over all host cores) on the same
graph, same queries, same L_search. See DESIGN.md "Measurement".
"""
import argparse
import hashlib
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG", "WARN")
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # NCCL logs to stdout by default: rank 0's stdout is ONE JSON line

N_BASE = int(os.environ.get("B200ANN_BENCH_N", 1000000))
DIM = 128
NQ = int(os.environ.get("B200ANN_BENCH_NQ", 10000))
TOPK = 10
KNN_K = int(os.environ.get("B200ANN_BENCH_KNN", 200))
NSG_L, NSG_R, NSG_C = 40, 50, 500
RECALL_TARGET = 0.95
L_SWEEP = [10, 12, 14, 16, 18, 20, 24, 28, 32, 36, 40, 48, 56, 64, 80, 100, 120, 160, 200, 256]
WORKLOAD = "NSG %dx%d SIFT-shape synthetic, exact kNN K=%d, NSG L=%d R=%d C=%d, %d-query batch k=%d" % (
    N_BASE, DIM, KNN_K, NSG_L, NSG_R, NSG_C, NQ, TOPK)
# C5: hybrid hnsw_nsg on Deep-shaped rows in 8 row shards
C5_ON = int(os.environ.get("B200ANN_BENCH_C5", 1))
C5_SHARDS = 8
C5_ROWS = int(os.environ.get("B200ANN_BENCH_C5_N", 1250000))
C5_DIM = 96
C5_M, C5_EFC = 16, 200
C5_WORKLOAD = "hybrid hnsw_nsg, %d shards x %dx%d Deep-shape synthetic (HNSW M=%d efC=%d levels + NSG from exact kNN K=%d, L=%d R=%d C=%d), %d-query batch k=%d" % (
    C5_SHARDS, C5_ROWS, C5_DIM, C5_M, C5_EFC, KNN_K, NSG_L, NSG_R, NSG_C, NQ, TOPK)


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def recall_at_k(found, truth):
    hit = 0
    for f, t in zip(found, truth):
        hit += len(set(f.tolist()) & set(t.tolist()))
    return hit / float(truth.size)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None
        self.nvml = None
        self.samples = []
        self.stop_flag = False
        self.thread = None

    def _nvml_loop(self):
        # NVML queries take microseconds: a short timed region (tens of ms) still gets dozens of samples,
        # which nvidia-smi -lms 100 (one process start + 100 ms period) cannot deliver
        n = self.nvml
        hd = n.nvmlDeviceGetHandleByIndex(self.index)
        while not self.stop_flag:
            try:
                sm = n.nvmlDeviceGetClockInfo(hd, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(hd, n.NVML_CLOCK_SM)
                rs = n.nvmlDeviceGetCurrentClocksEventReasons(hd) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else n.nvmlDeviceGetCurrentClocksThrottleReasons(hd)
                self.samples.append((sm, mx, rs))
            except Exception:
                break
            time.sleep(0.002)

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._nvml_loop, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            n = self.nvml
            bits = {"hw_slowdown": getattr(n, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                    "hw_thermal_slowdown": getattr(n, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                    "sw_thermal_slowdown": getattr(n, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                    "sw_power_cap": getattr(n, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
            if self.samples:
                reasons = sorted(nm for nm, b in bits.items() if any(r & b for _, _, r in self.samples))
                out = {"sm_mhz": float(np.median([x[0] for x in self.samples])),
                       "sm_max_mhz": float(max(x[1] for x in self.samples)), "reasons": reasons,
                       "samples": len(self.samples), "source": "nvml"}
            return out
        if not self.proc:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 6:
                    continue
                try:
                    sm.append(float(f[0])); mx.append(float(f[1]))
                except ValueError:
                    continue
                for nm, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "samples": len(sm), "source": "nvidia-smi"}
        return out


def synth():
    spec = importlib.util.spec_from_file_location("b200ann_synth", os.path.join(ROOT, "hnsw-nsg_b200", "synth.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def make_data(rank):
    """the C2 base is the same on every rank (seed 1000: every rank holds a replica); rank r searches its own slice of
    the query stream (seed 2 + r; rank 0's batch is the round-1 batch)"""
    S = synth()
    t = time.time()
    base = S.sift_shape(N_BASE, DIM, seed=1000)
    q = S.sift_shape(NQ, DIM, seed=2 + rank)
    log("rank %d: data %dx%d + %d queries in %.1fs" % (rank, N_BASE, DIM, NQ, time.time() - t))
    return base, q


def cache_path():
    """the NSG graph of the C2 workload, keyed by the workload only: written by whichever arm builds it first so that
    --impl reference can run without the product (it never feeds the product arm's build timings)"""
    key = "%s|%d|%d|%d|%d|%d|%d" % (WORKLOAD, N_BASE, KNN_K, NSG_L, NSG_R, NSG_C, NQ)
    return os.environ.get("B200ANN_BENCH_INDEX",
                          os.path.join(tempfile.gettempdir(), "b200ann_bench_c2_%s.npz" % hashlib.md5(key.encode()).hexdigest()[:12]))


def write_nsg(path, graph, ep, width, invalid=0xFFFFFFFF):
    """IndexNSG::Save format (CNNS/src/nsg/index_nsg.cpp:37-49): [u32 width][u32 ep] then per node [u32 k][u32 x k]"""
    n, gs = graph.shape
    valid = graph != invalid
    deg = valid.sum(1).astype(np.uint32)
    out = np.empty(2 + n + int(deg.sum()), np.uint32)
    out[0], out[1] = width, ep
    starts = 2 + np.arange(n, dtype=np.int64) + np.concatenate([[0], np.cumsum(deg[:-1], dtype=np.int64)])
    out[starts] = deg
    cols = np.broadcast_to(np.arange(gs), (n, gs))
    pos = (starts[:, None] + 1 + cols)[valid]
    out[pos] = graph[valid]
    out.tofile(path)


def dist_env():
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    return rank, local, world


def allmax(x, world):
    """max over ranks of a python float"""
    if world == 1:
        return float(x)
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(x)], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def build_c2(B, h, base, rank, world):
    """Upload + exact kNN graph (K=200) + NSG build, timed fresh in this process (no cache is ever read here).
    world > 1: the kNN graph is partitioned by row blocks - every rank computes the rows of its block
    (b200ann_knn_graph_rows), one NCCL all-gather makes the graph whole on every rank - and every rank then builds
    its replica of the NSG. Seconds are the max over ranks."""
    import torch
    import torch.distributed as dist
    info = {}
    torch.cuda.synchronize()
    t = time.time()
    ix = B.Index(h, base, B.L2)
    info["upload_s"] = time.time() - t
    knn = torch.empty((N_BASE, KNN_K), dtype=torch.int32, device="cuda")
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = time.time()
    if world == 1:
        ix.knn_graph_raw(KNN_K, knn)
        kernel_ms = h.last_kernel_ms
        gather_s = 0.0
    else:
        per = (N_BASE + world - 1) // world
        r0, r1 = min(N_BASE, rank * per), min(N_BASE, (rank + 1) * per)
        pad = torch.empty((world * per, KNN_K), dtype=torch.int32, device="cuda")
        ix.knn_graph_rows(r0, r1 - r0, KNN_K, out=pad[rank * per:rank * per + (r1 - r0)])
        kernel_ms = h.last_kernel_ms
        torch.cuda.synchronize()
        tg = time.time()
        dist.all_gather_into_tensor(pad.view(-1), pad[rank * per:(rank + 1) * per].reshape(-1).clone())
        knn.copy_(pad[:N_BASE])
        torch.cuda.synchronize()
        gather_s = time.time() - tg
        del pad
    info["knn_graph_s"] = allmax(time.time() - t, world)
    info["knn_graph_kernel_ms"] = allmax(kernel_ms, world)
    info["knn_graph_allgather_s"] = allmax(gather_s, world)
    st = h.knn_last_stats
    info["knn_graph_stats"] = st
    rows = N_BASE if world == 1 else (N_BASE + world - 1) // world
    flops = 2.0 * rows * N_BASE * DIM   # this rank's block of the n x n contraction
    info["knn_graph_roofline"] = {
        "bound": "tensor", "kernel": "tc::sweep_kernel (tcgen05 bf16)" if st["rows"] else "bf_knn_kernel (fp32 SIMT)",
        "achieved": flops * max(1, st["passes"]) / (kernel_ms * 1e-3) / 1e12, "useful": flops / (kernel_ms * 1e-3) / 1e12,
        "peak": measured_peaks().get("bf16_tflops_sustained"), "unit": "TFLOP/s", "rows_per_gpu": rows}
    if info["knn_graph_roofline"]["peak"]:
        info["knn_graph_roofline"]["frac"] = info["knn_graph_roofline"]["achieved"] / info["knn_graph_roofline"]["peak"]
    if st["rows"]:
        # the sweep's second bound: every 128 x 256 fp32 accumulator tile (128 KB) is read out of TMEM by tcgen05.ld at
        # 64 B/clk per SM (B300_MICROARCH.md "LDTM throughput"), however short the contraction is
        sm_clk = 1.965e9
        try:
            sm_clk = torch.cuda.get_device_properties(0).clock_rate * 1e3
        except Exception:
            pass
        n_sm = torch.cuda.get_device_properties(0).multi_processor_count
        tmem_bytes = float(st["tiles"]) * 128 * 256 * 4
        info["knn_graph_roofline"]["tmem_read"] = {
            "bytes": tmem_bytes, "achieved_B_per_clk_per_sm": tmem_bytes / (kernel_ms * 1e-3) / sm_clk / n_sm,
            "peak_B_per_clk_per_sm": 64.0, "frac": tmem_bytes / (kernel_ms * 1e-3) / sm_clk / n_sm / 64.0,
            "note": "tiles counts the main pass; the sampled pre-pass adds 1/16"}
    log("rank %d: exact kNN graph K=%d in %.2fs (kernel %.0f ms, all-gather %.3fs)" % (
        rank, KNN_K, info["knn_graph_s"], kernel_ms, gather_s))
    torch.cuda.synchronize()
    t = time.time()
    bi = ix.nsg_build((knn, KNN_K), NSG_L, NSG_R, NSG_C, seed=1)
    info["nsg_build_s"] = allmax(time.time() - t, world)
    info["nsg_build_device_ms"] = h.last_kernel_ms
    info["nsg"] = bi
    info["build_s"] = info["knn_graph_s"] + info["nsg_build_s"]
    log("rank %d: NSG build in %.2fs %s" % (rank, info["nsg_build_s"], bi))
    return ix, knn, info


def sweep_L(search, gt, what):
    """smallest L of L_SWEEP whose recall@TOPK against gt reaches the target"""
    sweep = []
    for L in L_SWEEP:
        if L < TOPK:
            continue
        r = recall_at_k(search(L), gt)
        sweep.append((L, round(r, 4)))
        if r >= RECALL_TARGET:
            return L, r, sweep
    raise RuntimeError("%s: recall@%d >= %.2f not reached in sweep %s" % (what, TOPK, RECALL_TARGET, sweep))


def other_builds(B, h, ix, base, knn_exact, q, gt, info):
    """rank 0, N = 1: the other construction paths of SURVEY 8(a) on the same rows, reported under "build":
    NN-descent (efanna IndexGraph::Build; C2 as BASELINE.json writes it - "efanna kNN K=200" - with the efanna README's
    SIFT1M parameters L=200 iter=10 S=10 R=100), the NSG built FROM that graph with its own L_search / recall, and the
    HNSW addPoint loop (M=16 efC=200). Each also gets the UNMODIFIED reference timed on a bounded row sample."""
    import torch
    n = base.shape[0]
    K, L, IT, S, R_ = KNN_K, 200, 10, 10, 100
    out = torch.empty((n, K), dtype=torch.int32, device="cuda")
    t = time.time()
    _, ninfo = ix.nndescent(K, L, IT, S, R_, seed=1, out=out)
    info["nndescent_s"] = time.time() - t
    info["nndescent_device_ms"] = h.last_kernel_ms
    a, b = out[:20000].cpu().numpy(), knn_exact[:20000].cpu().numpy()
    rec = float(np.mean([len(set(a[i].tolist()) & set(b[i].tolist())) / K for i in range(0, 20000, 10)]))
    info["nndescent"] = dict(ninfo, K=K, L=L, iter=IT, S=S, R=R_, recall_vs_exact=round(rec, 4),
                             join_gather_GBps=ninfo["dist_evals"] * 2 * DIM * 4 / max(1e-9, ninfo["join_us"] * 1e-6) / 1e9)
    log("NN-descent K=%d in %.2fs (%s)" % (K, info["nndescent_s"], info["nndescent"]))
    # C2 as written: NSG from the NN-descent graph, on a second index over the same rows
    try:
        ix2 = B.Index(h, base, B.L2)
        t = time.time()
        ix2.nsg_build((out, K), NSG_L, NSG_R, NSG_C, seed=1)
        t_nsg = time.time() - t
        L2, r2, sw2 = sweep_L(lambda LL: ix2.nsg_search(q, TOPK, LL, seed=1)[0], gt, "NSG from NN-descent graph")
        ms = []
        for _ in range(5):
            ix2.nsg_search(q, TOPK, L2, seed=1)
            ms.append(h.last_kernel_ms)
        info["nsg_from_nndescent"] = {"nsg_build_s": t_nsg, "L_search": L2, "recall_at_10": round(r2, 4),
                                      "kernel_qps": len(q) / (float(np.median(ms)) * 1e-3), "sweep": sw2}
        log("NSG from the NN-descent graph: %s" % info["nsg_from_nndescent"])
        ix2.close()
    except Exception as e:
        log("NSG from NN-descent graph skipped: %r" % (e,))
    del out
    t = time.time()
    hinfo = ix.hnsw_build(16, 200, seed=100)
    info["hnsw_build_s"] = time.time() - t
    info["hnsw_build_device_ms"] = h.last_kernel_ms
    bytes_ = (hinfo["beam_evals"] + hinfo["heuristic_evals"]) * DIM * 4 + hinfo["hops"] * 33 * 4
    info["hnsw"] = dict(hinfo, M=16, efc=200, gather_GBps=bytes_ / (h.last_kernel_ms * 1e-3) / 1e9)
    log("HNSW build M=16 efC=200 in %.2fs (%s)" % (info["hnsw_build_s"], info["hnsw"]))
    try:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import refutil as R
        if R.have_ref():
            m = int(os.environ.get("B200ANN_BENCH_CPU_BUILD_ROWS", 100000))
            sub = np.ascontiguousarray(base[:m])
            threads = os.cpu_count() or 1
            t = time.time()
            R.ref_nndescent(sub, K, L, IT, S, R_, os.path.join(tempfile.mkdtemp(), "g"), seed=1, threads=0, lib="efanna_fast")
            t_nd = time.time() - t
            ref = R.RefHnsw(0, DIM, m, 16, 200, lib="hnsw_fast")
            t = time.time()
            ref.add(sub, label0=0, threads=0)
            t_h = time.time() - t
            ref.close()
            sx = B.Index(h, sub, B.L2)
            sx.nndescent(K, L, IT, S, R_, seed=1)
            g_nd = h.last_kernel_ms * 1e-3
            sx.hnsw_build(16, 200)
            g_h = h.last_kernel_ms * 1e-3
            sx.close()
            info["build_cpu_reference"] = {
                "sample": "first %d rows, unmodified reference from oracle/_ref (-O3/-Ofast builds), OpenMP" % m,
                "cores": threads, "rows": m, "nndescent_s": t_nd, "nndescent_gpu_same_rows_s": g_nd,
                "hnsw_addpoint_s": t_h, "hnsw_gpu_same_rows_s": g_h,
                "note": "NN-descent work grows ~linearly with rows at fixed K/L/iter, the addPoint loop ~n log n: "
                        "x%d rows gives the full-size CPU estimate" % (n // m)}
            log("reference builds on %d rows: %s" % (m, info["build_cpu_reference"]))
    except Exception as e:  # the build baselines are extras; the headline never depends on them
        log("build cpu reference skipped: %r" % (e,))


def load_sharded():
    spec = importlib.util.spec_from_file_location("b200ann_sharded", os.path.join(ROOT, "hnsw-nsg_b200", "sharded.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def run_c5(B, h, rank, world, steps, warmup):
    """BASELINE.json configs[4]: hybrid hnsw_nsg over 8 row shards spread over the ranks (strong scaling: the same 8
    shards at every N, so the merged result is the same at every N). Per step every rank runs b200ann_hybrid_search on
    each shard it owns, writing (distance, global id) lists straight into its block of the send buffer (ids made global
    in the kernel epilogue, b200ann_set_id_base), ONE NCCL all-gather exchanges the blocks, b200ann_merge_topk_gathered
    merges. Everything is enqueued on the library's stream; no host synchronisation inside a step.
    Returns the dict reported as "c5" (rank 0) or None."""
    import torch
    import torch.distributed as dist
    S = synth()
    SH = load_sharded()
    own = SH.owned_shards(rank, world, C5_SHARDS)
    lp = len(own)
    nq, K = NQ, TOPK
    q = S.deep_shape(nq, C5_DIM, seed=5)
    shards = []
    tb = {"knn_graph_s": 0.0, "nsg_build_s": 0.0, "hnsw_build_s": 0.0, "data_s": 0.0}
    gts = []
    for s in own:
        t = time.time()
        base = S.deep_shape(C5_ROWS, C5_DIM, seed=3000 + s)
        tb["data_s"] += time.time() - t
        ix = B.Index(h, base, B.L2)
        gi, gd = ix.bf_knn(q, K)
        gts.append((gi.astype(np.int64) + s * C5_ROWS, gd))
        knn = torch.empty((C5_ROWS, KNN_K), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize(); t = time.time()
        ix.knn_graph_raw(KNN_K, knn)
        tb["knn_graph_s"] += time.time() - t
        t = time.time()
        ix.nsg_build((knn, KNN_K), NSG_L, NSG_R, NSG_C, seed=1)
        tb["nsg_build_s"] += time.time() - t
        del knn
        t = time.time()
        ix.hnsw_build(C5_M, C5_EFC, seed=100)
        tb["hnsw_build_s"] += time.time() - t
        shards.append(ix)
        del base
        log("rank %d: C5 shard %d built (cumulative %s)" % (rank, s, {k: round(v, 2) for k, v in tb.items()}))
    torch.cuda.empty_cache()
    build_s = allmax(tb["knn_graph_s"] + tb["nsg_build_s"] + tb["hnsw_build_s"], world)

    # global exact ground truth = merge of the per-shard exact lists (host, untimed)
    gi = np.stack([g[0] for g in gts]); gd = np.stack([g[1] for g in gts])   # [lp, nq, K]
    if world > 1:
        ti = torch.from_numpy(gi).cuda(); td = torch.from_numpy(gd).cuda()
        ai = torch.empty((world,) + ti.shape, dtype=ti.dtype, device="cuda")
        ad = torch.empty((world,) + td.shape, dtype=td.dtype, device="cuda")
        dist.all_gather_into_tensor(ai.view(-1), ti.view(-1)); dist.all_gather_into_tensor(ad.view(-1), td.view(-1))
        gi = ai.view(-1, nq, K).cpu().numpy(); gd = ad.view(-1, nq, K).cpu().numpy()
    alld = np.transpose(gd, (1, 0, 2)).reshape(nq, -1); alli = np.transpose(gi, (1, 0, 2)).reshape(nq, -1)
    order = np.argsort(alld, axis=1, kind="stable")[:, :K]
    truth = np.take_along_axis(alli, order, axis=1)

    ext = torch.cuda.ExternalStream(h.stream, device=torch.cuda.current_device())
    q_dev = torch.from_numpy(q).cuda()
    block, bd, bi = SH.alloc_block(lp, nq, K, "cuda")
    gathered = torch.empty((world, block.numel()), dtype=torch.int32, device="cuda")
    out_i = torch.empty((nq, K), dtype=torch.int32, device="cuda")
    out_d = torch.empty((nq, K), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    h.set_async(True)
    state = {"L": 40, "comm": True}

    def search_only():
        for l, s in enumerate(own):
            h.set_id_base(s * C5_ROWS)
            shards[l].hybrid_search_raw(q_dev, nq, C5_DIM, K, state["L"], 1, bi[l], bd[l], None, None)
        h.set_id_base(0)

    def step():
        search_only()
        if not state["comm"]:
            return
        with torch.cuda.stream(ext):
            g = SH.exchange(block, world, gathered)
        h.merge_topk_gathered(g, world, lp, nq, K, K, out_d, out_i)

    def found(L):
        state["L"] = L
        step()
        torch.cuda.synchronize()
        return out_i.cpu().numpy().view(np.uint32).astype(np.int64)

    L, rec, sweep = sweep_L(found, truth, "C5 hybrid")   # the same on every rank: same shards, same queries
    state["L"] = L
    log("rank %d: C5 L_search=%d recall@%d=%.4f (sweep %s)" % (rank, L, K, rec, sweep))

    def timed(fn):
        for _ in range(warmup):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = h.launch_count
        t0 = time.perf_counter()
        e0.record(ext)
        for _ in range(steps):
            fn()
        e1.record(ext)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        return allmax(e0.elapsed_time(e1), world) / steps, allmax(wall, world) / steps, h.launch_count - l0

    step_ms, wall_ms, launches = timed(step)
    state["comm"] = False
    search_ms, _, _ = timed(search_only)
    state["comm"] = True
    res_i = found(L)
    rec_final = recall_at_k(res_i, truth)
    # work counters of one shard search (device stats) for the roofline of the hybrid kernel on 96-d rows
    st_dev = torch.zeros((nq, 2), dtype=torch.int32, device="cuda")
    h.set_async(False)
    shards[0].hybrid_search_raw(q_dev, nq, C5_DIM, K, L, 1, bi[0], bd[0], None, st_dev)
    kms = []
    for _ in range(3):
        shards[0].hybrid_search_raw(q_dev, nq, C5_DIM, K, L, 1, bi[0], bd[0], None, None)
        kms.append(h.last_kernel_ms)
    st = st_dev.cpu().numpy().astype(np.float64)
    g0, _, _ = shards[0].get_graph()
    gs = (g0.shape[1] + 7) // 8 * 8
    alg = st[:, 0].sum() * C5_DIM * 4 + st[:, 1].sum() * gs * 4 + nq * C5_DIM * 4
    peak, _ = measured_peak_gbs()
    out = None
    if rank == 0:
        out = {"workload": C5_WORKLOAD, "metric": "QPS @ recall@10>=0.95", "value": nq / (step_ms * 1e-3), "unit": "queries/s",
               "scaling": "strong", "n_gpus": world, "shards": C5_SHARDS, "shards_per_gpu": lp, "total_rows": C5_SHARDS * C5_ROWS,
               "L_search": L, "recall_at_10": round(rec_final, 4), "ms_per_step": step_ms, "ms_per_step_wall": wall_ms,
               "search_only_ms": search_ms, "exchange_and_merge_ms": step_ms - search_ms,
               "comm_share": (step_ms - search_ms) / step_ms if step_ms > 0 else None,
               "exchange": "none (one GPU)" if world == 1 else "1 ncclAllGather of %d B per rank + merge kernel" % (block.numel() * 4),
               "gpu_launches": launches, "steps": steps,
               "build_s": build_s, "build": {k: allmax(v, 1) for k, v in tb.items()},
               "shard_kernel": {"kernel": "beam_search2_kernel (hybrid: HNSW levels + SearchFromEnterpoint)", "ms": float(np.mean(kms)),
                                "algorithmic_bytes": alg, "achieved_GBps": alg / (float(np.mean(kms)) * 1e-3) / 1e9,
                                "frac_of_hbm_peak": alg / (float(np.mean(kms)) * 1e-3) / 1e9 / peak,
                                "dist_evals_per_query": st[:, 0].mean(), "hops_per_query": st[:, 1].mean()}}
    for ix in shards:
        ix.close()
    torch.cuda.empty_cache()
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank, local, world = dist_env()
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("--gpus %d needs torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = importlib.import_module("hnsw-nsg_b200")
    h = B.Handle(local)
    h2 = B.Handle(local)     # second stream: a second batch in flight
    base, q = make_data(rank)
    ix, knn, info = build_c2(B, h, base, rank, world)
    gt, _ = ix.bf_knn(q, TOPK)
    L, rec0, sweep = sweep_L(lambda LL: ix.nsg_search(q, TOPK, LL, seed=1)[0], gt, "C2 NSG")
    if world > 1:  # every replica searches with the largest L any rank's batch needs
        t = torch.tensor([L], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        L = int(t.item())
    info["L_search"], info["sweep"] = L, sweep
    log("rank %d: L_search=%d recall@%d=%.4f (sweep %s)" % (rank, L, TOPK, rec0, sweep))
    if rank == 0 and world == 1 and not args.no_extras:
        other_builds(B, h, ix, base, knn, q, gt, info)
    del knn
    torch.cuda.empty_cache()
    if rank == 0:   # the graph as a file for --impl reference (never read by this arm)
        try:
            g, ep, width = ix.get_graph()
            np.savez(cache_path(), graph=g, ep=ep, width=width, L=L)
        except Exception as e:
            log("index cache write failed: %s" % e)

    nq, K = q.shape[0], TOPK
    hs = [h, h2]
    exts = [torch.cuda.ExternalStream(x.stream, device=local) for x in hs]
    q_dev = torch.from_numpy(q).cuda()
    ids_dev = [torch.empty((nq, K), dtype=torch.int32, device="cuda") for _ in hs]
    q_pin = [torch.from_numpy(q).pin_memory() for _ in hs]
    ids_pin = [torch.empty((nq, K), dtype=torch.int32).pin_memory() for _ in hs]
    torch.cuda.synchronize()

    def search(hh, qbuf, out):
        # the C ABI call a user makes; the handle decides the stream
        hh._check(B.lib().b200ann_nsg_search(hh._h, ix._ix, B._ptr(qbuf), nq, DIM, K, L, 1, B._ptr(out), None, None))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_device(n_flight, steps, warmup):
        """K steps, device-resident buffers, n_flight handles used alternately (stream-ordered calls, no host sync)"""
        for x in hs:
            x.set_async(True)
        for i in range(warmup):
            search(hs[i % n_flight], q_dev, ids_dev[i % n_flight])
        barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(n_flight)]
        l0 = sum(x.launch_count for x in hs)
        t0 = time.perf_counter()
        e0.record(exts[0])
        for j in range(1, n_flight):
            exts[j].wait_event(e0)   # nothing of the timed region starts before e0
        for i in range(steps):
            search(hs[i % n_flight], q_dev, ids_dev[i % n_flight])
        for j in range(n_flight):
            ends[j].record(exts[j])
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        dev = max(e0.elapsed_time(e) for e in ends)
        for x in hs:
            x.set_async(False)
        return allmax(dev, world), allmax(wall, world), sum(x.launch_count for x in hs) - l0

    def timed_e2e(n_threads, steps, warmup):
        """K steps through the synchronous C-ABI call with pinned HOST buffers (H2D of the queries, kernel, D2H of the
        ids inside every call), n_threads host threads with a handle each; wall clock around all of them"""
        import threading
        per = [steps // n_threads + (1 if t < steps % n_threads else 0) for t in range(n_threads)]
        for t in range(n_threads):
            for _ in range(max(1, warmup // n_threads)):
                search(hs[t], q_pin[t], ids_pin[t])
        barrier()
        start = threading.Barrier(n_threads + 1)

        def worker(t):
            start.wait()
            for _ in range(per[t]):
                search(hs[t], q_pin[t], ids_pin[t])

        th = [threading.Thread(target=worker, args=(t,)) for t in range(n_threads)]
        for x in th:
            x.start()
        start.wait()
        t0 = time.perf_counter()
        for x in th:
            x.join()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        if world > 1:
            dist.barrier()
        return allmax(wall, world)

    clocks = ClockSampler(local)
    clocks.start()
    dev2_ms, wall2_ms, launches = timed_device(2, args.steps, args.warmup)
    clk = clocks.stop()
    dev1_ms, wall1_ms, _ = timed_device(1, args.steps, args.warmup)
    e2e2_ms = timed_e2e(2, args.steps, args.warmup)
    e2e1_ms = timed_e2e(1, args.steps, args.warmup)
    step_ms, serial_ms = dev2_ms / args.steps, dev1_ms / args.steps

    # ---- correctness of what was timed + roofline accounting (untimed)
    rec = recall_at_k(ids_dev[0].cpu().numpy().view(np.uint32), gt)
    rec_b = recall_at_k(ids_dev[1].cpu().numpy().view(np.uint32), gt)
    assert rec == rec_b, "the two in-flight batches disagree"
    assert (ids_pin[0].numpy() == ids_dev[0].cpu().numpy()).all(), "host-buffer results differ from device-buffer results"
    _, _, st = ix.nsg_search(q, K, L, seed=1, want_stats=True)
    h.set_visited_bits(13)    # a visited set that forgets nothing: the distance evaluations the algorithm itself needs
    _, _, st_exact = ix.nsg_search(q, K, L, seed=1, want_stats=True)
    h.set_visited_bits(0)
    g, ep, width = ix.get_graph()
    gs = (g.shape[1] + 7) // 8 * 8
    n_dist = float(st_exact[:, 0].astype(np.float64).sum()); n_hops = float(st_exact[:, 1].astype(np.float64).sum())
    n_dist_run = float(st[:, 0].astype(np.float64).sum())
    alg_bytes = n_dist * DIM * 4 + n_hops * gs * 4 + nq * DIM * 4
    kms = []
    for _ in range(max(5, min(args.steps, 20))):   # the search kernel alone (CUDA events inside the library around its launch)
        ix.nsg_search_raw(q_dev, nq, DIM, K, L, 1, ids_dev[0], None, None)
        kms.append(h.last_kernel_ms)
    kernel_ms = float(np.mean(kms))
    peak, peak_src = measured_peak_gbs()
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "search_kernel_traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    except Exception:
        pass
    probe = None
    try:
        probe = max(ix.probe_gather(8 << 20, infl, 8, 8)[0] for infl in (1, 2) for _ in range(2))
    except Exception as e:
        log("probe_gather skipped: %r" % (e,))

    line = None
    if rank == 0:
        line = {
            "metric": "QPS @ recall@10>=0.95", "value": world * nq / (step_ms * 1e-3), "unit": "queries/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "L_search": L,
                       "parallelism": "1 GPU" if world == 1 else "%d replicas (kNN graph built by row blocks + NCCL all-gather), query stream split, no collective in the search path" % world,
                       "batches_in_flight": 2,
                       "l2": "inputs larger than L2 (vectors %d MB + graph %d MB per GPU; no flush)" % (
                           N_BASE * DIM * 4 >> 20, N_BASE * gs * 4 >> 20)},
            "recall_at_10": round(rec, 4), "rows_per_gpu": N_BASE, "queries_per_step": world * nq,
            "serial": {"value": world * nq / (serial_ms * 1e-3), "unit": "queries/s", "ms_per_step": serial_ms,
                       "what": "one batch in flight: every step's kernel starts after the previous one has drained"},
            "wall_ms_per_step": wall2_ms / args.steps,
            "e2e": {"value": world * nq / (e2e2_ms / args.steps * 1e-3), "unit": "queries/s", "ms_per_step": e2e2_ms / args.steps,
                    "h2d_bytes_per_step": nq * DIM * 4, "d2h_bytes_per_step": nq * K * 4,
                    "how": "synchronous b200ann_nsg_search calls with pinned host buffers from 2 host threads, a handle each",
                    "serial": {"value": world * nq / (e2e1_ms / args.steps * 1e-3), "ms_per_step": e2e1_ms / args.steps,
                               "how": "the same from 1 host thread"}},
            "gpu_launches": launches,
            "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "beam_search2_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "dram_frac": (traffic / (kernel_ms * 1e-3) / 1e9 / peak) if traffic else None,
                         "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes,
                         "dist_evals_exact_per_query": n_dist / nq, "dist_evals_run_per_query": n_dist_run / nq,
                         "hops_per_query": n_hops / nq,
                         "random_gather_probe_GBps": probe,
                         "pipelined_frac": alg_bytes / (step_ms * 1e-3) / 1e9 / peak},
            "build": {k: info.get(k) for k in ("build_s", "upload_s", "knn_graph_s", "knn_graph_kernel_ms", "knn_graph_allgather_s",
                                               "knn_graph_stats", "knn_graph_roofline", "nsg_build_s", "nsg_build_device_ms",
                                               "nsg", "nndescent_s", "nndescent_device_ms", "nndescent", "nsg_from_nndescent",
                                               "hnsw_build_s", "hnsw_build_device_ms", "hnsw", "build_cpu_reference", "sweep")},
        }
        line["build"]["from_cache"] = False
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        qps, rids, secs = cpu_reference_qps(base, q, g, ep, width, L, threads)
        line["cpu_baseline"] = {"value": qps, "unit": "queries/s", "cores": threads, "kind": "reference",
                                "sample": "all %d queries, 1 pass after a 2000-query warm-up (%.2f s), same graph and L_search; "
                                          "IndexNSG::Search from oracle/_ref/libref_nsg_fast, OpenMP" % (nq, secs),
                                "recall_at_10": round(recall_at_k(rids, gt), 4),
                                "ids_equal_gpu": float(np.mean(rids == ids_dev[0].cpu().numpy().view(np.uint32)))}
    ix.close()
    del base
    if C5_ON:
        try:
            c5 = run_c5(B, h, rank, world, max(3, args.steps // 5), args.warmup)
            if rank == 0:
                line["c5"] = c5
        except Exception as e:
            log("C5 skipped: %r" % (e,))
            if rank == 0:
                line["c5"] = {"error": repr(e)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def ref_nsg(base, graph, ep, width):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refutil as R
    fd, path = tempfile.mkstemp(suffix=".nsg")
    os.close(fd)
    write_nsg(path, graph, ep, width)
    ref = R.RefNsg(base, path, lib="nsg_fast")
    os.unlink(path)
    return ref


def cpu_reference_qps(base, q, graph, ep, width, L, threads, passes=1):
    """UNMODIFIED reference IndexNSG::Search (oracle/_ref/libref_nsg_fast*.so built from
    /root/reference/CNNS/src/nsg/index_nsg.cpp), OpenMP over `threads` host cores, on the same graph.
    Returns (qps, ids, seconds)."""
    ref = ref_nsg(base, graph, ep, width)
    ref.search(q[:min(len(q), 2000)], TOPK, L, seed=1, threads=threads)  # warm
    best, ids = None, None
    for _ in range(passes):
        t = time.perf_counter()
        ids = ref.search(q, TOPK, L, seed=1, threads=threads)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return len(q) / best, ids, best


def run_reference(args):
    """the reference's own CPU implementation of the timed path (IndexNSG::Search, unmodified sources compiled into
    oracle/_ref) on the host cores: same workload, graph, queries and L_search as the product arm. The index is an INPUT
    of the timed path (the reference's own CPU build at 1M x 128, K=200 takes tens of minutes): it is read from the file
    either arm leaves in /tmp (or B200ANN_BENCH_INDEX), so this arm runs without the product; only when no such file
    exists is the graph built - untimed - by b200ann_nsg_build and handed over in IndexNSG::Load's own format.
    N > 1: rank 0 alone runs (the CPU has no replicas to add); its line carries the N the driver asked for."""
    rank, local, world = dist_env()
    if rank != 0:
        return
    base, q = make_data(0)
    cp = cache_path()
    src = None
    if os.path.exists(cp):
        z = np.load(cp)
        g, ep, width, L = z["graph"], int(z["ep"]), int(z["width"]), int(z["L"])
        src = "read from %s (written by an earlier run of either arm)" % cp
    else:
        import torch
        torch.cuda.set_device(0)
        B = importlib.import_module("hnsw-nsg_b200")
        h = B.Handle(0)
        ix, knn, info = build_c2(B, h, base, 0, 1)
        del knn
        gt, _ = ix.bf_knn(q, TOPK)
        L, _, _ = sweep_L(lambda LL: ix.nsg_search(q, TOPK, LL, seed=1)[0], gt, "C2 NSG")
        g, ep, width = ix.get_graph()
        ix.close(); h.close()
        try:
            np.savez(cp, graph=g, ep=ep, width=width, L=L)
        except Exception:
            pass
        src = "built untimed by b200ann_nsg_build, loaded through IndexNSG::Load"
    ref = ref_nsg(base, g, ep, width)
    threads = os.cpu_count() or 1
    for _ in range(args.warmup):
        ref.search(q, TOPK, L, seed=1, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ids = ref.search(q, TOPK, L, seed=1, threads=threads)
    dt = (time.perf_counter() - t0) / args.steps
    qps = len(q) / dt
    # exact ground truth on the host for the recall of what was timed (bounded: 500 queries)
    sub = np.random.default_rng(0).choice(len(q), min(200, len(q)), replace=False)
    d2 = (q[sub] ** 2).sum(1)[:, None] - 2.0 * q[sub] @ base.T + (base ** 2).sum(1)[None, :]
    truth = np.argsort(d2, axis=1, kind="stable")[:, :TOPK]
    rec = recall_at_k(ids[sub], truth)
    line = {
        "impl": "reference", "metric": "QPS @ recall@10>=0.95", "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "L_search": L},
        "recall_at_10": round(rec, 4), "recall_sample": "%d of the %d queries against an exact host scan" % (len(sub), len(q)),
        "index": src,
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": threads, "kind": "reference",
                         "sample": "each step = all %d queries through IndexNSG::Search (oracle/_ref/libref_nsg_fast, "
                                   "OpenMP over %d threads)" % (len(q), threads)},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the other construction paths (NN-descent, HNSW, CPU builds)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

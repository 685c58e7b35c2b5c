#!/usr/bin/env python
"""bench.py — the hot path of Cookiecoolkid/hnsw-nsg on B200, BASELINE.json configs[1]:
NSG on 1M x 128 SIFT-shaped synthetic data (exact kNN graph K=200 -> NSG L=40 R=50 C=500),
10k-query batches, k=10, metric = QPS at recall@10 >= 0.95.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the graph search over one 10k-query batch at the smallest L_search whose
recall@10 (against exact brute-force ground truth) is >= 0.95.
  value  device-resident queries/results, CUDA events on the library's stream
  e2e    the same call through the C ABI with HOST buffers (pinned), copies inside the timed region
  N > 1  one process per GPU (torchrun); every rank owns a 1M-row shard with its own sub-index, all
         ranks search the same query batch, per-shard top-k lists are exchanged with one NCCL
         all-gather and merged on the GPU (weak scaling: the base grows with N).
--impl reference times the UNMODIFIED reference IndexNSG::Search (oracle/_ref, OpenMP over all host
cores) on the same graph, same queries, same L_search. See DESIGN.md "Measurement".
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
os.environ.setdefault("NCCL_DEBUG", "WARN")   # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line

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
    """rank r's shard = rows of seed 1000+r; queries are the same on every rank (seed 2)"""
    S = synth()
    t = time.time()
    base = S.sift_shape(N_BASE, DIM, seed=1000 + rank)
    q = S.sift_shape(NQ, DIM, seed=2)
    log("rank %d: data %dx%d + %d queries in %.1fs" % (rank, N_BASE, DIM, NQ, time.time() - t))
    return base, q


def cache_path(rank):
    lib = os.path.join(ROOT, "hnsw-nsg_b200", "libb200ann.so")
    key = "%s|%d|%d|%d|%d|%d|%d|%d|%s" % (WORKLOAD, rank, N_BASE, KNN_K, NSG_L, NSG_R, NSG_C, NQ,
                                          os.path.getmtime(lib) if os.path.exists(lib) else 0)
    return os.path.join(tempfile.gettempdir(), "b200ann_bench_%s.npz" % hashlib.md5(key.encode()).hexdigest()[:12])


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


def other_builds(B, h, ix, base, knn_exact, info):
    """the two other construction paths of SURVEY 8(a), timed on the same 1M x 128 rows (reported under "build"):
    NN-descent (efanna IndexGraph::Build, the README's SIFT1M parameters K=200 L=200 iter=10 S=10 R=100) and the
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
    # algorithmic bytes of the local join: two rows gathered per distance evaluation
    info["nndescent"] = dict(ninfo, K=K, L=L, iter=IT, S=S, R=R_, recall_vs_exact=round(rec, 4),
                             join_gather_GBps=ninfo["dist_evals"] * 2 * DIM * 4 / max(1e-9, ninfo["join_us"] * 1e-6) / 1e9)
    del out
    log("rank 0: NN-descent K=%d in %.2fs (%s)" % (K, info["nndescent_s"], info["nndescent"]))
    t = time.time()
    hinfo = ix.hnsw_build(16, 200, seed=100)
    info["hnsw_build_s"] = time.time() - t
    info["hnsw_build_device_ms"] = h.last_kernel_ms
    bytes_ = (hinfo["beam_evals"] + hinfo["heuristic_evals"]) * DIM * 4 + hinfo["hops"] * 33 * 4
    info["hnsw"] = dict(hinfo, M=16, efc=200, gather_GBps=bytes_ / (h.last_kernel_ms * 1e-3) / 1e9)
    log("rank 0: HNSW build M=16 efC=200 in %.2fs (%s)" % (info["hnsw_build_s"], info["hnsw"]))
    try:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import refutil as R
        import tempfile
        if R.have_ref() and int(os.environ.get("WORLD_SIZE", "1")) == 1:   # other ranks spin in NCCL: no clean host cores
            m = 50000
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
            info["build_cpu_reference"] = {"sample": "first %d rows, unmodified reference from oracle/_ref (-O3/-Ofast builds), OpenMP" % m,
                                           "cores": threads, "nndescent_s": t_nd, "nndescent_gpu_same_rows_s": g_nd,
                                           "hnsw_addpoint_s": t_h, "hnsw_gpu_same_rows_s": g_h}
            log("rank 0: reference builds on %d rows: %s" % (m, info["build_cpu_reference"]))
    except Exception as e:  # the build baselines are extras; the headline never depends on them
        log("build cpu reference skipped: %r" % (e,))


def build_index(B, h, base, q, rank):
    """upload, exact ground truth, exact kNN graph (K=200), NSG build, L_search sweep.
    Returns (ix, info). The graph is cached in /tmp so that the second arm on the same box reuses it."""
    import torch
    info = {}
    ix = B.Index(h, base, B.L2)
    gt, _ = ix.bf_knn(q, TOPK)
    info["gt_ms"] = h.last_kernel_ms
    cp = cache_path(rank)
    if os.path.exists(cp):
        z = np.load(cp)
        ix.set_graph(z["graph"], int(z["ep"]))
        info.update(json.loads(str(z["info"])))
        info["graph_from_cache"] = True
        log("rank %d: graph loaded from %s" % (rank, cp))
    else:
        knn = torch.empty((N_BASE, KNN_K), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        t = time.time()
        ix.knn_graph_raw(KNN_K, knn)
        info["knn_graph_s"] = time.time() - t
        info["knn_graph_kernel_ms"] = h.last_kernel_ms
        st = h.knn_last_stats
        info["knn_graph_stats"] = st
        # tensor roofline of the kNN-graph build: 2*n*n*dim useful flops (x passes issued) over the device time
        flops = 2.0 * N_BASE * N_BASE * DIM
        info["knn_graph_roofline"] = {
            "bound": "tensor", "kernel": "tc::sweep_kernel (tcgen05 bf16)" if st["rows"] else "bf_knn_kernel (fp32 SIMT)",
            "achieved": flops * max(1, st["passes"]) / (h.last_kernel_ms * 1e-3) / 1e12, "useful": flops / (h.last_kernel_ms * 1e-3) / 1e12,
            "peak": measured_peaks().get("bf16_tflops_sustained"), "unit": "TFLOP/s"}
        if info["knn_graph_roofline"]["peak"]:
            info["knn_graph_roofline"]["frac"] = info["knn_graph_roofline"]["achieved"] / info["knn_graph_roofline"]["peak"]
        log("rank %d: exact kNN graph K=%d in %.2fs" % (rank, KNN_K, info["knn_graph_s"]))
        t = time.time()
        bi = ix.nsg_build((knn, KNN_K), NSG_L, NSG_R, NSG_C, seed=1)
        info["nsg_build_s"] = time.time() - t
        info["nsg_build_device_ms"] = h.last_kernel_ms
        info["nsg"] = bi
        log("rank %d: NSG build in %.2fs %s" % (rank, info["nsg_build_s"], bi))
        if rank == 0:
            other_builds(B, h, ix, base, knn, info)
        del knn
        torch.cuda.empty_cache()
        g, ep, width = ix.get_graph()
        try:
            np.savez(cp, graph=g, ep=ep, info=json.dumps({k: info[k] for k in info if k != "gt_ms"}))
        except Exception as e:  # cache is best effort
            log("cache write failed: %s" % e)
    # smallest L_search reaching the recall target
    chosen = None
    sweep = []
    for L in L_SWEEP:
        if L < TOPK:
            continue
        ids, _ = ix.nsg_search(q, TOPK, L, seed=1)
        r = recall_at_k(ids, gt)
        sweep.append((L, round(r, 4)))
        if r >= RECALL_TARGET:
            chosen = (L, r)
            break
    if chosen is None:
        raise RuntimeError("recall@%d >= %.2f not reached in sweep %s" % (TOPK, RECALL_TARGET, sweep))
    info["L_search"], info["recall"] = chosen[0], chosen[1]
    info["sweep"] = sweep
    log("rank %d: L_search=%d recall@%d=%.4f (sweep %s)" % (rank, chosen[0], TOPK, chosen[1], sweep))
    return ix, gt, info


def cpu_reference_qps(base, q, graph, ep, width, L, threads, passes=1):
    """UNMODIFIED reference IndexNSG::Search (oracle/_ref/libref_nsg_fast*.so built from
    /root/reference/CNNS/src/nsg/index_nsg.cpp), OpenMP over `threads` host cores, on the same graph.
    Returns (qps, recall ids)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refutil as R
    fd, path = tempfile.mkstemp(suffix=".nsg")
    os.close(fd)
    write_nsg(path, graph, ep, width)
    ref = R.RefNsg(base, path, lib="nsg_fast")
    os.unlink(path)
    ref.search(q[:min(len(q), 2000)], TOPK, L, seed=1, threads=threads)  # warm
    best = None
    ids = None
    for _ in range(passes):
        t = time.perf_counter()
        ids = ref.search(q, TOPK, L, seed=1, threads=threads)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return len(q) / best, ids, best


def dist_env():
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    return rank, local, world


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank, local, world = dist_env()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = importlib.import_module("hnsw-nsg_b200")
    h = B.Handle(local)
    base, q = make_data(rank)
    ix, gt, info = build_index(B, h, base, q, rank)
    L = info["L_search"]
    if world > 1:  # every shard searches with the largest L any shard needs
        t = torch.tensor([L], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        L = int(t.item())
    nq, K = q.shape[0], TOPK
    ext = torch.cuda.ExternalStream(h.stream, device=local)

    # ---- device-resident buffers
    q_dev = torch.from_numpy(q).cuda()
    ids_dev = torch.empty((nq, K), dtype=torch.int32, device="cuda")
    d_dev = torch.empty((nq, K), dtype=torch.float32, device="cuda")
    gathered_i = torch.empty((world * nq, K), dtype=torch.int32, device="cuda") if world > 1 else None
    gathered_d = torch.empty((world * nq, K), dtype=torch.float32, device="cuda") if world > 1 else None
    out_i = torch.empty((nq, K), dtype=torch.int32, device="cuda")
    out_d = torch.empty((nq, K), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()

    if world > 1:
        h.set_async(True)   # search and merge are enqueued on the library's stream; no host sync inside a step

    def step_device():
        if world == 1:
            ix.nsg_search_raw(q_dev, nq, DIM, K, L, 1, ids_dev, None, None)
            return
        # one stream end to end: search kernel -> local->global ids -> NCCL all-gathers -> merge kernel
        with torch.cuda.stream(ext):
            ix.nsg_search_raw(q_dev, nq, DIM, K, L, 1, ids_dev, d_dev, None)
            ids_dev.add_(rank * N_BASE)  # local -> global row ids (u32 stored as i32 bits)
            dist.all_gather_into_tensor(gathered_d, d_dev)
            dist.all_gather_into_tensor(gathered_i, ids_dev)
            h.merge_topk_raw(gathered_d, gathered_i, world, nq, K, K, out_d, out_i)

    # ---- pinned host buffers for the end-to-end leg
    q_pin = torch.from_numpy(q).pin_memory()
    ids_pin = torch.empty((nq, K), dtype=torch.int32).pin_memory()
    d_pin = torch.empty((nq, K), dtype=torch.float32).pin_memory()

    def step_e2e():
        if world == 1:
            ix.nsg_search_raw(q_pin, nq, DIM, K, L, 1, ids_pin, None, None)
        else:
            with torch.cuda.stream(ext):
                q_dev.copy_(q_pin, non_blocking=True)
            step_device()
            with torch.cuda.stream(ext):
                ids_pin.copy_(out_i, non_blocking=True)
                d_pin.copy_(out_d, non_blocking=True)
            ext.synchronize()   # the caller holds its results in host memory when the step returns

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step, steps, warmup):
        for _ in range(warmup):
            step()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = h.launch_count
        t0 = time.perf_counter()
        e0.record(ext)
        for _ in range(steps):
            step()
        e1.record(ext)
        barrier()
        wall = time.perf_counter() - t0
        dev_ms = e0.elapsed_time(e1)
        ms = torch.tensor([dev_ms, wall * 1e3], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0]), float(ms[1]), h.launch_count - l0

    clocks = ClockSampler(local)
    clocks.start()
    dev_ms, wall_ms, launches = timed(step_device, args.steps, args.warmup)
    clk = clocks.stop()
    # single GPU: device events bracket exactly the kernels; multi GPU: the collective runs on torch's
    # stream, so the step time is the barrier-to-barrier wall clock (max over ranks)
    step_ms = (dev_ms if world == 1 else wall_ms) / args.steps
    e2e_dev_ms, e2e_wall_ms, _ = timed(step_e2e, args.steps, args.warmup)
    e2e_ms = e2e_wall_ms / args.steps
    if world > 1:
        torch.cuda.synchronize()
        h.set_async(False)   # the accounting below reads b200ann_last_kernel_ms

    # ---- correctness of what was timed + roofline accounting (untimed)
    if world == 1:
        ids_final = ids_dev.cpu().numpy().view(np.uint32)
        rec = recall_at_k(ids_final, gt)
    else:
        # global ground truth: merge of the per-shard exact lists
        gtd_i, gtd_d = ix.bf_knn(q, K)
        gi = torch.from_numpy(gtd_i.view(np.int32)).cuda() + rank * N_BASE
        gd = torch.from_numpy(gtd_d).cuda()
        all_i = torch.empty((world * nq, K), dtype=torch.int32, device="cuda")
        all_d = torch.empty((world * nq, K), dtype=torch.float32, device="cuda")
        dist.all_gather_into_tensor(all_i, gi); dist.all_gather_into_tensor(all_d, gd)
        all_i = all_i.view(world, nq, K); all_d = all_d.view(world, nq, K)
        order = torch.argsort(all_d.permute(1, 0, 2).reshape(nq, -1), dim=1, stable=True)[:, :K]
        truth = torch.gather(all_i.permute(1, 0, 2).reshape(nq, -1), 1, order).cpu().numpy()
        rec = recall_at_k(out_i.cpu().numpy(), truth)
    _, _, st = ix.nsg_search(q, K, L, seed=1, want_stats=True)
    g, ep, width = ix.get_graph()
    gs = (g.shape[1] + 7) // 8 * 8
    n_dist = float(st[:, 0].astype(np.float64).sum()); n_hops = float(st[:, 1].astype(np.float64).sum())
    alg_bytes = n_dist * DIM * 4 + n_hops * gs * 4 + nq * DIM * 4
    # the search kernel alone (CUDA events inside the library around its launch)
    kms = []
    for _ in range(max(3, args.steps)):
        ix.nsg_search_raw(q_dev, nq, DIM, K, L, 1, ids_dev, None, None)
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

    line = None
    if rank == 0:
        line = {
            "metric": "QPS @ recall@10>=0.95", "value": world * nq / (step_ms * 1e-3) if world > 1 else nq / (step_ms * 1e-3),
            "unit": "queries/s" if world == 1 else "shard-queries/s (each query searched on every shard; global_qps = value / n_gpus)",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "L_search": L, "recall_at_10": round(rec, 4),
                       "rows_per_gpu": N_BASE, "total_rows": N_BASE * world,
                       "l2": "inputs larger than L2 (vectors %d MB + graph %d MB per GPU; no flush)" % (
                           N_BASE * DIM * 4 >> 20, N_BASE * gs * 4 >> 20),
                       "parallelism": "1 GPU" if world == 1 else "base sharded x%d, NCCL all-gather + GPU top-k merge" % world},
            "global_qps": nq / (step_ms * 1e-3),
            "e2e": {"value": (world if world > 1 else 1) * nq / (e2e_ms * 1e-3), "unit": "queries/s" if world == 1 else "shard-queries/s",
                    "ms_per_step": e2e_ms, "h2d_bytes_per_step": nq * DIM * 4,
                    "d2h_bytes_per_step": nq * K * 4 * (1 if world == 1 else 2)},
            "gpu_launches": launches,
            "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "beam_search_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes,
                         "dist_evals_per_query": n_dist / nq, "hops_per_query": n_hops / nq},
            "build": {k: info.get(k) for k in ("knn_graph_s", "knn_graph_kernel_ms", "knn_graph_stats", "knn_graph_roofline", "nsg_build_s", "nsg_build_device_ms",
                                               "nsg", "nndescent_s", "nndescent_device_ms", "nndescent", "hnsw_build_s",
                                               "hnsw_build_device_ms", "hnsw", "build_cpu_reference", "graph_from_cache", "gt_ms")},
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            qps, rids, secs = cpu_reference_qps(base, q, g, ep, width, L, threads)
            line["cpu_baseline"] = {"value": qps, "unit": "queries/s", "cores": threads, "kind": "reference",
                                    "sample": "all %d queries, 1 pass after a 2000-query warm-up (%.2f s), same graph and L_search; "
                                              "IndexNSG::Search from oracle/_ref/libref_nsg_fast, OpenMP" % (nq, secs),
                                    "recall_at_10": round(recall_at_k(rids, gt), 4)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_reference(args):
    """the reference's own CPU implementation of the timed path (IndexNSG::Search, unmodified sources
    compiled into oracle/_ref) on the host cores, same config / graph / queries / L_search.
    The index is an INPUT of the timed path; at 1M x 128 with K=200 the reference's own CPU build
    (NN-descent + NSG Link with its serial InterInsert) takes tens of minutes, so the graph is
    built (untimed setup, cached in /tmp) by b200ann_nsg_build and handed to the reference through
    its own IndexNSG::Load file format."""
    rank, local, world = dist_env()
    if rank != 0:
        return
    import torch
    torch.cuda.set_device(0)
    B = importlib.import_module("hnsw-nsg_b200")
    h = B.Handle(0)
    base, q = make_data(0)
    ix, gt, info = build_index(B, h, base, q, 0)
    L = info["L_search"]
    g, ep, width = ix.get_graph()
    ix.close(); h.close()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refutil as R
    fd, path = tempfile.mkstemp(suffix=".nsg")
    os.close(fd)
    write_nsg(path, g, ep, width)
    ref = R.RefNsg(base, path, lib="nsg_fast")
    os.unlink(path)
    threads = os.cpu_count() or 1
    for _ in range(args.warmup):
        ref.search(q, TOPK, L, seed=1, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ids = ref.search(q, TOPK, L, seed=1, threads=threads)
    dt = (time.perf_counter() - t0) / args.steps
    qps = len(q) / dt
    rec = recall_at_k(ids, gt)
    line = {
        "impl": "reference", "metric": "QPS @ recall@10>=0.95", "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "L_search": L, "recall_at_10": round(rec, 4),
                   "index": "built untimed by b200ann_nsg_build, loaded through IndexNSG::Load"},
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
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

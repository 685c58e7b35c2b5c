"""Test-side helpers: ctypes bindings for the oracle (oracle/liboracle.so) and for the compiled
UNMODIFIED reference (oracle/_ref/libref_*.so), seeded synthetic data and parsers
for the reference's on-disk graph formats. TEST INFRASTRUCTURE ONLY — the product never imports this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
INVALID = 0xFFFFFFFF

# oracle distance modes (oracle/oracle.c)
(L2_SCALAR, IP_SCALAR, L2_HNSW_AVX, IP_HNSW_AVX, L2_EFANNA_AVX, L2_GPU, IP_GPU, FASTL2_EFANNA_AVX, L2_SEQ_FMA,
 IP_SEQ_FMA, L2_HYBRID_AVX, L2_U8) = range(12)

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def _build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s", os.path.join(ORACLE_DIR, "liboracle.so")])
    return so


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(_build_oracle())
        lib.orc_dist.restype = C.c_float
        lib.orc_dist.argtypes = [C.c_int, f32p, f32p, C.c_size_t]
        lib.orc_set_gpu_team.argtypes = [C.c_int]
        lib.orc_seed_hash.restype = C.c_uint32
        lib.orc_seed_hash.argtypes = [C.c_uint32] * 3
        lib.orc_bf_knn.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, f32p, C.c_size_t, C.c_size_t,
                                   C.c_size_t, C.c_int, u32p, f32p]
        lib.orc_bf_knn_u8.argtypes = [u8p, C.c_size_t, C.c_size_t, C.c_size_t, u8p, C.c_size_t, C.c_size_t,
                                      C.c_size_t, u32p, i32p]
        lib.orc_insert_into_pool.argtypes = [u32p, f32p, i32p, C.c_uint, C.c_uint32, C.c_float]
        lib.orc_nsg_search.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, C.c_uint32,
                                       f32p, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_int, C.c_int,
                                       C.c_uint32, u32p, C.c_void_p, C.c_void_p]
        lib.orc_hnsw_search.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, i32p, u32p,
                                        u32p, C.c_uint, C.c_void_p, C.c_uint32, C.c_int, f32p, C.c_size_t,
                                        C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int, u64p, f32p,
                                        C.c_void_p]
        lib.orc_nsg_init_ep.restype = C.c_uint32
        lib.orc_nsg_init_ep.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, C.c_uint32, C.c_uint,
                                        C.c_int, C.c_int, C.c_uint32, C.c_void_p]
        lib.orc_nsg_link.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, C.c_uint32, C.c_uint,
                                     C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_uint32, u32p, f32p, C.c_void_p]
        lib.orc_nsg_interinsert.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, C.c_int, C.c_int,
                                            u32p, f32p, C.c_void_p]
        lib.orc_nsg_cut_to_graph.argtypes = [u32p, f32p, C.c_size_t, C.c_uint, u32p, C.c_uint]
        lib.orc_nsg_tree_grow.restype = C.c_long
        lib.orc_nsg_tree_grow.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, C.c_uint32,
                                          C.c_uint, C.c_int, C.c_int, C.c_uint32, C.c_uint, C.POINTER(C.c_uint)]
        lib.orc_nsg_search_from.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, u32p,
                                            f32p, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_int, C.c_int,
                                            C.c_uint32, u32p, C.c_void_p, C.c_void_p]
        lib.orc_hybrid_search.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, i32p, u32p, u32p, C.c_uint,
                                          C.c_void_p, C.c_uint32, C.c_int, u32p, C.c_uint, f32p, C.c_size_t,
                                          C.c_size_t, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_uint32,
                                          u32p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.orc_merge_topk.argtypes = [f32p, u32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, f32p, u32p]
        lib.orc_mt19937.argtypes = [C.c_uint32, u32p, C.c_size_t]
        lib.orc_gen_random.argtypes = [C.c_int, C.c_uint32, C.c_uint32, u32p, C.c_uint, C.c_uint]
        lib.orc_nndescent_ref.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_uint,
                                          C.c_uint, C.c_uint, C.c_int, u32p, C.c_void_p]
        lib.orc_nndescent_batch.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_uint,
                                            C.c_uint, C.c_uint, C.c_uint32, C.c_int, u32p, C.c_void_p, C.c_void_p]
        lib.orc_nsg_search_qidx.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, C.c_uint32, f32p,
                                            C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_int, C.c_uint32, u32p, u32p, f32p]
        lib.orc_hnsw_search_filtered.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, i32p, u32p,
                                                 u32p, C.c_uint, C.c_void_p, C.c_uint32, C.c_int, u8p, f32p,
                                                 C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int,
                                                 C.c_int, u64p, f32p, C.c_void_p, C.c_void_p]
        lib.orc_hnsw_search_epsilon.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, i32p, u32p,
                                                u32p, C.c_uint, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, f32p,
                                                C.c_size_t, C.c_size_t, C.c_float, C.c_size_t, C.c_size_t, C.c_int,
                                                u64p, f32p, u32p, C.c_void_p]
        lib.orc_hnsw_update.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_int, C.c_int, i32p,
                                        u32p, u32p, u32p, C.c_size_t, C.c_uint32, C.c_int, u32p, f32p, C.c_size_t, C.c_void_p]
        lib.orc_hnsw_search_multivector.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint, i32p, u32p, u32p,
                                                    C.c_uint, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, u32p, f32p,
                                                    C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int,
                                                    u64p, f32p, u32p, C.c_void_p]
        lib.orc_hnsw_levels.argtypes = [C.c_size_t, C.c_uint, C.c_uint32, i32p]
        lib.orc_hnsw_batch_end.restype = C.c_size_t
        lib.orc_hnsw_batch_end.argtypes = [C.c_size_t, C.c_size_t, i32p, C.c_int, C.c_uint, C.c_uint]
        lib.orc_hnsw_build.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_int, C.c_int,
                                       C.c_uint, C.c_uint, i32p, u32p, u32p, u32p, C.POINTER(C.c_uint32),
                                       C.POINTER(C.c_int), C.c_void_p]
        _oracle = lib
    return _oracle


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


# ------------------------------------------------------------------ oracle wrappers
def orc_bf_knn(base, q, k, mode):
    n, dim = base.shape
    nq = q.shape[0]
    ids = np.empty((nq, k), np.uint32)
    d = np.empty((nq, k), np.float32)
    oracle().orc_bf_knn(base, n, dim, base.shape[1], q, nq, q.shape[1], k, mode, ids, d)
    return ids, d


def orc_bf_knn_u8(base, q, k):
    n, dim = base.shape
    nq = q.shape[0]
    ids = np.empty((nq, k), np.uint32)
    d = np.empty((nq, k), np.int32)
    oracle().orc_bf_knn_u8(base, n, dim, base.shape[1], q, nq, q.shape[1], k, ids, d)
    return ids, d


def orc_nsg_search(base, graph, ep, q, K, L, mode, rng_mode=1, seed=0, want_stats=False):
    n, dim = base.shape
    nq = q.shape[0]
    ids = np.empty((nq, K), np.uint32)
    d = np.empty((nq, K), np.float32)
    st = np.zeros((nq, 2), np.uint64) if want_stats else None
    rc = oracle().orc_nsg_search(base, n, dim, dim, graph, graph.shape[1], ep, q, nq, q.shape[1], K, L, mode,
                                 rng_mode, seed, ids, _vp(d), _vp(st))
    assert rc == 0
    return (ids, d, st) if want_stats else (ids, d)


def u8_rows(a):
    """[n, dim] uint8 -> float32 view [n, ceil(dim/4)] of the rows zero-padded to 4 bytes: how the oracle takes byte
    rows in mode L2_U8 (strides in 4-byte units; pass dim=<byte dim> to the orc_hnsw_* wrappers). Distances come back
    as int32 bit patterns in the float32 arrays (.view(np.int32))."""
    a = np.ascontiguousarray(a, np.uint8)
    n, dim = a.shape
    rb = (dim + 3) // 4 * 4
    out = np.zeros((n, rb), np.uint8)
    out[:, :dim] = a
    return out.view(np.float32)


def orc_hnsw_search(base, h, q, k, ef, mode, variant=1, want_stats=False, dim=None):
    """h: dict with links0, levels, upper_ptr, upper_links, labels, ep, maxlevel (see parse_hnsw)."""
    n, sdim = base.shape
    dim = dim or sdim
    nq = q.shape[0]
    lab = np.empty((nq, k), np.uint64)
    d = np.empty((nq, k), np.float32)
    st = np.zeros((nq, 2), np.uint64) if want_stats else None
    up = h["upper_links"]
    if up.size == 0:
        up = np.full((1, max(1, h["maxM"])), INVALID, np.uint32)
    oracle().orc_hnsw_search(base, n, dim, sdim, h["links0"], h["links0"].shape[1], h["levels"], h["upper_ptr"],
                             up, up.shape[1], _vp(h["labels"]), h["ep"], h["maxlevel"], q, nq, q.shape[1], k, ef,
                             mode, variant, lab, d, _vp(st))
    return (lab, d, st) if want_stats else (lab, d)


def orc_hnsw_search_filtered(base, h, allow, q, k, ef, mode, variant=0, ccap=0, want_stats=False, dim=None):
    """allow [n] u8 by internal id (0 = filtered out or deleted). variant 0 = the reference's two heaps, 1 = the CUDA
    kernel's bounded arrays (ccap). Returns (labels, d, inexact[, stats])."""
    n, sdim = base.shape
    dim = dim or sdim
    nq = q.shape[0]
    lab = np.empty((nq, k), np.uint64)
    d = np.empty((nq, k), np.float32)
    bad = np.zeros(nq, np.uint8)
    st = np.zeros((nq, 2), np.uint64) if want_stats else None
    up = h["upper_links"]
    if up.size == 0:
        up = np.full((1, max(1, h["maxM"])), INVALID, np.uint32)
    rc = oracle().orc_hnsw_search_filtered(base, n, dim, sdim, h["links0"], h["links0"].shape[1], h["levels"],
                                           h["upper_ptr"], up, up.shape[1], _vp(h["labels"]), h["ep"], h["maxlevel"],
                                           np.ascontiguousarray(allow, np.uint8), q, nq, q.shape[1], k, ef, ccap, mode,
                                           variant, lab, d, _vp(st), _vp(bad))
    assert rc == 0
    return (lab, d, bad, st) if want_stats else (lab, d, bad)


def orc_hnsw_search_epsilon(base, h, q, eps, min_cand, max_cand, mode, allow=None, internal_ids=True, want_stats=False,
                            dim=None):
    """searchStopConditionClosest with EpsilonSearchStopCondition; allow [n] u8 by internal id or None.
    Returns (ids or labels [nq,max], d, counts[, stats])."""
    n, sdim = base.shape
    dim = dim or sdim
    nq = q.shape[0]
    lab = np.empty((nq, max_cand), np.uint64)
    d = np.empty((nq, max_cand), np.float32)
    cnt = np.zeros(nq, np.uint32)
    st = np.zeros((nq, 2), np.uint64) if want_stats else None
    up = h["upper_links"]
    if up.size == 0:
        up = np.full((1, max(1, h["maxM"])), INVALID, np.uint32)
    a = None if allow is None else np.ascontiguousarray(allow, np.uint8)
    rc = oracle().orc_hnsw_search_epsilon(base, n, dim, sdim, h["links0"], h["links0"].shape[1], h["levels"],
                                          h["upper_ptr"], up, up.shape[1], None if internal_ids else _vp(h["labels"]),
                                          h["ep"], h["maxlevel"], _vp(a), q, nq, q.shape[1], eps, min_cand, max_cand,
                                          mode, lab, d, cnt, _vp(st))
    assert rc == 0
    return (lab, d, cnt, st) if want_stats else (lab, d, cnt)


def orc_hnsw_search_multivector(base, h, doc, q, num_docs, ef_collection, max_results, mode, allow=None, internal_ids=True,
                                want_stats=False):
    """searchStopConditionClosest with MultiVectorSearchStopCondition; doc [n] u32 by internal id.
    Returns (ids or labels [nq,max_results], d, counts[, stats])."""
    n, dim = base.shape
    nq = q.shape[0]
    lab = np.empty((nq, max_results), np.uint64)
    d = np.empty((nq, max_results), np.float32)
    cnt = np.zeros(nq, np.uint32)
    st = np.zeros((nq, 2), np.uint64) if want_stats else None
    up = h["upper_links"]
    if up.size == 0:
        up = np.full((1, max(1, h["maxM"])), INVALID, np.uint32)
    a = None if allow is None else np.ascontiguousarray(allow, np.uint8)
    rc = oracle().orc_hnsw_search_multivector(base, n, dim, dim, h["links0"], h["links0"].shape[1], h["levels"],
                                              h["upper_ptr"], up, up.shape[1], None if internal_ids else _vp(h["labels"]),
                                              h["ep"], h["maxlevel"], _vp(a), np.ascontiguousarray(doc, np.uint32), q, nq,
                                              q.shape[1], num_docs, ef_collection, max_results, mode, lab, d, cnt, _vp(st))
    assert rc == 0, rc
    return (lab, d, cnt, st) if want_stats else (lab, d, cnt)


def orc_nsg_search_from(base, graph, entry, q, K, L, mode, rng_mode=1, seed=0, want_stats=False):
    n, dim = base.shape
    nq = q.shape[0]
    ids = np.empty((nq, K), np.uint32)
    d = np.empty((nq, K), np.float32)
    st = np.zeros((nq, 2), np.uint64) if want_stats else None
    rc = oracle().orc_nsg_search_from(base, n, dim, dim, graph, graph.shape[1], np.ascontiguousarray(entry, np.uint32),
                                      q, nq, q.shape[1], K, L, mode, rng_mode, seed, ids, _vp(d), _vp(st))
    assert rc == 0
    return (ids, d, st) if want_stats else (ids, d)


def orc_hybrid_search(base, h, graph, q, K, L, mode_hnsw, mode_nsg, rng_mode=1, seed=0, want_stats=False):
    """h: parse_hnsw dict (upper layers + labels are used); graph: flat NSG adjacency."""
    n, dim = base.shape
    nq = q.shape[0]
    ids = np.empty((nq, K), np.uint32)
    d = np.empty((nq, K), np.float32)
    entry = np.empty(nq, np.uint32)
    st = np.zeros((nq, 2), np.uint64) if want_stats else None
    up = h["upper_links"]
    if up.size == 0:
        up = np.full((1, max(1, h["maxM"])), INVALID, np.uint32)
    rc = oracle().orc_hybrid_search(base, n, dim, dim, h["levels"], h["upper_ptr"], up, up.shape[1], _vp(h["labels"]),
                                    h["ep"], h["maxlevel"], graph, graph.shape[1], q, nq, q.shape[1], K, L,
                                    mode_hnsw, mode_nsg, rng_mode, seed, ids, _vp(d), _vp(entry), _vp(st))
    assert rc == 0
    return (ids, d, entry, st) if want_stats else (ids, d, entry)


def pad_graph(knn, gstride=None):
    """[n,K] ids -> flat adjacency [n,gstride] (gstride multiple of 8) padded with INVALID"""
    n, K = knn.shape
    gs = gstride or (K + 7) // 8 * 8
    g = np.full((n, gs), INVALID, np.uint32)
    g[:, :K] = knn
    return g


def orc_nsg_init_ep(base, graph, ep0, L, mode, rng_mode=1, seed=0):
    n, dim = base.shape
    center = np.empty(dim, np.float32)
    ep = oracle().orc_nsg_init_ep(base, n, dim, dim, graph, graph.shape[1], ep0, L, mode, rng_mode, seed, _vp(center))
    return int(ep), center


def orc_nsg_link(base, knn_graph, ep, L, R_, C_, mode, rng_mode=1, seed=0, want_stats=False):
    n, dim = base.shape
    ci = np.empty((n, R_), np.uint32)
    cd = np.empty((n, R_), np.float32)
    st = np.zeros((n, 3), np.uint64) if want_stats else None
    oracle().orc_nsg_link(base, n, dim, dim, knn_graph, knn_graph.shape[1], ep, L, R_, C_, mode, rng_mode, seed, ci, cd,
                          _vp(st))
    return (ci, cd, st) if want_stats else (ci, cd)


def orc_nsg_interinsert(base, ci, cd, mode, variant):
    n, dim = base.shape
    ci, cd = ci.copy(), cd.copy()
    ne = C.c_uint64(0)
    oracle().orc_nsg_interinsert(base, n, dim, dim, ci.shape[1], mode, variant, ci, cd, C.byref(ne))
    return ci, cd, ne.value


def orc_nsg_cut_to_graph(ci, cd, gstride):
    n, R_ = ci.shape
    g = np.empty((n, gstride), np.uint32)
    oracle().orc_nsg_cut_to_graph(ci, cd, n, R_, g, gstride)
    return g


def orc_nsg_tree_grow(base, graph, ep, L, mode, width_in, rng_mode=1, seed=0):
    n, dim = base.shape
    g = graph.copy()
    w = C.c_uint(0)
    added = oracle().orc_nsg_tree_grow(base, n, dim, dim, g, g.shape[1], ep, L, mode, rng_mode, seed, width_in, C.byref(w))
    return g, int(added), w.value


def orc_nndescent_ref(base, K, L, it, S, R_, mode, seed=1):
    """IndexGraph::Build restated statement for statement (one thread, libc rand() seeded with `seed`)."""
    n, dim = base.shape
    libc().srand(seed)
    knn = np.zeros((n, K), np.uint32)
    info = np.zeros(4, np.uint64)
    rc = oracle().orc_nndescent_ref(base, n, dim, dim, K, L, it, S, R_, mode, knn, _vp(info))
    assert rc >= 0
    return knn, info


def orc_nndescent_batch(base, K, L, it, S, R_, mode, seed=1):
    """the order-independent batch form the CUDA path computes; returns ids, distances, info."""
    n, dim = base.shape
    knn = np.zeros((n, K), np.uint32)
    d = np.zeros((n, K), np.float32)
    info = np.zeros(4, np.uint64)
    rc = oracle().orc_nndescent_batch(base, n, dim, dim, K, L, it, S, R_, seed, mode, knn, _vp(d), _vp(info))
    assert rc >= 0
    return knn, d, info


def orc_kdtree_graph(base, K, n_trees, m_level, seed=0):
    """IndexKDtree::Build in the batch form the CUDA path computes -> (knn [n,K], d, info[4])"""
    n, dim = base.shape
    knn = np.empty((n, K), np.uint32)
    d = np.empty((n, K), np.float32)
    info = np.zeros(4, np.uint64)
    lib = oracle()
    lib.orc_kdtree_graph.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_uint, C.c_uint32, u32p,
                                     C.c_void_p, C.c_void_p]
    rc = lib.orc_kdtree_graph(base, n, dim, dim, K, n_trees, m_level, seed, knn, _vp(d), _vp(info))
    assert rc == 0
    return knn, d, info


def orc_nndescent_refine(base, init_graph, K, L, it, S, R_, mode, seed=1):
    """IndexGraph::RefineGraph in batch form: NN-descent whose pools start from init_graph [n, init_k]"""
    n, dim = base.shape
    g = np.ascontiguousarray(init_graph, np.uint32)
    knn = np.empty((n, K), np.uint32)
    d = np.empty((n, K), np.float32)
    info = np.zeros(4, np.uint64)
    lib = oracle()
    lib.orc_nndescent_refine.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_uint, C.c_uint,
                                         C.c_uint, C.c_uint32, C.c_int, u32p, C.c_uint, u32p, C.c_void_p, C.c_void_p]
    rc = lib.orc_nndescent_refine(base, n, dim, dim, K, L, it, S, R_, seed, mode, g, g.shape[1], knn, _vp(d), _vp(info))
    assert rc >= 0
    return knn, d, info


def orc_nsg_search_qidx(base, graph, ep, q, qidx, K, L, mode, seed=0):
    n, dim = base.shape
    nq = q.shape[0]
    ids = np.empty((nq, K), np.uint32)
    d = np.empty((nq, K), np.float32)
    rc = oracle().orc_nsg_search_qidx(base, n, dim, dim, np.ascontiguousarray(graph), graph.shape[1], ep, q, nq, q.shape[1],
                                      K, L, mode, seed, np.ascontiguousarray(qidx, np.uint32), ids, d)
    assert rc == 0
    return ids, d


def orc_hnsw_levels(n, M, seed=100):
    lv = np.zeros(n, np.int32)
    oracle().orc_hnsw_levels(n, M, seed, lv)
    return lv


def orc_hnsw_build(base, M, efc, mode, variant, ratio=16, bmax=4096, seed=100, label0=0, dim=None):
    """hnswlib addPoint loop over rows 0..n-1 -> the flat dict parse_hnsw returns. variant 0 = the reference one
    point at a time, 1 = the CUDA path's batch form."""
    n, sdim = base.shape
    dim = dim or sdim
    lv = orc_hnsw_levels(n, M, seed)
    rows = int(lv.sum())
    links0 = np.zeros((n, 2 * M), np.uint32)
    upper_ptr = np.zeros(n, np.uint32)
    upper = np.zeros((max(rows, 1), M), np.uint32)
    ep, ml = C.c_uint32(0), C.c_int(0)
    info = np.zeros(4, np.uint64)
    rc = oracle().orc_hnsw_build(base, n, dim, sdim, M, efc, mode, variant, ratio, bmax, lv, links0, upper_ptr, upper,
                                 C.byref(ep), C.byref(ml), _vp(info))
    assert rc == 0
    return dict(n=n, dim=dim, M=M, maxM=M, maxM0=2 * M, efc=max(efc, M), ep=ep.value, maxlevel=ml.value, links0=links0,
                levels=lv, upper_ptr=upper_ptr, upper_links=upper[:rows], labels=np.arange(n, dtype=np.uint64) + label0,
                info=info)


def orc_hnsw_update(base, h, ids, newvec, mode, variant, efc=None, dim=None):
    """updatePoint for the elements `ids` (internal ids) with the vectors `newvec`: returns (new base, new graph dict).
    variant 0 = the reference one element at a time, 1 = the CUDA path's batch form. h: dict as parse_hnsw returns."""
    n, sdim = base.shape
    dim = dim or sdim
    data = np.ascontiguousarray(base).copy()
    links0 = np.ascontiguousarray(h["links0"]).copy()
    up = np.ascontiguousarray(h["upper_links"]).copy()
    rows = up.shape[0] if up.ndim == 2 and up.size else 0
    M = h["maxM"]
    if rows == 0:
        up = np.full((1, max(1, M)), INVALID, np.uint32)
    info = np.zeros(2, np.uint64)
    ids = np.ascontiguousarray(ids, np.uint32)
    newvec = np.ascontiguousarray(newvec, base.dtype)
    # the batch form moves at most 1/32 of the index per pass (b200ann_hnsw_update_batch, capi.cu): a larger call is a
    # sequence of passes, each on the graph the previous one left
    sub = len(ids) if variant == 0 else max(1, n // 32)
    for s in range(0, len(ids), sub):
        one = np.zeros(2, np.uint64)
        rc = oracle().orc_hnsw_update(data, n, dim, sdim, M, efc if efc is not None else h.get("efc", 200), mode, variant,
                                      np.ascontiguousarray(h["levels"], np.int32), links0,
                                      np.ascontiguousarray(h["upper_ptr"], np.uint32), up, rows, h["ep"], h["maxlevel"],
                                      np.ascontiguousarray(ids[s:s + sub]), np.ascontiguousarray(newvec[s:s + sub]),
                                      len(ids[s:s + sub]), _vp(one))
        assert rc == 0
        info += one
    g = dict(h)
    g.update(links0=links0, upper_links=up[:rows] if rows else h["upper_links"], info=info)
    return data, g


def orc_merge_topk(d, ids, kout):
    parts, nq, k = d.shape
    od = np.empty((nq, kout), np.float32)
    oi = np.empty((nq, kout), np.uint32)
    oracle().orc_merge_topk(np.ascontiguousarray(d, np.float32), np.ascontiguousarray(ids, np.uint32), parts, nq, k, kout, od, oi)
    return od, oi


def libc():
    return C.CDLL("libc.so.6")


# ------------------------------------------------------------------ reference on-disk formats
def parse_hnsw(path, elem="f32"):
    """hnswlib saveIndex blob (hnswlib/hnswlib/hnswalg.h:685-713) -> flat arrays (oracle.c layout).
    elem = "u8": a HierarchicalNSW<int> / L2SpaceI file (rows of dim unsigned bytes)."""
    raw = np.fromfile(path, np.uint8)
    hdr = raw[:96]
    q = lambda off: int(hdr[off:off + 8].view(np.uint64)[0])
    offL0, maxel, cur, sdpe, label_off, off_data = [q(8 * i) for i in range(6)]
    maxlevel = int(hdr[48:52].view(np.int32)[0])
    ep = int(hdr[52:56].view(np.uint32)[0])
    maxM, maxM0, M = q(56), q(64), q(72)
    efc = q(88)
    n = cur
    blob = raw[96:96 + n * sdpe].reshape(n, sdpe)
    cnt = blob[:, offL0:offL0 + 2].copy().view(np.uint16).reshape(n).astype(np.int64)
    l0 = blob[:, offL0 + 4:offL0 + 4 + 4 * maxM0].copy().view(np.uint32).reshape(n, maxM0)
    links0 = np.where(np.arange(maxM0)[None, :] < cnt[:, None], l0, INVALID).astype(np.uint32)
    if elem == "u8":
        dim = label_off - off_data
        vec = blob[:, off_data:off_data + dim].copy()
    else:
        dim = (label_off - off_data) // 4
        vec = blob[:, off_data:off_data + 4 * dim].copy().view(np.float32).reshape(n, dim)
    labels = blob[:, label_off:label_off + 8].copy().view(np.uint64).reshape(n)
    pos = 96 + n * sdpe
    levels = np.zeros(n, np.int32)
    upper_ptr = np.zeros(n, np.uint32)
    rows = []
    slpe = maxM * 4 + 4
    for i in range(n):
        sz = int(raw[pos:pos + 4].view(np.uint32)[0])
        pos += 4
        if sz:
            lv = sz // slpe
            levels[i] = lv
            upper_ptr[i] = len(rows)
            for l in range(lv):
                b = raw[pos + l * slpe: pos + (l + 1) * slpe]
                c = int(b[:2].view(np.uint16)[0])
                r = np.full(maxM, INVALID, np.uint32)
                r[:c] = b[4:4 + 4 * c].view(np.uint32)
                rows.append(r)
            pos += sz
    upper = np.stack(rows) if rows else np.zeros((0, maxM), np.uint32)
    return dict(n=n, dim=dim, M=M, maxM=maxM, maxM0=maxM0, efc=efc, ep=ep, maxlevel=maxlevel,
                links0=np.ascontiguousarray(links0), levels=levels, upper_ptr=upper_ptr,
                upper_links=np.ascontiguousarray(upper), labels=np.ascontiguousarray(labels), vectors=vec)


def parse_nsg(path, n):
    """IndexNSG::Save (CNNS/src/nsg/index_nsg.cpp:37-49): [u32 width][u32 ep] then per node [u32 k][u32 x k]."""
    raw = np.fromfile(path, np.uint32)
    width, ep = int(raw[0]), int(raw[1])
    rows, pos = [], 2
    for _ in range(n):
        k = int(raw[pos])
        rows.append(raw[pos + 1:pos + 1 + k])
        pos += 1 + k
    assert pos == raw.size
    gs = max(8, (max(len(r) for r in rows) + 7) // 8 * 8)
    g = np.full((n, gs), INVALID, np.uint32)
    for i, r in enumerate(rows):
        g[i, :len(r)] = r
    return g, ep, width


def write_knn_graph(path, knn):
    """efanna kNN-graph file (IndexGraph::Save, efanna_graph/src/index_graph.cpp:348-357; read by
    IndexNSG::Load_nn_graph, index_nsg.cpp:128-148): per node [u32 K][u32 x K]."""
    n, K = knn.shape
    out = np.empty((n, K + 1), np.uint32)
    out[:, 0] = K
    out[:, 1:] = knn
    out.tofile(path)


def read_knn_graph(path):
    raw = np.fromfile(path, np.uint32)
    K = int(raw[0])
    return np.ascontiguousarray(raw.reshape(-1, K + 1)[:, 1:])


# ------------------------------------------------------------------ seeded synthetic data
def data_uniform(n, dim, seed):
    return np.random.default_rng(seed).random((n, dim), dtype=np.float32)


def data_sift_like(n, dim, seed, n_clusters=64, centers_seed=12345):
    """integer-valued fp32 in [0,218], clustered, heavy-tailed: every L2 partial sum is an integer
    < 2^24, so fp32 distances are exact under ANY summation order (CPU SIMD, GPU, scalar)."""
    cent = np.abs(np.random.default_rng(centers_seed).normal(0, 1, (n_clusters, dim))) * 40
    rng = np.random.default_rng(seed)
    a = rng.integers(0, n_clusters, n)
    x = cent[a] + rng.normal(0, 1, (n, dim)) * 18
    return np.floor(np.clip(np.abs(x), 0, 218)).astype(np.float32)


def _synth():
    import importlib.util
    spec = importlib.util.spec_from_file_location("b200ann_synth", os.path.join(ROOT, "hnsw-nsg_b200", "synth.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def data_sift_shape(n, dim, seed, latent=16):
    """bench.py's SIFT-shaped generator (hnsw-nsg_b200/synth.py)"""
    return _synth().sift_shape(n, dim, seed, latent)


def data_glove_like(n, dim, seed, n_clusters=100, centers_seed=54321):
    cent = np.random.default_rng(centers_seed).normal(0, 1, (n_clusters, dim))
    rng = np.random.default_rng(seed)
    x = cent[rng.integers(0, n_clusters, n)] * 0.7 + rng.normal(0, 1, (n, dim))
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x.astype(np.float32)


def recall(found, truth):
    """mean |found_row ∩ truth_row| / k"""
    k = truth.shape[1]
    hit = 0
    for f, t in zip(found, truth):
        hit += len(set(f.tolist()) & set(t.tolist()))
    return hit / (truth.shape[0] * k)


# ------------------------------------------------------------------ compiled reference (oracle/_ref)
def have_ref():
    return os.path.exists(os.path.join(REF_DIR, "libref_hnsw.so"))


def host_has_avx512():
    try:
        return "avx512f" in open("/proc/cpuinfo").read()
    except OSError:
        return False


_reflibs = {}


def reflib(name):
    """name in {hnsw, hnsw_avx, hnsw_fast, nsg, nsg_fast, efanna, efanna_fast}; *_fast resolves to the
    AVX-512 build when the host supports it."""
    if name.endswith("_fast") and host_has_avx512():
        name += "512"
    if name in _reflibs:
        return _reflibs[name]
    lib = C.CDLL(os.path.join(REF_DIR, "libref_%s.so" % name))
    if name.startswith("hnsw"):
        lib.ref_dist.restype = C.c_float
        lib.ref_dist.argtypes = [C.c_int, f32p, f32p, C.c_size_t]
        lib.ref_bf_knn.argtypes = [C.c_int, f32p, C.c_size_t, C.c_size_t, f32p, C.c_size_t, C.c_size_t, u64p,
                                   f32p, C.c_int]
        lib.ref_bf_knn_u8.argtypes = [u8p, C.c_size_t, C.c_size_t, u8p, C.c_size_t, C.c_size_t, u64p, i32p, C.c_int]
        lib.ref_hnsw_new.restype = C.c_void_p
        lib.ref_hnsw_new.argtypes = [C.c_int] + [C.c_size_t] * 5
        lib.ref_hnsw_load.restype = C.c_void_p
        lib.ref_hnsw_load.argtypes = [C.c_int, C.c_size_t, C.c_char_p]
        lib.ref_hnsw_free.argtypes = [C.c_void_p]
        lib.ref_hnsw_add.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_size_t, C.c_int]
        lib.ref_hnsw_search.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_size_t, C.c_size_t, u64p, f32p,
                                        C.c_int, C.c_void_p, C.c_void_p]
        lib.ref_hnsw_save.argtypes = [C.c_void_p, C.c_char_p]
        lib.ref_hnsw_search_filtered.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, u64p,
                                                 f32p, C.c_int]
        lib.ref_hnsw_mark_deleted.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
        lib.ref_hnsw_search_epsilon.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_float, C.c_size_t, C.c_size_t,
                                                C.c_void_p, u64p, f32p, u32p, C.c_int]
        lib.ref_hnswmv_new.restype = C.c_void_p
        lib.ref_hnswmv_new.argtypes = [C.c_int] + [C.c_size_t] * 5
        lib.ref_hnswmv_free.argtypes = [C.c_void_p]
        lib.ref_hnswmv_add.argtypes = [C.c_void_p, f32p, u32p, C.c_size_t, C.c_size_t]
        lib.ref_hnswmv_search.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, u64p,
                                          f32p, u32p, C.c_int]
        lib.ref_hnswmv_save.argtypes = [C.c_void_p, C.c_char_p]
        lib.ref_hnswi_new.restype = C.c_void_p
        lib.ref_hnswi_new.argtypes = [C.c_size_t] * 5
        lib.ref_hnswi_load.restype = C.c_void_p
        lib.ref_hnswi_load.argtypes = [C.c_size_t, C.c_char_p]
        lib.ref_hnswi_free.argtypes = [C.c_void_p]
        lib.ref_hnswi_add.argtypes = [C.c_void_p, u8p, C.c_size_t, C.c_size_t, C.c_int]
        lib.ref_hnswi_search.argtypes = [C.c_void_p, u8p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, u64p, i32p, C.c_int]
        lib.ref_hnswi_search_epsilon.argtypes = [C.c_void_p, u8p, C.c_size_t, C.c_float, C.c_size_t, C.c_size_t, u64p, i32p,
                                                 u32p, C.c_int]
        lib.ref_hnswi_save.argtypes = [C.c_void_p, C.c_char_p]
    elif name.startswith("nsg"):
        lib.ref_nsg_dist_l2.restype = C.c_float
        lib.ref_nsg_dist_l2.argtypes = [f32p, f32p, C.c_uint]
        lib.ref_nsg_build.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_char_p, C.c_uint, C.c_uint, C.c_uint,
                                      C.c_char_p, C.c_uint, C.c_int]
        lib.ref_nsg_load.restype = C.c_void_p
        lib.ref_nsg_load.argtypes = [C.c_size_t, C.c_size_t, C.c_char_p]
        lib.ref_nsg_free.argtypes = [C.c_void_p]
        lib.ref_nsg_search.argtypes = [C.c_void_p, f32p, f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint,
                                       u32p, C.c_uint, C.c_int]
        lib.ref_nsg_optimize.argtypes = [C.c_void_p, f32p]
        lib.ref_nsg_search_opt.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint, u32p,
                                           C.c_uint, C.c_int]
    elif name.startswith("hybrid"):
        lib.ref_hybrid_dist.restype = C.c_float
        lib.ref_hybrid_dist.argtypes = [f32p, f32p, C.c_uint]
        lib.ref_hybrid_build.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_char_p, C.c_uint, C.c_uint, C.c_uint,
                                         C.c_uint, C.c_uint, C.c_char_p, C.c_uint, C.c_int]
        lib.ref_hybrid_load.restype = C.c_void_p
        lib.ref_hybrid_load.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_char_p, C.c_uint]
        lib.ref_hybrid_free.argtypes = [C.c_void_p]
        lib.ref_hybrid_search.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.c_uint]
        lib.ref_hybrid_search_from.argtypes = [C.c_void_p, f32p, f32p, C.c_size_t, C.c_size_t, C.c_uint, u32p, u32p,
                                               C.c_uint]
    elif name.startswith("efanna"):
        lib.ref_nndescent_build.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_uint, C.c_uint,
                                            C.c_uint, C.c_char_p, C.c_uint, C.c_int]
    _reflibs[name] = lib
    return lib


class _quiet:
    """the reference prints a line per node in IndexNSG::Build (index_nsg.cpp:497): silence fd 1."""

    def __enter__(self):
        import sys
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *a):
        os.dup2(self.saved, 1)
        os.close(self.null)
        os.close(self.saved)


def ref_bf_knn(base, q, k, metric, lib="hnsw", threads=0):
    n, dim = base.shape
    lab = np.empty((q.shape[0], k), np.uint64)
    d = np.empty((q.shape[0], k), np.float32)
    reflib(lib).ref_bf_knn(metric, base, n, dim, q, q.shape[0], k, lab, d, threads)
    return lab, d


def ref_bf_knn_u8(base, q, k, lib="hnsw", threads=0):
    n, dim = base.shape
    lab = np.empty((q.shape[0], k), np.uint64)
    d = np.empty((q.shape[0], k), np.int32)
    reflib(lib).ref_bf_knn_u8(base, n, dim, q, q.shape[0], k, lab, d, threads)
    return lab, d


class RefHnswMV:
    """the unmodified HierarchicalNSW<float> over MultiVectorL2Space / MultiVectorInnerProductSpace<unsigned>
    (oracle/ref_hnsw_capi.cpp ref_hnswmv_*): rows = vector + document id"""

    def __init__(self, metric, dim, max_elements, M=16, efc=200, seed=100):
        self.lib = reflib("hnsw")
        self.dim = dim
        self.h = self.lib.ref_hnswmv_new(metric, dim, max_elements, M, efc, seed)
        assert self.h

    def add(self, data, doc, label0=0):
        self.lib.ref_hnswmv_add(self.h, np.ascontiguousarray(data, np.float32), np.ascontiguousarray(doc, np.uint32),
                                data.shape[0], label0)

    def search(self, q, num_docs, ef_collection, max_results, allow_by_label=None, threads=0):
        ids = np.empty((q.shape[0], max_results), np.uint64)
        d = np.empty((q.shape[0], max_results), np.float32)
        cnt = np.zeros(q.shape[0], np.uint32)
        a = None if allow_by_label is None else np.ascontiguousarray(allow_by_label, np.uint8)
        rc = self.lib.ref_hnswmv_search(self.h, q, q.shape[0], num_docs, ef_collection, max_results, _vp(a), ids, d, cnt, threads)
        assert rc == 0
        return ids, d, cnt

    def save(self, path):
        assert self.lib.ref_hnswmv_save(self.h, path.encode()) == 0

    def close(self):
        if self.h:
            self.lib.ref_hnswmv_free(self.h)
            self.h = None


class RefHnswI:
    """the unmodified HierarchicalNSW<int> over L2SpaceI (oracle/ref_hnsw_capi.cpp ref_hnswi_*)"""

    def __init__(self, dim, max_elements=0, M=16, efc=200, seed=100, path=None):
        self.lib = reflib("hnsw")
        self.dim = dim
        self.h = self.lib.ref_hnswi_new(dim, max_elements, M, efc, seed) if path is None else self.lib.ref_hnswi_load(dim, path.encode())
        assert self.h

    def add(self, data, label0=0, threads=1):
        data = np.ascontiguousarray(data, np.uint8)
        self.lib.ref_hnswi_add(self.h, data, data.shape[0], label0, threads)

    def search(self, q, k, ef, allow_by_label=None, threads=0):
        q = np.ascontiguousarray(q, np.uint8)
        lab = np.empty((q.shape[0], k), np.uint64)
        d = np.empty((q.shape[0], k), np.int32)
        a = None if allow_by_label is None else np.ascontiguousarray(allow_by_label, np.uint8)
        self.lib.ref_hnswi_search(self.h, q, q.shape[0], k, ef, _vp(a), lab, d, threads)
        return lab, d

    def search_epsilon(self, q, eps, min_cand, max_cand, threads=0):
        q = np.ascontiguousarray(q, np.uint8)
        ids = np.empty((q.shape[0], max_cand), np.uint64)
        d = np.empty((q.shape[0], max_cand), np.int32)
        cnt = np.zeros(q.shape[0], np.uint32)
        self.lib.ref_hnswi_search_epsilon(self.h, q, q.shape[0], eps, min_cand, max_cand, ids, d, cnt, threads)
        return ids, d, cnt

    def save(self, path):
        assert self.lib.ref_hnswi_save(self.h, path.encode()) == 0

    def close(self):
        if self.h:
            self.lib.ref_hnswi_free(self.h)
            self.h = None


class RefHnsw:
    def __init__(self, metric, dim, max_elements=0, M=16, efc=200, seed=100, lib="hnsw", path=None):
        self.lib = reflib(lib)
        self.dim = dim
        if path is None:
            self.h = self.lib.ref_hnsw_new(metric, dim, max_elements, M, efc, seed)
        else:
            self.h = self.lib.ref_hnsw_load(metric, dim, path.encode())
        assert self.h

    def add(self, data, label0=0, threads=1):
        self.lib.ref_hnsw_add(self.h, data, data.shape[0], label0, threads)

    def search(self, q, k, ef, threads=0):
        lab = np.empty((q.shape[0], k), np.uint64)
        d = np.empty((q.shape[0], k), np.float32)
        self.lib.ref_hnsw_search(self.h, q, q.shape[0], k, ef, lab, d, threads, None, None)
        return lab, d

    def search_filtered(self, q, k, ef, allow_by_label=None, threads=0):
        """searchKnn with a filter functor over a label bitmap (None = deleted marks only); short rows: ~0 / -1"""
        lab = np.empty((q.shape[0], k), np.uint64)
        d = np.empty((q.shape[0], k), np.float32)
        a = None if allow_by_label is None else np.ascontiguousarray(allow_by_label, np.uint8)
        self.lib.ref_hnsw_search_filtered(self.h, q, q.shape[0], k, ef, _vp(a), lab, d, threads)
        return lab, d

    def search_epsilon(self, q, eps, min_cand, max_cand, allow_by_label=None, threads=0):
        """searchStopConditionClosest(q, EpsilonSearchStopCondition(eps, min, max), filter): (internal ids, d, counts)"""
        ids = np.empty((q.shape[0], max_cand), np.uint64)
        d = np.empty((q.shape[0], max_cand), np.float32)
        cnt = np.zeros(q.shape[0], np.uint32)
        a = None if allow_by_label is None else np.ascontiguousarray(allow_by_label, np.uint8)
        self.lib.ref_hnsw_search_epsilon(self.h, q, q.shape[0], eps, min_cand, max_cand, _vp(a), ids, d, cnt, threads)
        return ids, d, cnt

    def mark_deleted(self, label, on=True):
        assert self.lib.ref_hnsw_mark_deleted(self.h, int(label), 1 if on else 0) == 0

    def save(self, path):
        self.lib.ref_hnsw_save(self.h, path.encode())

    def close(self):
        if self.h:
            self.lib.ref_hnsw_free(self.h)
            self.h = None

    __del__ = close


def ref_nsg_build(base, knn_path, L, R, C_, out_path, seed=1, threads=1, lib="nsg"):
    n, dim = base.shape
    with _quiet():
        reflib(lib).ref_nsg_build(base, n, dim, knn_path.encode(), L, R, C_, out_path.encode(), seed, threads)


class RefNsg:
    def __init__(self, base, path, lib="nsg"):
        self.lib = reflib(lib)
        self.base = base
        self.n, self.dim = base.shape
        self.h = self.lib.ref_nsg_load(self.n, self.dim, path.encode())
        self.opt = False

    def search(self, q, K, L, seed=1, threads=1):
        assert not self.opt
        ids = np.empty((q.shape[0], K), np.uint32)
        self.lib.ref_nsg_search(self.h, self.base, q, q.shape[0], self.dim, K, L, ids, seed, threads)
        return ids

    def optimize(self):
        self.lib.ref_nsg_optimize(self.h, self.base)
        self.opt = True

    def search_opt(self, q, K, L, seed=1, threads=1):
        assert self.opt
        ids = np.empty((q.shape[0], K), np.uint32)
        self.lib.ref_nsg_search_opt(self.h, q, q.shape[0], self.dim, K, L, ids, seed, threads)
        return ids

    def close(self):
        if self.h:
            self.lib.ref_nsg_free(self.h)
            self.h = None

    __del__ = close


def ref_hybrid_build(base, knn_path, L, R, C_, M, efc, prefix, seed=1, threads=1, lib="hybrid"):
    """HNSW_NSG build sequence of hnsw_nsg/tests/test_hnsw_nsg_search.cpp:318-351 -> prefix_hnsw.bin / prefix_nsg.bin"""
    n, dim = base.shape
    with _quiet():
        reflib(lib).ref_hybrid_build(base, n, dim, knn_path.encode(), L, R, C_, M, efc, prefix.encode(), seed, threads)


class RefHybrid:
    def __init__(self, base, prefix, L_search, lib="hybrid"):
        self.lib = reflib(lib)
        self.base = base
        self.n, self.dim = base.shape
        with _quiet():
            self.h = self.lib.ref_hybrid_load(base, self.n, self.dim, prefix.encode(), L_search)
        assert self.h

    def search(self, q, k, ef, seed=1):
        kout = max(k, ef)
        ids = np.empty((q.shape[0], kout), np.uint32)
        with _quiet():
            self.lib.ref_hybrid_search(self.h, q, q.shape[0], k, ef, ids, seed)
        return ids

    def search_from(self, q, K, L, entry, seed=1):
        ids = np.empty((q.shape[0], K), np.uint32)
        self.lib.ref_hybrid_search_from(self.h, self.base, q, q.shape[0], K, L, np.ascontiguousarray(entry, np.uint32),
                                        ids, seed)
        return ids

    def close(self):
        if self.h:
            self.lib.ref_hybrid_free(self.h)
            self.h = None

    __del__ = close


def ref_nndescent(base, K, L, it, S, R, out_path, seed=1, threads=0, lib="efanna"):
    n, dim = base.shape
    with _quiet():
        reflib(lib).ref_nndescent_build(base, n, dim, K, L, it, S, R, out_path.encode(), seed, threads)
    return read_knn_graph(out_path)

"""hnsw-nsg_b200 — Python face of the C ABI in include/b200ann.h (ctypes over libb200ann.so).

The library is hand-written sm_100a CUDA; there is no CPU fallback. Importing this module without the
built library raises; creating a Handle without a CUDA device raises.

Import with ``importlib.import_module("hnsw-nsg_b200")`` (the directory name carries a hyphen).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ann.so")
INVALID = 0xFFFFFFFF
L2, IP, L2_U8 = 0, 1, 2   # L2_U8: unsigned-byte rows, int squared L2 (hnswlib L2SpaceI); HNSW entry points only

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "hnsw-nsg_b200: %s is missing — build it with `make -C hnsw-nsg_b200/csrc` "
        "(or __graft_entry__.build()); there is no CPU fallback." % LIB_PATH)

_lib = C.CDLL(LIB_PATH)
_vp = C.c_void_p
_sz = C.c_size_t
_lib.b200ann_create.argtypes = [C.c_int, C.POINTER(_vp)]
_lib.b200ann_destroy.argtypes = [_vp]
_lib.b200ann_last_error.restype = C.c_char_p
_lib.b200ann_last_error.argtypes = [_vp]
_lib.b200ann_stream.restype = _vp
_lib.b200ann_stream.argtypes = [_vp]
_lib.b200ann_launch_count.restype = C.c_uint64
_lib.b200ann_launch_count.argtypes = [_vp]
_lib.b200ann_last_kernel_ms.restype = C.c_float
_lib.b200ann_last_kernel_ms.argtypes = [_vp]
_lib.b200ann_set_visited_bits.argtypes = [_vp, C.c_int]
_lib.b200ann_index_create.argtypes = [_vp, _vp, _sz, _sz, _sz, C.c_int, C.POINTER(_vp)]
_lib.b200ann_index_destroy.argtypes = [_vp]
_lib.b200ann_index_size.restype = _sz
_lib.b200ann_index_size.argtypes = [_vp]
_lib.b200ann_index_dim.restype = _sz
_lib.b200ann_index_dim.argtypes = [_vp]
_lib.b200ann_index_set_graph.argtypes = [_vp, _vp, _sz, C.c_uint32]
_lib.b200ann_index_get_graph.argtypes = [_vp, _vp, _sz, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
_lib.b200ann_index_set_hnsw.argtypes = [_vp, _vp, _sz, _vp, _vp, _vp, _sz, _sz, _vp, C.c_uint32, C.c_int]
_lib.b200ann_bf_knn.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _vp, _vp]
_lib.b200ann_knn_graph.argtypes = [_vp, _vp, _sz, _vp]
_lib.b200ann_knn_graph_rows.argtypes = [_vp, _vp, _sz, _sz, _sz, _vp]
_lib.b200ann_bf_knn_u8.argtypes = [_vp, _vp, _sz, _sz, _sz, _vp, _sz, _sz, _sz, _vp, _vp]
_lib.b200ann_hnsw_build.argtypes = [_vp, _vp, _sz, _sz, C.c_uint32, _sz, _sz, _vp, _vp]
_lib.b200ann_hnsw_insert_batch.argtypes = [_vp, _vp, _vp, _sz, _vp, _sz, _sz, _vp]
_lib.b200ann_index_hnsw_shape.argtypes = [_vp, _vp]
_lib.b200ann_index_get_hnsw.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp]
_lib.b200ann_nndescent.argtypes = [_vp, _vp, _sz, _sz, _sz, _sz, _sz, C.c_uint32, _vp, _vp, _vp]
_lib.b200ann_nsg_search.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, C.c_uint32, _vp, _vp, _vp]
_lib.b200ann_hnsw_search.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, _vp, _vp, _vp]
_lib.b200ann_hnsw_search_multivector.argtypes = [_vp, _vp, _vp, _sz, _sz, _vp, _sz, _sz, _sz, _vp, C.c_int, _vp, _vp, _vp, _vp]
_lib.b200ann_kdtree_graph.argtypes = [_vp, _vp, _sz, _sz, _sz, C.c_uint32, _vp, _vp, _vp]
_lib.b200ann_nndescent_refine.argtypes = [_vp, _vp, _sz, _sz, _sz, _sz, _sz, C.c_uint32, _vp, _sz, _vp, _vp, _vp]
_lib.b200ann_index_set_labels.argtypes = [_vp, _vp, _vp, _sz]
_lib.b200ann_hnsw_update_batch.argtypes = [_vp, _vp, _vp, _vp, _sz, _sz, _vp]
_lib.b200ann_index_create_u8.argtypes = [_vp, _vp, _sz, _sz, _sz, _vp]
_lib.b200ann_bf_knn_noalloc.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _vp, _vp]
NOALLOC_SCRATCH_BYTES = 16 << 20   # B200ANN_NOALLOC_SCRATCH_BYTES, include/b200ann.h
_lib.b200ann_hnsw_search_filtered.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, _vp, _vp, _vp, _vp, _vp]
_lib.b200ann_hnsw_search_epsilon.argtypes = [_vp, _vp, _vp, _sz, _sz, C.c_float, _sz, _sz, _vp, C.c_int, _vp, _vp,
                                             _vp, _vp]
_u32 = C.c_uint32
_lib.b200ann_nsg_search_from.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, _u32, _vp, _vp, _vp, _vp]
_lib.b200ann_nsg_search_clustered.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, _u32, _vp, _vp, _vp, _vp, _vp, _vp]
_lib.b200ann_hybrid_search.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, _u32, _vp, _vp, _vp, _vp]
_lib.b200ann_set_link_visited_bits.argtypes = [_vp, C.c_int]
_lib.b200ann_nsg_build.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, _u32, _u32, _vp]
_lib.b200ann_nsg_init_ep.argtypes = [_vp, _vp, _vp, _sz, _sz, _u32, _u32, C.POINTER(_u32), _vp]
_lib.b200ann_nsg_link.argtypes = [_vp, _vp, _vp, _sz, _u32, _sz, _sz, _sz, _u32, _vp, _vp, _vp]
_lib.b200ann_nsg_interinsert.argtypes = [_vp, _vp, _sz, _vp, _vp, C.POINTER(C.c_uint64)]
_lib.b200ann_merge_topk.argtypes = [_vp, _vp, _vp, _sz, _sz, _sz, _sz, _vp, _vp]
_lib.b200ann_merge_topk_gathered.argtypes = [_vp, _vp, _sz, _sz, _sz, _sz, _sz, _vp, _vp]
_lib.b200ann_set_knn_mode.argtypes = [_vp, C.c_int, C.c_int]
_lib.b200ann_set_async.argtypes = [_vp, C.c_int]
_lib.b200ann_cluster_search.argtypes = [_vp, _vp, _vp, _sz, _vp, _vp, _vp, _sz, _vp, _vp, _sz, _sz, _sz, _sz, _sz, _sz, _u32,
                                        C.c_int, _vp, _vp, _vp]
_lib.b200ann_knn_last_stats.argtypes = [_vp, _vp]
_lib.b200ann_set_search_options.argtypes = [_vp, C.c_int, C.c_int, C.c_int]
_lib.b200ann_set_query_order.argtypes = [_vp, _vp, _sz]
_lib.b200ann_set_search_kernel.argtypes = [_vp, C.c_int, C.c_int, C.c_int]
_lib.b200ann_set_id_base.argtypes = [_vp, _u32]
_lib.b200ann_probe_gather.argtypes = [_vp, _vp, C.c_uint64, C.c_int, C.c_int, C.c_int, _u32, C.POINTER(C.c_float)]
_lib.b200ann_nsg_finish.argtypes = [_vp, _vp, _vp, _vp, _sz, _u32, _sz, _u32, _sz, C.POINTER(_u32)]

EXPORTS = [
    "b200ann_create", "b200ann_destroy", "b200ann_last_error", "b200ann_stream", "b200ann_launch_count",
    "b200ann_last_kernel_ms", "b200ann_set_visited_bits", "b200ann_index_create", "b200ann_index_destroy",
    "b200ann_index_size", "b200ann_index_dim", "b200ann_index_set_graph", "b200ann_index_get_graph",
    "b200ann_index_set_hnsw", "b200ann_bf_knn", "b200ann_knn_graph", "b200ann_nsg_search", "b200ann_hnsw_search",
    "b200ann_hnsw_search_filtered",
    "b200ann_bf_knn_noalloc",
    "b200ann_index_create_u8",
    "b200ann_hnsw_update_batch",
    "b200ann_index_set_labels",
    "b200ann_nndescent_refine",
    "b200ann_kdtree_graph",
    "b200ann_hnsw_search_multivector",
    "b200ann_hnsw_search_epsilon",
    "b200ann_set_link_visited_bits", "b200ann_nsg_build", "b200ann_nsg_init_ep", "b200ann_nsg_link",
    "b200ann_nsg_interinsert", "b200ann_nsg_finish", "b200ann_merge_topk", "b200ann_nsg_search_from",
    "b200ann_hybrid_search", "b200ann_set_knn_mode", "b200ann_knn_last_stats", "b200ann_nndescent",
    "b200ann_hnsw_build", "b200ann_index_hnsw_shape", "b200ann_index_get_hnsw", "b200ann_nsg_search_clustered",
    "b200ann_knn_graph_rows", "b200ann_hnsw_insert_batch", "b200ann_merge_topk_gathered", "b200ann_set_search_options", "b200ann_set_search_kernel", "b200ann_set_id_base", "b200ann_set_query_order", "b200ann_probe_gather",
    "b200ann_set_async", "b200ann_cluster_search", "b200ann_bf_knn_u8", "b200ann_upload_u8", "b200ann_free_device",
]


class B200AnnError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b200ann error %d: %s" % (code, msg))
        self.code = code


def dist_team(dim):
    """lanes per row used by the search kernels (csrc/dist.cuh dist_team_for_nch)"""
    return 8 if (dim + 3) // 4 <= 64 else 32


def _ptr(a):
    """numpy array -> host pointer; torch CUDA/CPU tensor -> data_ptr; int -> raw device pointer"""
    if a is None:
        return None
    if isinstance(a, int):
        return _vp(a)
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data_as(_vp)
    if hasattr(a, "data_ptr"):
        assert a.is_contiguous()
        return _vp(a.data_ptr())
    raise TypeError(type(a))


class Handle:
    """one per GPU and host thread: device, stream, last error (include/b200ann.h)"""

    def __init__(self, device=0):
        h = _vp()
        rc = _lib.b200ann_create(device, C.byref(h))
        if rc != 0:
            raise B200AnnError(rc, "b200ann_create(device=%d) failed: no usable CUDA device (no CPU fallback)" % device)
        self._h = h
        self.device = device

    def _check(self, rc):
        if rc != 0:
            raise B200AnnError(rc, _lib.b200ann_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            _lib.b200ann_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def stream(self):
        return _lib.b200ann_stream(self._h)

    @property
    def launch_count(self):
        return int(_lib.b200ann_launch_count(self._h))

    @property
    def last_kernel_ms(self):
        return float(_lib.b200ann_last_kernel_ms(self._h))

    def set_knn_mode(self, mode, slack=-1):
        """0 auto, 1 fp32 SIMT (csrc/bf_knn.cu), 2 tcgen05 tensor cores + fp32 re-rank (csrc/knn_tc.cu)"""
        self._check(_lib.b200ann_set_knn_mode(self._h, mode, slack))

    @property
    def knn_last_stats(self):
        out = np.zeros(8, np.uint64)
        self._check(_lib.b200ann_knn_last_stats(self._h, out.ctypes.data_as(_vp)))
        keys = ["rows", "fallback_rows", "reranked", "passes", "splits", "shortlist", "tiles", "a_resident"]
        return dict(zip(keys, (int(x) for x in out)))

    def set_async(self, on):
        """device-to-device nsg_search / merge_topk calls stop synchronizing the stream (see b200ann.h)"""
        self._check(_lib.b200ann_set_async(self._h, 1 if on else 0))

    def set_search_options(self, prefetch=-1, visited_compact=-1, ctas_per_sm=-1):
        """scheduling knobs of the search kernels (never change results); -1 = leave as is"""
        self._check(_lib.b200ann_set_search_options(self._h, prefetch, visited_compact, ctas_per_sm))

    def set_search_kernel(self, version=-1, ring_stages=-1, mode=-1):
        """version 0 = auto (default: 2 for L <= 64, else 1), 2 = TMA row ring + register pool, 1 = round-1 kernel;
        mode 0 auto / 1 query per warp / 2 query per block; results are identical"""
        self._check(_lib.b200ann_set_search_kernel(self._h, version, ring_stages, mode))

    def set_id_base(self, id_base):
        """added to every valid id of the NSG-family searches (row shards: local -> global rows)"""
        self._check(_lib.b200ann_set_id_base(self._h, id_base))

    def set_query_order(self, order):
        """order: device u32/i32 permutation (torch tensor / raw pointer) or None; see b200ann.h"""
        n = 0 if order is None else int(order.numel()) if hasattr(order, "numel") else int(len(order))
        self._order_keepalive = order
        self._check(_lib.b200ann_set_query_order(self._h, _ptr(order), n))

    def set_visited_bits(self, bits):
        self._check(_lib.b200ann_set_visited_bits(self._h, bits))

    def merge_topk(self, d, ids, kout):
        """d/ids: numpy [parts, nq, k] -> (out_d, out_i) [nq, kout]"""
        d = np.ascontiguousarray(d, np.float32)
        ids = np.ascontiguousarray(ids, np.uint32)
        parts, nq, k = d.shape
        od = np.empty((nq, kout), np.float32)
        oi = np.empty((nq, kout), np.uint32)
        self._check(_lib.b200ann_merge_topk(self._h, _ptr(d), _ptr(ids), parts, nq, k, kout, _ptr(od), _ptr(oi)))
        return od, oi

    def merge_topk_raw(self, d, ids, parts, nq, k, kout, out_d, out_i):
        self._check(_lib.b200ann_merge_topk(self._h, _ptr(d), _ptr(ids), parts, nq, k, kout, _ptr(out_d), _ptr(out_i)))

    def merge_topk_gathered(self, buf, world, lp, nq, k, kout, out_d, out_i):
        """buf: device buffer filled by one all-gather of per-rank blocks [d: lp x nq x k][ids: lp x nq x k]"""
        self._check(_lib.b200ann_merge_topk_gathered(self._h, _ptr(buf), world, lp, nq, k, kout, _ptr(out_d), _ptr(out_i)))

    def bf_knn_u8(self, base, q, k):
        """exact kNN over u8 rows with int squared-L2 (hnswlib L2SpaceI + BruteforceSearch<int>);
        base [n,dim], q [nq,dim] uint8 numpy -> (ids u32 [nq,k], dists i32 [nq,k])"""
        base = np.ascontiguousarray(base, np.uint8)
        q = np.ascontiguousarray(q, np.uint8)
        ids = np.empty((q.shape[0], k), np.uint32)
        d = np.empty((q.shape[0], k), np.int32)
        self._check(_lib.b200ann_bf_knn_u8(self._h, _ptr(base), base.shape[0], base.shape[1], base.shape[1], _ptr(q),
                                           q.shape[0], q.shape[1], k, _ptr(ids), _ptr(d)))
        return ids, d

    def bf_knn_u8_raw(self, base, n, dim, bstride, q, nq, qstride, k, ids, dists):
        """device (torch tensor / raw pointer) or host buffers, strides in bytes"""
        self._check(_lib.b200ann_bf_knn_u8(self._h, _ptr(base), n, dim, bstride, _ptr(q), nq, qstride, k, _ptr(ids),
                                           _ptr(dists)))

    def set_link_visited_bits(self, bits):
        self._check(_lib.b200ann_set_link_visited_bits(self._h, bits))


class Index:
    """vectors (+ graphs) resident in HBM"""

    def __init__(self, handle, data, metric=L2, n=None, dim=None, stride=None):
        self.h = handle
        if n is None:
            n, dim = data.shape
            stride = dim
        ix = _vp()
        if metric == L2_U8:   # data: uint8 rows, stride in bytes
            handle._check(_lib.b200ann_index_create_u8(handle._h, _ptr(data), n, dim, stride, C.byref(ix)))
        else:
            handle._check(_lib.b200ann_index_create(handle._h, _ptr(data), n, dim, stride, metric, C.byref(ix)))
        self._ix = ix
        self.n, self.dim, self.metric = n, dim, metric
        self._qt = np.uint8 if metric == L2_U8 else np.float32    # element type of rows and queries
        self._dt = np.int32 if metric == L2_U8 else np.float32    # type of the distances the searches return

    def close(self):
        if getattr(self, "_ix", None):
            _lib.b200ann_index_destroy(self._ix)
            self._ix = None

    __del__ = close

    def probe_gather(self, rows, inflight=2, warps_per_block=8, blocks_per_sm=8, seed=1):
        """random whole-row gather rate of this index in GB/s (b200ann_probe_gather)"""
        ms = C.c_float(0)
        self.h._check(_lib.b200ann_probe_gather(self.h._h, self._ix, rows, inflight, warps_per_block, blocks_per_sm, seed,
                                                C.byref(ms)))
        return rows * self.dim * 4 / (ms.value * 1e-3) / 1e9, ms.value

    # ---- graphs
    def set_graph(self, graph, ep):
        graph = np.ascontiguousarray(graph, np.uint32)
        self.h._check(_lib.b200ann_index_set_graph(self._ix, _ptr(graph), graph.shape[1], ep))

    def get_graph(self):
        ep, width = C.c_uint32(), C.c_uint32()
        self.h._check(_lib.b200ann_index_get_graph(self._ix, None, 0, C.byref(ep), C.byref(width)))
        g = np.empty((self.n, max(1, width.value)), np.uint32)
        self.h._check(_lib.b200ann_index_get_graph(self._ix, _ptr(g), g.shape[1], C.byref(ep), C.byref(width)))
        return g, ep.value, width.value

    def set_hnsw(self, links0, levels, upper_ptr, upper_links, labels, ep, maxlevel):
        links0 = np.ascontiguousarray(links0, np.uint32)
        levels = np.ascontiguousarray(levels, np.int32)
        upper_ptr = np.ascontiguousarray(upper_ptr, np.uint32)
        upper_links = np.ascontiguousarray(upper_links, np.uint32)
        rows = upper_links.shape[0] if upper_links.ndim == 2 else 0
        maxM = upper_links.shape[1] if upper_links.ndim == 2 else 0
        labels = None if labels is None else np.ascontiguousarray(labels, np.uint64)
        self.h._check(_lib.b200ann_index_set_hnsw(self._ix, _ptr(links0), links0.shape[1], _ptr(levels), _ptr(upper_ptr),
                                                  _ptr(upper_links) if rows else None, rows, maxM, _ptr(labels), ep,
                                                  maxlevel))

    # ---- exact brute force
    def bf_knn(self, q, k):
        q = np.ascontiguousarray(q, np.float32)
        ids = np.empty((q.shape[0], k), np.uint32)
        d = np.empty((q.shape[0], k), np.float32)
        self.h._check(_lib.b200ann_bf_knn(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1], k, _ptr(ids), _ptr(d)))
        return ids, d

    def knn_graph(self, K):
        knn = np.empty((self.n, K), np.uint32)
        self.h._check(_lib.b200ann_knn_graph(self.h._h, self._ix, K, _ptr(knn)))
        return knn

    def knn_graph_rows(self, row0, nrows, K, out=None):
        """rows [row0, row0+nrows) of the exact kNN graph; out: optional device buffer [nrows,K]"""
        knn = out if out is not None else np.empty((nrows, K), np.uint32)
        self.h._check(_lib.b200ann_knn_graph_rows(self.h._h, self._ix, row0, nrows, K, _ptr(knn)))
        return knn

    def knn_graph_raw(self, K, out):
        """out: device buffer [n,K] u32/i32 (torch tensor or raw pointer)"""
        self.h._check(_lib.b200ann_knn_graph(self.h._h, self._ix, K, _ptr(out)))

    def hnsw_build(self, M=16, efc=200, seed=100, ratio=0, bmax=0, labels=None):
        """HierarchicalNSW(space, n, M, efc, seed) + addPoint(row i, label i) for every row, on the GPU."""
        info = np.zeros(8, np.uint64)
        lab = None if labels is None else np.ascontiguousarray(labels, np.uint64)
        self.h._check(_lib.b200ann_hnsw_build(self.h._h, self._ix, M, efc, seed, ratio, bmax, _ptr(lab), _ptr(info)))
        keys = ("beam_evals", "hops", "heuristic_evals", "batches", "upper_rows", "maxlevel", "ep", "largest_group")
        return dict(zip(keys, (int(x) for x in info)))

    def hnsw_insert_batch(self, data, labels=None, efc=0):
        """nb addPoint calls on the live graph: appends the rows of `data` and links them in place"""
        data = np.ascontiguousarray(data, self._qt)
        info = np.zeros(8, np.uint64)
        lab = None if labels is None else np.ascontiguousarray(labels, np.uint64)
        self.h._check(_lib.b200ann_hnsw_insert_batch(self.h._h, self._ix, _ptr(data), data.shape[1], _ptr(lab), data.shape[0],
                                                     efc, _ptr(info)))
        self.n += data.shape[0]
        keys = ("beam_evals", "hops", "heuristic_evals", "batches", "upper_rows", "maxlevel", "ep", "largest_group")
        return dict(zip(keys, (int(x) for x in info)))

    def hnsw_update_batch(self, ids, data):
        """updatePoint for the elements `ids` (distinct internal ids) with the new vectors `data`, in place"""
        ids = np.ascontiguousarray(ids, np.uint32)
        data = np.ascontiguousarray(data, self._qt)
        assert data.shape[0] == ids.shape[0]
        info = np.zeros(8, np.uint64)
        self.h._check(_lib.b200ann_hnsw_update_batch(self.h._h, self._ix, _ptr(ids), _ptr(data), data.shape[1], ids.shape[0],
                                                     _ptr(info)))
        keys = ("dist_evals", "hops", "heuristic_evals", "chunks", "lists_rewritten", "maxlevel", "ep", "largest_group")
        return dict(zip(keys, (int(x) for x in info)))

    def set_labels(self, ids, labels):
        """new labels for single elements (internal ids) of an index that holds an HNSW graph"""
        ids = np.ascontiguousarray(ids, np.uint32)
        labels = np.ascontiguousarray(labels, np.uint64)
        self.h._check(_lib.b200ann_index_set_labels(self._ix, _ptr(ids), _ptr(labels), ids.shape[0]))

    def get_hnsw(self):
        """the HNSW graph in the flat layout set_hnsw takes (dict like tests/refutil.parse_hnsw)"""
        sh = np.zeros(5, np.uint64)
        self.h._check(_lib.b200ann_index_hnsw_shape(self._ix, _ptr(sh)))
        maxM0, maxM, rows, maxlevel, ep = (int(x) for x in sh)
        links0 = np.empty((self.n, maxM0), np.uint32)
        levels = np.empty(self.n, np.int32)
        upper_ptr = np.empty(self.n, np.uint32)
        upper = np.empty((rows, maxM), np.uint32)
        labels = np.empty(self.n, np.uint64)
        self.h._check(_lib.b200ann_index_get_hnsw(self._ix, _ptr(links0), _ptr(levels), _ptr(upper_ptr),
                                                  _ptr(upper) if rows else None, _ptr(labels)))
        return dict(n=self.n, maxM=maxM, maxM0=maxM0, ep=ep, maxlevel=maxlevel, links0=links0, levels=levels,
                    upper_ptr=upper_ptr, upper_links=upper, labels=labels)

    def kdtree_graph(self, K, n_trees, m_level, seed=0, want_dists=False):
        """IndexKDtree::Build: initial kNN graph [n,K] from n_trees randomized KD-trees merged below depth m_level"""
        knn = np.empty((self.n, K), np.uint32)
        d = np.empty((self.n, K), np.float32) if want_dists else None
        info = np.zeros(8, np.uint64)
        self.h._check(_lib.b200ann_kdtree_graph(self.h._h, self._ix, K, n_trees, m_level, seed, _ptr(knn), _ptr(d), _ptr(info)))
        inf = dict(zip(("dist_evals", "nodes", "leaves", "deepest", "levels"), (int(x) for x in info)))
        return (knn, d, inf) if want_dists else (knn, inf)

    def nndescent_refine(self, init_graph, K, L, it, S, R, seed=0, want_dists=False):
        """IndexGraph::RefineGraph: NN-descent started from the rows of init_graph [n, init_k] (u32, INVALID padding)"""
        g = np.ascontiguousarray(init_graph, np.uint32)
        assert g.shape[0] == self.n
        knn = np.empty((self.n, K), np.uint32)
        d = np.empty((self.n, K), np.float32) if want_dists else None
        info = np.zeros(8, np.uint64)
        self.h._check(_lib.b200ann_nndescent_refine(self.h._h, self._ix, K, L, it, S, R, seed, _ptr(g), g.shape[1], _ptr(knn),
                                                    _ptr(d), _ptr(info)))
        keys = ("dist_evals", "offers", "inserts", "iterations", "join_us", "update_us")
        inf = dict(zip(keys, (int(x) for x in info)))
        return (knn, d, inf) if want_dists else (knn, inf)

    def nndescent(self, K, L, it, S, R, seed=0, want_dists=False, out=None):
        """IndexGraph::Build (NN-descent) on the GPU: returns knn [n,K] (+ dists) and the info dict.
        out: optional device buffer [n,K] (torch tensor / raw pointer) instead of a numpy result."""
        knn = out if out is not None else np.empty((self.n, K), np.uint32)
        d = np.empty((self.n, K), np.float32) if want_dists else None
        info = np.zeros(8, np.uint64)
        self.h._check(_lib.b200ann_nndescent(self.h._h, self._ix, K, L, it, S, R, seed, _ptr(knn), _ptr(d), _ptr(info)))
        keys = ("dist_evals", "offers", "inserts", "iterations", "join_us", "update_us")
        inf = dict(zip(keys, (int(x) for x in info)))
        return (knn, d, inf) if want_dists else (knn, inf)

    # ---- NSG construction (IndexNSG::Build and its phases)
    def nsg_build(self, knn, L, R, C_, seed=0, ep0=INVALID):
        """knn: numpy [n,kstride] u32, or (device pointer / torch tensor, kstride)"""
        kptr, kstride = (knn if isinstance(knn, tuple) else (np.ascontiguousarray(knn, np.uint32), knn.shape[1]))
        info = np.zeros(8, np.uint64)
        self.h._check(_lib.b200ann_nsg_build(self.h._h, self._ix, _ptr(kptr), kstride, L, R, C_, seed, ep0, _ptr(info)))
        keys = ("ep", "width", "tree_grow_edges", "beam_evals", "hops", "prune_evals", "interinsert_evals")
        return dict(zip(keys, (int(x) for x in info)))

    def nsg_init_ep(self, knn, L, seed=0, ep0=INVALID):
        knn = np.ascontiguousarray(knn, np.uint32)
        ep = _u32()
        center = np.empty(self.dim, np.float32)
        self.h._check(_lib.b200ann_nsg_init_ep(self.h._h, self._ix, _ptr(knn), knn.shape[1], L, seed, ep0, C.byref(ep),
                                               _ptr(center)))
        return ep.value, center

    def nsg_link(self, knn, ep, L, R, C_, seed=0, want_stats=False):
        knn = np.ascontiguousarray(knn, np.uint32)
        ci = np.empty((self.n, R), np.uint32)
        cd = np.empty((self.n, R), np.float32)
        st = np.zeros((self.n, 3), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_nsg_link(self.h._h, self._ix, _ptr(knn), knn.shape[1], ep, L, R, C_, seed, _ptr(ci),
                                            _ptr(cd), _ptr(st)))
        return (ci, cd, st) if want_stats else (ci, cd)

    def nsg_interinsert(self, cut_ids, cut_d):
        ci = np.ascontiguousarray(cut_ids, np.uint32).copy()
        cd = np.ascontiguousarray(cut_d, np.float32).copy()
        ne = C.c_uint64(0)
        self.h._check(_lib.b200ann_nsg_interinsert(self.h._h, self._ix, ci.shape[1], _ptr(ci), _ptr(cd), C.byref(ne)))
        return ci, cd, ne.value

    def nsg_finish(self, cut_ids, cut_d, ep, L, seed=0, slack=14):
        ci = np.ascontiguousarray(cut_ids, np.uint32)
        cd = np.ascontiguousarray(cut_d, np.float32)
        added = _u32()
        self.h._check(_lib.b200ann_nsg_finish(self.h._h, self._ix, _ptr(ci), _ptr(cd), ci.shape[1], ep, L, seed, slack,
                                              C.byref(added)))
        return added.value

    # ---- graph search
    def nsg_search(self, q, K, L, seed=0, want_stats=False):
        q = np.ascontiguousarray(q, np.float32)
        ids = np.empty((q.shape[0], K), np.uint32)
        d = np.empty((q.shape[0], K), np.float32)
        st = np.zeros((q.shape[0], 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_nsg_search(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1], K, L, seed,
                                              _ptr(ids), _ptr(d), _ptr(st)))
        return (ids, d, st) if want_stats else (ids, d)

    def hnsw_search(self, q, k, ef, want_stats=False):
        q = np.ascontiguousarray(q, self._qt)
        lab = np.empty((q.shape[0], k), np.uint64)
        d = np.empty((q.shape[0], k), self._dt)
        st = np.zeros((q.shape[0], 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_hnsw_search(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1], k, ef, _ptr(lab),
                                               _ptr(d), _ptr(st)))
        return (lab, d, st) if want_stats else (lab, d)

    def hnsw_search_filtered(self, q, k, ef, allow, want_stats=False):
        """searchKnn(query, k, isIdAllowed) batched. allow: bool / 0-1 array [n] by internal id (row), True = the
        element may be returned (passes the filter and is not deleted). -> (labels, dists, inexact[, stats]); short
        rows are padded with ~0 / inf; inexact[q] = 1: the bounded candidate set overflowed for that query."""
        q = np.ascontiguousarray(q, self._qt)
        a = np.asarray(allow).astype(bool)
        assert a.shape == (self.n,)
        bits = np.packbits(a, bitorder="little")
        words = np.zeros(((self.n + 31) // 32) * 4, np.uint8)
        words[:len(bits)] = bits
        words = words.view(np.uint32)
        lab = np.empty((q.shape[0], k), np.uint64)
        d = np.empty((q.shape[0], k), self._dt)
        bad = np.zeros(q.shape[0], np.uint8)
        st = np.zeros((q.shape[0], 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_hnsw_search_filtered(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1], k, ef,
                                                        _ptr(words), _ptr(lab), _ptr(d), _ptr(st), _ptr(bad)))
        return (lab, d, bad, st) if want_stats else (lab, d, bad)

    def hnsw_search_epsilon(self, q, epsilon, min_num_candidates, max_num_candidates, allow=None, internal_ids=False,
                            want_stats=False):
        """searchStopConditionClosest(query, EpsilonSearchStopCondition(epsilon, min, max), filter) batched
        (hnswalg.h:1327-1378, stop_condition.h:219-275). -> (labels [nq,max], dists [nq,max], counts [nq][, stats]);
        row q holds counts[q] results ascending by distance, all within epsilon; the rest is ~0 / inf.
        internal_ids=True returns internal ids like the reference does."""
        q = np.ascontiguousarray(q, self._qt)
        words = None
        if allow is not None:
            a = np.asarray(allow).astype(bool)
            assert a.shape == (self.n,)
            bits = np.packbits(a, bitorder="little")
            words = np.zeros(((self.n + 31) // 32) * 4, np.uint8)
            words[:len(bits)] = bits
            words = words.view(np.uint32)
        mx = int(max_num_candidates)
        lab = np.empty((q.shape[0], mx), np.uint64)
        d = np.empty((q.shape[0], mx), self._dt)
        cnt = np.zeros(q.shape[0], np.uint32)
        st = np.zeros((q.shape[0], 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_hnsw_search_epsilon(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1],
                                                       float(epsilon), int(min_num_candidates), mx, _ptr(words),
                                                       1 if internal_ids else 0, _ptr(lab), _ptr(d), _ptr(cnt), _ptr(st)))
        return (lab, d, cnt, st) if want_stats else (lab, d, cnt)

    def hnsw_search_multivector(self, q, doc_ids, num_docs, ef_collection, max_results, allow=None, internal_ids=False,
                                want_stats=False):
        """searchStopConditionClosest(query, MultiVectorSearchStopCondition(space, num_docs, ef_collection), filter)
        batched (stop_condition.h:146-215). doc_ids [n] u32 by internal id. -> (labels [nq,max_results], dists,
        counts[, stats]); row q holds counts[q] results ascending by distance covering num_docs documents."""
        q = np.ascontiguousarray(q, self._qt)
        doc = np.ascontiguousarray(doc_ids, np.uint32)
        assert doc.shape == (self.n,)
        words = None
        if allow is not None:
            a = np.asarray(allow).astype(bool)
            bits = np.packbits(a, bitorder="little")
            words = np.zeros(((self.n + 31) // 32) * 4, np.uint8)
            words[:len(bits)] = bits
            words = words.view(np.uint32)
        mx = int(max_results)
        lab = np.empty((q.shape[0], mx), np.uint64)
        d = np.empty((q.shape[0], mx), self._dt)
        cnt = np.zeros(q.shape[0], np.uint32)
        st = np.zeros((q.shape[0], 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_hnsw_search_multivector(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1], _ptr(doc),
                                                           int(num_docs), int(ef_collection), mx, _ptr(words),
                                                           1 if internal_ids else 0, _ptr(lab), _ptr(d), _ptr(cnt), _ptr(st)))
        return (lab, d, cnt, st) if want_stats else (lab, d, cnt)

    def nsg_search_from(self, q, entry, K, L, seed=0, want_stats=False):
        """IndexNSG::SearchFromEnterpoint batched: query i starts at entry[i]"""
        q = np.ascontiguousarray(q, np.float32)
        entry = np.ascontiguousarray(entry, np.uint32)
        ids = np.empty((q.shape[0], K), np.uint32)
        d = np.empty((q.shape[0], K), np.float32)
        st = np.zeros((q.shape[0], 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_nsg_search_from(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1], K, L, seed,
                                                   _ptr(entry), _ptr(ids), _ptr(d), _ptr(st)))
        return (ids, d, st) if want_stats else (ids, d)

    def nsg_search_clustered(self, q, ep, base, size, K, L, seed=0, want_stats=False):
        """query i searches the cluster (ep[i], base[i], size[i]) of a block-diagonal graph; size 0 = skip"""
        q = np.ascontiguousarray(q, np.float32)
        nq = q.shape[0]
        ep, base, size = (np.ascontiguousarray(a, np.uint32) for a in (ep, base, size))
        ids = np.empty((nq, K), np.uint32)
        d = np.empty((nq, K), np.float32)
        st = np.zeros((nq, 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_nsg_search_clustered(self.h._h, self._ix, _ptr(q), nq, q.shape[1], K, L, seed, _ptr(ep),
                                                        _ptr(base), _ptr(size), _ptr(ids), _ptr(d), _ptr(st)))
        return (ids, d, st) if want_stats else (ids, d)

    def hybrid_search(self, q, K, L, seed=0, want_stats=False):
        """HNSW_NSG::searchKnn batched (K = max(k, ef), L = L_search) -> ids, dists, entry labels"""
        q = np.ascontiguousarray(q, np.float32)
        ids = np.empty((q.shape[0], K), np.uint32)
        d = np.empty((q.shape[0], K), np.float32)
        entry = np.empty(q.shape[0], np.uint32)
        st = np.zeros((q.shape[0], 2), np.uint32) if want_stats else None
        self.h._check(_lib.b200ann_hybrid_search(self.h._h, self._ix, _ptr(q), q.shape[0], q.shape[1], K, L, seed,
                                                 _ptr(ids), _ptr(d), _ptr(entry), _ptr(st)))
        return (ids, d, entry, st) if want_stats else (ids, d, entry)

    def hybrid_search_raw(self, q, nq, qstride, K, L, seed, ids, dists=None, entry_out=None, stats=None):
        self.h._check(_lib.b200ann_hybrid_search(self.h._h, self._ix, _ptr(q), nq, qstride, K, L, seed, _ptr(ids),
                                                 _ptr(dists), _ptr(entry_out), _ptr(stats)))

    # raw (pointer-level) entry for device-resident buffers: torch tensors or ints
    def nsg_search_raw(self, q, nq, qstride, K, L, seed, ids, dists=None, stats=None):
        self.h._check(_lib.b200ann_nsg_search(self.h._h, self._ix, _ptr(q), nq, qstride, K, L, seed, _ptr(ids),
                                              _ptr(dists), _ptr(stats)))


def lib():
    return _lib

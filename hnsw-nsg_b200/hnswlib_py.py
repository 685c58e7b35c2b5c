"""Python face with the names and behaviour of the reference's hnswlib Python bindings, over
the C ABI: `from hnsw-nsg_b200 import hnswlib_py as hnswlib` is the whole migration for a script that used

    hnswlib.Index(space, dim) / init_index / add_items / knn_query / set_ef / save_index / load_index / get_items
    hnswlib.BFIndex(space, dim) / init_index / add_items / knn_query / delete_vector / save_index / load_index

Reference: hnswlib/python_bindings/bindings.cpp:151-700 (Index<float>), :723-910 (BFIndex<float>), :914-1011 (the
pybind11 names, argument names and defaults mirrored here). Same argument meaning, return shapes / dtypes
(labels uint64 [n,k], distances float32 [n,k]) and error behaviour (RuntimeError with the reference's messages).

Everything that computes runs on the GPU: add_items stages vectors on the host (the reference copies them too,
hnswalg.h:1204); the graph is built by b200ann_hnsw_build (same level sequence as HierarchicalNSW for the seed,
batched insertion) on the first knn_query / save_index after an add; knn_query is ONE b200ann_hnsw_search launch
for the whole query batch. `num_threads` arguments are accepted and ignored. There is no CPU path: without the
CUDA library every compute call raises.

knn_query(filter=...) and mark_deleted / unmark_deleted run the reference's searchBaseLayerST<false> on the GPU
(b200ann_hnsw_search_filtered): the Python callable is evaluated once per call over all labels into a bitmap.
add_items(replace_deleted=True) reuses the slots of deleted elements (hnswalg.h:954-992, needs
init_index(allow_replace_deleted=True)); the graph is rebuilt on the next query."""
import importlib
import pickle  # noqa: F401  (Index/BFIndex define __getstate__/__setstate__)

import numpy as np

_SPACES = ("l2", "ip", "cosine")
_handles = {}


def _B():
    return importlib.import_module(__package__)


def _handle(device=0):
    """one library handle (stream + scratch) per device, shared by every index of this module"""
    h = _handles.get(device)
    if h is None:
        h = _handles[device] = _B().Handle(device)
    return h


def _as_rows(data, dim):
    """bindings.cpp:83-99 get_input_array_shapes: 1-D input is one row; anything else must be [rows, dim]"""
    a = np.asarray(data, dtype=np.float32)
    if a.ndim == 1:
        a = a[None, :]
    if a.ndim != 2:
        raise RuntimeError("Input vector data wrong shape. Number of dimensions %d. Data must be a 1D or 2D array." % a.ndim)
    if a.shape[1] != dim:
        raise RuntimeError("Wrong dimensionality of the vectors")
    return np.ascontiguousarray(a)


def _as_ids(ids, rows, first):
    """bindings.cpp:102-127 get_input_ids_and_check_shapes; default labels are cur_l + row, cur_l = rows added so far"""
    if ids is None:
        return np.arange(first, first + rows, dtype=np.uint64)
    a = np.asarray(ids)
    if a.ndim == 0 and rows == 1:
        a = a[None]
    if a.ndim != 1 or a.shape[0] != rows:
        raise RuntimeError("The input label shape %d does not match the input data vector shape %d" % (a.ndim, 1))
    return a.astype(np.uint64)


def _normalize(a):
    """bindings.cpp:240-247 normalize_vector, row-wise in fp32 like the reference"""
    norm = np.sqrt((a * a).sum(1, dtype=np.float32)).astype(np.float32)
    return (a * (np.float32(1.0) / (norm + np.float32(1e-30)))[:, None]).astype(np.float32)


# ---------------------------------------------------------------- HierarchicalNSW::saveIndex / loadIndex file format
# hnswlib/hnswlib/hnswalg.h:685-713 (save), :716-822 (load). 96-byte header, then cur_element_count rows of
# [u16 count, 2 pad][maxM0 x u32 links][dim x f32][u64 label], then per element [u32 bytes][levels x (u16 count, 2 pad,
# maxM x u32)].
_INVALID = 0xFFFFFFFF


def _write_hnsw_file(path, g, vec, max_elements, M, efc, deleted=None):
    """g: the flat dict Index.get_hnsw() returns; vec [n,dim] fp32 as stored (normalised for cosine); deleted [n] bool:
    markDelete's DELETE_MARK, bit 0 of byte 2 of the level-0 list header (hnswalg.h:21, :864-878)"""
    n, dim = vec.shape
    maxM, maxM0 = int(g["maxM"]), int(g["maxM0"])
    sl0 = maxM0 * 4 + 4
    sdpe = sl0 + dim * 4 + 8
    hdr = np.zeros(96, np.uint8)
    for i, v in enumerate((0, max_elements, n, sdpe, sl0 + dim * 4, sl0)):
        hdr[8 * i:8 * i + 8] = np.array([v], np.uint64).view(np.uint8)
    hdr[48:52] = np.array([g["maxlevel"] if n else -1], np.int32).view(np.uint8)
    hdr[52:56] = np.array([g["ep"] if n else 0], np.uint32).view(np.uint8)
    for off, v in ((56, maxM), (64, maxM0), (72, M), (88, efc)):
        hdr[off:off + 8] = np.array([v], np.uint64).view(np.uint8)
    hdr[80:88] = np.array([1.0 / np.log(float(M))], np.float64).view(np.uint8)
    blob = np.zeros((n, sdpe), np.uint8)
    if n:
        l0 = np.ascontiguousarray(g["links0"], np.uint32)
        cnt = (l0 != _INVALID).sum(1).astype(np.uint16)           # rows are packed first, INVALID padded
        blob[:, 0:2] = cnt.view(np.uint8).reshape(n, 2)
        if deleted is not None:
            blob[:, 2] = np.asarray(deleted, bool).astype(np.uint8)
        blob[:, 4:sl0] = np.where(l0 == _INVALID, 0, l0).astype(np.uint32).view(np.uint8).reshape(n, maxM0 * 4)
        blob[:, sl0:sl0 + dim * 4] = np.ascontiguousarray(vec, np.float32).view(np.uint8).reshape(n, dim * 4)
        blob[:, sl0 + dim * 4:] = np.ascontiguousarray(g["labels"], np.uint64).view(np.uint8).reshape(n, 8)
    slpe = maxM * 4 + 4
    levels = np.asarray(g["levels"], np.int64)
    tail = np.zeros(4 * n + int(levels.sum()) * slpe, np.uint8)
    if n:
        pos = 4 * np.arange(n, dtype=np.int64) + slpe * (np.cumsum(levels) - levels)
        sz = (levels * slpe).astype(np.uint32).view(np.uint8).reshape(n, 4)
        tail[pos[:, None] + np.arange(4)] = sz
        up = np.ascontiguousarray(g["upper_links"], np.uint32)
        if up.shape[0]:
            owner = np.repeat(np.arange(n), levels)               # upper rows are stored element by element, level 1 first
            lvl = np.arange(len(owner)) - np.repeat(np.cumsum(levels) - levels, levels)
            src = np.asarray(g["upper_ptr"], np.int64)[owner] + lvl
            rows = up[src]
            rb = np.zeros((len(owner), slpe), np.uint8)
            rb[:, 0:2] = (rows != _INVALID).sum(1).astype(np.uint16).view(np.uint8).reshape(-1, 2)
            rb[:, 4:] = np.where(rows == _INVALID, 0, rows).astype(np.uint32).view(np.uint8).reshape(-1, maxM * 4)
            off = pos[owner] + 4 + lvl * slpe
            tail[off[:, None] + np.arange(slpe)] = rb
    with open(path, "wb") as f:
        f.write(hdr.tobytes())
        f.write(blob.tobytes())
        f.write(tail.tobytes())


def _read_hnsw_file(path, dim):
    try:
        raw = np.fromfile(path, np.uint8)
    except (FileNotFoundError, OSError):
        raise RuntimeError("Cannot open file")                    # hnswalg.h:720
    q = lambda off: int(raw[off:off + 8].view(np.uint64)[0])
    off_l0, maxel, n, sdpe, label_off, off_data = [q(8 * i) for i in range(6)]
    maxlevel = int(raw[48:52].view(np.int32)[0])
    ep = int(raw[52:56].view(np.uint32)[0])
    maxM, maxM0, M, efc = q(56), q(64), q(72), q(88)
    if (label_off - off_data) // 4 != dim:
        raise RuntimeError("index file dimension does not match the space")
    blob = raw[96:96 + n * sdpe].reshape(n, sdpe)
    cnt = blob[:, off_l0:off_l0 + 2].copy().view(np.uint16).reshape(n).astype(np.int64)
    deleted = (blob[:, off_l0 + 2] & 1).astype(bool)             # hnswalg.h:811-815
    l0 = blob[:, off_l0 + 4:off_l0 + 4 + 4 * maxM0].copy().view(np.uint32).reshape(n, maxM0)
    links0 = np.where(np.arange(maxM0)[None, :] < cnt[:, None], l0, _INVALID).astype(np.uint32)
    vec = blob[:, off_data:off_data + 4 * dim].copy().view(np.float32).reshape(n, dim)
    labels = blob[:, label_off:label_off + 8].copy().view(np.uint64).reshape(n)
    slpe = maxM * 4 + 4
    levels = np.zeros(n, np.int32)
    starts = np.zeros(n, np.int64)
    pos = 96 + n * sdpe
    for i in range(n):                                            # sizes chain: each element's tail length is in the file
        sz = int(raw[pos]) | int(raw[pos + 1]) << 8 | int(raw[pos + 2]) << 16 | int(raw[pos + 3]) << 24
        starts[i] = pos + 4
        levels[i] = sz // slpe
        pos += 4 + sz
    lv64 = levels.astype(np.int64)
    upper_ptr = (np.cumsum(lv64) - lv64).astype(np.uint32)
    owner = np.repeat(np.arange(n), lv64)
    if len(owner):
        lvl = np.arange(len(owner)) - np.repeat(upper_ptr.astype(np.int64), lv64)
        off = starts[owner] + lvl * slpe
        rb = raw[off[:, None] + np.arange(slpe)]
        c = rb[:, 0:2].copy().view(np.uint16).reshape(-1).astype(np.int64)
        r = rb[:, 4:].copy().view(np.uint32).reshape(-1, maxM)
        upper = np.where(np.arange(maxM)[None, :] < c[:, None], r, _INVALID).astype(np.uint32)
    else:
        upper = np.zeros((0, maxM), np.uint32)
    upper_ptr = np.where(levels > 0, upper_ptr, 0).astype(np.uint32)
    return dict(n=n, M=M, maxM=maxM, maxM0=maxM0, efc=efc, ep=ep, maxlevel=maxlevel, max_elements=maxel,
                links0=np.ascontiguousarray(links0), levels=levels, upper_ptr=upper_ptr,
                upper_links=np.ascontiguousarray(upper), labels=np.ascontiguousarray(labels),
                vectors=np.ascontiguousarray(vec), deleted=deleted)


class Index:
    """hnswlib.Index (bindings.cpp:151-700): HNSW over float32 vectors, built and searched on the GPU."""

    def __init__(self, space, dim, device=0):
        if space not in _SPACES:
            raise RuntimeError("Space name must be one of l2, ip, or cosine.")   # bindings.cpp:172
        self.space, self.dim = space, int(dim)
        self._device = device
        self._normalize = space == "cosine"                      # :165-170: cosine = inner product over unit vectors
        self._inited = False
        self._ef = 10                                             # default_ef, :181
        self.num_threads = 1
        self._ix = None
        self._dirty = True
        self._graph = None                                        # flat HNSW dict when loaded from a file
        self._built = 0                                           # rows linked into the device graph
        self._upd, self._relab = set(), set()                     # linked rows whose vector / label changed since

    # ---- construction
    def init_index(self, max_elements, M=16, ef_construction=200, random_seed=100, allow_replace_deleted=False):
        if self._inited:
            raise RuntimeError("The index is already initiated.")  # :198
        self._max = int(max_elements)
        self._M = min(int(M), 10000)                              # hnswalg.h:98-104
        self._efc = max(int(ef_construction), self._M)            # :107
        self._seed = int(random_seed)
        self._allow_replace = bool(allow_replace_deleted)
        self._vec = np.zeros((0, self.dim), np.float32)
        self._labels = np.zeros(0, np.uint64)
        self._deleted = np.zeros(0, bool)
        self._lookup = {}
        self._cur_l = 0
        self._inited = True

    def _need_init(self):
        if not self._inited:
            raise RuntimeError("Index is not initialized")

    def add_items(self, data, ids=None, num_threads=-1, replace_deleted=False):
        """bindings.cpp:250-330. A label that is already present replaces that element's vector (the reference's
        updatePoint, hnswalg.h:1158-1175). On the next query new rows are linked into the live device graph
        (b200ann_hnsw_insert_batch) and moved rows re-linked in place (b200ann_hnsw_update_batch); nothing is rebuilt."""
        self._need_init()
        if self._graph is not None and self._ix is None:
            self._build()                                         # a loaded graph goes to the device before it is extended
        if replace_deleted and not self._allow_replace:           # hnswalg.h:955-957
            raise RuntimeError("Replacement of deleted elements is disabled in constructor")
        # replace_deleted (hnswalg.h:954-992): a new point takes the slot (internal id, vector, label) of an element
        # marked deleted; the reference picks an unspecified one (unordered_set::begin), here the smallest internal id
        vacant = np.flatnonzero(self._deleted).tolist() if replace_deleted else []
        vacant.reverse()
        rows = _as_rows(data, self.dim)
        labels = _as_ids(ids, len(rows), self._cur_l)
        self._cur_l += len(rows)                                  # :328
        if self._normalize:
            rows = _normalize(rows)
        new_rows, new_labels = [], []
        try:
            for r, lab in enumerate(labels.tolist()):
                if vacant:
                    at = vacant.pop()
                    self._lookup.pop(int(self._labels[at]), None)    # :982-987
                    self._labels[at] = lab
                    self._lookup[lab] = at
                    self._deleted[at] = False                     # unmarkDeletedInternal, :989
                    self._vec[at] = rows[r]
                    if at < self._built:                          # :989-991 updatePoint on the reused slot
                        self._upd.add(at); self._relab.add(at)
                    continue
                at = self._lookup.get(lab)
                if at is not None:
                    if self._allow_replace and at < len(self._vec) and self._deleted[at]:   # :1163-1167
                        raise RuntimeError("Can't use addPoint to update deleted elements if replacement of deleted "
                                           "elements is enabled.")
                    if at < len(self._vec):
                        self._vec[at] = rows[r]
                        self._deleted[at] = False                 # hnswalg.h:1166-1168: an update un-deletes
                        if at < self._built:
                            self._upd.add(at)
                    else:
                        new_rows[at - len(self._vec)] = r
                    continue
                if len(self._labels) + len(new_labels) >= self._max:
                    raise RuntimeError("The number of elements exceeds the specified limit")   # hnswalg.h:1187
                self._lookup[lab] = len(self._labels) + len(new_labels)
                new_rows.append(r)
                new_labels.append(lab)
        finally:                                                  # rows accepted before an error stay, like the reference
            if new_rows:
                self._vec = np.concatenate([self._vec, rows[new_rows]])
                self._labels = np.concatenate([self._labels, np.array(new_labels, np.uint64)])
                self._deleted = np.concatenate([self._deleted, np.zeros(len(new_labels), bool)])
            self._dirty = True
            self._graph = None

    def _build(self):
        """vectors -> device, graph built by b200ann_hnsw_build (or installed from a loaded file)"""
        self._need_init()
        if not self._dirty and self._ix is not None:
            return
        B = _B()
        if self._ix is not None and self._built > 0:              # a live graph: append and update in place
            n = len(self._vec)
            if n > self._built:
                self._ix.hnsw_insert_batch(self._vec[self._built:], labels=self._labels[self._built:], efc=self._efc)
            if self._upd:
                ids = np.array(sorted(self._upd), np.uint32)
                self._ix.hnsw_update_batch(ids, self._vec[ids])
                if self._relab:
                    rl = np.array(sorted(self._relab), np.uint32)
                    self._ix.set_labels(rl, self._labels[rl])
            self._built = n
            self._upd, self._relab = set(), set()
            self._dirty = False
            return
        if self._ix is not None:
            self._ix.close()
            self._ix = None
        self._built = 0
        self._upd, self._relab = set(), set()
        if len(self._vec) == 0:
            self._dirty = False
            return
        metric = B.L2 if self.space == "l2" else B.IP
        ix = B.Index(_handle(self._device), self._vec, metric)
        if self._graph is not None:
            g = self._graph
            ix.set_hnsw(g["links0"], g["levels"], g["upper_ptr"], g["upper_links"], g["labels"], g["ep"], g["maxlevel"])
        else:
            ix.hnsw_build(self._M, self._efc, self._seed, labels=self._labels)
        self._ix = ix
        self._built = len(self._vec)
        self._dirty = False

    # ---- search
    def knn_query(self, data, k=1, num_threads=-1, filter=None):
        """bindings.cpp:612-698 -> (labels uint64 [rows,k], distances float32 [rows,k]), nearest first"""
        self._need_init()
        q = _as_rows(data, self.dim)
        if self._normalize:
            q = _normalize(q)
        self._build()
        k = int(k)
        short = "Cannot return the results in a contiguous 2D array. Probably ef or M is too small"
        if self._ix is None or k > len(self._labels):
            raise RuntimeError(short)
        if filter is None and not self._deleted.any():            # hnswalg.h:1306 bare_bone_search
            try:
                lab, d = self._ix.hnsw_search(q, k, max(self._ef, k))  # searchKnn uses max(ef_, k), :1309
            except _B().B200AnnError as e:
                if e.code == -4:                                  # B200ANN_ERR_FEWER: the reference's message, :646-648
                    raise RuntimeError(short)
                raise
        else:
            allow = ~self._deleted
            if filter is not None:                                # CustomFilterFunctor, bindings.cpp:129-142
                allow &= np.fromiter((bool(filter(l)) for l in self._labels.tolist()), bool, len(self._labels))
            lab, d, bad = self._ix.hnsw_search_filtered(q, k, max(self._ef, k), allow)
            if bad.any():
                raise RuntimeError("knn_query: the filter is too selective for the graph walk on %d queries (candidate "
                                   "set overflow); use BFIndex.knn_query(filter=...) for an exact scan" % int(bad.sum()))
        if (lab[:, k - 1] == np.uint64(0xFFFFFFFFFFFFFFFF)).any():
            raise RuntimeError(short)
        return lab, d

    def set_ef(self, ef):
        self._ef = int(ef)

    ef = property(lambda self: self._ef, lambda self, v: self.set_ef(v))

    def set_num_threads(self, num_threads):
        self.num_threads = int(num_threads)

    # ---- data access
    def get_items(self, ids=None, return_type="numpy"):
        """bindings.cpp:333-365: stored vectors (normalised ones for cosine) of the given labels"""
        if return_type not in ("numpy", "list"):
            raise ValueError("Invalid return type")
        self._need_init()
        if ids is None:
            out = np.zeros((0, self.dim), np.float32)
        else:
            a = np.asarray(ids)
            if a.ndim == 0:
                raise ValueError("ids must be an array, not a scalar")
            try:
                rows = [self._lookup[int(x)] for x in a.reshape(-1).tolist()]
            except KeyError:
                raise RuntimeError("Label not found")             # hnswalg.h:831
            if any(r < len(self._deleted) and self._deleted[r] for r in rows):
                raise RuntimeError("Label not found")             # :832 a deleted element is not returned either
            out = self._vec[rows].copy()
        return out if return_type == "numpy" else out.tolist()

    def get_ids_list(self):
        return [int(x) for x in self._labels.tolist()]

    def get_max_elements(self):
        return self._max

    def get_current_count(self):
        return len(self._labels)

    def resize_index(self, new_size):
        """hnswalg.h:640-675 resizeIndex"""
        self._need_init()
        if new_size < len(self._labels):
            raise RuntimeError("Cannot resize, max element is less than the current number of elements")
        self._max = int(new_size)

    def _row_of(self, label):
        self._need_init()
        at = self._lookup.get(int(label))
        if at is None:
            raise RuntimeError("Label not found")                 # hnswalg.h:859
        return at

    def mark_deleted(self, label):
        """hnswalg.h:853-885 markDelete: the element stays in the graph, is walked through and never returned"""
        at = self._row_of(label)
        if self._deleted[at]:
            raise RuntimeError("The requested to delete element is already deleted")
        self._deleted[at] = True

    def unmark_deleted(self, label):
        """hnswalg.h:893-922 unmarkDelete"""
        at = self._row_of(label)
        if not self._deleted[at]:
            raise RuntimeError("The requested to undelete element is not deleted")
        self._deleted[at] = False

    M = property(lambda self: self._M if self._inited else 0)
    ef_construction = property(lambda self: self._efc if self._inited else 0)
    max_elements = property(lambda self: self._max if self._inited else 0)
    element_count = property(lambda self: len(self._labels) if self._inited else 0)

    # ---- files
    def save_index(self, path_to_index):
        """HierarchicalNSW::saveIndex format: the reference (CPU) loads what this writes and vice versa"""
        self._build()
        if self._ix is None:
            g = dict(maxM=self._M, maxM0=2 * self._M, maxlevel=-1, ep=0, links0=np.zeros((0, 2 * self._M), np.uint32),
                     levels=np.zeros(0, np.int32), upper_ptr=np.zeros(0, np.uint32),
                     upper_links=np.zeros((0, self._M), np.uint32), labels=self._labels)
        else:
            g = self._ix.get_hnsw()
        _write_hnsw_file(path_to_index, g, self._vec, self._max, self._M, self._efc, self._deleted)

    def index_file_size(self):
        self._build()
        n = len(self._labels)
        lv = int(self._ix.get_hnsw()["levels"].sum()) if self._ix is not None else 0
        return 96 + n * (2 * self._M * 4 + 4 + self.dim * 4 + 8) + 4 * n + lv * (self._M * 4 + 4)

    def load_index(self, path_to_index, max_elements=0, allow_replace_deleted=False):
        g = _read_hnsw_file(path_to_index, self.dim)
        if self._ix is not None:
            self._ix.close()
            self._ix = None
        self._max = max(int(max_elements), int(g["max_elements"]))
        self._M, self._efc = int(g["M"]), int(g["efc"])
        self._seed = 100
        self._allow_replace = bool(allow_replace_deleted)
        self._vec, self._labels = g["vectors"], g["labels"]
        self._deleted = g["deleted"].copy()
        self._lookup = {int(l): i for i, l in enumerate(self._labels.tolist())}
        self._graph = g
        self._cur_l = len(self._labels)                           # :236
        self._inited = True
        self._dirty = True

    # ---- pickle (bindings.cpp:380-480 keeps the same parameters + the serialised index)
    def __getstate__(self):
        st = dict(space=self.space, dim=self.dim, device=self._device, ef=self._ef, num_threads=self.num_threads,
                  inited=self._inited)
        if self._inited:
            self._build()
            st.update(cur_l=self._cur_l, max=self._max, M=self._M, efc=self._efc, seed=self._seed, vec=self._vec, labels=self._labels, deleted=self._deleted,
                      allow_replace_deleted=self._allow_replace,   # bindings.cpp:491, :567 pickle it too
                      graph=self._ix.get_hnsw() if self._ix is not None else None)
        return st

    def __setstate__(self, st):
        Index.__init__(self, st["space"], st["dim"], st["device"])
        self._ef, self.num_threads = st["ef"], st["num_threads"]
        if st["inited"]:
            self.init_index(st["max"], st["M"], st["efc"], st["seed"], st.get("allow_replace_deleted", False))
            self._vec, self._labels, self._deleted = st["vec"], st["labels"], st["deleted"]
            self._lookup = {int(l): i for i, l in enumerate(self._labels.tolist())}
            self._graph = st["graph"]
            self._cur_l = st["cur_l"]

    def __repr__(self):
        return "<hnswlib.Index(space='%s', dim=%d)>" % (self.space, self.dim)

    def close(self):
        if self._ix is not None:
            self._ix.close()
            self._ix = None


class BFIndex:
    """hnswlib.BFIndex (bindings.cpp:725-910) over hnswlib::BruteforceSearch (bruteforce.h:10-172): exact scan with
    top-k on the GPU (b200ann_bf_knn). Like the reference, cosine normalises the stored vectors only (:795-803);
    knn_query passes the queries through as they are (:874-876)."""

    def __init__(self, space, dim, device=0):
        if space not in _SPACES:
            raise RuntimeError("Space name must be one of l2, ip, or cosine.")
        self.space, self.dim = space, int(dim)
        self._device = device
        self._normalize = space == "cosine"
        self._inited = False
        self.num_threads = 1
        self._ix = None
        self._dirty = True

    def init_index(self, max_elements):
        if self._inited:
            raise RuntimeError("The index is already initiated.")
        self._max = int(max_elements)
        self._vec = np.zeros((self._max, self.dim), np.float32)
        self._labels = np.zeros(self._max, np.uint64)
        self._count = 0
        self._cur_l = 0
        self._lookup = {}
        self._inited = True

    def _need_init(self):
        if not self._inited:
            raise RuntimeError("Index is not initialized")

    def add_items(self, data, ids=None):
        self._need_init()
        rows = _as_rows(data, self.dim)
        labels = _as_ids(ids, len(rows), self._cur_l)
        if self._normalize:
            rows = _normalize(rows)
        for r, lab in enumerate(labels.tolist()):                 # bruteforce.h:61-83 addPoint
            at = self._lookup.get(lab)
            if at is None:
                if self._count >= self._max:
                    raise RuntimeError("The number of elements exceeds the specified limit\n")
                at = self._count
                self._count += 1
                self._lookup[lab] = at
            self._labels[at] = lab
            self._vec[at] = rows[r]
        self._cur_l += len(rows)
        self._dirty = True

    def delete_vector(self, label):
        """bruteforce.h:86-103 removePoint: the last element takes the hole; an unknown label is ignored"""
        self._need_init()
        at = self._lookup.pop(int(label), None)
        if at is None:
            return
        last = self._count - 1
        if at != last:
            self._vec[at] = self._vec[last]
            self._labels[at] = self._labels[last]
            self._lookup[int(self._labels[at])] = at
        self._count -= 1
        self._dirty = True

    def _upload(self):
        B = _B()
        if not self._dirty and self._ix is not None:
            return
        if self._ix is not None:
            self._ix.close()
            self._ix = None
        if self._count:
            self._ix = B.Index(_handle(self._device), self._vec[:self._count], B.L2 if self.space == "l2" else B.IP)
        self._dirty = False

    def knn_query(self, data, k=1, num_threads=-1, filter=None):
        """bindings.cpp:851-905. filter(label) -> bool keeps a label (bruteforce.h:110-135): the scan then runs over the
        kept rows only. Slots past the number of available elements hold label 0 / distance inf (the reference leaves
        them uninitialised)."""
        self._need_init()
        q = _as_rows(data, self.dim)
        k = int(k)
        lab = np.zeros((len(q), k), np.uint64)
        d = np.full((len(q), k), np.inf, np.float32)
        if filter is None:
            self._upload()
            ix, labels, n = self._ix, self._labels, self._count
        else:
            B = _B()
            keep = np.array([i for i in range(self._count) if filter(int(self._labels[i]))], np.int64)
            n = len(keep)
            labels = self._labels[keep]
            ix = B.Index(_handle(self._device), np.ascontiguousarray(self._vec[keep]),
                         B.L2 if self.space == "l2" else B.IP) if n else None
        kk = min(k, n)
        if kk:
            ids, dd = ix.bf_knn(q, kk)
            lab[:, :kk] = labels[ids]
            d[:, :kk] = dd
        if filter is not None and ix is not None:
            ix.close()
        return lab, d

    def set_num_threads(self, num_threads):
        self.num_threads = int(num_threads)

    def get_max_elements(self):
        return self._max

    def get_current_count(self):
        return self._count

    def save_index(self, path_to_index):
        """bruteforce.h:139-151: [u64 maxelements][u64 size_per_element][u64 count] + maxelements x (vector, u64 label)"""
        self._need_init()
        spe = self.dim * 4 + 8
        blob = np.zeros((self._max, spe), np.uint8)
        blob[:, :self.dim * 4] = self._vec.view(np.uint8).reshape(self._max, self.dim * 4)
        blob[:, self.dim * 4:] = self._labels.view(np.uint8).reshape(self._max, 8)
        with open(path_to_index, "wb") as f:
            f.write(np.array([self._max, spe, self._count], np.uint64).tobytes())
            f.write(blob.tobytes())

    def load_index(self, path_to_index, max_elements=0):
        """bruteforce.h:154-171"""
        try:
            raw = np.fromfile(path_to_index, np.uint8)
        except (FileNotFoundError, OSError):
            raise RuntimeError("Cannot open file")
        mx, spe, cnt = (int(v) for v in raw[:24].view(np.uint64))
        if spe != self.dim * 4 + 8:
            raise RuntimeError("index file dimension does not match the space")
        blob = raw[24:24 + mx * spe].reshape(mx, spe)
        self._max, self._count, self._cur_l = mx, cnt, cnt
        self._vec = np.ascontiguousarray(blob[:, :self.dim * 4]).view(np.float32).reshape(mx, self.dim).copy()
        self._labels = np.ascontiguousarray(blob[:, self.dim * 4:]).view(np.uint64).reshape(mx).copy()
        self._lookup = {int(self._labels[i]): i for i in range(cnt)}
        self._inited = True
        self._dirty = True

    def __repr__(self):
        return "<hnswlib.BFIndex(space='%s', dim=%d)>" % (self.space, self.dim)

    def close(self):
        if self._ix is not None:
            self._ix.close()
            self._ix = None

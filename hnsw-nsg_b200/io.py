"""fvecs / ivecs / centroid files in the reference's layouts (CNNS/src/utils/aux_util.cpp:8-84), numpy only.
fvecs: per row [i32 dim][f32 x dim]; ivecs (ground truth, kNN graphs): per row [u32 k][u32 x k];
centroids.fvecs: [i32 n_clusters][i32 m][u32 dim] then n_clusters*(m+1) fvecs rows."""
import numpy as np


def load_fvecs(path):
    raw = np.fromfile(path, np.uint8)
    if raw.size == 0:
        return np.zeros((0, 0), np.float32)
    dim = int(raw[:4].view(np.int32)[0])
    rows = raw.size // ((dim + 1) * 4)          # the reference ignores a ragged tail the same way (aux_util.cpp:19)
    a = raw[:rows * (dim + 1) * 4].view(np.float32).reshape(rows, dim + 1)
    return np.ascontiguousarray(a[:, 1:])


def save_fvecs(path, x):
    x = np.ascontiguousarray(x, np.float32)
    n, dim = x.shape
    out = np.empty((n, dim + 1), np.float32)
    out[:, 0] = np.array([dim], np.int32).view(np.float32)[0]
    out[:, 1:] = x
    out.tofile(path)


def load_ivecs(path):
    """equal-length rows -> [n,k] uint32 array; ragged rows (loadGT accepts them) -> list of arrays"""
    raw = np.fromfile(path, np.uint32)
    rows, pos = [], 0
    while pos < raw.size:
        k = int(raw[pos])
        if pos + 1 + k > raw.size:
            break                                 # truncated last row: dropped, like loadGT (aux_util.cpp:45-46)
        rows.append(raw[pos + 1:pos + 1 + k])
        pos += 1 + k
    if rows and all(len(r) == len(rows[0]) for r in rows):
        return np.ascontiguousarray(np.stack(rows))
    return rows


def save_ivecs(path, ids):
    ids = np.ascontiguousarray(ids, np.uint32)
    n, k = ids.shape
    out = np.empty((n, k + 1), np.uint32)
    out[:, 0] = k
    out[:, 1:] = ids
    out.tofile(path)


def load_centroids(path):
    """-> (points [n_clusters*(m+1), dim], n_clusters, m): each cluster's centroid followed by its m samples"""
    raw = np.fromfile(path, np.uint8)
    n_clusters, m = (int(v) for v in raw[:8].view(np.int32))
    dim = int(raw[8:12].view(np.uint32)[0])
    total = n_clusters * (m + 1)
    a = raw[12:12 + total * (dim + 1) * 4].view(np.float32).reshape(total, dim + 1)
    if not (a[:, 0].view(np.uint32) == dim).all():
        raise ValueError("Dimension mismatch in centroids file")
    return np.ascontiguousarray(a[:, 1:]), n_clusters, m


def save_centroids(path, points, n_clusters, m):
    points = np.ascontiguousarray(points, np.float32)
    dim = points.shape[1]
    assert points.shape[0] == n_clusters * (m + 1)
    with open(path, "wb") as f:
        np.array([n_clusters, m], np.int32).tofile(f)
        np.array([dim], np.uint32).tofile(f)
        out = np.empty((points.shape[0], dim + 1), np.float32)
        out[:, 0] = np.array([dim], np.uint32).view(np.float32)[0]
        out[:, 1:] = points
        out.tofile(f)

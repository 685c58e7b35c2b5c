"""world_size-2 gloo test of the sharded-search host logic (hnsw-nsg_b200/sharded.py): shard rows,
search each shard, all-gather the per-shard lists, merge == brute force over the whole base.
CPU only: the per-shard search and the merge are the oracle here; on GPUs they are b200ann kernels."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import refutil as R


def _sharded():
    spec = importlib.util.spec_from_file_location("b200ann_sharded", os.path.join(R.ROOT, "hnsw-nsg_b200", "sharded.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _worker(rank, world, port, n_per, dim, nq, k, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S = _sharded()
    base = R.data_uniform(n_per * world, dim, 1)
    q = R.data_uniform(nq, dim, 2)
    shard = np.ascontiguousarray(base[rank * n_per:(rank + 1) * n_per])
    ids, d = R.orc_bf_knn(shard, q, k, R.L2_SCALAR)            # local ids of this shard
    if rank == 1:
        ids[0, -1] = R.INVALID; d[0, -1] = np.inf                # an unfilled slot must stay INVALID

    def merge(gd, gi, kout):
        od, oi = R.orc_merge_topk(gd.numpy(), gi.numpy().view(np.uint32), kout)
        return od, oi

    od, oi = S.sharded_topk(torch.from_numpy(d), torch.from_numpy(ids.view(np.int32)), rank, n_per, k, merge)
    tid, td = R.orc_bf_knn(base, q, k, R.L2_SCALAR)
    ok = (oi[1:] == tid[1:]).all() and (od[1:] == td[1:]).all() and R.INVALID not in oi
    out[rank] = bool(ok)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_topk_world2():
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        port = 29500 + (os.getpid() % 2000)
        mp.spawn(_worker, args=(world, port, 1500, 16, 40, 10, out), nprocs=world, join=True)
        assert out[0] and out[1]


def test_merge_oracle_matches_numpy():
    rng = np.random.default_rng(0)
    parts, nq, k = 3, 20, 7
    d = np.sort(rng.random((parts, nq, k), dtype=np.float32), axis=2)
    ids = rng.permutation(parts * nq * k).astype(np.uint32).reshape(parts, nq, k)
    od, oi = R.orc_merge_topk(d, ids, 10)
    for q in range(nq):
        alld = d[:, q].ravel(); alli = ids[:, q].ravel()
        o = np.lexsort((alli, alld))[:10]
        assert (oi[q] == alli[o]).all() and (od[q] == alld[o]).all()

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


SHARDS = 4


def _run(rank, world, n_per, dim, nq, k):
    """the sharded search of bench.py's C5 step with the oracle in place of the kernels: -> (out_d, out_i)"""
    S = _sharded()
    base = R.data_uniform(n_per * SHARDS, dim, 1)
    q = R.data_uniform(nq, dim, 2)
    own = S.owned_shards(rank, world, SHARDS)
    lp = len(own)
    block, bd, bi = S.alloc_block(lp, nq, k, "cpu")
    for l, s in enumerate(own):
        shard = np.ascontiguousarray(base[s * n_per:(s + 1) * n_per])
        ids, d = R.orc_bf_knn(shard, q, k, R.L2_SCALAR)          # local ids of this shard
        ids = ids.astype(np.int64) + s * n_per                    # what b200ann_set_id_base does in the kernel epilogue
        if s == 1:
            ids[0, -1] = R.INVALID; d[0, -1] = np.inf             # an unfilled slot must stay INVALID
        bd[l] = torch.from_numpy(d)
        bi[l] = torch.from_numpy(ids.astype(np.uint32).view(np.int32))
    g = S.exchange(block, world)
    gd, gi = S.split_gathered(g, lp, nq, k)
    od, oi = R.orc_merge_topk(gd.numpy(), gi.numpy().view(np.uint32), k)
    return base, q, od, oi


def _worker(rank, world, port, n_per, dim, nq, k, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    base, q, od, oi = _run(rank, world, n_per, dim, nq, k)
    tid, td = R.orc_bf_knn(base, q, k, R.L2_SCALAR)
    ok = (oi[1:] == tid[1:]).all() and (od[1:] == td[1:]).all() and R.INVALID not in oi
    out[rank] = (bool(ok), oi.tobytes(), od.tobytes())
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_topk_world2_equals_world1():
    """4 shards over 2 ranks (gloo): the merged lists equal brute force over the whole base AND the lists one rank
    holding all 4 shards produces - the result does not depend on how many GPUs the shards are spread over"""
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        port = 29500 + (os.getpid() % 2000)
        mp.spawn(_worker, args=(world, port, 900, 16, 40, 10, out), nprocs=world, join=True)
        assert out[0][0] and out[1][0]
        assert out[0][1] == out[1][1] and out[0][2] == out[1][2]
        _, _, od, oi = _run(0, 1, 900, 16, 40, 10)
        assert oi.tobytes() == out[0][1] and od.tobytes() == out[0][2]


def test_merge_oracle_matches_numpy():
    rng = np.random.default_rng(0)
    parts, nq, k = 3, 20, 7
    d = np.sort(rng.random((parts, nq, k), dtype=np.float32), axis=2)
    ids = rng.permutation(parts * nq * k).astype(np.uint32).reshape(parts, nq, k)
    od, oi = R.orc_merge_topk(d, ids, 10)
    for q in range(nq):
        alld = d[:, q].ravel(); alli = ids[:, q].ravel()
        o = np.argsort(alld, kind="stable")[:10]   # (distance, part, position)
        assert (oi[q] == alli[o]).all() and (od[q] == alld[o]).all()

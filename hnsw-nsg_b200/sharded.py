"""Row-sharded search across ranks (one process per GPU): every rank searches its own sub-index,
per-shard [nq,k] (distance, global id) lists are exchanged with ONE all-gather and merged.
Backend agnostic (`nccl` on GPUs, `gloo` in the CPU tests); the merge itself is injected:
on a GPU it is Handle.merge_topk_raw (csrc/merge.cu).
Reference analogue: the per-cluster search + host merge of CNNS/tests/cluster_hnsw_nsg_search.cpp:189-270."""
import torch
import torch.distributed as dist


def to_global(local_ids, rank, rows_per_rank, invalid=-1):
    """local row ids (int32 bit pattern of u32) -> global ids; INVALID (0xFFFFFFFF == -1) stays INVALID"""
    return torch.where(local_ids == invalid, local_ids, local_ids + rank * rows_per_rank)


def gather_lists(d, ids):
    """d, ids [nq,k] on every rank -> ([world,nq,k], [world,nq,k]) on every rank, part r = rank r"""
    world = dist.get_world_size()
    nq = d.shape[0]
    gd = torch.empty((world * nq,) + tuple(d.shape[1:]), dtype=d.dtype, device=d.device)
    gi = torch.empty((world * nq,) + tuple(ids.shape[1:]), dtype=ids.dtype, device=ids.device)
    dist.all_gather_into_tensor(gd, d.contiguous())   # concatenation along dim 0 == [world, nq, k]
    dist.all_gather_into_tensor(gi, ids.contiguous())
    return gd.view((world,) + tuple(d.shape)), gi.view((world,) + tuple(ids.shape))


def sharded_topk(d, local_ids, rank, rows_per_rank, kout, merge):
    """merge(gd, gi, kout) -> (out_d, out_i): the kout smallest (distance,id) of the union per query"""
    gi_local = to_global(local_ids, rank, rows_per_rank)
    gd, gi = gather_lists(d, gi_local)
    return merge(gd, gi, kout)

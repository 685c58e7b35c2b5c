"""Row-sharded search across ranks (one process per GPU): the base is cut into `shards` row blocks, rank r owns the
lp = shards / world consecutive shards r*lp .. r*lp + lp - 1 (so block order == shard order == row order and the merged
result, ties included, is the same for every world size); every rank searches the same query batch on the shards it owns, writing (distance, global id)
lists straight into its BLOCK of a send buffer; ONE all-gather exchanges the blocks and every rank merges them.

Block layout (32-bit words): [distances: lp x nq x k float32][ids: lp x nq x k uint32], lp = shards per rank - what
b200ann_merge_topk_gathered (csrc/merge.cu) reads after the collective. Backend agnostic (`nccl` on GPUs, `gloo` in
the CPU tests). The merged result does not depend on the number of ranks the shards are spread over.
Reference analogue: the per-cluster search + host merge of CNNS/tests/cluster_hnsw_nsg_search.cpp:189-270."""
import torch
import torch.distributed as dist


def owned_shards(rank, world, shards):
    """shards of rank `rank`: rank*lp .. rank*lp + lp - 1 with lp = shards / world (shards % world must be 0)"""
    assert shards % world == 0, "the shard count must be a multiple of the world size"
    lp = shards // world
    return list(range(rank * lp, (rank + 1) * lp))


def alloc_block(lp, nq, k, device):
    """-> (block int32 [2*lp*nq*k], d float32 view [lp,nq,k], ids int32 view [lp,nq,k] (u32 bit patterns))"""
    block = torch.empty(2 * lp * nq * k, dtype=torch.int32, device=device)
    half = lp * nq * k
    d = block[:half].view(torch.float32).view(lp, nq, k)
    ids = block[half:].view(lp, nq, k)
    return block, d, ids


def exchange(block, world, out=None):
    """all-gather of every rank's block -> [world, len(block)] (world == 1: the block itself, no collective)"""
    if world == 1:
        return block.view(1, -1)
    if out is None:
        out = torch.empty((world, block.numel()), dtype=block.dtype, device=block.device)
    dist.all_gather_into_tensor(out.view(-1), block)
    return out


def split_gathered(g, lp, nq, k):
    """gathered blocks [world, 2*lp*nq*k] -> (d [world*lp,nq,k] float32, ids [world*lp,nq,k] int32), part = rank*lp + l"""
    world = g.shape[0]
    half = lp * nq * k
    d = g[:, :half].contiguous().view(torch.float32).view(world * lp, nq, k)
    ids = g[:, half:].contiguous().view(world * lp, nq, k)
    return d, ids

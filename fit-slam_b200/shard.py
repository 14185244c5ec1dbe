"""Candidate sharding across ranks: the host-side half of the multi-GPU path.

The costmap, the map points and the frontier list are replicated (extraction is deterministic, every rank finds the same
list); what shards is the scoring of the (frontier, yaw) candidates.

In a cycle (fitslam_cycle_begin / _end) rank r owns the frontiers f with f % nranks == r: ownership is known the moment
a frontier exists, so a rank scores its share while the search is still running.  Every rank writes 16-byte score
records {int32 information, int32 max_index, float32 fisher, uint32 flags} for its frontiers into a zeroed buffer of one
record per frontier, and `allreduce_records_` (integer sum) leaves record f at index f on every rank.

The stand-alone scoring calls on a caller's list (fitslam_score_arrival / fitslam_score_fisher under fitslam_set_shard) use
contiguous blocks instead: rank r owns [r * B, min(F, (r + 1) * B)), B = ceil(F / nranks) -- the same arithmetic as
`shard_block` / `shard_range` in csrc/ctx.h -- with the records in its block of an (nranks * B)-record buffer, merged by
the in-place all-gather `allgather_records_`.
There is no other exchange on the path (the reference has none at all: it is single-process, DEP/src/ExplorationBT.cpp).
"""
import numpy as np

RECORD_DTYPE = np.dtype([("information", np.int32), ("max_index", np.int32), ("fisher", np.float32), ("flags", np.uint32)])
FLAG_ACHIEVABLE, FLAG_FOOTPRINT, FLAG_RAY_FAILED, FLAG_FISHER_VALID, FLAG_FISHER_SAFE = 1, 2, 4, 8, 16


def shard_block(n, nranks):
    """Records per rank: ceil(n / nranks)."""
    return (int(n) + int(nranks) - 1) // int(nranks)


def shard_range(n, rank, nranks):
    """(first, count) of the frontiers rank `rank` scores; count is 0 for ranks past the end."""
    b = shard_block(n, nranks)
    lo = rank * b
    if lo >= n:
        return n, 0
    return lo, min(lo + b, n) - lo


def pack_records(arrival, fisher=None, fisher_safe=None):
    """fitslam_arrival fields (+ optional Fisher information per frontier) -> the 16-byte records."""
    n = len(arrival["information"])
    rec = np.zeros(n, RECORD_DTYPE)
    rec["information"] = arrival["information"]
    rec["max_index"] = arrival["max_index"]
    flags = (np.asarray(arrival["achievable"], np.uint32) & 1) * FLAG_ACHIEVABLE
    flags |= (np.asarray(arrival["footprint_in_lethal"], np.uint32) & 1) * FLAG_FOOTPRINT
    flags |= (np.asarray(arrival["ray_failed"], np.uint32) & 1) * FLAG_RAY_FAILED
    if fisher is not None:
        rec["fisher"] = fisher
        flags |= FLAG_FISHER_VALID
        if fisher_safe is not None:
            flags |= (np.asarray(fisher_safe, np.uint32) & 1) * FLAG_FISHER_SAFE
    rec["flags"] = flags
    return rec


def allgather_records_(rec, rank, world, group=None):
    """In-place all-gather of the record buffer `rec` (a torch tensor of world * B rows, any device): every rank
    contributes rows [rank * B, (rank + 1) * B).  One collective per cycle; NCCL on the GPUs, gloo in the CPU tests."""
    import torch.distributed as dist
    if world == 1:
        return rec
    assert rec.shape[0] % world == 0, "the record buffer holds world * ceil(F / world) rows"
    b = rec.shape[0] // world
    dist.all_gather_into_tensor(rec, rec[rank * b:(rank + 1) * b], group=group)
    return rec


def owner(f, nranks):
    """Rank that scores frontier f in a sharded cycle (fitslam_cycle_begin): interleaved, so that ownership is known the
    moment a frontier exists and a rank can score its share while the search is still running."""
    return f % nranks


def allreduce_records_(rec, world, group=None):
    """Merge of the cycle's record buffer `rec` (a torch int32 tensor of F rows x 4, any device): every rank has written
    the rows of the frontiers it owns into a zeroed buffer, so an integer sum is a merge (the float field travels as its
    bit pattern).  One collective per cycle; NCCL on the GPUs, gloo in the CPU tests."""
    import torch.distributed as dist
    if world == 1:
        return rec
    dist.all_reduce(rec, op=dist.ReduceOp.SUM, group=group)
    return rec

"""Read-batch sharding across the GPUs of one box: contiguous slices of the query / pair batch per
rank, the index replicated, results concatenated on the host. No collective on the data path; torch.distributed is
used only to replicate the index once and to gather results where one process wants them all."""
import numpy as np


def shard_bounds(n, world, rank):
    """[begin, end) of rank's contiguous slice: ceil(n / world) items per rank, the last ranks may be short or empty"""
    per = (n + world - 1) // world
    b = min(n, rank * per)
    return b, min(n, b + per)


def shard_string_set(symbols, offsets, world, rank):
    """slice of a ConcatenatedStringSet (8-bit symbols + count+1 offsets): rebased offsets, symbol sub-array"""
    n = len(offsets) - 1
    b, e = shard_bounds(n, world, rank)
    off = np.asarray(offsets[b:e + 1], dtype=np.uint32)
    if len(off) == 0:
        return np.zeros(0, dtype=np.uint8), np.zeros(1, dtype=np.uint32), (b, e)
    sym = np.asarray(symbols[int(off[0]):int(off[-1])])
    return sym, (off - off[0]).astype(np.uint32), (b, e)


def concat_results(parts):
    """host-side concatenation of per-rank result arrays in rank order"""
    parts = [p for p in parts if p is not None and len(p)]
    if not parts:
        return np.zeros(0)
    return np.concatenate(parts, axis=0)


def gather_to_rank0(local, group=None):
    """gathers variable-length per-rank numpy results on rank 0 (gloo or nccl); returns the concatenation there, None elsewhere"""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    parts = [None] * world if rank == 0 else None
    dist.gather_object(local, parts, dst=0, group=group)
    return concat_results(parts) if rank == 0 else None

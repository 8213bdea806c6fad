"""Multi-GPU host logic: one process per GPU, pictures (or streams) sharded, no data-path collective.

A PAME job needs only SOURCE-picture pyramids (Source/Lib/Codec/me_process.c:248-262 resolves
`ref_pa_pic_ptr_array` to pa-reference objects, never to reconstructed pictures), so every non-intra
picture of a sequence is an independent job once the source frames exist. The reference runs them
concurrently on its ME thread pool; here they are dealt round-robin in decode order to the ranks
(SURVEY.md §8e). Each rank builds the pyramids it needs from the source frames itself — building is
~1 % of a step, cheaper than any exchange — and only result tables travel (to rank 0, host side).

Works with any torch.distributed backend (`nccl` on the GPU box, `gloo` in the CPU tests).
"""
from typing import Callable, Dict, List, Sequence, Tuple

import numpy as np


def rank_plan(n_jobs: int, world_size: int, rank: int) -> List[int]:
    """Indices (decode order) of the jobs rank `rank` owns: i % world_size == rank."""
    if not (0 <= rank < world_size):
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    return list(range(rank, n_jobs, world_size))


def pictures_for(specs: Sequence[Tuple[object, int, Dict[Tuple[int, int], int]]], idx: Sequence[int]) -> List[int]:
    """Picture numbers whose pyramids must be resident on the rank that runs jobs `idx`."""
    need = set()
    for i in idx:
        _, cur, refs = specs[i]
        need.add(cur)
        need.update(refs.values())
    return sorted(need)


def stream_plan(n_streams: int, world_size: int, rank: int) -> List[int]:
    """Stream sharding (BASELINE configs[4]): stream s -> rank s % world_size."""
    return list(range(rank, n_streams, world_size))


def run_sharded(specs, run_local: Callable[[List[int]], Dict[int, Dict[str, np.ndarray]]], dist=None):
    """Run the jobs of `specs` sharded over the initialised process group.

    run_local(indices) -> {job index: {array name: ndarray}} computes this rank's share.
    Returns on rank 0 the merged {job index: arrays} of ALL jobs (None on other ranks).
    """
    if dist is None or not dist.is_initialized():
        world, rank = 1, 0
    else:
        world, rank = dist.get_world_size(), dist.get_rank()
    mine = rank_plan(len(specs), world, rank)
    local = run_local(mine)
    if set(local) != set(mine):
        raise RuntimeError("rank %d returned jobs %s, owns %s" % (rank, sorted(local), mine))
    if world == 1:
        return local
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local, gathered, dst=0)
    if rank != 0:
        return None
    merged: Dict[int, Dict[str, np.ndarray]] = {}
    for part in gathered:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError("job %d computed twice" % k)
            merged[k] = v
    if sorted(merged) != list(range(len(specs))):
        raise RuntimeError("jobs missing after gather")
    return merged


def max_over_ranks(value_ms: float, dist=None, device=None) -> float:
    """Device time of a multi-GPU step = max over ranks (never wall clock)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value_ms)
    import torch
    t = torch.tensor([float(value_ms)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

"""Multi-GPU host logic: one process per GPU, pictures (or streams) sharded; the only data-path collective is
the optional broadcast of source-picture pyramids between GPUs (share_pyramids).

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


# ---------------------------------------------------------------- pyramids shared over NVLink
def pyramid_owner(picture_number: int, world_size: int) -> int:
    """The rank that uploads picture `picture_number` and builds its pyramid: picture % world_size."""
    return int(picture_number) % int(world_size)


def sharing_plan(needed_by_rank: Sequence[Sequence[int]], world_size: int) -> List[Tuple[int, int, List[int]]]:
    """needed_by_rank[r] = pictures whose pyramids rank r must hold. Returns, in picture order, the
    broadcasts to run: (picture, owner rank, ranks that need it besides the owner). Pictures nobody but the
    owner needs are left out. Every rank computes the same plan (collectives must be issued in one order)."""
    plan = []
    for pic in sorted({p for need in needed_by_rank for p in need}):
        owner = pyramid_owner(pic, world_size)
        users = [r for r in range(world_size) if r != owner and pic in set(needed_by_rank[r])]
        if users:
            plan.append((pic, owner, users))
    return plan


class _DeviceBytes:
    """A span of device memory as a CUDA-array-interface object (torch.as_tensor aliases it, no copy)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def share_pyramids(ses, plan, rank: int, dist, torch, slot_bytes: int, cache=None):
    """Run the transfers of `plan` (sharing_plan) on the initialised NCCL group: the owner's pyramid (the three
    padded planes of one ring slot, contiguous in HBM) goes straight into the slot the referencing ranks adopted
    for the picture — point-to-point `ncclSend` / `ncclRecv` over NVLink / NVSwitch, all pictures of the step in ONE
    NCCL group (one launch; 60 separate broadcasts cost 2 ms of launch latency at 1080p), no host copy, no
    rebuild. The caller orders this after the owners' pyramid builds (ses.sync()); the transfers are enqueued
    behind the current torch stream and the launches that follow wait for them. `cache` (a dict the caller keeps
    while the pictures stay resident) remembers the slot tensors: a ring slot does not move. Returns the bytes
    this rank sent + received."""
    ops, moved = [], 0
    cache = {} if cache is None else cache
    for pic, owner, users in plan:
        if rank != owner and rank not in users:
            continue
        t = cache.get(pic)
        if t is None:
            if rank != owner:
                ses.adopt_picture(pic)
            p0, n0 = ses.plane_device_ptr(pic, 0)
            p1, n1 = ses.plane_device_ptr(pic, 1)
            p2, n2 = ses.plane_device_ptr(pic, 2)
            if p1 != p0 + n0 or p2 != p1 + n1 or n0 + n1 + n2 != int(slot_bytes):
                raise RuntimeError("pyramid planes of a slot are expected to be contiguous")
            t = cache[pic] = torch.as_tensor(_DeviceBytes(p0, slot_bytes), device="cuda")
        if rank == owner:
            for u in users:
                ops.append(dist.P2POp(dist.isend, t, u))
                moved += slot_bytes
        else:
            ops.append(dist.P2POp(dist.irecv, t, owner))
            moved += slot_bytes
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return moved


# ---------------------------------------------------------------- pyramids all-gathered in place (pipelined form)
def gather_groups(pictures: Sequence[int], world_size: int) -> List[int]:
    """Picture p lives in ring slot p on every rank and is built by rank p % N: the N pictures of group g =
    [g N, (g + 1) N) sit in N consecutive slots, one built by each rank, and ONE in-place ncclAllGather over that
    span gives every rank the whole group. Returns the groups that hold `pictures`, in order of first appearance."""
    seen, order = set(), []
    for p in pictures:
        g = int(p) // world_size
        if g not in seen:
            seen.add(g)
            order.append(g)
    return order


def all_gather_groups(groups: Sequence[int], rank: int, world: int, dist, torch, pool_ptr: int, slot_bytes: int, cache=None,
                      pool=None) -> int:
    """In-place all-gather (NCCL over NVLink / NVSwitch) of the ring slots of every group in `groups`, enqueued behind
    the CURRENT torch stream; the caller brackets the call with MeSession.external_write_begin / _end on that stream,
    so nothing synchronises with the host. `pool`: a 1-D uint8 tensor that IS the ring (CPU tests with gloo); otherwise
    the ring is the device memory at pool_ptr. Returns the bytes this rank received."""
    cache = {} if cache is None else cache
    moved = 0
    for g in groups:
        t = cache.get(g)
        if t is None:
            if pool is not None:
                t = pool[g * world * slot_bytes:(g + 1) * world * slot_bytes]
            else:
                t = torch.as_tensor(_DeviceBytes(pool_ptr + g * world * slot_bytes, world * slot_bytes), device="cuda")
            cache[g] = t
        dist.all_gather_into_tensor(t, t[rank * slot_bytes:(rank + 1) * slot_bytes])
        moved += (world - 1) * slot_bytes
    return moved

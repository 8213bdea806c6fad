"""N > 1 host logic on CPU: two processes, gloo, 127.0.0.1. The per-rank compute is the oracle
(this is a test of sharding / gathering / timing reduction, not of the kernels)."""
import ctypes as C
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
PKG = "svt-av1-mod-by-patman_b200"
shard = importlib.import_module(PKG + ".shard")
gop = importlib.import_module(PKG + ".gop")

W, H, FRAMES = 192, 128, 7


def _specs():
    import _util as U
    lib = U.product_lib()
    specs = []
    for j in gop.ra_jobs(FRAMES, 2, 2, 1):
        d = U.derive_input(8, W, H, hierarchical_levels=2, temporal_layer_index=j.temporal_layer, is_ref=int(j.is_ref),
                           ref_list0_count_try=j.n_l0, ref_list1_count_try=j.n_l1, ref_count_used_list0=2,
                           ref_count_used_list1=1)
        p = U.abi.SvtMeParams()
        assert lib.svtme_derive_params(C.byref(d), C.byref(p)) == 0
        U.finish_params(p, d, j.picture_number, j.refs)
        specs.append((p, j.picture_number, j.refs))
    return specs


def _run_local(specs, seq, idx):
    import _util as U
    need = shard.pictures_for(specs, idx)
    pyr = {n: U.CpuPyramid("orc", seq[n], n) for n in need}
    out = {}
    for i in idx:
        p, cur, refs = specs[i]
        arrays = U.cpu_me_picture("orc", p, pyr[cur], {k: pyr[v] for k, v in refs.items()}, deep=False)
        out[i] = {k: np.array(v, copy=True) for k, v in arrays.items()}
    return out


def _worker(rank, world, port, q):
    import torch.distributed as dist
    import _util as U
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        specs = _specs()
        seq = U.synth_sequence(W, H, FRAMES, seed=3)
        merged = shard.run_sharded(specs, lambda idx: _run_local(specs, seq, idx), dist)
        t = shard.max_over_ranks(10.0 + rank, dist)
        if rank == 0:
            q.put(("ok", merged, t, shard.rank_plan(len(specs), world, rank)))
        else:
            assert merged is None
            q.put(("ok1", None, t, shard.rank_plan(len(specs), world, rank)))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_rank_plan_partitions_jobs():
    for world in (1, 2, 3, 8):
        seen = []
        for r in range(world):
            seen += shard.rank_plan(31, world, r)
        assert sorted(seen) == list(range(31))
    with pytest.raises(ValueError):
        shard.rank_plan(4, 2, 2)
    assert shard.stream_plan(64, 8, 3) == list(range(3, 64, 8))


def test_pictures_for_lists_current_and_references():
    specs = [(None, 4, {(0, 0): 0, (1, 0): 8}), (None, 2, {(0, 0): 0, (1, 0): 4})]
    assert shard.pictures_for(specs, [0]) == [0, 4, 8]
    assert shard.pictures_for(specs, [0, 1]) == [0, 2, 4, 8]


@pytest.mark.timeout(300)
def test_two_ranks_gloo_picture_sharding_matches_single_process():
    import torch.multiprocessing as mp
    import _util as U
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    r0 = next(g for g in got if g[0] == "ok")
    r1 = next(g for g in got if g[0] == "ok1")
    merged = r0[1]
    assert r0[2] == r1[2] == 11.0  # max over ranks of (10 + rank)
    assert sorted(r0[3] + r1[3]) == list(range(len(merged)))
    # single-process answer
    specs = _specs()
    seq = U.synth_sequence(W, H, FRAMES, seed=3)
    single = _run_local(specs, seq, list(range(len(specs))))
    assert sorted(merged) == sorted(single)
    for i in single:
        for k in single[i]:
            assert np.array_equal(single[i][k], merged[i][k]), (i, k)


def _gather_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        slot_bytes, n_slots = 48, 6  # three groups of two pictures; picture p sits in slot p and is built by rank p % 2
        pool = torch.zeros(n_slots * slot_bytes, dtype=torch.uint8)
        for p in range(n_slots):
            if shard.pyramid_owner(p, world) == rank:
                pool[p * slot_bytes:(p + 1) * slot_bytes] = torch.arange(slot_bytes, dtype=torch.uint8) + 7 * p
        groups = shard.gather_groups([4, 5, 0, 3], world)  # first-need order: groups 2, 0, 1
        moved = shard.all_gather_groups(groups, rank, world, dist, torch, 0, slot_bytes, pool=pool)
        q.put((rank, groups, moved, pool.numpy().copy()))
    finally:
        dist.destroy_process_group()


def test_gather_groups_order_and_membership():
    assert shard.gather_groups([32, 0, 16, 8, 4, 33, 7], 8) == [4, 0, 2, 1]
    assert shard.gather_groups([], 4) == []


@pytest.mark.timeout(300)
def test_two_ranks_gloo_in_place_all_gather_of_ring_slot_groups():
    """The pyramid exchange of the pipelined picture-sharded run, on CPU tensors: every rank owns one slot of each
    group of N consecutive slots; after the in-place all-gather of a group every rank holds all of its slots."""
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gather_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = np.concatenate([(np.arange(48) + 7 * p).astype(np.uint8) for p in range(6)])
    for rank, groups, moved, pool in got:
        assert groups == [2, 0, 1] and moved == 3 * 48
        assert np.array_equal(pool, want), rank

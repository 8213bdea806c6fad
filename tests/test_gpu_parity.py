"""GPU parity tests (-m gpu): everything goes through the C ABI of libsvtme_b200.so and is compared
bit-for-bit with the oracle restatement / the committed golden vectors. Nothing here reads
/root/reference."""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import _util as U
import make_golden as G

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
me = importlib.import_module(U.PKG + ".me")
shard = importlib.import_module(U.PKG + ".shard")


def gpu_stage(seq, p, cur, refs, W, H, deep=True):
    with me.MeSession(0, W, H, capacity=len(set(refs.values())) + 1) as ses:
        for n in set(refs.values()) | {cur}:
            ses.put_picture(n, seq[n])
        return ses.run_picture(p, cur, refs, deep=deep)


def derive(ses_or_lib, d):
    p = U.abi.SvtMeParams()
    assert U.product_lib().svtme_derive_params(C.byref(d), C.byref(p)) == 0
    return p


@pytest.mark.parametrize("case", G.STAGE_CASES, ids=[c[0] for c in G.STAGE_CASES])
def test_stage_matches_golden_and_oracle(case):
    name, W, H, preset, tl, refs, cur, kw = case
    seq, d = G.stage_case(*case)
    p = U.finish_params(derive(None, d), d, cur, refs)
    got = gpu_stage(seq, p, cur, refs, W, H)
    want = np.load(os.path.join(GOLD, "stage_%s.npz" % name))
    for k in want.files:
        assert np.array_equal(want[k], got[k]), "golden " + k
    pyr = {n: U.CpuPyramid("orc", seq[n], n) for n in set(refs.values()) | {cur}}
    orc = U.cpu_me_picture("orc", p, pyr[cur], {k: pyr[v] for k, v in refs.items()})
    for k in orc:
        assert np.array_equal(orc[k], got[k]), "oracle " + k


@pytest.mark.parametrize("case", [c for c in G.STAGE_CASES if c[3] in (4, 8)], ids=lambda c: c[0])
def test_stage_generic_engine_for_refinement_rounds(case, monkeypatch):
    """HME level 1 / 2 rounds through the generic staged engine instead of the warp-per-window
    refinement kernel (the path presets with wide refinement windows take): same golden result."""
    monkeypatch.setenv("SVTME_NO_REFINE_KERNEL", "1")
    monkeypatch.setenv("SVTME_NO_L0_STRIP", "1")
    name, W, H, preset, tl, refs, cur, kw = case
    seq, d = G.stage_case(*case)
    p = U.finish_params(derive(None, d), d, cur, refs)
    got = gpu_stage(seq, p, cur, refs, W, H)
    want = np.load(os.path.join(GOLD, "stage_%s.npz" % name))
    for k in want.files:
        assert np.array_equal(want[k], got[k]), "golden " + k


def test_strip_mode_segment_splitting_and_tiled_windows():
    """SVTME_STRIP_ARENA_LIMIT (read once when the library loads, hence a child process) shrinks the arena
    the strip kernels may fill: segments are halved until their union fits and single windows larger than
    the limit take the tiled path. Results must not change."""
    import subprocess
    code = r"""
import os, sys, ctypes as C, importlib, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r); sys.path.insert(0, %r)
import _util as U, make_golden as G
me = importlib.import_module(U.PKG + ".me")
for case in [c for c in G.STAGE_CASES if c[3] in (4, 8)]:
    name, W, H, preset, tl, refs, cur, kw = case
    seq, d = G.stage_case(*case)
    p = U.abi.SvtMeParams()
    assert U.product_lib().svtme_derive_params(C.byref(d), C.byref(p)) == 0
    p = U.finish_params(p, d, cur, refs)
    with me.MeSession(0, W, H, capacity=len(set(refs.values())) + 1) as ses:
        for n in set(refs.values()) | {cur}:
            ses.put_picture(n, seq[n])
        got = ses.run_picture(p, cur, refs, deep=True)
    want = np.load(os.path.join(%r, "stage_%%s.npz" %% name))
    for k in want.files:
        assert np.array_equal(want[k], got[k]), (name, k)
print("ok")
""" % (os.path.dirname(os.path.abspath(__file__)), GOLD, os.path.dirname(os.path.dirname(os.path.abspath(__file__))), GOLD)
    for limit in ("1024", "4096"):
        env = dict(os.environ, SVTME_STRIP_ARENA_LIMIT=limit)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "ok" in r.stdout, (limit, r.stdout[-500:], r.stderr[-2000:])


def test_10bit_sources_give_the_msb_pyramid():
    """svtme_put_pictures_10bit (host and device sources, aligned and odd strides): all three planes incl. pads equal
    the pyramid of (sample >> 2), which is what the oracle's unpack + pyramid give."""
    import torch
    W, H = 328, 200
    rng = np.random.default_rng(17)
    orc = U.oracle_lib()
    for stride in (W, W + 8, W + 3):
        src = rng.integers(0, 1024, size=(2, H, stride), dtype=np.uint16)
        src[1] |= rng.integers(0, 64, size=(H, stride), dtype=np.uint16) << 10  # junk above bit 9 wraps like the reference's cast
        msb = np.zeros((2, H, W), dtype=np.uint8)
        for n in range(2):
            orc.orc_unpack_msb(src[n].ctypes.data_as(C.c_void_p), C.c_uint32(stride), msb[n].ctypes.data_as(C.c_void_p), C.c_uint32(W),
                               C.c_uint32(W), C.c_uint32(H))
        dev = torch.from_numpy(src).cuda()
        with me.MeSession(0, W, H, capacity=6) as ses:
            ses.put_pictures([0, 1], [src[0].ctypes.data, src[1].ctypes.data], stride, bits=10)
            ses.put_pictures([2, 3], [dev[0].data_ptr(), dev[1].data_ptr()], stride, device=True, bits=10)
            ses.put_pictures([4, 5], [msb[0].ctypes.data, msb[1].ctypes.data], W)
            for n in range(2):
                pyr = U.CpuPyramid("orc", msb[n], n)
                for lvl in range(3):
                    want = ses.get_plane(4 + n, lvl)
                    assert np.array_equal(ses.get_plane(n, lvl), want), (stride, n, lvl, "host")
                    assert np.array_equal(ses.get_plane(2 + n, lvl), want), (stride, n, lvl, "device")
                    assert np.array_equal(pyr.plane(lvl)[:, :want.shape[1]], want), (stride, n, lvl, "oracle")


def test_pyramid_matches_golden_and_oracle():
    seq = U.synth_sequence(328, 200, 1, seed=5)
    want = np.load(os.path.join(GOLD, "pyramid.npz"))
    with me.MeSession(0, 328, 200, capacity=1) as ses:
        ses.put_picture(0, seq[0])
        for lvl, k in enumerate(("full", "quarter", "sixteenth")):
            got = ses.get_plane(0, lvl)
            assert np.array_equal(want[k][:, :got.shape[1]], got), k


@pytest.mark.parametrize("size", [(1920, 1080), (856, 480), (64, 64)])
def test_pyramid_sizes_vs_oracle(size):
    W, H = size
    img = U.random_frames(W, H, 1, seed=W)[0]
    pyr = U.CpuPyramid("orc", img, 0)
    with me.MeSession(0, W, H, capacity=1) as ses:
        ses.put_picture(0, img)
        for lvl in range(3):
            got = ses.get_plane(0, lvl)
            assert np.array_equal(pyr.plane(lvl)[:, :got.shape[1]], got), lvl


@pytest.mark.parametrize("size", [(856, 480), (328, 200)])
def test_batched_pyramid_build_host_and_device_sources(size):
    """svtme_put_pictures / svtme_put_pictures_device (2 launches per batch; aligned device rows are
    read in place, misaligned ones go through a copy) == the oracle pyramid of every picture."""
    import torch
    W, H = size
    n = 3
    imgs = U.random_frames(W, H, n, seed=H)
    want = [U.CpuPyramid("orc", imgs[i], i) for i in range(n)]
    with me.MeSession(0, W, H, capacity=3 * n) as ses:
        host = np.ascontiguousarray(imgs)
        ses.put_pictures(list(range(n)), [host[i].ctypes.data for i in range(n)], W, device=False)
        stride = (W + 15) // 16 * 16  # W % 16 != 0 for 856 / 328: pad the rows to keep them aligned
        dev = torch.zeros((n, H, stride), dtype=torch.uint8, device="cuda")
        dev[:, :, :W] = torch.from_numpy(host).cuda()
        ses.put_pictures([10 + i for i in range(n)], [dev[i].data_ptr() for i in range(n)], stride, device=True)
        mis = torch.zeros((n, H, stride + 3), dtype=torch.uint8, device="cuda")  # odd stride: copy path
        mis[:, :, :W] = torch.from_numpy(host).cuda()
        ses.put_pictures([20 + i for i in range(n)], [mis[i].data_ptr() for i in range(n)], stride + 3, device=True)
        torch.cuda.synchronize()
        for base in (0, 10, 20):
            for i in range(n):
                for lvl in range(3):
                    got = ses.get_plane(base + i, lvl)
                    assert np.array_equal(want[i].plane(lvl)[:, :got.shape[1]], got), (base, i, lvl)


@pytest.mark.parametrize("preset,tl,refs,cur", [
    (8, 3, {(0, 0): 2, (0, 1): 1, (0, 2): 0, (1, 0): 4, (1, 1): 5}, 3),
    (4, 2, {(0, 0): 2, (0, 1): 1, (0, 2): 0, (0, 3): 0, (1, 0): 4, (1, 1): 5, (1, 2): 6}, 3),
    (0, 1, {(0, 0): 0, (1, 0): 4}, 2),
    (13, 1, {(0, 0): 1, (1, 0): 3}, 2),
])
def test_stage_edges_and_ragged_sizes_vs_oracle(preset, tl, refs, cur):
    """Incomplete b64 rows/columns (height 1080-like remainder 56, width remainder 24) and motion
    that pushes search windows into the picture padding."""
    for (W, H), motion in [((472, 312), (9, -7)), ((216, 120), (-12, 10))]:
        seq = U.synth_sequence(W, H, 7, seed=preset + W, motion=motion, noise=3.0)
        l0 = 1 + max(ri for (li, ri) in refs if li == 0)
        l1 = 1 + max([ri for (li, ri) in refs if li == 1], default=-1)
        d = U.derive_input(preset, W, H, temporal_layer_index=tl, ref_list0_count_try=l0, ref_list1_count_try=l1,
                           ref_count_used_list0=l0, ref_count_used_list1=l1)
        p = U.finish_params(derive(None, d), d, cur, refs)
        got = gpu_stage(seq, p, cur, refs, W, H)
        pyr = {n: U.CpuPyramid("orc", seq[n], n) for n in set(refs.values()) | {cur}}
        orc = U.cpu_me_picture("orc", p, pyr[cur], {k: pyr[v] for k, v in refs.items()})
        for k in orc:
            assert np.array_equal(orc[k], got[k]), (W, H, k)


def test_dependent_references_with_small_areas_vs_oracle():
    """enable_me_sr_adjustment == 2 makes the full-pel area of every other reference depend on the 64x64 result of
    (list 0, ref 0) (motion_estimation.c:1345-1350). With small areas all references of a b64 run in ONE CTA of the
    lean kernel, in the reference's loop order; with large areas the two-round path is taken. Both against the oracle
    (the parameter set with the small area is synthetic: the rtc screen-content preset with its full-pel area cut to
    preset 8's)."""
    W, H = 472, 312
    refs, cur = {(0, 0): 2, (0, 1): 1, (0, 2): 0, (1, 0): 4, (1, 1): 5}, 3
    seq = U.synth_sequence(W, H, 7, seed=31, motion=(5, -3), noise=3.0)
    pyr = {n: U.CpuPyramid("orc", seq[n], n) for n in set(refs.values()) | {cur}}
    for small in (True, False):
        d = U.derive_input(10, W, H, temporal_layer_index=2, ref_list0_count_try=3, ref_list1_count_try=2, ref_count_used_list0=3,
                           ref_count_used_list1=2, rtc_tune=1, sc_class1=1)
        p = U.finish_params(derive(None, d), d, cur, refs)
        assert p.enable_me_sr_adjustment == 2
        if small:
            p.me_sa.sa_min.width, p.me_sa.sa_min.height, p.me_sa.sa_max.width, p.me_sa.sa_max.height = 9, 3, 9, 5
        got = gpu_stage(seq, p, cur, refs, W, H)
        orc = U.cpu_me_picture("orc", p, pyr[cur], {k: pyr[v] for k, v in refs.items()})
        for k in orc:
            assert np.array_equal(orc[k], got[k]), (small, k)


def test_stage_flat_and_random_pictures_tie_breaks():
    """All-equal pictures make every position tie (first-in-raster-order must win); i.i.d. random
    pictures give non-trivial minima everywhere (SadTest RANDOM pattern)."""
    W, H = 256, 192
    refs, cur = {(0, 0): 0, (0, 1): 1, (1, 0): 3}, 2
    for kind in ("flat", "random"):
        seq = np.full((4, H, W), 128, dtype=np.uint8) if kind == "flat" else U.random_frames(W, H, 4)
        for preset in (4, 9):
            d = U.derive_input(preset, W, H, temporal_layer_index=1, ref_list0_count_try=2, ref_list1_count_try=1,
                               ref_count_used_list0=2, ref_count_used_list1=1)
            p = U.finish_params(derive(None, d), d, cur, refs)
            got = gpu_stage(seq, p, cur, refs, W, H)
            pyr = {n: U.CpuPyramid("orc", seq[n], n) for n in range(4)}
            orc = U.cpu_me_picture("orc", p, pyr[cur], {k: pyr[v] for k, v in refs.items()})
            for k in orc:
                assert np.array_equal(orc[k], got[k]), (kind, preset, k)


def test_batched_jobs_equal_single_jobs():
    W, H = 320, 192
    seq = U.synth_sequence(W, H, 6, seed=11)
    d = U.derive_input(8, W, H, temporal_layer_index=1, ref_list0_count_try=1, ref_list1_count_try=1)
    with me.MeSession(0, W, H, capacity=6) as ses:
        for n in range(6):
            ses.put_picture(n, seq[n])
        specs = []
        for cur in (1, 2, 3, 4):
            p = U.finish_params(derive(None, d), d, cur, {(0, 0): cur - 1, (1, 0): cur + 1})
            specs.append((p, cur, {(0, 0): cur - 1, (1, 0): cur + 1}))
        jobs, outs = ses.make_jobs(specs, deep=True)
        ses.run_jobs(jobs)
        for (p, cur, refs), batch in zip(specs, outs):
            single = ses.run_picture(p, cur, refs, deep=True)
            for k in single:
                assert np.array_equal(single[k], batch[k]), (cur, k)


def test_queued_batches_equal_synchronous_runs_and_reuploads_are_ordered():
    """svtme_submit_pictures / svtme_collect_results: several batches in flight return, in submission
    order, exactly what the synchronous call returns; a slot re-uploaded with DIFFERENT content while
    earlier batches that read the old content are still queued must not disturb them (the upload stream
    waits for the slot's readers)."""
    W, H = 320, 192
    seqs = [U.synth_sequence(W, H, 6, seed=s) for s in (11, 12, 13)]
    d = U.derive_input(8, W, H, temporal_layer_index=1, ref_list0_count_try=1, ref_list1_count_try=1)
    specs = []
    for cur in (1, 2, 3, 4):
        p = U.finish_params(derive(None, d), d, cur, {(0, 0): cur - 1, (1, 0): cur + 1})
        specs.append((p, cur, {(0, 0): cur - 1, (1, 0): cur + 1}))
    import torch
    hosts = [torch.from_numpy(s).pin_memory() for s in seqs]  # pinned: the uploads are truly asynchronous
    with me.MeSession(0, W, H, capacity=6) as ses:
        want = []
        for seq in seqs:
            for n in range(6):
                ses.put_picture(n, seq[n])
            jobs, outs = ses.make_jobs(specs, deep=True)
            ses.run_jobs(jobs)
            want.append(outs)
        queued = []
        for rnd, hst in enumerate(hosts):  # three rounds back to back, same slots, different pictures
            ses.put_pictures(list(range(6)), [hst.data_ptr() + n * W * H for n in range(6)], W)
            a, oa = ses.make_jobs(specs[:2], deep=True)
            b, ob = ses.make_jobs(specs[2:], deep=True)
            ses.submit_jobs(a)
            ses.submit_jobs(b)
            queued += [(a, oa, rnd, 0), (b, ob, rnd, 2)]
        assert ses.pending_batches() == 6
        with pytest.raises(me.SvtMeError):
            ses.collect_jobs(queued[1][0] if len(queued[1][0]) != len(queued[0][0]) else ses.make_jobs(specs[:1])[0])
        for jobs, outs, rnd, first in queued:
            ses.collect_jobs(jobs)
            for i, o in enumerate(outs):
                for k in o:
                    assert np.array_equal(o[k], want[rnd][first + i][k]), (rnd, first + i, k)
        assert ses.pending_batches() == 0
        with pytest.raises(me.SvtMeError):
            ses.collect_jobs(queued[0][0])  # nothing submitted


def test_error_behaviour():
    lib = U.product_lib()
    h = C.c_void_p()
    assert lib.svtme_open(0, 100, 100, 2, C.byref(h)) == -1  # not a multiple of 8
    with me.MeSession(0, 128, 128, capacity=1) as ses:
        img = np.zeros((128, 128), dtype=np.uint8)
        ses.put_picture(0, img)
        with pytest.raises(me.SvtMeError):
            ses.put_picture(1, img)  # ring full
        d = U.derive_input(8, 128, 128)
        p = U.finish_params(derive(None, d), d, 0, {(0, 0): 5, (1, 0): 6})
        with pytest.raises(me.SvtMeError):
            ses.run_picture(p, 0, {(0, 0): 5, (1, 0): 6})  # references not resident
        with pytest.raises(me.SvtMeError):
            ses.release_picture(9)


def test_put_is_transactional_when_the_ring_fills_up():
    """a batched put that fails at picture i > 0 (ring full, null picture) must leave no trace: the
    earlier pictures of the call are NOT resident (a later launch on them reports NOTFOUND instead of running on
    empty pyramids) and pictures that were resident before stay usable."""
    W, H = 128, 128
    imgs = U.random_frames(W, H, 4, seed=3)
    host = np.ascontiguousarray(imgs)
    d = U.derive_input(8, W, H)
    with me.MeSession(0, W, H, capacity=3) as ses:
        ses.put_pictures([0, 1], [host[0].ctypes.data, host[1].ctypes.data], W)
        with pytest.raises(me.SvtMeError):  # 2 new pictures, 1 free slot
            ses.put_pictures([2, 3], [host[2].ctypes.data, host[3].ctypes.data], W)
        with pytest.raises(me.SvtMeError):  # null picture at i = 1
            ses.put_pictures([2, 3], [host[2].ctypes.data, None], W)
        p = U.finish_params(derive(None, d), d, 2, {(0, 0): 0, (1, 0): 1})
        with pytest.raises(me.SvtMeError):
            ses.run_picture(p, 2, {(0, 0): 0, (1, 0): 1})  # picture 2 must not be resident
        ses.put_picture(2, imgs[2])  # the slot is still free
        got = ses.run_picture(p, 2, {(0, 0): 0, (1, 0): 1}, deep=True)
        with pytest.raises(me.SvtMeError):
            ses.run_picture(p, 2, {(0, 0): 0, (1, 0): 3})  # NOTFOUND half-way through a launch ...
        again = ses.run_picture(p, 2, {(0, 0): 0, (1, 0): 1}, deep=True)  # ... leaves the handle usable
        ses.put_picture(1, imgs[3])  # re-upload into a slot the failed launch named: no stale reader event
        ses.put_picture(1, imgs[1])
        third = ses.run_picture(p, 2, {(0, 0): 0, (1, 0): 1}, deep=True)
    pyr = {n: U.CpuPyramid("orc", imgs[n], n) for n in range(3)}
    orc = U.cpu_me_picture("orc", p, pyr[2], {(0, 0): pyr[0], (1, 0): pyr[1]})
    for k in orc:
        assert np.array_equal(orc[k], got[k]) and np.array_equal(orc[k], again[k]) and np.array_equal(orc[k], third[k]), k


def test_two_handles_from_two_threads_and_on_two_devices():
    """The reference is ONE process with many ME threads: handles opened on different devices (when the box has
    them; else twice on device 0) and driven from two host threads must each give the single-handle result
    (kernel attributes are set per device, svtme_configure_stage_kernels)."""
    import threading
    import torch
    ndev = torch.cuda.device_count()
    W, H = 320, 192
    seq = U.synth_sequence(W, H, 3, seed=21, motion=(4, 2))
    d = U.derive_input(4, W, H, temporal_layer_index=1, ref_list0_count_try=1, ref_list1_count_try=1)
    p = U.finish_params(derive(None, d), d, 1, {(0, 0): 0, (1, 0): 2})
    want = gpu_stage(seq, p, 1, {(0, 0): 0, (1, 0): 2}, W, H)
    res, err = {}, []

    def work(dev):
        try:
            with me.MeSession(dev % ndev, W, H, capacity=3) as ses:
                for _ in range(3):
                    for n in range(3):
                        ses.put_picture(n, seq[n])
                    res[dev] = ses.run_picture(p, 1, {(0, 0): 0, (1, 0): 2}, deep=True)
        except Exception as e:  # noqa: BLE001
            err.append(repr(e))

    ths = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not err, err
    for dev in range(2):
        for k in want:
            assert np.array_equal(want[k], res[dev][k]), (dev, k)


def test_share_picture_between_handles_and_external_writes():
    """svtme_share_picture (peer copy of a pyramid into another handle's ring: another device when the box has two, else
    device 0 again) and svtme_external_write_begin / _end (pyramids written on the caller's stream, here a device copy
    standing in for an NCCL receive): ME on the receiving handle equals ME on the handle that built the pyramids, and the
    ordering holds without host synchronisation across repeated rounds that overwrite the same slots."""
    import torch
    ndev = torch.cuda.device_count()
    W, H = 320, 192
    refs, cur = {(0, 0): 0, (1, 0): 2}, 1
    d = U.derive_input(6, W, H, temporal_layer_index=1, ref_list0_count_try=1, ref_list1_count_try=1)
    p = U.finish_params(derive(None, d), d, cur, refs)
    seqs = [U.synth_sequence(W, H, 3, seed=40 + k, motion=(2 + k, -1)) for k in range(3)]
    with me.MeSession(0, W, H, capacity=3) as a, me.MeSession(1 % ndev, W, H, capacity=3) as b, me.MeSession(0, W, H, capacity=3) as c:
        xs = torch.cuda.Stream(device=0)
        for rnd, seq in enumerate(seqs):  # back to back, no host sync in between: slots are overwritten while launches are queued
            a.put_pictures([0, 1, 2], [np.ascontiguousarray(seq[n]).ctypes.data for n in range(3)], W)
            ja, oa = a.make_jobs([(p, cur, refs)], deep=True)
            a.submit_jobs(ja)
            for n in range(3):
                a.share_picture_to(b, n)
            jb, ob = b.make_jobs([(p, cur, refs)], deep=True)
            b.submit_jobs(jb)
            # "receive" a's pyramids into c on the caller's stream
            c.external_write_begin([0, 1, 2], xs.cuda_stream)
            a.external_write_begin([0, 1, 2], xs.cuda_stream)  # sender side: wait for a's builds
            with torch.cuda.stream(xs):
                for n in range(3):
                    nbytes = sum(a.plane_device_ptr(n, l)[1] for l in range(3))
                    src = torch.as_tensor(shard._DeviceBytes(a.plane_device_ptr(n, 0)[0], nbytes), device="cuda:0")
                    dst = torch.as_tensor(shard._DeviceBytes(c.plane_device_ptr(n, 0)[0], nbytes), device="cuda:0")
                    dst.copy_(src, non_blocking=True)
            c.external_write_end(xs.cuda_stream)
            a.external_write_end(xs.cuda_stream)  # a's next upload into these slots waits for the copies that read them
            jc, oc = c.make_jobs([(p, cur, refs)], deep=True)
            c.submit_jobs(jc)
            a.collect_jobs(ja)
            b.collect_jobs(jb)
            c.collect_jobs(jc)
            pyr = {n: U.CpuPyramid("orc", seq[n], n) for n in range(3)}
            want = U.cpu_me_picture("orc", p, pyr[cur], {k: pyr[v] for k, v in refs.items()})
            for k in want:
                assert np.array_equal(want[k], oa[0][k]), (rnd, "builder", k)
                assert np.array_equal(want[k], ob[0][k]), (rnd, "peer copy", k)
                assert np.array_equal(want[k], oc[0][k]), (rnd, "external write", k)


def test_queued_results_land_directly_in_page_locked_caller_arrays():
    """svtme_pin_host_memory: jobs whose three big tables are page-locked get them by direct download, jobs with pageable
    arrays through the pinned mirror; mixed in one batch, both equal the synchronous run."""
    W, H = 320, 192
    seq = U.synth_sequence(W, H, 6, seed=11)
    d = U.derive_input(4, W, H, temporal_layer_index=1, ref_list0_count_try=1, ref_list1_count_try=1)
    specs = []
    for cur in (1, 2, 3, 4):
        p = U.finish_params(derive(None, d), d, cur, {(0, 0): cur - 1, (1, 0): cur + 1})
        specs.append((p, cur, {(0, 0): cur - 1, (1, 0): cur + 1}))
    with me.MeSession(0, W, H, capacity=6) as ses:
        for n in range(6):
            ses.put_picture(n, seq[n])
        jw, want = ses.make_jobs(specs, deep=True)
        ses.run_jobs(jw)
        ja, oa = ses.make_jobs(specs[:2], deep=False, pin=True)
        jb, ob = ses.make_jobs(specs[2:], deep=True, pin=False)
        jc, oc = ses.make_jobs(specs, deep=True, pin=True)
        mixed = (U.abi.SvtMeJob * 4)(ja[0], jb[0], ja[1], jb[1])  # pinned / pageable alternating in one batch
        for rnd in range(2):
            for o in oa + ob + oc:
                for a in o.values():
                    a[...] = 0xAB if a.dtype == np.uint8 else 0
            ses.submit_jobs(mixed)
            ses.submit_jobs(jc)
            ses.collect_jobs(mixed)
            ses.collect_jobs(jc)
            for got, w in ((oa[0], want[0]), (ob[0], want[2]), (oa[1], want[1]), (ob[1], want[3])):
                for k in got:
                    assert np.array_equal(got[k], w[k]), (rnd, "mixed", k)
            for got, w in zip(oc, want):
                for k in got:
                    assert np.array_equal(got[k], w[k]), (rnd, "all pinned", k)

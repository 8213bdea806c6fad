"""Temporal-filter ME (ME_MCTF): svt_aom_motion_estimation_b64 as called by
produce_temporally_filtered_pic (Source/Lib/Codec/temporal_filtering.c:3075) — FULL_SAD HME, one reference, unscaled
picture distance, no pruning, early return after HME under tf_me_exit_th, no candidate arrays. CPU: parameter derivation
and the oracle against golden fixtures generated from the compiled reference (tests/golden/make_golden_tf.py) and, where
oracle/_ref is present, against the reference itself. GPU (-m gpu): the CUDA path through the C ABI against both."""
import ctypes as C
import gzip
import importlib
import json
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import _util as U
import make_golden_tf as T

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def product_tf_params(d, centre, frame):
    p = U.abi.SvtMeParams()
    assert U.product_lib().svtme_derive_params_tf(C.byref(d), C.byref(p)) == 0
    return T.finish_tf(p, centre, frame)


def assert_tf_equal(want, got, p, what):
    """Blocks that returned early keep whatever the tables held before (nothing reads them): only their HME results count."""
    n = want["hme_sad"].reshape(-1, 2, 4).shape[0]
    early = want["hme_sad"].reshape(n, 2, 4)[:, 0, 0] < p.tf_me_exit_th
    if "tf_early_exit" in want:
        assert np.array_equal(early, np.asarray(want["tf_early_exit"]).astype(bool))
    for k in T.TF_KEYS:
        w, g = np.asarray(want[k]).reshape(n, 2, 4, -1)[:, 0, 0], np.asarray(got[k]).reshape(n, 2, 4, -1)[:, 0, 0]
        if k in ("p_sb_best_sad", "p_sb_best_mv"):
            w, g = w[~early], g[~early]
        assert np.array_equal(w, g), (what, k)


def test_tf_parameter_derivation_matches_reference_grid():
    lib = U.product_lib()
    grid = json.load(gzip.open(os.path.join(GOLD, "tf_params.json.gz"), "rt"))
    assert len(grid) == 15 * 5 * 5 * 2
    for row in grid:
        d = U.abi.SvtMeTfDeriveInput()
        for k, v in row["in"].items():
            setattr(d, k, v)
        p = U.abi.SvtMeParams()
        assert lib.svtme_derive_params_tf(C.byref(d), C.byref(p)) == 0
        assert p.as_dict() == row["out"], row["in"]
    bad = T.tf_input(3, 320, 192, 5, 0, 0)
    assert lib.svtme_derive_params_tf(C.byref(bad), C.byref(U.abi.SvtMeParams())) == -1


@pytest.mark.parametrize("case", T.TF_CASES, ids=[c[0] for c in T.TF_CASES])
def test_oracle_tf_me_matches_golden_and_reference(case):
    name, W, H, preset, level, exit_th, centre, frame, qp_opt = case
    seq, d = T.tf_case(*case)
    p = product_tf_params(d, centre, frame)
    want = np.load(os.path.join(GOLD, "tf_stage_%s.npz" % name))
    for which in ["orc"] + (["ref"] if U.ref_lib() is not None else []):
        for simd in ((0, 1) if which == "ref" else (0,)):
            if which == "ref":
                U.ref_lib().ref_set_simd(simd)
            pyr = {n: U.CpuPyramid(which, seq[n], n) for n in (centre, frame)}
            got = U.cpu_me_picture(which, p, pyr[centre], {(0, 0): pyr[frame]})
            assert_tf_equal(want, got, p, (which, simd))
    if U.ref_lib() is not None:
        U.ref_lib().ref_set_simd(0)
    assert 0 < int(np.asarray(want["tf_early_exit"]).sum()) < len(want["tf_early_exit"]) or exit_th == 0


@pytest.mark.gpu
@pytest.mark.parametrize("case", T.TF_CASES, ids=[c[0] for c in T.TF_CASES])
def test_gpu_tf_me_matches_golden_and_oracle(case):
    me = importlib.import_module(U.PKG + ".me")
    name, W, H, preset, level, exit_th, centre, frame, qp_opt = case
    seq, d = T.tf_case(*case)
    p = product_tf_params(d, centre, frame)
    with me.MeSession(0, W, H, capacity=2) as ses:
        ses.put_picture(centre, seq[centre])
        ses.put_picture(frame, seq[frame])
        got = ses.run_picture(p, centre, {(0, 0): frame}, deep=True)
    assert_tf_equal(np.load(os.path.join(GOLD, "tf_stage_%s.npz" % name)), got, p, "golden")
    pyr = {n: U.CpuPyramid("orc", seq[n], n) for n in (centre, frame)}
    assert_tf_equal(U.cpu_me_picture("orc", p, pyr[centre], {(0, 0): pyr[frame]}), got, p, "oracle")
    for k in ("total_me_candidate_index", "me_mv_array", "me_candidate_array", "me_64x64_distortion"):
        assert not np.any(got[k]), k  # nothing is exported per PU for the temporal filter


@pytest.mark.gpu
def test_gpu_tf_me_batch_of_neighbour_frames_at_1080p():
    """One centre picture against its six neighbours in ONE batch (what produce_temporally_filtered_pic loops over),
    1920x1080, against the CPU checker."""
    me = importlib.import_module(U.PKG + ".me")
    W, H, centre = 1920, 1080, 3
    seq = U.synth_sequence(W, H, 7, seed=5, motion=(2, 1), noise=1.5)
    d = T.tf_input(8, W, H, 2, 16 * 16, 1)
    specs = [(product_tf_params(d, centre, f), centre, {(0, 0): f}) for f in range(7) if f != centre]
    with me.MeSession(0, W, H, capacity=7) as ses:
        host = np.ascontiguousarray(seq)
        ses.put_pictures(list(range(7)), [host[i].ctypes.data for i in range(7)], W)
        jobs, outs = ses.make_jobs(specs, deep=True)
        ses.run_jobs(jobs)
    which = U.best_cpu_checker()
    pyrs = U.CpuPyramidSet(which).build({n: seq[n] for n in range(7)}, 8)
    cj, want = U.cpu_mt_jobs(which, specs, pyrs, 510, deep=True)
    U.cpu_me_pictures_mt(which, cj, 8)
    for (p, _, _), w, g in zip(specs, want, outs):
        assert_tf_equal(w, g, p, "1080p")
    pyrs.free()

"""Dynamic-GOP detector HME (dg_detector_hme_level0, pd_process.c:491-589). CPU: the oracle
restatement against the reference's own function compiled from its sources (C and AVX2 svt_sad_loop_kernel) and against a
golden fixture. GPU (-m gpu): the CUDA kernel through the C ABI against the CPU checker. Nothing reads /root/reference."""
import ctypes as C
import importlib
import json
import os

import numpy as np
import pytest

import _util as U

me = importlib.import_module(U.PKG + ".me")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dg_golden.json")

# (width, height, input_resolution): window 16 (<= 360p range), 64 (480p range), 128 (above); incomplete last b64 row /
# column; pictures smaller than the window (everything is clipped)
CASES = [(192, 136, 1), (328, 200, 2), (328, 200, 4), (640, 360, 1), (640, 360, 4), (128, 64, 6)]


def pair(w, h, seed, kind):
    if kind == "pan":  # global translation + noise: most vectors equal, sum_in_vectors cancels left / right
        seq = U.synth_sequence(w, h, 2, seed=seed, motion=(5, -3), noise=2.0)
        return seq[1], seq[0]
    if kind == "zoom":  # centre-symmetric motion: vectors point inwards
        rng = np.random.default_rng(seed)
        base = rng.integers(0, 256, size=(h // 8 + 2, w // 8 + 2), dtype=np.uint8)
        big = np.kron(base, np.ones((8, 8), np.uint8))
        a = big[4:4 + h, 4:4 + w]
        ys, xs = np.mgrid[0:h, 0:w]
        b = big[np.clip(4 + ys + (ys - h // 2) // 16, 0, big.shape[0] - 1), np.clip(4 + xs + (xs - w // 2) // 16, 0, big.shape[1] - 1)]
        return np.ascontiguousarray(a), np.ascontiguousarray(b)
    rng = np.random.default_rng(seed)  # unrelated pictures: large SADs (tot_cplx)
    return rng.integers(0, 256, size=(h, w), dtype=np.uint8), rng.integers(0, 256, size=(h, w), dtype=np.uint8)


@pytest.mark.parametrize("case", CASES, ids=lambda c: "%dx%d_res%d" % c)
def test_oracle_matches_compiled_reference(case):
    if U.ref_lib() is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    w, h, res = case
    for simd in (0, 1):
        U.ref_lib().ref_set_simd(simd)
        for kind in ("pan", "zoom", "random"):
            a, b = pair(w, h, w + res, kind)
            want = U.cpu_dg_detector_hme("ref", U.CpuPyramid("ref", a, 1), U.CpuPyramid("ref", b, 0), res)
            got = U.cpu_dg_detector_hme("orc", U.CpuPyramid("orc", a, 1), U.CpuPyramid("orc", b, 0), res)
            assert want == got, (kind, simd, want, got)
    U.ref_lib().ref_set_simd(0)


def test_oracle_matches_golden_fixture():
    """tests/golden/dg_golden.json: outputs of the compiled reference (tests/golden/make_golden_dg.py)."""
    gold = json.load(open(GOLD))
    assert len(gold) >= 12
    for g in gold:
        a, b = pair(g["w"], g["h"], g["seed"], g["kind"])
        got = U.cpu_dg_detector_hme("orc", U.CpuPyramid("orc", a, 1), U.CpuPyramid("orc", b, 0), g["res"])
        assert list(got) == g["metrics"], g


def test_identical_pictures_give_zero_vectors():
    a, _ = pair(256, 192, 3, "random")  # complete b64 only: a block inside the replicated pad would tie along the pad direction
    m = U.cpu_dg_detector_hme("orc", U.CpuPyramid("orc", a, 1), U.CpuPyramid("orc", a, 0), 4)
    assert m == (0, 0, 0, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES + [(1920, 1080, 4)], ids=lambda c: "%dx%d_res%d" % c)
def test_gpu_dg_detector_matches_the_cpu_checker(case):
    w, h, res = case
    which = U.best_cpu_checker()
    with me.MeSession(0, w, h, capacity=2) as ses:
        for kind in ("pan", "zoom", "random"):
            a, b = pair(w, h, w + res, kind)
            a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
            ses.put_pictures([1, 0], [a.ctypes.data, b.ctypes.data], w)
            m = U.abi.SvtMeDgMetrics()
            assert ses.lib.svtme_dg_detector_hme(ses.h, 1, 0, res, C.byref(m)) == 0, ses.lib.svtme_last_error()
            want = U.cpu_dg_detector_hme(which, U.CpuPyramid(which, a, 1), U.CpuPyramid(which, b, 0), res)
            assert (int(m.tot_dist), int(m.tot_cplx), int(m.tot_active), int(m.sum_in_vectors)) == want, kind
        m = U.abi.SvtMeDgMetrics()
        assert ses.lib.svtme_dg_detector_hme(ses.h, 1, 9, res, C.byref(m)) != 0  # not resident
        assert ses.lib.svtme_dg_detector_hme(ses.h, 1, 0, 7, C.byref(m)) != 0  # not an EbInputResolution value

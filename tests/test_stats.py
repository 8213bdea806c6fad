"""Picture statistics (svt_aom_gathering_picture_statistics, pic_analysis_process.c:1560).
CPU part: the oracle restatement against the reference's own function compiled from its sources (C kernels and the
AVX2 / SSE2 kernels its RTCD installs) and against golden fixtures generated from that build. GPU part (-m gpu): the CUDA
kernels through the C ABI against the CPU checker, bit for bit. Nothing here reads /root/reference at run time."""
import ctypes as C
import importlib
import os

import numpy as np
import pytest

import _util as U

me = importlib.import_module(U.PKG + ".me")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# (width, height): complete b64 only; incomplete last row / column (1080 = 16 * 64 + 56); widths whose 1/16 plane does not
# divide by the region count
SIZES = [(128, 64), (192, 136), (328, 200), (640, 360)]
CONFIGS = [dict(), dict(prec_sub=0), dict(all_sizes=0), dict(scd=1), dict(regions=(1, 1)), dict(regions=(3, 2), scd=1),
           dict(calc_hist=0), dict(calculate_variance=0)]


def pictures(w, h, seed):
    rng = np.random.default_rng(seed)
    yield "random", rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    yield "synthetic", U.synth_sequence(w, h, 2, seed=seed)[1]
    flat = np.full((h, w), 255, np.uint8)  # the largest sums: mean^2 and E[x^2] at the top of their range
    flat[::2, ::3] = 0
    yield "extreme", flat


def same(a, b, what):
    for k in a:
        assert np.array_equal(a[k], b[k]), (what, k, np.argwhere(a[k] != b[k])[:4])


@pytest.mark.parametrize("size", SIZES, ids=lambda s: "%dx%d" % s)
def test_oracle_matches_compiled_reference(size):
    if U.ref_lib() is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    w, h = size
    for simd in (0, 1):
        U.ref_lib().ref_set_simd(simd)
        for name, luma in pictures(w, h, seed=w + h):
            ref_pyr, orc_pyr = U.CpuPyramid("ref", luma, 0), U.CpuPyramid("orc", luma, 0)
            for cfg in CONFIGS:
                p = U.stats_params(**cfg)
                same(U.cpu_picture_statistics("ref", ref_pyr, p), U.cpu_picture_statistics("orc", orc_pyr, p), (name, cfg, simd))
    U.ref_lib().ref_set_simd(0)


def test_oracle_matches_golden_fixture():
    """tests/golden/stats_golden.npz: outputs of the compiled reference (tests/golden/make_golden_stats.py)."""
    g = np.load(os.path.join(GOLD, "stats_golden.npz"))
    n = int(g["n_cases"])
    assert n >= 8
    for i in range(n):
        luma, cfg = g["luma_%d" % i], [int(x) for x in g["cfg_%d" % i]]
        p = U.stats_params(calc_hist=cfg[0], calculate_variance=cfg[1], prec_sub=cfg[2], all_sizes=cfg[3], scd=cfg[4], regions=(cfg[5], cfg[6]))
        got = U.cpu_picture_statistics("orc", U.CpuPyramid("orc", luma, 0), p)
        for k in got:
            assert np.array_equal(got[k], g["%s_%d" % (k, i)]), (i, k)


def test_variance_of_a_flat_picture_is_zero_and_of_a_checkerboard_is_known():
    w, h = 128, 128
    flat = np.full((h, w), 77, np.uint8)
    s = U.cpu_picture_statistics("orc", U.CpuPyramid("orc", flat, 0), U.stats_params(prec_sub=0))
    assert not s["variance"].any() and int(s["avg_luma"][0]) == 77 and int(s["pic_avg_variance"][0]) == 0
    assert (s["average_intensity_per_region"] == 77).all()
    chk = np.zeros((h, w), np.uint8)
    chk[:, ::2] = 200  # columns alternate 200 / 0: mean 100, E[x^2] = 20000, variance 10000 in every block size
    s = U.cpu_picture_statistics("orc", U.CpuPyramid("orc", chk, 0), U.stats_params(prec_sub=0))
    assert (s["variance"] == 10000).all() and int(s["pic_avg_variance"][0]) == 10000
    # histogram: bins start at 1 and are scaled by 16 * decim^2 (decim 4 without scene-change detection)
    hist = s["picture_histogram"]
    assert hist.shape == (4, 4, 256) and (hist % 256 == 0).all() and (hist >= 256).all()


# ---------------------------------------------------------------- GPU
def gpu_stats(ses, pic, params, n_b64):
    out, arrays = U.abi.alloc_stats_out(n_b64, params)
    rc = ses.lib.svtme_picture_statistics(ses.h, pic, C.byref(params), C.byref(out))
    assert rc == 0, ses.lib.svtme_last_error()
    return arrays


@pytest.mark.gpu
@pytest.mark.parametrize("size", SIZES + [(1920, 1080)], ids=lambda s: "%dx%d" % s)
def test_gpu_statistics_match_the_cpu_checker(size):
    w, h = size
    which = U.best_cpu_checker()
    n_b64 = ((w + 63) // 64) * ((h + 63) // 64)
    with me.MeSession(0, w, h, capacity=4) as ses:
        for k, (name, luma) in enumerate(pictures(w, h, seed=w + h)):
            luma = np.ascontiguousarray(luma)
            ses.put_pictures([k], [luma.ctypes.data], w)
            pyr = U.CpuPyramid(which, luma, k)
            for cfg in CONFIGS if w <= 640 else CONFIGS[:4]:
                p = U.stats_params(**cfg)
                same(U.cpu_picture_statistics(which, pyr, p), gpu_stats(ses, k, p, n_b64), (name, cfg))


@pytest.mark.gpu
def test_gpu_statistics_of_10bit_sources_and_errors():
    w, h = 640, 360
    rng = np.random.default_rng(9)
    luma8 = U.synth_sequence(w, h, 1, seed=4)[0]
    luma16 = np.ascontiguousarray((luma8.astype(np.uint16) << 2) | rng.integers(0, 4, size=luma8.shape, dtype=np.uint16))
    p = U.stats_params()
    with me.MeSession(0, w, h, capacity=2) as ses:
        ses.put_pictures([5], [luma16.ctypes.data], w, bits=10)  # the statistics see the 8 MSBs, like the reference's 8-bit buffer
        same(U.cpu_picture_statistics("orc", U.CpuPyramid("orc", luma8, 5), p), gpu_stats(ses, 5, p, 60), "10-bit")
        out, _ = U.abi.alloc_stats_out(60, p)
        assert ses.lib.svtme_picture_statistics(ses.h, 77, C.byref(p), C.byref(out)) != 0  # not resident
        bad = U.stats_params(regions=(5, 1))
        assert ses.lib.svtme_picture_statistics(ses.h, 5, C.byref(bad), C.byref(out)) != 0


@pytest.mark.gpu
def test_gpu_statistics_batch_of_pictures_in_one_call():
    w, h = 328, 200
    n_b64 = 6 * 4
    seq = U.synth_sequence(w, h, 5, seed=21, noise=5.0)
    p = U.stats_params(scd=1, regions=(2, 3))
    with me.MeSession(0, w, h, capacity=5) as ses:
        ses.put_pictures(list(range(5)), [np.ascontiguousarray(f).ctypes.data for f in seq], w)
        outs = (U.abi.SvtMeStatsOut * 5)()
        keep = []
        for i in range(5):
            o, arrays = U.abi.alloc_stats_out(n_b64, p)
            outs[i] = o
            keep.append(arrays)
        nums = (C.c_uint64 * 5)(4, 2, 0, 1, 3)
        assert ses.lib.svtme_pictures_statistics(ses.h, 5, nums, C.byref(p), outs) == 0, ses.lib.svtme_last_error()
        for i, pic in enumerate([4, 2, 0, 1, 3]):
            same(U.cpu_picture_statistics("orc", U.CpuPyramid("orc", seq[pic], pic), p), keep[i], ("batch", pic))

"""CPU tests (-m "not gpu"): the oracle restatement against the committed golden vectors (generated
from the compiled reference by tests/golden/make_golden.py) and, where oracle/_ref is available,
against the compiled reference itself on fresh seeded inputs."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import _util as U
import make_golden as G

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
needs_ref = pytest.mark.skipif(U.ref_lib() is None, reason="oracle/_ref/libsvtref.so not built (no /root/reference)")


def run_stage(which, case, params=None):
    name, W, H, preset, tl, refs, cur, kw = case
    seq, d = G.stage_case(*case)
    p = params if params is not None else U.ref_params(d, cur, refs)
    pyr = {n: U.CpuPyramid(which, seq[n], n) for n in set(refs.values()) | {cur}}
    return U.cpu_me_picture(which, p, pyr[cur], {k: pyr[v] for k, v in refs.items()})


def golden_params(case):
    """SvtMeParams for a golden stage case without needing oracle/_ref: rebuilt from params of the
    product-independent derivation stored next to the fixture."""
    name, W, H, preset, tl, refs, cur, kw = case
    seq, d = G.stage_case(*case)
    if U.ref_lib() is not None:
        return U.ref_params(d, cur, refs)
    pytest.skip("needs oracle/_ref for parameter derivation")


@pytest.mark.parametrize("case", G.STAGE_CASES, ids=[c[0] for c in G.STAGE_CASES])
def test_oracle_stage_matches_golden(case):
    want = np.load(os.path.join(GOLD, "stage_%s.npz" % case[0]))
    got = run_stage("orc", case, golden_params(case))
    for k in want.files:
        assert np.array_equal(want[k], got[k]), k


def test_oracle_pyramid_matches_golden():
    seq = U.synth_sequence(328, 200, 1, seed=5)
    pyr = U.CpuPyramid("orc", seq[0], 0)
    want = np.load(os.path.join(GOLD, "pyramid.npz"))
    for lvl, k in enumerate(("full", "quarter", "sixteenth")):
        assert np.array_equal(want[k], pyr.plane(lvl)), k


def test_oracle_sad_loop_matches_golden():
    rows = np.load(os.path.join(GOLD, "kernels.npz"))["sad_loop"]
    rng = np.random.default_rng(13596)
    src = rng.integers(0, 256, size=(128, 128), dtype=np.uint8)
    refb = rng.integers(0, 256, size=(288, 288), dtype=np.uint8)
    lib = U.oracle_lib()
    for bw, bh, sw, sh, skip, best_w, x_w, y_w in rows.tolist():
        best, x, y = C.c_uint64(0), C.c_int16(0), C.c_int16(0)
        lib.orc_sad_loop_kernel(src.ctypes.data_as(C.c_void_p), C.c_uint32(128), refb.ctypes.data_as(C.c_void_p),
                                C.c_uint32(288), C.c_uint32(bh), C.c_uint32(bw), C.byref(best), C.byref(x), C.byref(y),
                                C.c_uint32(288), C.c_uint8(skip), C.c_int16(sw), C.c_int16(sh))
        assert (best.value, x.value, y.value) == (best_w, x_w, y_w), (bw, bh, sw, sh, skip)


@needs_ref
@pytest.mark.parametrize("simd", [0, 1])
def test_oracle_stage_matches_reference_fresh(simd):
    """Fresh seeded inputs, reference C and AVX2 kernels (the reference's own differential idea)."""
    U.ref_lib().ref_set_simd(simd)
    try:
        W, H = 384, 216
        seq = U.synth_sequence(W, H, 6, seed=99 + simd, motion=(-4, 3), noise=3.0)
        for preset, tl, refs, cur in [(7, 2, {(0, 0): 1, (0, 1): 0, (1, 0): 3, (1, 1): 5}, 2), (3, 1, {(0, 0): 0, (1, 0): 4}, 2)]:
            l0 = 1 + max(ri for (li, ri) in refs if li == 0)
            l1 = 1 + max(ri for (li, ri) in refs if li == 1)
            d = U.derive_input(preset, W, H, temporal_layer_index=tl, ref_list0_count_try=l0, ref_list1_count_try=l1,
                               ref_count_used_list0=l0, ref_count_used_list1=l1)
            p = U.ref_params(d, cur, refs)
            res = {}
            for which in ("orc", "ref"):
                pyr = {n: U.CpuPyramid(which, seq[n], n) for n in set(refs.values()) | {cur}}
                res[which] = U.cpu_me_picture(which, p, pyr[cur], {k: pyr[v] for k, v in refs.items()})
            for k in res["ref"]:
                assert np.array_equal(res["orc"][k], res["ref"][k]), (preset, k)
    finally:
        U.ref_lib().ref_set_simd(0)


@needs_ref
def test_oracle_kernels_match_reference_c():
    """SadTest.cc shapes: RANDOM / REF_MAX / SRC_MAX patterns, sub_sad on and off."""
    ref, orc = U.ref_lib(), U.oracle_lib()
    rng = np.random.default_rng(7)
    for pattern in ("random", "ref_max", "src_max"):
        src = rng.integers(0, 256, size=(64, 64), dtype=np.uint8)
        rf = rng.integers(0, 256, size=(64, 80), dtype=np.uint8)
        if pattern == "ref_max":
            rf[:] = 255
        if pattern == "src_max":
            src[:] = 255
        for sub in (0, 1):
            outs = []
            for lib, fn in ((ref, "svt_ext_all_sad_calculation_8x8_16x16_c"), (orc, "orc_ext_all_sad_calculation_8x8_16x16")):
                bs8 = np.full(64, 0x7FFFFFFE, dtype=np.uint32); bs16 = np.full(16, 0x7FFFFFFE, dtype=np.uint32)
                bm8 = np.zeros(64, dtype=np.uint32); bm16 = np.zeros(16, dtype=np.uint32)
                e16 = np.zeros((16, 8), dtype=np.uint32); e8 = np.zeros((64, 8), dtype=np.uint32)
                getattr(lib, fn)(src.ctypes.data_as(C.c_void_p), C.c_uint32(64), rf.ctypes.data_as(C.c_void_p), C.c_uint32(80),
                                 C.c_uint32(0x00030005), bs8.ctypes.data_as(C.c_void_p), bs16.ctypes.data_as(C.c_void_p),
                                 bm8.ctypes.data_as(C.c_void_p), bm16.ctypes.data_as(C.c_void_p), e16.ctypes.data_as(C.c_void_p),
                                 e8.ctypes.data_as(C.c_void_p), C.c_uint8(sub))
                outs.append((bs8, bs16, bm8, bm16, e16))
            for a, b in zip(*outs):
                assert np.array_equal(a, b), (pattern, sub)
    # downsample_2d: PictureOperatorTest.cc:230-232 sizes / steps
    for (w, h) in [(176, 144), (88, 72), (960, 540)]:
        img = rng.integers(0, 256, size=(h, w), dtype=np.uint8)
        for step in (2, 4, 8):
            o = []
            for lib, fn in ((ref, "svt_aom_downsample_2d_c"), (orc, "orc_downsample_2d")):
                out = np.zeros((h // step + 1, w // step + 1), dtype=np.uint8)
                getattr(lib, fn)(img.ctypes.data_as(C.c_void_p), C.c_uint32(w), C.c_uint32(w), C.c_uint32(h),
                                 out.ctypes.data_as(C.c_void_p), C.c_uint32(out.shape[1]), C.c_uint32(step))
                o.append(out)
            assert np.array_equal(o[0], o[1]), (w, h, step)


def test_unpack_msb_oracle_vs_reference():
    """10-bit source -> 8-bit ME plane: the restatement against svt_aom_un_pack2d of the compiled reference, on
    sizes that take both of its branches (multiple-of-4 width with even height, and not) and on values with the
    upper six bits set (the cast to uint8_t wraps)."""
    ref, orc = U.ref_lib(), U.oracle_lib()
    if ref is None:
        pytest.skip("compiled reference (oracle/_ref) not available")
    rng = np.random.default_rng(3)
    for (w, h) in [(64, 64), (136, 72), (35, 21), (66, 33)]:
        for hi in (1024, 65536):
            src = rng.integers(0, hi, size=(h, w + 5), dtype=np.uint16)
            a = np.zeros((h, w + 3), dtype=np.uint8)
            b = np.zeros((h, w + 3), dtype=np.uint8)
            ref.ref_unpack_msb(src.ctypes.data_as(C.c_void_p), C.c_uint32(w + 5), a.ctypes.data_as(C.c_void_p), C.c_uint32(w + 3),
                               C.c_uint32(w), C.c_uint32(h))
            orc.orc_unpack_msb(src.ctypes.data_as(C.c_void_p), C.c_uint32(w + 5), b.ctypes.data_as(C.c_void_p), C.c_uint32(w + 3),
                               C.c_uint32(w), C.c_uint32(h))
            assert np.array_equal(a, b), (w, h, hi)
            assert np.array_equal(b[:, :w], (src[:, :w] >> 2).astype(np.uint8))


@needs_ref
def test_thread_pool_drivers_equal_the_single_picture_calls():
    """bench.py's all-cores CPU arm (oracle/mt_pool.h: (picture, b64-row) tasks on a pthread pool; pyramids one
    picture per task, AVX2 downsample_2d) gives exactly what the single-threaded single-picture calls give, for the
    compiled reference (C and AVX2 kernels) and for the oracle port — and the pyramids refilled in place equal
    fresh ones."""
    W, H = 472, 312
    seq = U.synth_sequence(W, H, 7, seed=9, motion=(5, -3), noise=2.0)
    refs = {(0, 0): 2, (0, 1): 1, (0, 2): 0, (1, 0): 4, (1, 1): 5}
    specs = []
    for preset, tl, cur in ((8, 3, 3), (4, 2, 3), (6, 0, 6)):
        r = refs if cur == 3 else {(0, 0): 5, (0, 1): 3, (1, 0): 5, (1, 1): 3}
        l0, l1 = 1 + max(ri for (li, ri) in r if li == 0), 1 + max(ri for (li, ri) in r if li == 1)
        d = U.derive_input(preset, W, H, temporal_layer_index=tl, ref_list0_count_try=l0, ref_list1_count_try=l1,
                           ref_count_used_list0=3, ref_count_used_list1=2)
        specs.append((U.ref_params(d, cur, r), cur, r))
    n_b64 = ((W + 63) // 64) * ((H + 63) // 64)
    single = {n: U.CpuPyramid("orc", seq[n], n) for n in range(7)}
    want = [U.cpu_me_picture("orc", p, single[cur], {k: single[v] for k, v in r.items()}) for p, cur, r in specs]
    for which, simd in (("ref", 0), ("ref", 1), ("orc", 0)):
        if which == "ref":
            U.ref_lib().ref_set_simd(simd)
        pyrs = U.CpuPyramidSet(which)
        for rebuild in range(2):  # the second build refills the planes in place
            pyrs.build({n: seq[n] for n in range(7)}, 3)
            for n in range(7):
                for lvl, name in enumerate(("full", "quarter", "sixteenth")):
                    pl = getattr(pyrs[n], name)
                    rows = pl.height + 2 * pl.org_y
                    got = np.ctypeslib.as_array(C.cast(pl.buffer_y, C.POINTER(C.c_uint8)), shape=(rows, pl.stride_y))
                    assert np.array_equal(got, single[n].plane(lvl)), (which, simd, n, name)
        for threads, seg_rows in ((1, 1), (3, 1), (4, 2), (16, 5)):
            jobs, outs = U.cpu_mt_jobs(which, specs, pyrs, n_b64, deep=True)
            U.cpu_me_pictures_mt(which, jobs, threads, seg_rows)
            for w, o in zip(want, outs):
                for k in w:
                    assert np.array_equal(w[k], o[k]), (which, simd, threads, seg_rows, k)
        pyrs.free()
    U.ref_lib().ref_set_simd(0)

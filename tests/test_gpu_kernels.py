"""GPU kernel-level parity (-m gpu): the RTCD-signature wrappers of the kernel-test ABI against the
oracle, with the shapes of the reference's own unit tests (test/SadTest.cc, PictureOperatorTest.cc)."""
import ctypes as C

import numpy as np
import pytest

import _util as U

pytestmark = pytest.mark.gpu
vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731

# test/SadTest.cc:507-518 search areas (subset incl. the extremes) and block sizes used by HME
AREAS = [(1, 1), (8, 1), (7, 3), (8, 3), (16, 5), (24, 9), (32, 17), (48, 48), (64, 32), (96, 3), (8, 100), (160, 40), (640, 5)]
BLOCKS = [(16, 8), (16, 16), (32, 16), (64, 32), (64, 64), (8, 8), (24, 12), (48, 24), (6, 3)]


def patterns(rng, shape_src, shape_ref):
    for name in ("random", "ref_max", "src_max", "equal"):
        src = rng.integers(0, 256, size=shape_src, dtype=np.uint8)
        ref = rng.integers(0, 256, size=shape_ref, dtype=np.uint8)
        if name == "ref_max":
            ref[:] = 255
        elif name == "src_max":
            src[:] = 255
        elif name == "equal":
            src[:] = 7
            ref[:] = 7
        yield name, src, ref


def test_sad_loop_kernel_vs_oracle():
    lib, orc = U.product_lib(), U.oracle_lib()
    rng = np.random.default_rng(13596)
    n = 0
    for pname, src, ref in patterns(rng, (128, 128), (800, 800)):
        for (bw, bh) in BLOCKS:
            for (sw, sh) in AREAS if pname == "random" else AREAS[::3]:
                for skip in (0, 1):
                    for sub in (1, 2):  # the HME callers double both strides and halve the height
                        res = []
                        for fn in (lib.svt_sad_loop_kernel_b200, orc.orc_sad_loop_kernel):
                            best, x, y = C.c_uint64(0), C.c_int16(-3), C.c_int16(-5)
                            fn(vp(src), C.c_uint32(128 * sub), vp(ref), C.c_uint32(800 * sub), C.c_uint32(max(1, bh // sub)),
                               C.c_uint32(bw), C.byref(best), C.byref(x), C.byref(y), C.c_uint32(800), C.c_uint8(skip),
                               C.c_int16(sw), C.c_int16(sh))
                            res.append((best.value, x.value, y.value))
                        if res[1][0] == 0xffffff:  # nothing searched: x/y stay what the caller had
                            assert res[0][0] == 0xffffff
                        else:
                            assert res[0] == res[1], (pname, bw, bh, sw, sh, skip, sub, res)
                        n += 1
    assert n > 500


@pytest.mark.parametrize("sub_sad", [1, 0])
def test_fullpel_sblock_vs_oracle(sub_sad):
    lib, orc = U.product_lib(), U.oracle_lib()
    rng = np.random.default_rng(5)
    for pname, src, ref in patterns(rng, (64, 64), (400, 400)):
        for (sw, sh) in [(1, 1), (8, 3), (16, 5), (7, 4), (64, 32), (72, 33), (128, 40), (3, 70)]:
            outs = []
            for fn in (lib.svt_open_loop_me_fullpel_search_sblock_b200, orc.orc_fullpel_search_sblock):
                bs = np.full(85, U.abi.MAX_SAD_VALUE, dtype=np.uint32)
                bm = np.zeros(85, dtype=np.uint32)
                # pre-seed a few entries so that "existing entry wins ties / stays if better" is exercised
                bs[3], bm[3] = 0, 0x00110022
                fn(vp(src), C.c_uint32(64), vp(ref), C.c_uint32(400), C.c_int16(-9), C.c_int16(4), C.c_uint32(sw),
                   C.c_uint32(sh), C.c_uint8(sub_sad), vp(bs), vp(bm))
                outs.append((bs, bm))
            assert np.array_equal(outs[0][0], outs[1][0]), (pname, sw, sh, "sad")
            assert np.array_equal(outs[0][1], outs[1][1]), (pname, sw, sh, "mv")


@pytest.mark.parametrize("sub_sad", [0, 1])
def test_ext_all_and_eight_sad_vs_oracle(sub_sad):
    """Allsad8x8_CalculationTest / Allsad32x32_CalculationTest (SadTest.cc:796-1071)."""
    lib, orc = U.product_lib(), U.oracle_lib()
    rng = np.random.default_rng(21)
    for pname, src, ref in patterns(rng, (64, 64), (64, 96)):
        outs = []
        for all_fn, eight_fn in ((lib.svt_ext_all_sad_calculation_8x8_16x16_b200, lib.svt_ext_eight_sad_calculation_32x32_64x64_b200),
                                 (orc.orc_ext_all_sad_calculation_8x8_16x16, orc.orc_ext_eight_sad_calculation_32x32_64x64)):
            bs8 = np.full(64, 0x7FFFFFFE, dtype=np.uint32); bs16 = np.full(16, 0x7FFFFFFE, dtype=np.uint32)
            bm8 = np.zeros(64, dtype=np.uint32); bm16 = np.zeros(16, dtype=np.uint32)
            e16 = np.zeros((16, 8), dtype=np.uint32); e8 = np.zeros((64, 8), dtype=np.uint32)
            all_fn(vp(src), C.c_uint32(64), vp(ref), C.c_uint32(96), C.c_uint32(0xfffe0007), vp(bs8), vp(bs16), vp(bm8), vp(bm16),
                   vp(e16), vp(e8), C.c_uint8(sub_sad))
            bs32 = np.full(4, 0x7FFFFFFE, dtype=np.uint32); bs64 = np.full(1, 0x7FFFFFFE, dtype=np.uint32)
            bm32 = np.zeros(4, dtype=np.uint32); bm64 = np.zeros(1, dtype=np.uint32); s32 = np.zeros((4, 8), dtype=np.uint32)
            eight_fn(vp(e16), vp(bs32), vp(bs64), vp(bm32), vp(bm64), C.c_uint32(0xfffe0007), vp(s32))
            outs.append((bs8, bs16, bm8, bm16, e16, bs32, bs64, bm32, bm64, s32))
        for i, (a, b) in enumerate(zip(*outs)):
            assert np.array_equal(a, b), (pname, i)


def test_single_position_kernels_vs_oracle():
    """Extsad8x8_CalculationTest / Extsad32x32_CalculationTest / InitializeBuffer32 (SadTest.cc:1102-1413)."""
    lib, orc = U.product_lib(), U.oracle_lib()
    rng = np.random.default_rng(3)
    src = rng.integers(0, 256, size=(16, 64), dtype=np.uint8)
    ref = rng.integers(0, 256, size=(16, 80), dtype=np.uint8)
    for sub in (0, 1):
        outs = []
        for f8, f32 in ((lib.svt_ext_sad_calculation_8x8_16x16_b200, lib.svt_ext_sad_calculation_32x32_64x64_b200),
                        (orc.orc_ext_sad_calculation_8x8_16x16, orc.orc_ext_sad_calculation_32x32_64x64)):
            bs8 = np.full(4, 5000, dtype=np.uint32); bs16 = np.full(1, 20000, dtype=np.uint32)
            bm8 = np.zeros(4, dtype=np.uint32); bm16 = np.zeros(1, dtype=np.uint32)
            s16 = np.zeros(1, dtype=np.uint32); s8 = np.zeros(4, dtype=np.uint32)
            f8(vp(src), C.c_uint32(64), vp(ref), C.c_uint32(80), vp(bs8), vp(bs16), vp(bm8), vp(bm16), C.c_uint32(0x00010002), vp(s16),
               vp(s8), C.c_uint8(sub))
            sad16 = rng.integers(0, 60000, size=16).astype(np.uint32) if not outs else outs[0][-1]
            bs32 = np.full(4, 150000, dtype=np.uint32); bs64 = np.full(1, 400000, dtype=np.uint32)
            bm32 = np.zeros(4, dtype=np.uint32); bm64 = np.zeros(1, dtype=np.uint32); s32 = np.zeros(4, dtype=np.uint32)
            f32(vp(sad16), vp(bs32), vp(bs64), vp(bm32), vp(bm64), C.c_uint32(0x00050006), vp(s32))
            outs.append((bs8, bs16, bm8, bm16, s16, s8, bs32, bs64, bm32, bm64, s32, sad16))
        for i, (a, b) in enumerate(zip(*outs)):
            assert np.array_equal(a, b), (sub, i)
    buf = np.zeros(85, dtype=np.uint32)
    lib.svt_initialize_buffer_32bits_b200(vp(buf), C.c_uint32(21), C.c_uint32(1), C.c_uint32(U.abi.MAX_SAD_VALUE))
    assert (buf == U.abi.MAX_SAD_VALUE).all()
    lib.svt_nxm_sad_kernel_b200.restype = C.c_uint32
    orc.orc_nxm_sad_kernel.restype = C.c_uint32
    for (h, w) in [(32, 64), (8, 16), (7, 6)]:
        a = lib.svt_nxm_sad_kernel_b200(vp(src), C.c_uint32(64), vp(ref), C.c_uint32(80), C.c_uint32(min(h, 16)), C.c_uint32(w))
        b = orc.orc_nxm_sad_kernel(vp(src), C.c_uint32(64), vp(ref), C.c_uint32(80), C.c_uint32(min(h, 16)), C.c_uint32(w))
        assert a == b


def test_downsample_and_padding_vs_oracle():
    """Downsample2DTest (PictureOperatorTest.cc:225-335): 4 sizes x steps 2/4/8; padding of odd geometry."""
    lib, orc = U.product_lib(), U.oracle_lib()
    rng = np.random.default_rng(8)
    for (w, h) in [(1920, 1080), (960, 540), (176, 144), (88, 72)]:
        img = rng.integers(0, 256, size=(h, w), dtype=np.uint8)
        for step in (2, 4, 8):
            o = []
            for fn in (lib.svt_aom_downsample_2d_b200, orc.orc_downsample_2d):
                out = np.zeros((h // step + 1, w // step + 1), dtype=np.uint8)
                fn(vp(img), C.c_uint32(w), C.c_uint32(w), C.c_uint32(h), vp(out), C.c_uint32(out.shape[1]), C.c_uint32(step))
                o.append(out)
            assert np.array_equal(o[0], o[1]), (w, h, step)
    for (w, h, pw, ph) in [(40, 24, 16, 16), (33, 17, 5, 9), (64, 64, 72, 72)]:
        o = []
        for fn in (lib.svt_aom_generate_padding_b200, orc.orc_generate_padding):
            stride = w + 2 * pw
            buf = np.zeros((h + 2 * ph, stride), dtype=np.uint8)
            buf[ph:ph + h, pw:pw + w] = np.arange(h * w, dtype=np.uint32).reshape(h, w) % 251
            fn(vp(buf), C.c_uint32(stride), C.c_uint32(w), C.c_uint32(h), C.c_uint32(pw), C.c_uint32(ph))
            o.append(buf)
        assert np.array_equal(o[0], o[1]), (w, h, pw, ph)


def test_fullpel_sblock_batch_vs_oracle():
    """svt_open_loop_me_fullpel_search_sblock_b200_batch: n searches in one launch on device memory, every item with its own
    block, window position and area, tables read-modify-write — against the oracle, item by item."""
    import torch
    lib, orc = U.product_lib(), U.oracle_lib()
    rng = np.random.default_rng(77)
    S = 640
    src = rng.integers(0, 256, size=(256, S), dtype=np.uint8)
    ref = rng.integers(0, 256, size=(480, S), dtype=np.uint8)
    areas = [(16, 5), (8, 3), (64, 32), (24, 7), (1, 1), (7, 4), (72, 33), (3, 40)]
    n = len(areas)
    items = (U.abi.SvtMeFullpelItem * n)()
    want_s, want_m = [], []
    init_s = np.full((n, 85), U.abi.MAX_SAD_VALUE, dtype=np.uint32)
    init_m = np.zeros((n, 85), dtype=np.uint32)
    init_s[2, 3], init_m[2, 3] = 0, 0x00110022  # an existing better entry stays
    for i, (sw, sh) in enumerate(areas):
        sx, sy, rx, ry = 16 * (i % 4), 64 * (i % 3), 40 + 11 * i, 100 + 9 * i
        it = items[i]
        it.src_offset, it.ref_offset = sy * S + sx, ry * S + rx
        it.x_search_area_origin, it.y_search_area_origin, it.search_area_width, it.search_area_height = -9 + i, 4 - i, sw, sh
        bs, bm = init_s[i].copy(), init_m[i].copy()
        orc.orc_fullpel_search_sblock(C.c_void_p(src.ctypes.data + it.src_offset), C.c_uint32(S), C.c_void_p(ref.ctypes.data + it.ref_offset),
                                      C.c_uint32(S), C.c_int16(-9 + i), C.c_int16(4 - i), C.c_uint32(sw), C.c_uint32(sh), C.c_uint8(1), vp(bs), vp(bm))
        want_s.append(bs)
        want_m.append(bm)
    d_src, d_ref = torch.from_numpy(src).cuda(), torch.from_numpy(ref).cuda()
    d_items = torch.frombuffer(bytearray(bytes(items)), dtype=torch.uint8).cuda()
    d_s, d_m = torch.from_numpy(init_s.view(np.int32)).cuda(), torch.from_numpy(init_m.view(np.int32)).cuda()
    torch.cuda.synchronize()
    for _ in range(2):  # a second pass over its own results changes nothing (existing entries win ties)
        rc = lib.svt_open_loop_me_fullpel_search_sblock_b200_batch(None, n, d_src.data_ptr(), S, d_ref.data_ptr(), S, d_items.data_ptr(), 1,
                                                                   d_s.data_ptr(), d_m.data_ptr())
        assert rc == 0, lib.svtme_last_error()
        torch.cuda.synchronize()
        got_s, got_m = d_s.cpu().numpy().view(np.uint32), d_m.cpu().numpy().view(np.uint32)
        for i in range(n):
            assert np.array_equal(got_s[i], want_s[i]) and np.array_equal(got_m[i], want_m[i]), (areas[i], "item", i)

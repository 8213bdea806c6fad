"""Generate the committed golden fixtures from the UNMODIFIED reference compiled under
oracle/_ref (needs /root/reference; run in the build container, not on the GPU box):

    python tests/golden/make_golden.py

  params.json.gz: svt_aom_sig_deriv_me + friends for a grid of encoder settings
  stage_*.npz   : svt_aom_motion_estimation_b64 results of small synthetic pictures
  pyramid.npz   : padded full / quarter / sixteenth planes of one picture
  kernels.npz   : svt_sad_loop_kernel_c / svt_ext_all_sad_calculation_8x8_16x16_c known answers
Inputs are regenerated from seeds by tests/_util.py, so only outputs are stored.
"""
import ctypes as C
import gzip
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _util as U  # noqa: E402

STAGE_CASES = [
    # name, W, H, preset, temporal layer, refs{(list,idx): pic}, cur, extra derive kwargs
    ("p8_b_1p1", 320, 192, 8, 1, {(0, 0): 0, (1, 0): 4}, 2, {}),
    ("p8_b_3p2", 320, 184, 8, 3, {(0, 0): 2, (0, 1): 1, (0, 2): 0, (1, 0): 4, (1, 1): 5}, 3, {}),
    ("p4_b_4p3", 256, 136, 4, 2, {(0, 0): 2, (0, 1): 1, (0, 2): 0, (0, 3): 0, (1, 0): 4, (1, 1): 5, (1, 2): 6}, 3, {}),
    ("p4_base_2p2", 320, 192, 4, 0, {(0, 0): 0, (0, 1): 2, (1, 0): 0, (1, 1): 2}, 6, {}),
    ("p6_p_1", 344, 200, 6, 0, {(0, 0): 0}, 4, {}),
    ("p2_gm", 256, 128, 2, 1, {(0, 0): 0, (1, 0): 4}, 2, {"gm_level_enabled": 1, "gm_use_distance_based_active_th": 1}),
    ("p10_rtc_sc", 320, 192, 10, 1, {(0, 0): 1, (0, 1): 0}, 2, {"rtc_tune": 1, "sc_class1": 1}),
    ("p12", 320, 192, 12, 2, {(0, 0): 1, (1, 0): 3}, 2, {}),
]


def stage_case(name, W, H, preset, tl, refs, cur, kw):
    seq = U.synth_sequence(W, H, 7, seed=len(name) + W, motion=(3, -2))
    l0 = 1 + max(ri for (li, ri) in refs if li == 0)
    l1 = 1 + max([ri for (li, ri) in refs if li == 1], default=-1)
    d = U.derive_input(preset, W, H, temporal_layer_index=tl, ref_list0_count_try=l0, ref_list1_count_try=l1,
                       slice_type_b=1 if l1 else 0, ref_count_used_list0=max(l0, 1), ref_count_used_list1=l1, **kw)
    return seq, d


def main():
    ref = U.ref_lib()
    assert ref is not None, "oracle/_ref/libsvtref.so missing and /root/reference not mounted"
    ref.ref_set_simd(0)
    # ---- params
    grid = []
    for preset in range(-1, 14):
        for (w, h) in [(320, 192), (1280, 720), (1920, 1080), (3840, 2160), (7680, 4320)]:
            for kw in [dict(), dict(sc_class1=1), dict(rtc_tune=1), dict(rtc_tune=1, sc_class1=1),
                       dict(temporal_layer_index=0), dict(hierarchical_levels=3, qp=20), dict(qp=55, frame_rate_q16=10 << 16),
                       dict(safe_limit_nref=1, safe_limit_zz_th=60000)]:
                d = U.derive_input(preset, w, h, ref_list0_count_try=3, ref_list1_count_try=2, ref_count_used_list0=4,
                                   ref_count_used_list1=3, **kw)
                p = U.abi.SvtMeParams()
                assert ref.ref_derive_params(C.byref(d), C.byref(p)) == 0
                grid.append({"in": {k: getattr(d, k) for k, _ in d._fields_}, "out": p.as_dict()})
    with gzip.open(os.path.join(HERE, "params.json.gz"), "wt") as f:
        json.dump(grid, f, separators=(",", ":"))
    # ---- stage
    for case in STAGE_CASES:
        name, W, H, preset, tl, refs, cur, kw = case
        seq, d = stage_case(*case)
        p = U.ref_params(d, cur, refs)
        pyr = {n: U.CpuPyramid("ref", seq[n], n) for n in set(refs.values()) | {cur}}
        res = U.cpu_me_picture("ref", p, pyr[cur], {k: pyr[v] for k, v in refs.items()})
        np.savez_compressed(os.path.join(HERE, "stage_%s.npz" % name), **res)
    # ---- pyramid
    seq = U.synth_sequence(328, 200, 1, seed=5)
    pyr = U.CpuPyramid("ref", seq[0], 0)
    np.savez_compressed(os.path.join(HERE, "pyramid.npz"), full=pyr.plane(0), quarter=pyr.plane(1), sixteenth=pyr.plane(2))
    # ---- kernels: sad_loop + all_sad on seeded random buffers
    rng = np.random.default_rng(13596)
    src = rng.integers(0, 256, size=(128, 128), dtype=np.uint8)
    refb = rng.integers(0, 256, size=(288, 288), dtype=np.uint8)
    rows = []
    for (bw, bh) in [(16, 8), (16, 16), (32, 16), (64, 32), (64, 64), (8, 8), (24, 12), (48, 24)]:
        for (sw, sh) in [(8, 3), (16, 5), (48, 48), (8, 100), (96, 3), (1, 1), (7, 9), (64, 40)]:
            for skip in (0, 1):
                best = C.c_uint64(0)
                x, y = C.c_int16(0), C.c_int16(0)
                ref.svt_sad_loop_kernel_c(src.ctypes.data_as(C.c_void_p), C.c_uint32(128), refb.ctypes.data_as(C.c_void_p),
                                          C.c_uint32(288), C.c_uint32(bh), C.c_uint32(bw), C.byref(best), C.byref(x),
                                          C.byref(y), C.c_uint32(288), C.c_uint8(skip), C.c_int16(sw), C.c_int16(sh))
                rows.append([bw, bh, sw, sh, skip, best.value, x.value, y.value])
    np.savez_compressed(os.path.join(HERE, "kernels.npz"), sad_loop=np.array(rows, dtype=np.int64))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()

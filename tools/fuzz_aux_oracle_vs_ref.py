"""Differential fuzz on the CPU for the rows next to ME: temporal-filter ME (svtme_derive_params_tf against the reference's own
derivation, oracle against compiled reference incl. early exits), picture statistics and the DG-detector HME, on random
sizes / parameters / content; C and AVX2 kernels alternate. Needs oracle/_ref. Round 2: 190 configurations, no mismatch.
Usage: python tools/fuzz_aux_oracle_vs_ref.py [seed] [configs]"""
import ctypes as C
import os
import sys

import numpy as np

_T = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests")
sys.path.insert(0, _T)
sys.path.insert(0, os.path.join(_T, "golden"))
import _util as U, make_golden_tf as T
rng = np.random.default_rng(int(sys.argv[1])); N = int(sys.argv[2])
ref = U.ref_lib(); bad = 0
for it in range(N):
    ref.ref_set_simd(it & 1)
    W = int(rng.integers(2, 8)) * 64 + int(rng.choice([0, 8, 24, 40, 56])); H = int(rng.integers(2, 5)) * 64 + int(rng.choice([0, 8, 16, 32, 48]))
    seq = U.synth_sequence(W, H, 4, seed=int(rng.integers(1, 999)), motion=(int(rng.integers(-6, 7)), int(rng.integers(-4, 5))), noise=float(rng.choice([0.5, 2.0, 5.0])))
    # TF-ME
    preset, level = int(rng.integers(-1, 14)), int(rng.integers(0, 5))
    d = T.tf_input(preset, W, H, level, int(rng.choice([0, 256, 2000, 6000])), int(rng.integers(0, 2)), tl=int(rng.integers(0, 3)), is_ref=1)
    p = U.abi.SvtMeParams()
    if ref.ref_derive_params_tf(C.byref(d), C.byref(p)) == 0:
        q = U.abi.SvtMeParams()
        assert U.product_lib().svtme_derive_params_tf(C.byref(d), C.byref(q)) == 0
        if bytes(p) != bytes(q): bad += 1; print("TF PARAMS differ", preset, level, W, H)
        T.finish_tf(p, 1, 2)
        res = {}
        for which in ("orc", "ref"):
            pyr = {n: U.CpuPyramid(which, seq[n], n, hme_level1=int(p.enable_hme_level1_flag)) for n in (1, 2)}
            res[which] = U.cpu_me_picture(which, p, pyr[1], {(0, 0): pyr[2]})
        n = res["ref"]["hme_sad"].reshape(-1, 2, 4).shape[0]
        early = res["ref"]["hme_sad"].reshape(n, 2, 4)[:, 0, 0] < p.tf_me_exit_th
        for k in T.TF_KEYS:
            a, b = np.asarray(res["ref"][k]).reshape(n, 2, 4, -1)[:, 0, 0], np.asarray(res["orc"][k]).reshape(n, 2, 4, -1)[:, 0, 0]
            if k in ("p_sb_best_sad", "p_sb_best_mv"): a, b = a[~early], b[~early]
            if not np.array_equal(a, b): bad += 1; print("TF MISMATCH", k, preset, level, W, H)
    # statistics
    sp = U.stats_params(calc_hist=int(rng.random() < 0.9), calculate_variance=int(rng.random() < 0.9), prec_sub=int(rng.random() < 0.5), all_sizes=int(rng.random() < 0.6),
                        scd=int(rng.random() < 0.5), regions=(int(rng.integers(1, 5)), int(rng.integers(1, 5))))
    a = U.cpu_picture_statistics("ref", U.CpuPyramid("ref", seq[0], 0), sp); b = U.cpu_picture_statistics("orc", U.CpuPyramid("orc", seq[0], 0), sp)
    for k in a:
        if not np.array_equal(a[k], b[k]): bad += 1; print("STATS MISMATCH", k, W, H)
    # DG detector
    res_class = int(rng.integers(0, 7))
    x = U.cpu_dg_detector_hme("ref", U.CpuPyramid("ref", seq[3], 3), U.CpuPyramid("ref", seq[0], 0), res_class)
    y = U.cpu_dg_detector_hme("orc", U.CpuPyramid("orc", seq[3], 3), U.CpuPyramid("orc", seq[0], 0), res_class)
    if x != y: bad += 1; print("DG MISMATCH", W, H, res_class, x, y)
ref.ref_set_simd(0)
print("aux fuzz seed", sys.argv[1], "configs", N, "mismatching", bad)

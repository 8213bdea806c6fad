"""Generates tests/golden/stats_golden.npz from the compiled reference (oracle/_ref/libsvtref.so, built from
/root/reference by oracle/Makefile): svt_aom_gathering_picture_statistics on small seeded pictures. Run in the
container that has /root/reference; the fixture travels, the reference does not."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _util as U  # noqa: E402

CASES = [((192, 136), dict()), ((192, 136), dict(prec_sub=0)), ((328, 200), dict(all_sizes=0)), ((328, 200), dict(scd=1)),
         ((128, 64), dict(regions=(1, 1))), ((264, 136), dict(regions=(3, 2), scd=1)), ((264, 136), dict(prec_sub=0, regions=(2, 4))),
         ((328, 200), dict(calc_hist=0)), ((192, 136), dict(calculate_variance=0))]
out = {"n_cases": np.int64(len(CASES))}
assert U.ref_lib() is not None
U.ref_lib().ref_set_simd(0)
for i, ((w, h), cfg) in enumerate(CASES):
    luma = U.synth_sequence(w, h, 2, seed=11 + i, noise=6.0)[1] if i % 2 else np.random.default_rng(i).integers(0, 256, (h, w), dtype=np.uint8)
    p = U.stats_params(**cfg)
    res = U.cpu_picture_statistics("ref", U.CpuPyramid("ref", luma, 0), p)
    out["luma_%d" % i] = luma
    out["cfg_%d" % i] = np.array([p.calc_hist, p.calculate_variance, p.block_mean_prec_sub, p.all_block_sizes, p.scene_change_detection,
                                  p.regions_per_width, p.regions_per_height], np.int64)
    for k, v in res.items():
        out["%s_%d" % (k, i)] = v
np.savez_compressed(os.path.join(HERE, "stats_golden.npz"), **out)
print("wrote stats_golden.npz:", len(CASES), "cases")

"""Golden fixtures of the temporal-filter ME (ME_MCTF) from the UNMODIFIED reference compiled under oracle/_ref (needs
/root/reference; run in the build container):

    python tests/golden/make_golden_tf.py

  tf_params.json.gz : svt_aom_sig_deriv_me_tf (+ what produce_temporally_filtered_pic sets per block) for a grid
  tf_stage_*.npz    : svt_aom_motion_estimation_b64 with me_type = ME_MCTF on small synthetic pictures: the tables and HME
                      results the temporal filter reads (p_sb_best_sad / p_sb_best_mv, search centre, HME SAD)
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

# name, W, H, preset, tf hme_me_level, me_exit_th, centre picture, neighbour frame, qp_opt
TF_CASES = [
    # (exit thresholds chosen so that some blocks return early after HME and others run the full-pel search; the
    # encoder's own values are 16 * 16 ... for its level-based presets)
    ("tf_l0", 320, 192, 2, 0, 0, 3, 2, 0),
    ("tf_l1_exit", 328, 184, 5, 1, 3000, 3, 5, 0),
    ("tf_l2", 256, 136, 7, 2, 5000, 2, 3, 1),
    ("tf_l3_sub", 320, 192, 9, 3, 4800, 3, 1, 1),
    ("tf_l4_sub", 344, 200, 12, 4, 1000, 4, 3, 0),
]
TF_KEYS = ("p_sb_best_sad", "p_sb_best_mv", "hme_sc", "hme_sad", "do_ref")


def tf_input(preset, W, H, level, exit_th, qp_opt, tl=0, is_ref=1):
    d = U.abi.SvtMeTfDeriveInput()
    d.enc_mode, d.width, d.height, d.qp = preset, W, H, 35
    d.hme_me_level, d.qp_opt, d.me_exit_th = level, qp_opt, exit_th
    d.temporal_layer_index, d.is_ref = tl, is_ref
    return d


def tf_case(name, W, H, preset, level, exit_th, centre, frame, qp_opt):
    # little motion and noise: some blocks fall under tf_me_exit_th (early return after HME), others do not
    seq = U.synth_sequence(W, H, 7, seed=len(name) + W, motion=(1, 0), noise=1.0)
    return seq, tf_input(preset, W, H, level, exit_th, qp_opt)


def finish_tf(p, centre, frame):
    p.picture_number = centre
    p.ref_picture_number[0][0] = frame
    return p


def tf_outputs(res, p):
    """What the temporal filter reads: for blocks that returned early only the HME results of (list 0, ref 0)."""
    out = {k: np.array(res[k], copy=True) for k in TF_KEYS}
    early = out["hme_sad"].reshape(-1, 2, 4)[:, 0, 0] < p.tf_me_exit_th
    out["tf_early_exit"] = early.astype(np.uint8)
    return out


def main():
    ref = U.ref_lib()
    assert ref is not None, "oracle/_ref/libsvtref.so missing and /root/reference not mounted"
    ref.ref_set_simd(0)
    grid = []
    for preset in range(-1, 14):
        for (w, h) in [(320, 192), (856, 480), (1280, 720), (1920, 1080), (3840, 2160)]:
            for level in range(5):
                for qp_opt, exit_th in ((0, 0), (1, 16 * 16)):
                    d = tf_input(preset, w, h, level, exit_th, qp_opt, tl=level % 3, is_ref=1)
                    p = U.abi.SvtMeParams()
                    assert ref.ref_derive_params_tf(C.byref(d), C.byref(p)) == 0
                    grid.append({"in": {k: getattr(d, k) for k, _ in d._fields_ if not k.startswith("_")}, "out": p.as_dict()})
    with gzip.open(os.path.join(HERE, "tf_params.json.gz"), "wt") as f:
        json.dump(grid, f, separators=(",", ":"))
    for case in TF_CASES:
        name, W, H, preset, level, exit_th, centre, frame, qp_opt = case
        seq, d = tf_case(*case)
        p = U.abi.SvtMeParams()
        assert ref.ref_derive_params_tf(C.byref(d), C.byref(p)) == 0
        finish_tf(p, centre, frame)
        pyr = {n: U.CpuPyramid("ref", seq[n], n) for n in (centre, frame)}
        res = U.cpu_me_picture("ref", p, pyr[centre], {(0, 0): pyr[frame]})
        out = tf_outputs(res, p)
        print(name, "blocks", len(out["tf_early_exit"]), "early exits", int(out["tf_early_exit"].sum()))
        np.savez_compressed(os.path.join(HERE, "tf_stage_%s.npz" % name), **out)
    print("TF golden fixtures written to", HERE)


if __name__ == "__main__":
    main()

"""Shared test helpers: library loaders (oracle restatement, compiled reference, product) and
synthetic picture generators. The oracle libraries are CHECKERS only (see oracle/me_oracle.h)."""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "svt-av1-mod-by-patman_b200"
abi = importlib.import_module(PKG + ".abi")

ORACLE_SO = os.path.join(ROOT, "oracle", "libme_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libsvtref.so")
_libs = {}


def oracle_lib():
    if "orc" not in _libs:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], stdout=subprocess.DEVNULL)
        _libs["orc"] = C.CDLL(ORACLE_SO)
    return _libs["orc"]


def ref_lib():
    """Compiled reference (oracle/_ref). None when it was never built and /root/reference is absent."""
    if "ref" not in _libs:
        if not os.path.exists(REF_SO) and os.path.isdir("/root/reference/Source"):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], stdout=subprocess.DEVNULL)
        _libs["ref"] = C.CDLL(REF_SO) if os.path.exists(REF_SO) else None
    return _libs["ref"]


def product_lib():
    return abi.load_library()


# ---------------------------------------------------------------- synthetic pictures
def synth_sequence(width, height, n, seed=1, motion=(2, 1), noise=2.0, rects=16):
    """Config-2 style luma: band-limited noise translating `motion` px/frame,
    moving 64x64 rectangles, per-frame noise. Returns uint8 array [n, height, width]."""
    rng = np.random.default_rng(seed)
    big_h, big_w = height + 64 + abs(motion[1]) * n + 8, width + 64 + abs(motion[0]) * n + 8
    base = rng.integers(0, 256, size=(big_h // 8 + 2, big_w // 8 + 2)).astype(np.float32)
    base = np.kron(base, np.ones((8, 8), dtype=np.float32))[:big_h, :big_w]
    k = np.ones(5, dtype=np.float32) / 5
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, base)
    base = np.apply_along_axis(lambda c: np.convolve(c, k, mode="same"), 0, base)
    fine = rng.integers(-24, 25, size=(big_h, big_w)).astype(np.float32)
    base = np.clip(base + fine, 0, 255)
    rx = rng.integers(0, max(1, width - 64), size=rects)
    ry = rng.integers(0, max(1, height - 64), size=rects)
    rv = rng.integers(-6, 7, size=(rects, 2))
    rc = rng.integers(0, 256, size=rects)
    out = np.empty((n, height, width), dtype=np.uint8)
    for t in range(n):
        ox, oy = 32 + max(0, motion[0]) * t if motion[0] >= 0 else 32 - motion[0] * (n - t), 0
        ox = 32 + (motion[0] * t if motion[0] >= 0 else -motion[0] * (n - 1 - t))
        oy = 32 + (motion[1] * t if motion[1] >= 0 else -motion[1] * (n - 1 - t))
        f = base[oy:oy + height, ox:ox + width].copy()
        for i in range(rects):
            x = int(np.clip(rx[i] + rv[i, 0] * t, 0, max(0, width - 64)))
            y = int(np.clip(ry[i] + rv[i, 1] * t, 0, max(0, height - 64)))
            f[y:y + 64, x:x + 64] = rc[i]
        if noise > 0:
            f = f + rng.normal(0, noise, size=f.shape).astype(np.float32)
        out[t] = np.clip(np.rint(f), 0, 255).astype(np.uint8)
    return out


def random_frames(width, height, n, seed=13596):
    """Config-1 style: i.i.d. uniform u8 frames (the SadTest RANDOM pattern)."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n, height, width), dtype=np.uint8)


# ---------------------------------------------------------------- oracle / reference drivers
class CpuPyramid:
    """A pyramid built by the oracle restatement ('orc') or the compiled reference ('ref')."""

    def __init__(self, which, luma, picture_number, hme_level1=1):
        self.which = which
        self.lib = oracle_lib() if which == "orc" else ref_lib()
        self.pyr = abi.Pyramid()
        luma = np.ascontiguousarray(luma)
        h, w = luma.shape
        fn = getattr(self.lib, which + "_pyramid_build")
        rc = fn(C.byref(self.pyr), luma.ctypes.data_as(C.c_void_p), C.c_uint32(w), C.c_uint32(w), C.c_uint32(h),
                C.c_int(hme_level1), C.c_uint64(picture_number))
        assert rc == 0

    def plane(self, level):
        pl = [self.pyr.full, self.pyr.quarter, self.pyr.sixteenth][level]
        rows = pl.height + 2 * pl.org_y
        buf = (C.c_uint8 * (rows * pl.stride_y)).from_address(pl.buffer_y)
        return np.frombuffer(buf, dtype=np.uint8).reshape(rows, pl.stride_y)

    def __del__(self):
        try:
            getattr(self.lib, self.which + "_pyramid_free")(C.byref(self.pyr))
        except Exception:
            pass


def cpu_me_picture(which, params, cur, refs, deep=True, rows=None):
    """Run the whole-picture stage on the CPU checker. refs: dict {(list, idx): CpuPyramid}."""
    lib = oracle_lib() if which == "orc" else ref_lib()
    w, h = cur.pyr.full.width, cur.pyr.full.height
    n_b64 = ((w + 63) // 64) * ((h + 63) // 64)
    out, arrays = abi.alloc_out(n_b64, params, deep=deep)
    RefArr = (C.POINTER(abi.Pyramid) * abi.MAX_REFS) * abi.MAX_LISTS
    ra = RefArr()
    for (li, ri), pyr in refs.items():
        ra[li][ri] = C.pointer(pyr.pyr)
    if rows is None:
        rc = getattr(lib, which + "_me_picture")(C.byref(params), C.byref(cur.pyr), C.byref(ra), C.byref(out))
    else:
        rc = getattr(lib, which + "_me_picture_rows")(C.byref(params), C.byref(cur.pyr), C.byref(ra), C.byref(out),
                                                     C.c_uint32(rows[0]), C.c_uint32(rows[1]))
    assert rc == 0
    return arrays


def derive_input(enc_mode, width, height, **kw):
    d = abi.SvtMeDeriveInput()
    d.enc_mode, d.width, d.height = enc_mode, width, height
    d.qp = kw.get("qp", 35)
    d.hierarchical_levels = kw.get("hierarchical_levels", 5)
    d.temporal_layer_index = kw.get("temporal_layer_index", 1)
    d.is_ref = kw.get("is_ref", 1)
    d.slice_type_b = kw.get("slice_type_b", 1)
    d.ref_list0_count_try = kw.get("ref_list0_count_try", 1)
    d.ref_list1_count_try = kw.get("ref_list1_count_try", 1)
    d.frame_rate_q16 = kw.get("frame_rate_q16", 30 << 16)
    d.ref_count_used_list0 = kw.get("ref_count_used_list0", max(1, d.ref_list0_count_try))
    d.ref_count_used_list1 = kw.get("ref_count_used_list1", d.ref_list1_count_try)
    for k in ("rtc_tune", "sc_class1", "safe_limit_nref", "only_l_bwd", "use_best_me_unipred_cand_only",
              "similar_brightness_refs", "safe_limit_zz_th", "gm_level_enabled", "gm_use_distance_based_active_th"):
        setattr(d, k, kw.get(k, 0))
    return d


def finish_params(p, d, picture_number, ref_numbers):
    """Per-picture fields the encoder sets in me_process.c:219-227 / pd_process.c."""
    p.num_of_list_to_search = 2 if d.slice_type_b else 1
    p.num_of_ref_pic_to_search[0] = d.ref_list0_count_try
    p.num_of_ref_pic_to_search[1] = d.ref_list1_count_try if d.slice_type_b else 0
    p.temporal_layer_index = d.temporal_layer_index
    p.is_ref = d.is_ref
    p.hierarchical_levels = d.hierarchical_levels
    p.similar_brightness_refs = d.similar_brightness_refs
    p.max_number_of_pus_per_sb = 85
    p.only_l_bwd = d.only_l_bwd
    p.gm_enabled = d.gm_level_enabled
    p.gm_use_distance_based_active_th = d.gm_use_distance_based_active_th
    p.picture_number = picture_number
    for (li, ri), n in ref_numbers.items():
        p.ref_picture_number[li][ri] = n
    return p


def ref_params(d, picture_number, ref_numbers):
    """Parameters derived by the reference's own svt_aom_sig_deriv_me (oracle/_ref)."""
    p = abi.SvtMeParams()
    assert ref_lib().ref_derive_params(C.byref(d), C.byref(p)) == 0
    return finish_params(p, d, picture_number, ref_numbers)

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
me_host = importlib.import_module(PKG + ".me")

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


synth = importlib.import_module(PKG + ".synth")
synth_sequence, random_frames = synth.synth_sequence, synth.random_frames


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


def stats_params(calc_hist=1, calculate_variance=1, prec_sub=1, all_sizes=1, scd=0, regions=(4, 4)):
    p = abi.SvtMeStatsParams()
    p.calc_hist, p.calculate_variance, p.block_mean_prec_sub = calc_hist, calculate_variance, prec_sub
    p.all_block_sizes, p.scene_change_detection = all_sizes, scd
    p.regions_per_width, p.regions_per_height = regions
    return p


def cpu_picture_statistics(which, pyr, params):
    """svt_aom_gathering_picture_statistics on a CpuPyramid: the oracle restatement ('orc') or the compiled reference ('ref')."""
    lib = oracle_lib() if which == "orc" else ref_lib()
    w, h = pyr.pyr.full.width, pyr.pyr.full.height
    out, arrays = abi.alloc_stats_out(((w + 63) // 64) * ((h + 63) // 64), params)
    rc = getattr(lib, which + "_picture_statistics")(C.byref(pyr.pyr), C.byref(params), C.byref(out))
    assert rc == 0
    return arrays


def cpu_dg_detector_hme(which, src, ref, input_resolution):
    """dg_detector_hme_level0 over the whole picture between two CpuPyramids: (tot_dist, tot_cplx, tot_active, sum_in_vectors)."""
    lib = oracle_lib() if which == "orc" else ref_lib()
    m = abi.SvtMeDgMetrics()
    rc = getattr(lib, which + "_dg_detector_hme")(C.byref(src.pyr), C.byref(ref.pyr), C.c_uint32(input_resolution), C.byref(m))
    assert rc == 0
    return (int(m.tot_dist), int(m.tot_cplx), int(m.tot_active), int(m.sum_in_vectors))


derive_input, finish_params = me_host.derive_input, me_host.finish_params


def ref_params(d, picture_number, ref_numbers):
    """Parameters derived by the reference's own svt_aom_sig_deriv_me (oracle/_ref)."""
    p = abi.SvtMeParams()
    assert ref_lib().ref_derive_params(C.byref(d), C.byref(p)) == 0
    return finish_params(p, d, picture_number, ref_numbers)


# ---------------------------------------------------------------- all-cores drivers (bench.py's CPU arm, big parity cases)
class CpuMtJob(C.Structure):  # RefMtJob / OrcMtJob (oracle/ref_harness.c, oracle/me_oracle_stage.c)
    _fields_ = [("params", abi.SvtMeParams), ("cur", C.POINTER(abi.Pyramid)),
                ("refs", (C.POINTER(abi.Pyramid) * abi.MAX_REFS) * abi.MAX_LISTS), ("out", abi.SvtMeOut)]


class CpuPyramidSet:
    """Pyramids of several pictures built on a pthread pool; planes are recycled when the set is rebuilt."""

    def __init__(self, which):
        self.which = which
        self.lib = oracle_lib() if which == "orc" else ref_lib()
        self.nums, self.arr = [], None

    def build(self, frames, threads, hme_level1=1):
        """frames: {picture_number: 2-D uint8 array}, all of one size."""
        nums = sorted(frames)
        if nums != self.nums:
            self.free()
            self.nums, self.arr = nums, (abi.Pyramid * len(nums))()
        imgs = [np.ascontiguousarray(frames[n]) for n in nums]
        h, w = imgs[0].shape
        ptrs = (C.c_void_p * len(nums))(*[im.ctypes.data for im in imgs])
        pn = (C.c_uint64 * len(nums))(*nums)
        fn = getattr(self.lib, self.which + "_pyramids_build_mt")
        rc = fn(self.arr, ptrs, C.c_uint32(w), C.c_uint32(w), C.c_uint32(h), C.c_int(hme_level1), pn, C.c_uint32(len(nums)),
                C.c_uint32(threads))
        assert rc == 0
        return self

    def __getitem__(self, n):
        return self.arr[self.nums.index(n)]

    def free(self):
        if self.arr is not None:
            for i in range(len(self.nums)):
                getattr(self.lib, self.which + "_pyramid_free")(C.byref(self.arr[i]))
        self.nums, self.arr = [], None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def cpu_mt_jobs(which, specs, pyrs, n_b64, deep=False):
    """specs: [(params, cur, {(list, idx): picture number})]; pyrs: CpuPyramidSet. Returns (job array, [arrays])."""
    lib = oracle_lib() if which == "orc" else ref_lib()
    assert getattr(lib, which + "_sizeof_mt_job")() == C.sizeof(CpuMtJob)
    jobs = (CpuMtJob * len(specs))()
    outs = []
    for i, (p, cur, refs) in enumerate(specs):
        jobs[i].params = p
        jobs[i].cur = C.pointer(pyrs[cur])
        for (li, ri), n in refs.items():
            jobs[i].refs[li][ri] = C.pointer(pyrs[n])
        out, arrays = abi.alloc_out(n_b64, p, deep=deep)
        jobs[i].out = out
        outs.append(arrays)
    return jobs, outs


def cpu_me_pictures_mt(which, jobs, threads, seg_rows=1):
    lib = oracle_lib() if which == "orc" else ref_lib()
    rc = getattr(lib, which + "_me_pictures_mt")(jobs, C.c_uint32(len(jobs)), C.c_uint32(threads), C.c_uint32(seg_rows))
    assert rc == 0


def best_cpu_checker():
    """('ref', lib) when the compiled reference is available (AVX2 kernels on), else ('orc', lib)."""
    ref = ref_lib()
    if ref is not None:
        ref.ref_set_simd(1)
        return "ref"
    return "orc"

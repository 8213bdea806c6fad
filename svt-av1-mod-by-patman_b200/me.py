"""Host-side mirror of the reference's interface for the open-loop ME path.

The product is the C ABI (include/svtme_b200.h, csrc/libsvtme_b200.so); this module only wraps it
for Python callers (tests, bench): numpy arrays in, numpy arrays out, errors raised. It never
computes pixels itself and has no CPU fallback: without the CUDA library it raises.

Reference interface mirrored: the PAME branch of svt_aom_motion_estimation_kernel
(Source/Lib/Codec/me_process.c:126-313) -> MeSession.run_picture(s);
svt_aom_downsample_filtering_input_picture (pic_analysis_process.c:1922) -> MeSession.put_picture.
"""
import ctypes as C

import numpy as np

from . import abi


class SvtMeError(RuntimeError):
    pass


def _check(rc, lib):
    if rc != 0:
        raise SvtMeError("svtme error %d: %s" % (rc, lib.svtme_last_error().decode()))


def launch_count():
    return int(abi.load_library().svtme_launch_count())


class MeSession:
    """One handle = one GPU (one process per GPU when sharding pictures across GPUs)."""

    def __init__(self, device, width, height, capacity=16):
        self.lib = abi.load_library()
        self.h = C.c_void_p()
        self.width, self.height = width, height
        _check(self.lib.svtme_open(device, width, height, capacity, C.byref(self.h)), self.lib)
        self.n_b64 = int(self.lib.svtme_b64_count(self.h))
        self._pinned = []

    def close(self):
        if self.h:
            self.lib.svtme_sync(self.h)
            self.unpin_all()
            self.lib.svtme_close(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def derive_params(self, derive_input):
        p = abi.SvtMeParams()
        _check(self.lib.svtme_derive_params(C.byref(derive_input), C.byref(p)), self.lib)
        return p

    def put_picture(self, pic_num, luma):
        luma = np.ascontiguousarray(luma, dtype=np.uint8)
        assert luma.shape == (self.height, self.width)
        _check(self.lib.svtme_put_picture(self.h, pic_num, luma.ctypes.data, luma.strides[0]), self.lib)

    def put_picture_device(self, pic_num, dev_ptr, stride):
        _check(self.lib.svtme_put_picture_device(self.h, pic_num, dev_ptr, stride), self.lib)

    def release_picture(self, pic_num):
        _check(self.lib.svtme_release_picture(self.h, pic_num), self.lib)

    def put_pictures(self, picture_numbers, ptrs, stride, device=False, bits=8):
        """Batched pyramid build: ptrs = host (or device) addresses of the luma planes. bits=10: 16-bit samples,
        stride in samples; the ME planes get sample >> 2."""
        n = len(picture_numbers)
        nums = (C.c_uint64 * n)(*picture_numbers)
        arr = (C.c_void_p * n)(*ptrs)
        if bits == 10:
            fn = self.lib.svtme_put_pictures_10bit_device if device else self.lib.svtme_put_pictures_10bit
        else:
            fn = self.lib.svtme_put_pictures_device if device else self.lib.svtme_put_pictures
        _check(fn(self.h, n, nums, arr, stride), self.lib)

    def get_plane(self, pic_num, level, pad=None):
        pad = (72, 32, 16)[level] if pad is None else pad
        w, h = self.width >> level, self.height >> level
        out = np.zeros((h + 2 * pad, w + 2 * pad), dtype=np.uint8)
        _check(self.lib.svtme_get_plane(self.h, pic_num, level, out.ctypes.data, out.strides[0], pad), self.lib)
        return out

    def plane_device_ptr(self, pic_num, level):
        ptr, n = C.c_void_p(), C.c_size_t()
        _check(self.lib.svtme_plane_device_ptr(self.h, pic_num, level, C.byref(ptr), C.byref(n)), self.lib)
        return ptr.value, n.value

    def adopt_picture(self, pic_num):
        _check(self.lib.svtme_adopt_picture(self.h, pic_num), self.lib)

    def external_write_begin(self, picture_numbers, cuda_stream_ptr):
        """Ring slots for the pictures + `cuda_stream` waits for pending builds and for launches that still read the slots."""
        n = len(picture_numbers)
        _check(self.lib.svtme_external_write_begin(self.h, n, (C.c_uint64 * n)(*picture_numbers), cuda_stream_ptr), self.lib)

    def external_write_end(self, cuda_stream_ptr):
        """Later launches / uploads wait for everything enqueued on `cuda_stream` so far (the exchange)."""
        _check(self.lib.svtme_external_write_end(self.h, cuda_stream_ptr), self.lib)

    def share_picture_to(self, other, pic_num):
        """Peer copy of a pyramid into a handle on another device of the same process."""
        _check(self.lib.svtme_share_picture(self.h, other.h, pic_num), self.lib)

    BIG_TABLES = ("total_me_candidate_index", "me_mv_array", "me_candidate_array")

    def make_jobs(self, specs, deep=False, pin=False):
        """specs: list of (params, cur, {(list, idx): pic_num}). Returns (SvtMeJob array, [arrays]). pin: page-lock the three
        big result tables of every job (svtme_pin_host_memory) so that the queued form downloads straight into them."""
        jobs = (abi.SvtMeJob * len(specs))()
        outs = []
        for i, (p, cur, refs) in enumerate(specs):
            jobs[i].params = p
            jobs[i].cur = cur
            for (li, ri), n in refs.items():
                jobs[i].refs[li][ri] = n
            out, arrays = abi.alloc_out(self.n_b64, p, deep=deep)
            if pin:
                for name in self.BIG_TABLES:
                    _check(self.lib.svtme_pin_host_memory(arrays[name].ctypes.data, arrays[name].nbytes), self.lib)
                    self._pinned.append(arrays[name])
            jobs[i].out = out
            outs.append(arrays)
        return jobs, outs

    def unpin_all(self):
        for a in self._pinned:
            self.lib.svtme_unpin_host_memory(a.ctypes.data)
        self._pinned = []

    def run_jobs(self, jobs):
        _check(self.lib.svtme_run_pictures(self.h, jobs, len(jobs)), self.lib)

    def launch_jobs(self, jobs):
        _check(self.lib.svtme_launch_pictures(self.h, jobs, len(jobs)), self.lib)

    def fetch_jobs(self, jobs):
        _check(self.lib.svtme_fetch_results(self.h, jobs, len(jobs)), self.lib)

    def submit_jobs(self, jobs):
        """Queued form: enqueue the batch and its result download; several batches may be in flight."""
        _check(self.lib.svtme_submit_pictures(self.h, jobs, len(jobs)), self.lib)

    def collect_jobs(self, jobs):
        """Block until the OLDEST submitted batch is complete; fills the arrays made by make_jobs."""
        _check(self.lib.svtme_collect_results(self.h, jobs, len(jobs)), self.lib)

    def pending_batches(self):
        return int(self.lib.svtme_pending_batches(self.h))

    def sync(self):
        _check(self.lib.svtme_sync(self.h), self.lib)

    def set_stream(self, cuda_stream_ptr):
        _check(self.lib.svtme_set_stream(self.h, cuda_stream_ptr), self.lib)

    def last_work_ops(self):
        v = C.c_uint64()
        _check(self.lib.svtme_last_work_ops(self.h, C.byref(v)), self.lib)
        return int(v.value)

    PHASE_NAMES = ("state_clear", "zz_sad", "prehme", "hme_l0", "hme_l1", "hme_l2_centre", "integer_search", "finish")

    def last_phase_ms(self):
        """Device milliseconds per launch slot of the stage pipeline in the last launch."""
        v = (C.c_float * len(self.PHASE_NAMES))()
        _check(self.lib.svtme_last_phase_ms(self.h, v, len(self.PHASE_NAMES)), self.lib)
        return dict(zip(self.PHASE_NAMES, (float(x) for x in v)))

    def last_phase_work(self):
        """Algorithmic SAD byte-ops per launch slot of the stage pipeline in the last launch."""
        v = (C.c_uint64 * len(self.PHASE_NAMES))()
        _check(self.lib.svtme_last_phase_work(self.h, v, len(self.PHASE_NAMES)), self.lib)
        return dict(zip(self.PHASE_NAMES, (int(x) for x in v)))

    def set_search_timing(self, on=True):
        """Later launches also time the search launches of each slot on their own (svtme_last_search_ms)."""
        _check(self.lib.svtme_set_search_timing(self.h, int(bool(on))), self.lib)

    def last_search_ms(self):
        """Device milliseconds of the search launches of each slot (no bookkeeping launches) in the last launch."""
        v = (C.c_float * len(self.PHASE_NAMES))()
        _check(self.lib.svtme_last_search_ms(self.h, v, len(self.PHASE_NAMES)), self.lib)
        return dict(zip(self.PHASE_NAMES, (float(x) for x in v)))

    def last_kernel_ms(self):
        return float(self.lib.svtme_last_kernel_ms(self.h))

    def run_picture(self, params, cur, refs, deep=False):
        jobs, outs = self.make_jobs([(params, cur, refs)], deep=deep)
        self.run_jobs(jobs)
        return outs[0]


def derive_input(enc_mode, width, height, **kw):
    """Encoder-level settings -> SvtMeDeriveInput (defaults: CRF qp 35, random access, 30 fps)."""
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

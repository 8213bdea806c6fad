"""ctypes mirror of include/svtme_b200.h (the C ABI of the B200 open-loop ME path).

Struct layouts follow the header field by field; `check_layout(lib)` compares sizeof() with the
loaded library so that a drifted header fails loudly instead of corrupting memory.
Reference structures mirrored: MeContext (Source/Lib/Codec/me_context.h:366-509), MeSbResults
(me_sb_results.h:29-50).
"""
import ctypes as C
import os

MAX_LISTS, MAX_REFS, PU_COUNT = 2, 4, 85
MAX_SAD_VALUE = 128 * 128 * 255

u8, u16, u32, u64, i16, i32 = C.c_uint8, C.c_uint16, C.c_uint32, C.c_uint64, C.c_int16, C.c_int32


class SearchArea(C.Structure):
    _fields_ = [("width", u16), ("height", u16)]


class SearchAreaMinMax(C.Structure):
    _fields_ = [("sa_min", SearchArea), ("sa_max", SearchArea)]


class SvtMeParams(C.Structure):
    _fields_ = [
        ("me_sa", SearchAreaMinMax), ("hme_l0_sa", SearchAreaMinMax),
        ("hme_l1_sa", SearchArea), ("hme_l2_sa", SearchArea),
        ("num_hme_sa_w", u16), ("num_hme_sa_h", u16),
        ("enable_hme_flag", u8), ("enable_hme_level0_flag", u8), ("enable_hme_level1_flag", u8),
        ("enable_hme_level2_flag", u8),
        ("hme_search_method", u8), ("me_search_method", u8),
        ("reduce_hme_l0_sr_th_min", u8), ("reduce_hme_l0_sr_th_max", u8),
        ("prehme_enable", u8), ("prehme_skip_search_line", u8), ("prehme_l1_early_exit", u8), ("_pad0", u8),
        ("prehme_sa_cfg", SearchAreaMinMax * 2),
        ("enable_me_hme_ref_pruning", u8), ("_pad1", u8),
        ("prune_ref_if_hme_sad_dev_bigger_than_th", u16), ("prune_ref_if_me_sad_dev_bigger_than_th", u16),
        ("zz_sad_pct", u16), ("zz_sad_th", u32), ("phme_sad_th", u32), ("phme_sad_pct", u16),
        ("enable_me_sr_adjustment", u8), ("distance_based_hme_resizing", u8),
        ("reduce_me_sr_based_on_mv_length_th", u16), ("stationary_hme_sad_abs_th", u16),
        ("stationary_me_sr_divisor", u16), ("reduce_me_sr_based_on_hme_sad_abs_th", u16),
        ("me_sr_divisor_for_low_hme_sad", u16),
        ("me_8x8_var_enabled", u8), ("_pad2", u8),
        ("me_sr_div4_th", u32), ("me_sr_div2_th", u32), ("me_sr_mult2_th", u32),
        ("mv_sa_adj_enabled", u8), ("mv_sa_adj_nearest_ref_only", u8),
        ("mv_sa_adj_mv_size_th", u16), ("mv_sa_adj_sa_multiplier", u16),
        ("prune_me_candidates_th", i32),
        ("use_best_unipred_cand_only", u8), ("me_type", u8), ("tf_me_exit_th", C.c_uint16),
        ("me_early_exit_th", u32), ("me_safe_limit_zz_th", u32), ("prev_me_stage_based_exit_th", u32),
        ("num_of_list_to_search", u8), ("num_of_ref_pic_to_search", u8 * 2),
        ("temporal_layer_index", u8), ("is_ref", u8),
        ("hierarchical_levels", u8), ("similar_brightness_refs", u8),
        ("enable_me_8x8", u8), ("enable_me_16x16", u8), ("max_number_of_pus_per_sb", u8),
        ("only_l_bwd", u8), ("gm_enabled", u8), ("gm_use_distance_based_active_th", u8),
        ("input_resolution_le_480p", u8),
        ("max_cand", u8), ("max_refs", u8), ("max_l0", u8), ("_pad4", u8 * 3),
        ("picture_number", u64),
        ("ref_picture_number", (u64 * MAX_REFS) * MAX_LISTS),
    ]

    def as_dict(self):
        def conv(v):
            if isinstance(v, C.Structure):
                return {k: conv(getattr(v, k)) for k, _ in v._fields_}
            if isinstance(v, C.Array):
                return [conv(x) for x in v]
            return v
        return {k: conv(getattr(self, k)) for k, _ in self._fields_ if not k.startswith("_pad")}


class SvtMeDeriveInput(C.Structure):
    _fields_ = [
        ("enc_mode", i32), ("width", u32), ("height", u32), ("qp", u32),
        ("rtc_tune", u8), ("sc_class1", u8), ("hierarchical_levels", u8), ("temporal_layer_index", u8),
        ("is_ref", u8), ("slice_type_b", u8), ("ref_list0_count_try", u8), ("ref_list1_count_try", u8),
        ("frame_rate_q16", u32),
        ("safe_limit_nref", u8), ("only_l_bwd", u8), ("use_best_me_unipred_cand_only", u8),
        ("similar_brightness_refs", u8),
        ("safe_limit_zz_th", u32),
        ("ref_count_used_list0", u8), ("ref_count_used_list1", u8),
        ("gm_level_enabled", u8), ("gm_use_distance_based_active_th", u8),
    ]


class SvtMeFullpelItem(C.Structure):  # include/svtme_b200.h
    _fields_ = [("src_offset", u32), ("ref_offset", u32), ("x_search_area_origin", C.c_int16), ("y_search_area_origin", C.c_int16),
                ("search_area_width", C.c_uint16), ("search_area_height", C.c_uint16)]


class SvtMeTfDeriveInput(C.Structure):  # include/svtme_b200.h
    _fields_ = [("enc_mode", C.c_int32), ("width", u32), ("height", u32), ("qp", u32), ("hme_me_level", u8), ("qp_opt", u8),
                ("me_exit_th", C.c_uint16), ("temporal_layer_index", u8), ("is_ref", u8), ("_pad", u8 * 2)]


class SvtMeOut(C.Structure):
    _fields_ = [
        ("total_me_candidate_index", C.c_void_p), ("me_mv_array", C.c_void_p), ("me_candidate_array", C.c_void_p),
        ("me_64x64_distortion", C.c_void_p), ("me_32x32_distortion", C.c_void_p),
        ("me_16x16_distortion", C.c_void_p), ("me_8x8_distortion", C.c_void_p),
        ("me_8x8_cost_variance", C.c_void_p), ("rc_me_distortion", C.c_void_p),
        ("stationary_block_present_sb", C.c_void_p), ("rc_me_allow_gm", C.c_void_p),
        ("p_sb_best_sad", C.c_void_p), ("p_sb_best_mv", C.c_void_p),
        ("hme_sc", C.c_void_p), ("hme_sad", C.c_void_p), ("do_ref", C.c_void_p),
    ]


class SvtMeDgMetrics(C.Structure):  # include/svtme_b200.h
    _fields_ = [("tot_dist", u64), ("tot_cplx", u32), ("tot_active", u32), ("sum_in_vectors", C.c_int32)]


class SvtMeStatsParams(C.Structure):  # include/svtme_b200.h
    _fields_ = [("calc_hist", C.c_uint8), ("calculate_variance", C.c_uint8), ("block_mean_prec_sub", C.c_uint8),
                ("all_block_sizes", C.c_uint8), ("scene_change_detection", C.c_uint8), ("_pad", C.c_uint8 * 3),
                ("regions_per_width", u32), ("regions_per_height", u32)]


class SvtMeStatsOut(C.Structure):
    _fields_ = [("variance", C.c_void_p), ("pic_avg_variance", C.c_void_p), ("picture_histogram", C.c_void_p),
                ("average_intensity_per_region", C.c_void_p), ("avg_luma", C.c_void_p)]


def alloc_stats_out(n_b64, params):
    """Host arrays of one svtme_picture_statistics call (layouts of SvtMeStatsOut) and the struct pointing at them."""
    import numpy as np
    arrays = {
        "variance": np.zeros((n_b64, 85 if params.all_block_sizes else 1), np.uint16),
        "pic_avg_variance": np.zeros(1, np.uint16),
        "picture_histogram": np.zeros((4, 4, 256), np.uint32),
        "average_intensity_per_region": np.zeros((4, 4), np.uint64),
        "avg_luma": np.zeros(1, np.uint64),
    }
    out = SvtMeStatsOut()
    for name, a in arrays.items():
        setattr(out, name, a.ctypes.data)
    return out, arrays


class SvtMeJob(C.Structure):
    _fields_ = [("params", SvtMeParams), ("cur", u64), ("refs", (u64 * MAX_REFS) * MAX_LISTS), ("out", SvtMeOut)]


class Plane(C.Structure):  # OrcPlane / RefPlane (oracle side only)
    _fields_ = [("buffer_y", C.c_void_p), ("stride_y", u32), ("org_x", u32), ("org_y", u32),
                ("width", u32), ("height", u32)]


class Pyramid(C.Structure):  # OrcPyramid / RefPyramid (oracle side only)
    _fields_ = [("full", Plane), ("quarter", Plane), ("sixteenth", Plane), ("picture_number", u64)]


def exported_pu_count(p):
    """n_pu of MeSbResults: svt_aom_me_sb_results_ctor, Source/Lib/Codec/pcs.c:107-112."""
    if not p.enable_me_16x16:
        return 5
    return p.max_number_of_pus_per_sb if p.enable_me_8x8 else 21


OUT_SPECS = {  # name -> (numpy dtype, per-b64 element count as f(n_pu, params))
    "total_me_candidate_index": ("u1", lambda n, p: n),
    "me_mv_array": ("<u4", lambda n, p: n * p.max_refs),
    "me_candidate_array": ("u1", lambda n, p: n * p.max_cand),
    "me_64x64_distortion": ("<u4", lambda n, p: 1), "me_32x32_distortion": ("<u4", lambda n, p: 1),
    "me_16x16_distortion": ("<u4", lambda n, p: 1), "me_8x8_distortion": ("<u4", lambda n, p: 1),
    "me_8x8_cost_variance": ("<u4", lambda n, p: 1), "rc_me_distortion": ("<u4", lambda n, p: 1),
    "stationary_block_present_sb": ("u1", lambda n, p: 1), "rc_me_allow_gm": ("u1", lambda n, p: 1),
    "p_sb_best_sad": ("<u4", lambda n, p: 2 * 4 * 85), "p_sb_best_mv": ("<u4", lambda n, p: 2 * 4 * 85),
    "hme_sc": ("<i2", lambda n, p: 2 * 4 * 2), "hme_sad": ("<u8", lambda n, p: 2 * 4), "do_ref": ("u1", lambda n, p: 2 * 4),
}


def alloc_out(n_b64, params, deep=True):
    """Allocate zeroed numpy arrays for every SvtMeOut member; returns (SvtMeOut, dict of arrays)."""
    import numpy as np
    n_pu = exported_pu_count(params)
    out, arrays = SvtMeOut(), {}
    for name, (dt, cnt) in OUT_SPECS.items():
        if not deep and name in ("p_sb_best_sad", "p_sb_best_mv", "hme_sc", "hme_sad", "do_ref"):
            continue
        a = np.zeros((n_b64, max(1, cnt(n_pu, params))), dtype=dt)
        arrays[name] = a
        setattr(out, name, a.ctypes.data)
    return out, arrays


_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libsvtme_b200.so")
_lib = None


def load_library():
    """Load the CUDA C-ABI library. There is no fallback: a missing build is a hard error."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). The B200 ME path has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, cp = C.c_void_p, C.c_char_p
    P = C.POINTER
    sigs = {
        "svtme_derive_params": (C.c_int, [P(SvtMeDeriveInput), P(SvtMeParams)]),
        "svtme_derive_params_tf": (C.c_int, [P(SvtMeTfDeriveInput), P(SvtMeParams)]),
        "svtme_open": (C.c_int, [C.c_int, u32, u32, u32, P(vp)]),
        "svtme_close": (None, [vp]),
        "svtme_last_error": (cp, []),
        "svtme_b64_count": (u32, [vp]),
        "svtme_exported_pu_count": (u32, [P(SvtMeParams)]),
        "svtme_put_picture": (C.c_int, [vp, u64, vp, u32]),
        "svtme_put_picture_device": (C.c_int, [vp, u64, vp, u32]),
        "svtme_put_pictures": (C.c_int, [vp, u32, P(u64), P(vp), u32]),
        "svtme_put_pictures_device": (C.c_int, [vp, u32, P(u64), P(vp), u32]),
        "svtme_put_pictures_10bit": (C.c_int, [vp, u32, P(u64), P(vp), u32]),
        "svtme_put_pictures_10bit_device": (C.c_int, [vp, u32, P(u64), P(vp), u32]),
        "svtme_get_plane": (C.c_int, [vp, u64, C.c_int, vp, u32, u32]),
        "svtme_release_picture": (C.c_int, [vp, u64]),
        "svtme_plane_device_ptr": (C.c_int, [vp, u64, C.c_int, P(vp), P(C.c_size_t)]),
        "svtme_adopt_picture": (C.c_int, [vp, u64]),
        "svtme_external_write_begin": (C.c_int, [vp, u32, P(u64), vp]),
        "svtme_external_write_end": (C.c_int, [vp, vp]),
        "svtme_share_picture": (C.c_int, [vp, vp, u64]),
        "svtme_run_picture": (C.c_int, [vp, P(SvtMeParams), u64, P((u64 * MAX_REFS) * MAX_LISTS), P(SvtMeOut)]),
        "svtme_run_pictures": (C.c_int, [vp, P(SvtMeJob), u32]),
        "svtme_launch_pictures": (C.c_int, [vp, P(SvtMeJob), u32]),
        "svtme_fetch_results": (C.c_int, [vp, P(SvtMeJob), u32]),
        "svtme_submit_pictures": (C.c_int, [vp, P(SvtMeJob), u32]),
        "svtme_collect_results": (C.c_int, [vp, P(SvtMeJob), u32]),
        "svtme_pending_batches": (u32, [vp]),
        "svtme_pin_host_memory": (C.c_int, [vp, C.c_size_t]),
        "svtme_unpin_host_memory": (C.c_int, [vp]),
        "svtme_debug_mark": (C.c_int, [vp]),
        "svtme_debug_last_span": (C.c_int, [vp, P(C.c_float)]),
        "svtme_sync": (C.c_int, [vp]),
        "svtme_last_kernel_ms": (C.c_float, [vp]),
        "svtme_launch_count": (u64, []),
        "svt_open_loop_me_fullpel_search_sblock_b200_batch": (C.c_int, [vp, u32, vp, u32, vp, u32, vp, C.c_uint8, vp, vp]),
        "svtme_noalloc_setup": (C.c_int, [C.c_int]),
        "svt_sad_loop_kernel_b200_noalloc": (C.c_int, [vp, u32, vp, u32, C.c_size_t, vp, u32, C.c_size_t, u32, u32, u32, C.c_uint8,
                                                       C.c_int16, C.c_int16, vp, vp, vp]),
        "svtme_set_stream": (C.c_int, [vp, vp]),
        "svtme_last_work_ops": (C.c_int, [vp, P(u64)]),
        "svtme_last_phase_ms": (C.c_int, [vp, P(C.c_float), C.c_uint32]),
        "svtme_last_phase_work": (C.c_int, [vp, P(u64), C.c_uint32]),
        "svtme_dg_detector_hme": (C.c_int, [vp, u64, u64, u32, P(SvtMeDgMetrics)]),
        "svtme_picture_statistics": (C.c_int, [vp, u64, P(SvtMeStatsParams), P(SvtMeStatsOut)]),
        "svtme_pictures_statistics": (C.c_int, [vp, u32, P(u64), P(SvtMeStatsParams), P(SvtMeStatsOut)]),
        "svtme_last_statistics_ms": (C.c_int, [vp, P(C.c_float), P(C.c_float)]),
        "svtme_sizeof_stats_params": (u32, []),
        "svtme_set_search_timing": (C.c_int, [vp, C.c_int]),
        "svtme_last_search_ms": (C.c_int, [vp, P(C.c_float), C.c_uint32]),
        "svtme_measure_vabsdiff4_ginst_per_s": (C.c_double, []),
        "svtme_sizeof_params": (u32, []),
        "svtme_sizeof_out": (u32, []),
        "svtme_sizeof_job": (u32, []),
        "svtme_sizeof_derive_input": (u32, []),
        "svtme_sizeof_tf_derive_input": (u32, []),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    check_layout(lib)
    _lib = lib
    return lib


def check_layout(lib):
    pairs = [("svtme_sizeof_params", SvtMeParams), ("svtme_sizeof_out", SvtMeOut), ("svtme_sizeof_job", SvtMeJob),
             ("svtme_sizeof_derive_input", SvtMeDeriveInput), ("svtme_sizeof_tf_derive_input", SvtMeTfDeriveInput)]
    for fn, st in pairs:
        got, want = getattr(lib, fn)(), C.sizeof(st)
        if got != want:
            raise RuntimeError(f"ABI drift: {fn}() = {got} but ctypes mirror of {st.__name__} is {want} bytes")

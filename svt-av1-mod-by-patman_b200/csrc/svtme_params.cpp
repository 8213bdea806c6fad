// svtme_params.cpp — host-side derivation of the ME parameters from encoder-level settings.
// Mirrors, value for value, the reference's
//   svt_aom_sig_deriv_me            Source/Lib/Codec/enc_mode_config.c:684-823
//   set_me_search_params            :218-353     set_hme_search_params     :141-213
//   svt_aom_set_me_hme_ref_prune_ctrls :354-437  svt_aom_set_mv_based_sa_ctrls :439-459
//   svt_aom_set_me_sr_adjustment_ctrls :461-534  svt_aom_set_me_8x8_var_ctrls  :535-555
//   svt_aom_set_prehme_ctrls        :557-617     svt_aom_get_enable_me_8x8     :76-96
//   HME level flags                 :1798-1809   svt_aom_derive_input_resolution sequence_control_set.c:113-131
//   svt_aom_get_max_allocated_me_refs pcs.c:90-95
// Checked field-by-field against the compiled reference by tests/test_params.py.
#include <string.h>
#include "svtme_b200.h"

namespace {
enum { R240 = 0, R360, R480, R720, R1080, R4K, R8K };

int input_resolution(uint32_t size) {
    if (size < 0x28500) return R240;
    if (size < 0x4CE00) return R360;
    if (size < 0xA1400) return R480;
    if (size < 0x16DA00) return R720;
    if (size < 0x535200) return R1080;
    if (size < 0x140A000) return R4K;
    return R8K;
}
inline SvtMeSearchArea SA(uint16_t w, uint16_t h) { SvtMeSearchArea a = {w, h}; return a; }
inline int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
template <class T> inline T tmax(T a, T b) { return a > b ? a : b; }

void me_search_params(const SvtMeDeriveInput &in, int res, SvtMeParams &o) {
    const int m = in.enc_mode;
    uint8_t q_mult = 0;
    SvtMeSearchArea mn, mx;
    if (in.rtc_tune) {
        if (in.sc_class1) {
            if (m <= 8) { mn = SA(32, 32); mx = SA(96, 96); }
            else if (res < R1080) { mn = SA(16, 16); mx = SA(32, 16); }
            else if (m <= 10) { mn = SA(24, 24); mx = SA(24, 24); }
            else { mn = SA(16, 6); mx = SA(16, 9); }
        } else if (m <= 9) {
            if (res < R1080) { mn = SA(16, 16); mx = SA(32, 16); }
            else { mn = SA(16, 6); mx = SA(16, 9); }
        } else {
            if (res < R720) { mn = SA(8, 3); mx = SA(16, 9); }
            else if (res < R1080) { mn = SA(8, 1); mx = SA(16, 7); }
            else if (res < R4K) { mn = SA(8, 1); mx = SA(8, 7); }
            else { mn = SA(8, 1); mx = SA(8, 1); }
        }
    } else if (in.sc_class1) {
        if (m <= 1) { mn = SA(175, 175); mx = SA(750, 750); }
        else if (m <= 7) { mn = SA(48, 48); mx = SA(224, 224); }
        else if (m <= 8) { mn = SA(32, 32); mx = SA(164, 164); }
        else if (m <= 9) { mn = SA(32, 32); mx = SA(96, 96); }
        else if (m <= 10) { mn = SA(16, 16); mx = SA(96, 96); }
        else { mn = SA(8, 8); mx = SA(32, 32); }
    } else if (m <= 1) { mn = SA(64, 64); mx = SA(256, 256); }
    else if (m <= 2) { mn = SA(32, 32); mx = SA(128, 128); }
    else if (m <= 3) { mn = SA(24, 24); mx = SA(104, 104); }
    else if (m <= 6) { mn = SA(16, 16); mx = SA(64, 32); }
    else if (m <= 8) {
        if (in.hierarchical_levels <= 3) {
            if (res < R4K) { mn = SA(8, 5); mx = SA(16, 9); }
            else { mn = SA(8, 1); mx = SA(8, 1); }
        } else if (res < R1080) { mn = SA(16, 16); mx = SA(32, 16); }
        else { mn = SA(16, 6); mx = SA(16, 9); }
        q_mult = 4;
    } else if (m <= 10) {
        if (res < R1080) { mn = SA(16, 16); mx = SA(32, 16); }
        else { mn = SA(16, 6); mx = SA(16, 9); }
        q_mult = 4;
    } else { mn = SA(8, 3); mx = SA(8, 3); q_mult = 2; }
    if (q_mult) {
        const uint16_t qw = (uint16_t)clip3(500, 1000, (int)(q_mult * ((8 * (int)in.qp) - 125)));
        mn.width = (uint16_t)tmax(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)tmax(3, (mn.height * qw) / 1000);
        mx.width = (uint16_t)tmax(8, (mx.width * qw) / 1000);
        mx.height = (uint16_t)tmax(3, (mx.height * qw) / 1000);
    }
    if (in.frame_rate_q16 >> 16) { // "low_frame_rate_flag" exactly as written in the reference (:346-350)
        mn.width = (uint16_t)((mn.width * 3) >> 1);
        mn.height = (uint16_t)((mn.height * 3) >> 1);
    }
    o.me_sa.sa_min = mn;
    o.me_sa.sa_max = mx;
}

void hme_search_params(const SvtMeDeriveInput &in, int res, SvtMeParams &o) {
    const int m = in.enc_mode;
    o.num_hme_sa_w = o.num_hme_sa_h = 2;
    uint8_t q_mult = 0;
    SvtMeSearchArea mn, mx;
    if (m <= 1) {
        if (res < R4K) { mn = SA(32, 32); mx = SA(192, 192); }
        else { mn = SA(240, 240); mx = SA(480, 480); }
    } else if (m <= 3) { mn = SA(32, 32); mx = SA(192, 192); }
    else if (m <= 7) { mn = SA(32, 32); mx = SA(192, 192); q_mult = 3; }
    else if (!in.rtc_tune && m <= 8) {
        if (in.sc_class1 || res >= R4K) { mn = SA(32, 32); mx = SA(192, 192); }
        else { mn = SA(16, 16); mx = SA(192, 192); }
        q_mult = 3;
    } else if ((!in.rtc_tune && m <= 10) || (in.rtc_tune && m <= 8)) {
        if (in.sc_class1) { mn = SA(32, 32); mx = SA(192, 192); }
        else { mn = SA(16, 16); mx = SA(192, 192); }
        q_mult = 3;
    } else {
        if (in.sc_class1) { mn = SA(32, 32); mx = SA(192, 192); }
        else if (res < R4K) { mn = SA(8, 8); mx = SA(96, 96); }
        else { mn = SA(16, 16); mx = SA(96, 96); }
        q_mult = 3;
    }
    if (q_mult) {
        const uint16_t qw = (uint16_t)clip3(500, 1000, (int)(q_mult * ((8 * (int)in.qp) - 125)));
        mn.width = (uint16_t)tmax(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)tmax(8, (mn.height * qw) / 1000);
        mx.width = (uint16_t)tmax(96, (mx.width * qw) / 1000);
        mx.height = (uint16_t)tmax(96, (mx.height * qw) / 1000);
    }
    o.hme_l0_sa.sa_min = mn;
    o.hme_l0_sa.sa_max = mx;
    if (m <= 0) { o.hme_l1_sa = SA(16, 16); o.hme_l2_sa = SA(16, 16); }
    else { o.hme_l1_sa = SA(8, 3); o.hme_l2_sa = SA(8, 3); }
}

void prune_ctrls(int level, SvtMeParams &o) {
    static const struct { uint8_t en; uint16_t hme, me; uint32_t zz; uint16_t zzp; uint32_t ph; uint16_t php; } t[7] = {
        {0, 0xffff, 0xffff, 0, 0, 0, 0}, {1, 80, 0xffff, 0, 0, 0, 0}, {1, 50, 0xffff, 0, 0, 0, 0},
        {1, 30, 0xffff, 0, 0, 0, 0},     {1, 15, 60, 0, 0, 0, 0},     {1, 5, 60, 0, 0, 0, 0},
        {1, 5, 60, 20 * 64 * 64, 5, 10 * 64 * 64, 5}};
    o.enable_me_hme_ref_pruning = t[level].en;
    o.prune_ref_if_hme_sad_dev_bigger_than_th = t[level].hme;
    o.prune_ref_if_me_sad_dev_bigger_than_th = t[level].me;
    o.zz_sad_th = t[level].zz; o.zz_sad_pct = t[level].zzp;
    o.phme_sad_th = t[level].ph; o.phme_sad_pct = t[level].php;
}

void sr_adjustment_ctrls(int level, SvtMeParams &o) {
    static const struct { uint8_t en; uint16_t mvl, st, sdiv, lo, ldiv; uint8_t dist; } t[6] = {
        {0, 0, 0, 0, 0, 0, 0},           {1, 4, 12000, 8, 6000, 8, 0},   {1, 4, 12000, 8, 6000, 8, 1},
        {1, 4, 12000, 8, 12000, 8, 1},   {2, 16, 20000, 8, 20000, 8, 1}, {2, 20, 24000, 8, 24000, 8, 1}};
    o.enable_me_sr_adjustment = t[level].en;
    if (level) {
        o.reduce_me_sr_based_on_mv_length_th = t[level].mvl;
        o.stationary_hme_sad_abs_th = t[level].st;
        o.stationary_me_sr_divisor = t[level].sdiv;
        o.reduce_me_sr_based_on_hme_sad_abs_th = t[level].lo;
        o.me_sr_divisor_for_low_hme_sad = t[level].ldiv;
        o.distance_based_hme_resizing = t[level].dist;
    }
    if (o.enable_hme_level2_flag == 0) {
        const int d = o.enable_hme_level1_flag == 1 ? 4 : 16;
        o.stationary_hme_sad_abs_th = (uint16_t)(o.stationary_hme_sad_abs_th / d);
        o.reduce_me_sr_based_on_hme_sad_abs_th = (uint16_t)(o.reduce_me_sr_based_on_hme_sad_abs_th / d);
    }
}

void prehme_ctrls(int level, SvtMeParams &o) {
    o.prehme_enable = level != 0;
    if (!level) return;
    static const struct { uint16_t v[8]; uint8_t skip, l1; } t[5] = {
        {{0}, 0, 0},
        {{8, 144, 8, 496, 144, 3, 496, 3}, 0, 0},
        {{8, 100, 8, 400, 96, 3, 384, 3}, 0, 0},
        {{8, 100, 8, 350, 32, 7, 200, 7}, 1, 0},
        {{8, 100, 8, 350, 32, 7, 128, 7}, 1, 1}};
    const uint16_t *v = t[level].v;
    o.prehme_sa_cfg[0].sa_min = SA(v[0], v[1]); o.prehme_sa_cfg[0].sa_max = SA(v[2], v[3]);
    o.prehme_sa_cfg[1].sa_min = SA(v[4], v[5]); o.prehme_sa_cfg[1].sa_max = SA(v[6], v[7]);
    o.prehme_skip_search_line = t[level].skip;
    o.prehme_l1_early_exit = t[level].l1;
}
} // namespace

extern "C" int svtme_derive_params(const SvtMeDeriveInput *inp, SvtMeParams *out) {
    if (!inp || !out || inp->enc_mode < -1 || inp->enc_mode > 13) return SVTME_ERR_INVALID;
    const SvtMeDeriveInput &in = *inp;
    SvtMeParams &o = *out;
    memset(&o, 0, sizeof(o));
    const int m = in.enc_mode;
    const int res = input_resolution(in.width * in.height);
    const bool is_base = in.temporal_layer_index == 0;
    me_search_params(in, res, o);
    hme_search_params(in, res, o);
    o.enable_hme_flag = 1;
    o.enable_hme_level0_flag = 1;
    o.enable_hme_level1_flag = 1;
    o.enable_hme_level2_flag = (in.sc_class1 || m <= 6) ? 1 : 0;
    o.hme_search_method = 0; // SUB_SAD_SEARCH
    o.me_search_method = 0;
    if (in.rtc_tune) {
        if (in.sc_class1) { o.reduce_hme_l0_sr_th_min = 8; o.reduce_hme_l0_sr_th_max = 100; }
        else if (m <= 8) { o.reduce_hme_l0_sr_th_min = 0; o.reduce_hme_l0_sr_th_max = 0; }
        else { o.reduce_hme_l0_sr_th_min = 8; o.reduce_hme_l0_sr_th_max = 200; }
    }
    int prehme_level;
    if (in.sc_class1) prehme_level = in.rtc_tune ? 1 : 2;
    else if (in.rtc_tune) prehme_level = m <= 8 ? 2 : (m <= 10 ? 4 : 0);
    else prehme_level = m <= 8 ? 2 : 4;
    if (o.enable_hme_level1_flag == 0) prehme_level = 0;
    prehme_ctrls(prehme_level, o);
    int prune_level;
    if (in.sc_class1) prune_level = m <= 2 ? 1 : (m <= 8 ? 3 : 6);
    else if (m <= 0) prune_level = is_base ? 1 : 2;
    else if (m <= 1) prune_level = is_base ? 1 : 4;
    else if (m <= 5) prune_level = is_base ? 1 : 5;
    else if (m <= 10) prune_level = is_base ? 1 : 6;
    else prune_level = 6;
    prune_ctrls(prune_level, o);
    int sr_level;
    if (in.sc_class1) sr_level = m <= 8 ? 4 : 5;
    else if (m <= -1) sr_level = 0;
    else if (m <= 0) sr_level = 1;
    else sr_level = 3;
    sr_adjustment_ctrls(sr_level, o);
    if (m <= 3) {
        o.mv_sa_adj_enabled = 1; o.mv_sa_adj_nearest_ref_only = 1; o.mv_sa_adj_mv_size_th = 25; o.mv_sa_adj_sa_multiplier = 2;
    }
    o.me_8x8_var_enabled = 1; o.me_sr_div4_th = 80000; o.me_sr_div2_th = 150000; o.me_sr_mult2_th = 0xFFFFFFFFu;
    o.prune_me_candidates_th = m <= 5 ? 0 : 65;
    o.use_best_unipred_cand_only = in.use_best_me_unipred_cand_only;
    if (in.rtc_tune) o.me_early_exit_th = in.sc_class1 ? 64 * 64 : (m <= 10 ? 64 * 64 * 8 : 64 * 64 * 9);
    else o.me_early_exit_th = m <= 4 ? 0 : 64 * 64 * 8;
    o.me_safe_limit_zz_th = in.safe_limit_nref == 1 ? in.safe_limit_zz_th : 0;
    o.prev_me_stage_based_exit_th = (in.rtc_tune && in.sc_class1) ? 64 * 64 * 4 : 0;
    // svt_aom_get_enable_me_8x8 / _16x16
    if (in.rtc_tune) o.enable_me_8x8 = m <= 9;
    else o.enable_me_8x8 = m <= 5 ? 1 : (m <= 10 ? (res <= R720) : 0);
    o.enable_me_16x16 = 1;
    o.input_resolution_le_480p = res <= R480;
    o.max_refs = (uint8_t)(in.ref_count_used_list0 + in.ref_count_used_list1);
    o.max_cand = (uint8_t)(in.ref_count_used_list0 + in.ref_count_used_list1 +
                           in.ref_count_used_list0 * in.ref_count_used_list1 + (in.ref_count_used_list0 - 1) +
                           (in.ref_count_used_list1 == 3 ? 1 : 0));
    o.max_l0 = in.ref_count_used_list0;
    // per-picture fields (me_process.c:219-227)
    o.num_of_list_to_search = in.slice_type_b ? 2 : 1;
    o.num_of_ref_pic_to_search[0] = in.ref_list0_count_try;
    o.num_of_ref_pic_to_search[1] = in.slice_type_b ? in.ref_list1_count_try : 0;
    o.temporal_layer_index = in.temporal_layer_index;
    o.is_ref = in.is_ref;
    o.hierarchical_levels = in.hierarchical_levels;
    o.similar_brightness_refs = in.similar_brightness_refs;
    o.max_number_of_pus_per_sb = SVTME_PU_COUNT;
    o.only_l_bwd = in.only_l_bwd;
    o.gm_enabled = in.gm_level_enabled;
    o.gm_use_distance_based_active_th = in.gm_use_distance_based_active_th;
    return SVTME_OK;
}

extern "C" uint32_t svtme_exported_pu_count(const SvtMeParams *p) {
    if (!p->enable_me_16x16) return 5;
    return p->enable_me_8x8 ? p->max_number_of_pus_per_sb : 21;
}


// ---- temporal-filter ME (ME_MCTF): svt_aom_sig_deriv_me_tf (enc_mode_config.c:829-875), tf_set_me_hme_params_oq
// (:601-681), TF HME flags (:1810-1834), set_hme_search_params_mctf(ctx, 0) and the FULL_SAD HME search method of
// create_me_context_and_picture_control (temporal_filtering.c:2730-2749,315), per-block fields (:3046-3065).
extern "C" int svtme_derive_params_tf(const SvtMeTfDeriveInput *in, SvtMeParams *o) {
    if (!in || !o || in->hme_me_level > 4 || in->enc_mode < -1 || in->enc_mode > 13) return SVTME_ERR_INVALID;
    memset(o, 0, sizeof(*o));
    const int res = input_resolution(in->width * in->height);
    SvtMeSearchArea l0mn, l0mx, l1, l2, mn, mx;
    switch (in->hme_me_level) {
    case 0: l0mn = SA(30, 30); l0mx = SA(60, 60); l1 = SA(16, 16); l2 = SA(16, 16); mn = SA(60, 60); mx = SA(120, 120); break;
    case 1: l0mn = SA(16, 16); l0mx = SA(32, 32); l1 = SA(16, 16); l2 = SA(16, 16); mn = SA(16, 16); mx = SA(32, 32); break;
    case 2:
        if (res <= R360) { l0mn = SA(8, 8); l0mx = SA(8, 8); l1 = SA(8, 8); }
        else if (res <= R480) { l0mn = SA(8, 8); l0mx = SA(16, 16); l1 = SA(8, 8); }
        else { l0mn = SA(16, 16); l0mx = SA(32, 32); l1 = SA(16, 16); }
        l2 = SA(16, 16); mn = SA(8, 8); mx = SA(8, 8);
        break;
    case 3: l0mn = SA(8, 8); l0mx = SA(8, 8); l1 = SA(8, 8); l2 = SA(8, 8); mn = SA(8, 8); mx = SA(8, 8); break;
    default: l0mn = SA(4, 4); l0mx = SA(4, 4); l1 = SA(8, 8); l2 = SA(8, 8); mn = SA(8, 8); mx = SA(8, 8); break;
    }
    if (in->qp_opt) {
        const uint16_t qw = (uint16_t)clip3(250, 1000, (int)((8 * (int)in->qp) - 125));
        mn.width = (uint16_t)tmax(8, (mn.width * qw) / 1000); mn.height = (uint16_t)tmax(8, (mn.height * qw) / 1000);
        mx.width = (uint16_t)tmax(8, (mx.width * qw) / 1000); mx.height = (uint16_t)tmax(8, (mx.height * qw) / 1000);
    }
    o->num_hme_sa_w = o->num_hme_sa_h = 2;
    o->hme_l0_sa.sa_min = l0mn; o->hme_l0_sa.sa_max = l0mx; // set_hme_search_params_mctf(ctx, 0): the TF defaults
    o->hme_l1_sa = l1; o->hme_l2_sa = l2;
    o->me_sa.sa_min = mn; o->me_sa.sa_max = mx;
    o->enable_hme_flag = o->enable_hme_level0_flag = 1;
    o->enable_hme_level1_flag = in->hme_me_level <= 2;
    o->enable_hme_level2_flag = in->hme_me_level == 0;
    o->hme_search_method = 1;                          // FULL_SAD_SEARCH, temporal_filtering.c:315
    o->me_search_method = in->hme_me_level <= 2 ? 1 : 0;
    o->prune_ref_if_hme_sad_dev_bigger_than_th = o->prune_ref_if_me_sad_dev_bigger_than_th = (uint16_t)~0; // prune level 0
    o->me_early_exit_th = (in->enc_mode <= 6 || res <= R720) ? 0u : 64u * 64u * 4u;
    o->prev_me_stage_based_exit_th = (in->enc_mode <= 7 || res <= R720) ? 0u : 64u * 64u * 4u;
    o->num_of_list_to_search = 1;
    o->num_of_ref_pic_to_search[0] = 1;
    o->temporal_layer_index = in->temporal_layer_index;
    o->is_ref = in->is_ref;
    o->me_type = 1;
    o->tf_me_exit_th = in->me_exit_th;
    o->max_number_of_pus_per_sb = 85;
    o->enable_me_8x8 = o->enable_me_16x16 = 1; // (the TF reads the tables directly; nothing is exported per PU)
    o->max_refs = o->max_cand = o->max_l0 = 1;
    o->input_resolution_le_480p = res <= R480;
    return SVTME_OK;
}
extern "C" uint32_t svtme_sizeof_tf_derive_input(void) { return (uint32_t)sizeof(SvtMeTfDeriveInput); }

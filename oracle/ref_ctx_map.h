/*
 * ref_ctx_map.h — SvtMeParams <-> the reference's MeContext (Source/Lib/Codec/me_context.h:366-509), field by field.
 * TEST INFRASTRUCTURE ONLY (see me_oracle.h). Included by oracle/ref_harness.c (drives the compiled reference from
 * SvtMeParams) and by oracle/encoder_dump/svtme_dump.c (records what the real encoder handed to ME as SvtMeParams).
 * Needs the reference's me_context.h and include/svtme_b200.h to be included first.
 */
#ifndef REF_CTX_MAP_H
#define REF_CTX_MAP_H
static inline void params_to_ctx(const SvtMeParams *p, MeContext *c) {
#define SA(dst, src) do { (dst).width = (src).width; (dst).height = (src).height; } while (0)
    SA(c->me_sa.sa_min, p->me_sa.sa_min); SA(c->me_sa.sa_max, p->me_sa.sa_max);
    SA(c->hme_l0_sa.sa_min, p->hme_l0_sa.sa_min); SA(c->hme_l0_sa.sa_max, p->hme_l0_sa.sa_max);
    SA(c->hme_l1_sa, p->hme_l1_sa); SA(c->hme_l2_sa, p->hme_l2_sa);
    c->num_hme_sa_w = p->num_hme_sa_w; c->num_hme_sa_h = p->num_hme_sa_h;
    c->enable_hme_flag = p->enable_hme_flag; c->enable_hme_level0_flag = p->enable_hme_level0_flag;
    c->enable_hme_level1_flag = p->enable_hme_level1_flag; c->enable_hme_level2_flag = p->enable_hme_level2_flag;
    c->hme_search_method = p->hme_search_method; c->me_search_method = p->me_search_method;
    c->reduce_hme_l0_sr_th_min = p->reduce_hme_l0_sr_th_min; c->reduce_hme_l0_sr_th_max = p->reduce_hme_l0_sr_th_max;
    c->prehme_ctrl.enable = p->prehme_enable; c->prehme_ctrl.skip_search_line = p->prehme_skip_search_line;
    c->prehme_ctrl.l1_early_exit = p->prehme_l1_early_exit;
    for (int i = 0; i < 2; ++i) {
        SA(c->prehme_ctrl.prehme_sa_cfg[i].sa_min, p->prehme_sa_cfg[i].sa_min);
        SA(c->prehme_ctrl.prehme_sa_cfg[i].sa_max, p->prehme_sa_cfg[i].sa_max);
    }
    c->me_hme_prune_ctrls.enable_me_hme_ref_pruning = p->enable_me_hme_ref_pruning;
    c->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th = p->prune_ref_if_hme_sad_dev_bigger_than_th;
    c->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th = p->prune_ref_if_me_sad_dev_bigger_than_th;
    c->me_hme_prune_ctrls.zz_sad_th = p->zz_sad_th; c->me_hme_prune_ctrls.zz_sad_pct = p->zz_sad_pct;
    c->me_hme_prune_ctrls.phme_sad_th = p->phme_sad_th; c->me_hme_prune_ctrls.phme_sad_pct = p->phme_sad_pct;
    c->me_sr_adjustment_ctrls.enable_me_sr_adjustment = p->enable_me_sr_adjustment;
    c->me_sr_adjustment_ctrls.distance_based_hme_resizing = p->distance_based_hme_resizing;
    c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_mv_length_th = p->reduce_me_sr_based_on_mv_length_th;
    c->me_sr_adjustment_ctrls.stationary_hme_sad_abs_th = p->stationary_hme_sad_abs_th;
    c->me_sr_adjustment_ctrls.stationary_me_sr_divisor = p->stationary_me_sr_divisor;
    c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_hme_sad_abs_th = p->reduce_me_sr_based_on_hme_sad_abs_th;
    c->me_sr_adjustment_ctrls.me_sr_divisor_for_low_hme_sad = p->me_sr_divisor_for_low_hme_sad;
    c->me_8x8_var_ctrls.enabled = p->me_8x8_var_enabled; c->me_8x8_var_ctrls.me_sr_div4_th = p->me_sr_div4_th;
    c->me_8x8_var_ctrls.me_sr_div2_th = p->me_sr_div2_th; c->me_8x8_var_ctrls.me_sr_mult2_th = p->me_sr_mult2_th;
    c->mv_based_sa_adj.enabled = p->mv_sa_adj_enabled; c->mv_based_sa_adj.nearest_ref_only = p->mv_sa_adj_nearest_ref_only;
    c->mv_based_sa_adj.mv_size_th = p->mv_sa_adj_mv_size_th; c->mv_based_sa_adj.sa_multiplier = p->mv_sa_adj_sa_multiplier;
    c->prune_me_candidates_th = p->prune_me_candidates_th;
    c->use_best_unipred_cand_only = p->use_best_unipred_cand_only;
    c->me_early_exit_th = p->me_early_exit_th; c->me_safe_limit_zz_th = p->me_safe_limit_zz_th;
    c->prev_me_stage_based_exit_th = p->prev_me_stage_based_exit_th;
    c->num_of_list_to_search = p->num_of_list_to_search;
    c->num_of_ref_pic_to_search[0] = p->num_of_ref_pic_to_search[0];
    c->num_of_ref_pic_to_search[1] = p->num_of_ref_pic_to_search[1];
    c->temporal_layer_index = p->temporal_layer_index; c->is_ref = p->is_ref;
    c->me_type = p->me_type == 1 ? ME_MCTF : ME_OPEN_LOOP;
    c->tf_me_exit_th = p->tf_me_exit_th;
#undef SA
}

static inline void ctx_to_params(const MeContext *c, SvtMeParams *p) {
#define SA(dst, src) do { (dst).width = (src).width; (dst).height = (src).height; } while (0)
    SA(p->me_sa.sa_min, c->me_sa.sa_min); SA(p->me_sa.sa_max, c->me_sa.sa_max);
    SA(p->hme_l0_sa.sa_min, c->hme_l0_sa.sa_min); SA(p->hme_l0_sa.sa_max, c->hme_l0_sa.sa_max);
    SA(p->hme_l1_sa, c->hme_l1_sa); SA(p->hme_l2_sa, c->hme_l2_sa);
    p->num_hme_sa_w = c->num_hme_sa_w; p->num_hme_sa_h = c->num_hme_sa_h;
    p->enable_hme_flag = c->enable_hme_flag; p->enable_hme_level0_flag = c->enable_hme_level0_flag;
    p->enable_hme_level1_flag = c->enable_hme_level1_flag; p->enable_hme_level2_flag = c->enable_hme_level2_flag;
    p->hme_search_method = c->hme_search_method; p->me_search_method = c->me_search_method;
    p->reduce_hme_l0_sr_th_min = c->reduce_hme_l0_sr_th_min; p->reduce_hme_l0_sr_th_max = c->reduce_hme_l0_sr_th_max;
    p->prehme_enable = c->prehme_ctrl.enable;
    if (p->prehme_enable) {
        p->prehme_skip_search_line = c->prehme_ctrl.skip_search_line;
        p->prehme_l1_early_exit = c->prehme_ctrl.l1_early_exit;
        for (int i = 0; i < 2; ++i) {
            SA(p->prehme_sa_cfg[i].sa_min, c->prehme_ctrl.prehme_sa_cfg[i].sa_min);
            SA(p->prehme_sa_cfg[i].sa_max, c->prehme_ctrl.prehme_sa_cfg[i].sa_max);
        }
    }
    p->enable_me_hme_ref_pruning = c->me_hme_prune_ctrls.enable_me_hme_ref_pruning;
    p->prune_ref_if_hme_sad_dev_bigger_than_th = c->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th;
    p->prune_ref_if_me_sad_dev_bigger_than_th = c->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th;
    p->zz_sad_th = c->me_hme_prune_ctrls.zz_sad_th; p->zz_sad_pct = c->me_hme_prune_ctrls.zz_sad_pct;
    p->phme_sad_th = c->me_hme_prune_ctrls.phme_sad_th; p->phme_sad_pct = c->me_hme_prune_ctrls.phme_sad_pct;
    p->enable_me_sr_adjustment = c->me_sr_adjustment_ctrls.enable_me_sr_adjustment;
    if (p->enable_me_sr_adjustment) {
        p->distance_based_hme_resizing = c->me_sr_adjustment_ctrls.distance_based_hme_resizing;
        p->reduce_me_sr_based_on_mv_length_th = c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_mv_length_th;
        p->stationary_hme_sad_abs_th = c->me_sr_adjustment_ctrls.stationary_hme_sad_abs_th;
        p->stationary_me_sr_divisor = c->me_sr_adjustment_ctrls.stationary_me_sr_divisor;
        p->reduce_me_sr_based_on_hme_sad_abs_th = c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_hme_sad_abs_th;
        p->me_sr_divisor_for_low_hme_sad = c->me_sr_adjustment_ctrls.me_sr_divisor_for_low_hme_sad;
    }
    p->me_8x8_var_enabled = c->me_8x8_var_ctrls.enabled; p->me_sr_div4_th = c->me_8x8_var_ctrls.me_sr_div4_th;
    p->me_sr_div2_th = c->me_8x8_var_ctrls.me_sr_div2_th; p->me_sr_mult2_th = c->me_8x8_var_ctrls.me_sr_mult2_th;
    p->mv_sa_adj_enabled = c->mv_based_sa_adj.enabled;
    if (p->mv_sa_adj_enabled) {
        p->mv_sa_adj_nearest_ref_only = c->mv_based_sa_adj.nearest_ref_only;
        p->mv_sa_adj_mv_size_th = c->mv_based_sa_adj.mv_size_th; p->mv_sa_adj_sa_multiplier = c->mv_based_sa_adj.sa_multiplier;
    }
    p->prune_me_candidates_th = c->prune_me_candidates_th;
    p->use_best_unipred_cand_only = c->use_best_unipred_cand_only;
    p->me_early_exit_th = c->me_early_exit_th; p->me_safe_limit_zz_th = c->me_safe_limit_zz_th;
    p->prev_me_stage_based_exit_th = c->prev_me_stage_based_exit_th;
    p->me_type = c->me_type == ME_MCTF ? 1 : 0;
    p->tf_me_exit_th = c->me_type == ME_MCTF ? c->tf_me_exit_th : 0;
#undef SA
}

#endif

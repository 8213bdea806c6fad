/*
 * svtme_hooks_example.c — the encoder-side binding of libsvtme_b200.so, written against the reference's own headers
 * and syntax-checked against them (`make -C oracle integration-check`, gcc -fsyntax-only -I/root/reference/Source/...).
 * INTEGRATION.md quotes these functions; a maintainer adds the CALLS to them at the cited places. TEST INFRASTRUCTURE /
 * DOCUMENTATION ONLY: nothing here is compiled into the product, and no reference source is copied.
 *
 * State the patch adds to the reference's structs (shown here as parameters so that the file compiles against the
 * unmodified headers):
 *   EbEncHandle             : SvtMeHandle *svtme;                     (one per GPU)
 *   PictureParentControlSet : SvtmePcsExt  svtme;                     (flat result arrays, page-locked once per pcs)
 */
#include <string.h>
#include "pcs.h"
#include "sequence_control_set.h"
#include "reference_object.h"
#include "me_context.h"
#include "svt_log.h"
#include "svtme_b200.h"

typedef struct SvtmePcsExt {
    uint8_t  *total; /* [b64_total_count][n_pu]             MeSbResults.total_me_candidate_index, back to back */
    uint32_t *mv;    /* [b64_total_count][n_pu * max_refs]  MeSbResults.me_mv_array */
    uint8_t  *cand;  /* [b64_total_count][n_pu * max_cand]  MeSbResults.me_candidate_array */
} SvtmePcsExt;

/* H0 — Globals/enc_handle.c, right after svt_aom_setup_rtcd_internal() (:1476); svtme_close() at de-init. */
EbErrorType svtme_hook_open(SvtMeHandle **svtme, const SequenceControlSet *scs) {
    if (svtme_open(/*device*/ 0, scs->max_input_luma_width, scs->max_input_luma_height,
                   /*capacity: the PA reference window*/ scs->pa_reference_picture_buffer_init_count, svtme) != SVTME_OK) {
        SVT_ERROR("svtme_open: %s\n", svtme_last_error());
        return EB_ErrorInsufficientResources; /* no CPU fallback by design */
    }
    return EB_ErrorNone;
}

/* H1 — Codec/pic_analysis_process.c, svt_aom_picture_analysis_kernel: in place of svt_aom_pad_input_pictures (:2099)
 * and svt_aom_downsample_filtering_input_picture (:2126). The host copies of the 1/4 and 1/16 planes are read back only
 * while CPU stages still want them (svt_aom_gathering_picture_statistics :1560, the DG detector). */
EbErrorType svtme_hook_pyramid(SvtMeHandle *svtme, PictureParentControlSet *pcs, EbPaReferenceObject *pa) {
    EbPictureBufferDesc *in = pcs->enhanced_pic; /* 8-bit luma (the 8 MSBs of a 10-bit source) */
    if (svtme_put_picture(svtme, pcs->picture_number, in->buffer_y + in->org_x + in->org_y * in->stride_y, in->stride_y) != SVTME_OK)
        return EB_ErrorInsufficientResources;
    EbPictureBufferDesc *q = pa->quarter_downsampled_picture_ptr, *s = pa->sixteenth_downsampled_picture_ptr;
    /* dst = first byte of the padded plane; the planes' own pads (32 / 16) are copied with them */
    if (svtme_get_plane(svtme, pcs->picture_number, 1, q->buffer_y, q->stride_y, q->org_x) != SVTME_OK ||
        svtme_get_plane(svtme, pcs->picture_number, 2, s->buffer_y, s->stride_y, s->org_x) != SVTME_OK)
        return EB_ErrorInsufficientResources;
    return EB_ErrorNone;
}

/* H2 — Codec/me_process.c, PAME branch of svt_aom_motion_estimation_kernel, in place of the b64 loop (:174-289): the
 * first segment of a picture runs the whole picture, the other segments have nothing left to do for ME. */
EbErrorType svtme_hook_me_picture(SvtMeHandle *svtme, SequenceControlSet *scs, PictureParentControlSet *pcs, SvtmePcsExt *x) {
    SvtMeDeriveInput di;
    memset(&di, 0, sizeof(di));
    di.enc_mode = pcs->enc_mode;            di.width = scs->max_input_luma_width;  di.height = scs->max_input_luma_height;
    di.qp = scs->static_config.qp;          di.frame_rate_q16 = scs->frame_rate;
    di.rtc_tune = scs->static_config.pred_structure == SVT_AV1_PRED_LOW_DELAY_B;
    di.sc_class1 = pcs->sc_class1;          di.hierarchical_levels = pcs->hierarchical_levels;
    di.temporal_layer_index = pcs->temporal_layer_index;   di.is_ref = pcs->is_ref;
    di.slice_type_b = pcs->slice_type == B_SLICE;
    di.ref_list0_count_try = pcs->ref_list0_count_try;
    di.ref_list1_count_try = pcs->slice_type == B_SLICE ? pcs->ref_list1_count_try : 0;
    di.ref_count_used_list0 = pcs->pa_me_data->max_l0;     /* pd_process.c:3495-3511 */
    di.ref_count_used_list1 = (uint8_t)(pcs->pa_me_data->max_refs - pcs->pa_me_data->max_l0);
    di.safe_limit_nref = scs->mrp_ctrls.safe_limit_nref;   di.safe_limit_zz_th = scs->mrp_ctrls.safe_limit_zz_th;
    di.only_l_bwd = scs->mrp_ctrls.only_l_bwd;
    di.use_best_me_unipred_cand_only = pcs->use_best_me_unipred_cand_only;
    di.similar_brightness_refs = pcs->similar_brightness_refs;
    di.gm_level_enabled = pcs->gm_ctrls.enabled;
    di.gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    SvtMeParams prm;
    if (svtme_derive_params(&di, &prm) != SVTME_OK) /* == svt_aom_sig_deriv_me(scs, pcs, me_ctx), enc_mode_config.c:684 */
        return EB_ErrorBadParameter;
    /* per-picture fields, me_process.c:219-227 */
    prm.num_of_list_to_search = pcs->slice_type == P_SLICE ? 1 : 2;
    prm.num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
    prm.num_of_ref_pic_to_search[1] = pcs->slice_type == B_SLICE ? pcs->ref_list1_count_try : 0;
    prm.temporal_layer_index = pcs->temporal_layer_index;  prm.is_ref = pcs->is_ref;
    prm.hierarchical_levels = pcs->hierarchical_levels;    prm.similar_brightness_refs = pcs->similar_brightness_refs;
    prm.max_number_of_pus_per_sb = pcs->max_number_of_pus_per_sb;
    prm.only_l_bwd = scs->mrp_ctrls.only_l_bwd;
    prm.gm_enabled = pcs->gm_ctrls.enabled;
    prm.gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    prm.picture_number = pcs->picture_number;
    uint64_t refs[SVTME_MAX_LISTS][SVTME_MAX_REFS];
    memset(refs, 0, sizeof(refs));
    for (int l = 0; l < prm.num_of_list_to_search; l++)
        for (int r = 0; r < prm.num_of_ref_pic_to_search[l]; r++) { /* me_process.c:248-262 */
            EbPaReferenceObject *o = (EbPaReferenceObject *)pcs->ref_pa_pic_ptr_array[l][r]->object_ptr;
            refs[l][r] = prm.ref_picture_number[l][r] = o->picture_number;
        }
    SvtMeOut out;
    memset(&out, 0, sizeof(out));
    out.total_me_candidate_index = x->total;  out.me_mv_array = x->mv;  out.me_candidate_array = x->cand;
    out.me_64x64_distortion = pcs->me_64x64_distortion;   out.me_32x32_distortion = pcs->me_32x32_distortion;
    out.me_16x16_distortion = pcs->me_16x16_distortion;   out.me_8x8_distortion = pcs->me_8x8_distortion;
    out.me_8x8_cost_variance = pcs->me_8x8_cost_variance; out.rc_me_distortion = pcs->rc_me_distortion;
    out.stationary_block_present_sb = pcs->stationary_block_present_sb;
    out.rc_me_allow_gm = pcs->rc_me_allow_gm;
    if (svtme_run_picture(svtme, &prm, pcs->picture_number, refs, &out) != SVTME_OK) {
        SVT_ERROR("svtme_run_picture: %s\n", svtme_last_error());
        return EB_ErrorInsufficientResources;
    }
    /* into MeSbResults (me_sb_results.h:29-50; allocated in pcs.c:97-119) — or point MeSbResults at the flat arrays once */
    const uint32_t n_pu = svtme_exported_pu_count(&prm);
    for (uint32_t b = 0; b < pcs->b64_total_count; b++) {
        MeSbResults *r = pcs->pa_me_data->me_results[b];
        memcpy(r->total_me_candidate_index, x->total + (size_t)b * n_pu, n_pu);
        memcpy(r->me_mv_array, x->mv + (size_t)b * n_pu * prm.max_refs, 4u * n_pu * prm.max_refs);
        memcpy(r->me_candidate_array, x->cand + (size_t)b * n_pu * prm.max_cand, (size_t)n_pu * prm.max_cand);
    }
    pcs->me_processed_b64_count = pcs->b64_total_count; /* then the GM branch of :262-272 as before */
    return EB_ErrorNone;
}

/* H3 — where the PA reference object is released (svt_release_object(pcs->pa_ref_pic_wrapper), Codec/pd_process.c /
 * Codec/initial_rc_process.c). H4 — Codec/temporal_filtering.c:3737 pad_and_decimate_filtered_pic: the filtered picture
 * replaces the source, so svtme_hook_pyramid runs again with the same picture number (overwrites the ring slot). */
void svtme_hook_release(SvtMeHandle *svtme, const PictureParentControlSet *pcs) { (void)svtme_release_picture(svtme, pcs->picture_number); }

/* H5 — temporal-filter ME (Codec/temporal_filtering.c:3046-3080, one call per centre picture and neighbouring frame
 * instead of one svt_aom_motion_estimation_b64 call per block): the tables the filter reads come back as the deep arrays. */
EbErrorType svtme_hook_tf_me(SvtMeHandle *svtme, SequenceControlSet *scs, PictureParentControlSet *centre, uint64_t frame_picture_number,
                             uint32_t *p_sb_best_sad, uint32_t *p_sb_best_mv, int16_t *hme_sc, uint64_t *hme_sad) {
    SvtMeTfDeriveInput ti;
    memset(&ti, 0, sizeof(ti));
    ti.enc_mode = centre->enc_mode;  ti.width = scs->max_input_luma_width;  ti.height = scs->max_input_luma_height;
    ti.qp = scs->static_config.qp;   ti.hme_me_level = centre->tf_ctrls.hme_me_level;  ti.qp_opt = centre->tf_ctrls.qp_opt;
    ti.me_exit_th = centre->tf_ctrls.me_exit_th;
    ti.temporal_layer_index = centre->temporal_layer_index;  ti.is_ref = centre->is_ref;
    SvtMeParams prm;
    if (svtme_derive_params_tf(&ti, &prm) != SVTME_OK) /* == svt_aom_sig_deriv_me_tf + temporal_filtering.c:296,315,3046-3074 */
        return EB_ErrorBadParameter;
    prm.picture_number = centre->picture_number;
    prm.ref_picture_number[0][0] = frame_picture_number;
    uint64_t refs[SVTME_MAX_LISTS][SVTME_MAX_REFS];
    memset(refs, 0, sizeof(refs));
    refs[0][0] = frame_picture_number;
    SvtMeOut out;
    memset(&out, 0, sizeof(out));
    out.p_sb_best_sad = p_sb_best_sad;  out.p_sb_best_mv = p_sb_best_mv;  out.hme_sc = hme_sc;  out.hme_sad = hme_sad;
    /* per block: hme_sad[b][0][0] < prm.tf_me_exit_th  <=>  the reference returned early and set tf_use_pred_64x64_only_th = ~0 */
    return svtme_run_picture(svtme, &prm, centre->picture_number, refs, &out) == SVTME_OK ? EB_ErrorNone : EB_ErrorInsufficientResources;
}

/* H6 — picture statistics (Codec/pic_analysis_process.c:2137, right after the pyramid): in place of
 * svt_aom_gathering_picture_statistics(scs, pcs, input_padded_pic, sixteenth). With it the 1/16 plane no longer has to come
 * back to the host for this function (H1 keeps the read-back only while the DG detector wants the planes). `var_rows` /
 * `hist` are flat arrays the patch keeps next to pcs->variance / pcs->picture_histogram (those are arrays of row pointers). */
EbErrorType svtme_hook_statistics(SvtMeHandle *svtme, const SequenceControlSet *scs, PictureParentControlSet *pcs, uint16_t *var_rows,
                                  uint32_t *hist /* [4][4][256] */) {
    SvtMeStatsParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.calc_hist              = scs->calc_hist;
    sp.calculate_variance     = scs->calculate_variance;
    sp.block_mean_prec_sub    = scs->block_mean_calc_prec == BLOCK_MEAN_PREC_SUB;
    sp.all_block_sizes        = scs->static_config.enable_adaptive_quantization == 1 || scs->static_config.variance_octile;
    sp.scene_change_detection = scs->static_config.scene_change_detection != 0;
    sp.regions_per_width      = scs->picture_analysis_number_of_regions_per_width;
    sp.regions_per_height     = scs->picture_analysis_number_of_regions_per_height;
    uint64_t intensity[16];
    SvtMeStatsOut so;
    memset(&so, 0, sizeof(so));
    so.variance = var_rows;  so.pic_avg_variance = &pcs->pic_avg_variance;  so.picture_histogram = hist;
    so.average_intensity_per_region = intensity;  so.avg_luma = &pcs->avg_luma;
    if (svtme_picture_statistics(svtme, pcs->picture_number, &sp, &so) != SVTME_OK) return EB_ErrorInsufficientResources;
    const uint32_t n_var = sp.all_block_sizes ? 85 : 1;
    if (sp.calculate_variance)
        for (uint32_t b = 0; b < pcs->b64_total_count; ++b) memcpy(pcs->variance[b], var_rows + (size_t)b * n_var, n_var * sizeof(uint16_t));
    if (sp.calc_hist)
        for (uint32_t x = 0; x < sp.regions_per_width; ++x)
            for (uint32_t y = 0; y < sp.regions_per_height; ++y) {
                memcpy(pcs->picture_histogram[x][y], hist + ((size_t)x * 4 + y) * 256, 256 * sizeof(uint32_t));
                pcs->average_intensity_per_region[x][y] = intensity[x * 4 + y];
            }
    return EB_ErrorNone;
}

/* H7 — dynamic-GOP detector (Codec/me_process.c:327-332, TASK_DG_DETECTOR_HME; tasks posted by early_hme, pd_process.c:591-633):
 * in place of dg_detector_hme_level0(pcs, segment_index). The first segment's task runs the whole picture on the GPU and
 * completes the frame (metrics + semaphore), the other segments' tasks return at once. Both pictures are resident because
 * their pyramids went through H1. */
EbErrorType svtme_hook_dg_detector(SvtMeHandle *svtme, PictureParentControlSet *pcs, uint32_t segment_index) {
    if (segment_index != 0) return EB_ErrorNone;
    SvtMeDgMetrics m;
    if (svtme_dg_detector_hme(svtme, pcs->picture_number, pcs->dg_detector->ref_pic->picture_number, (uint32_t)pcs->input_resolution, &m) != SVTME_OK)
        return EB_ErrorInsufficientResources;
    svt_block_on_mutex(pcs->dg_detector->metrics_mutex);
    pcs->dg_detector->metrics.tot_dist += m.tot_dist;
    pcs->dg_detector->metrics.tot_cplx += m.tot_cplx;
    pcs->dg_detector->metrics.tot_active += m.tot_active;
    pcs->dg_detector->metrics.sum_in_vectors += m.sum_in_vectors;
    pcs->dg_detector->metrics.seg_completed = (uint16_t)(pcs->me_segments_column_count * pcs->me_segments_row_count);
    svt_post_semaphore(pcs->dg_detector->frame_done_sem); /* (:584-586: all segments done) */
    svt_release_mutex(pcs->dg_detector->metrics_mutex);
    return EB_ErrorNone;
}

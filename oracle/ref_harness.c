/*
 * ref_harness.c — thin driver around the UNMODIFIED reference sources, compiled where they lie
 * under /root/reference by oracle/Makefile into oracle/_ref/libsvtref.so (git-ignored).
 * TEST INFRASTRUCTURE ONLY: used to pin oracle/me_oracle*.c, to generate tests/golden/, and as
 * the "reference" CPU baseline of bench.py. No reference source is copied into this repo; this
 * file only includes the reference's headers and calls its functions:
 *   svt_aom_motion_estimation_b64      Source/Lib/Codec/motion_estimation.c:3146
 *   svt_aom_sig_deriv_me               Source/Lib/Codec/enc_mode_config.c:684
 *   svt_aom_downsample_filtering_input_picture  Source/Lib/Codec/pic_analysis_process.c:1922
 *   svt_aom_generate_padding           Source/Lib/Codec/pic_operators.c:338
 *   svt_aom_gathering_picture_statistics  Source/Lib/Codec/pic_analysis_process.c:1560
 *   dg_detector_hme_level0             Source/Lib/Codec/pd_process.c:491
 *   RTCD pointers                      Source/Lib/Codec/aom_dsp_rtcd.h:776,838-860
 */
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

/* Instantiate the RTCD function-pointer globals here (the reference does it in aom_dsp_rtcd.c /
 * common_dsp_rtcd.c, which drag in the whole library). */
#define AOM_RTCD_C
#define RTCD_C
#include "aom_dsp_rtcd.h"
#include "common_dsp_rtcd.h"
#include "pcs.h"
#include "sequence_control_set.h"
#include "me_context.h"
#include "motion_estimation.h"
#include "reference_object.h"
#include "pic_operators.h"
#include "pic_analysis_process.h"
#include "enc_mode_config.h"
#include "compute_sad_c.h"
#include "me_sad_calculation.h"
#include "pack_unpack_c.h"
#include "compute_mean.h"

#include "svtme_b200.h"
#include "mt_pool.h"

void svt_aom_get_max_allocated_me_refs(uint8_t ref_count_used_list0, uint8_t ref_count_used_list1,
                                       uint8_t *max_ref_to_alloc, uint8_t *max_cand_to_alloc); /* Codec/pcs.c:90 */

/* ---- symbols the pulled-in objects reference but this path never reaches ---- */
void svt_log(int level, const char *tag, const char *fmt, ...) { (void)level; (void)tag; (void)fmt; }
void svt_aom_assert_err(uint32_t condition, char *err_msg) {
    if (!condition) { fprintf(stderr, "ref assert: %s\n", err_msg); abort(); }
}

/* AVX2 entry points of Source/Lib/ASM_AVX2/compute_sad_intrin_avx2.c + picture_operators */
void svt_sad_loop_kernel_avx2_intrin(uint8_t *, uint32_t, uint8_t *, uint32_t, uint32_t, uint32_t, uint64_t *,
                                     int16_t *, int16_t *, uint32_t, uint8_t, int16_t, int16_t);
void svt_ext_all_sad_calculation_8x8_16x16_avx2(uint8_t *, uint32_t, uint8_t *, uint32_t, uint32_t, uint32_t *,
                                                uint32_t *, uint32_t *, uint32_t *, uint32_t[16][8],
                                                uint32_t[64][8], Bool);
void svt_ext_eight_sad_calculation_32x32_64x64_avx2(uint32_t[16][8], uint32_t *, uint32_t *, uint32_t *,
                                                    uint32_t *, uint32_t, uint32_t[4][8]);
void svt_ext_sad_calculation_8x8_16x16_avx2_intrin(uint8_t *, uint32_t, uint8_t *, uint32_t, uint32_t *,
                                                   uint32_t *, uint32_t *, uint32_t *, uint32_t, uint32_t *,
                                                   uint32_t *, Bool);
void svt_ext_sad_calculation_32x32_64x64_sse4_intrin(uint32_t *, uint32_t *, uint32_t *, uint32_t *, uint32_t *,
                                                     uint32_t, uint32_t *);
uint32_t svt_nxm_sad_kernel_helper_avx2(const uint8_t *, uint32_t, const uint8_t *, uint32_t, uint32_t, uint32_t);
/* Source/Lib/ASM_AVX2/transforms_intrin_avx2.c:171 (aom_dsp_rtcd.h:837) */
void svt_aom_downsample_2d_avx2(uint8_t *, uint32_t, uint32_t, uint32_t, uint8_t *, uint32_t, uint32_t);

static void *memcpy_shim(void *d, void const *s, size_t n) { return memcpy(d, s, n); }
static void memcpy_void(void *d, void const *s, size_t n) { memcpy(d, s, n); }

static int g_simd = -1;

/* What svt_aom_setup_rtcd_internal (aom_dsp_rtcd.c:187, x86 block :499-514) does for the ME
 * pointers: C first, overwritten by AVX2 when allowed. simd: 0 = C_DEFAULT, 1 = AVX2. */
int ref_set_simd(int simd) {
    (void)memcpy_shim;
    svt_memcpy                                 = memcpy_void;
    svt_sad_loop_kernel                        = svt_sad_loop_kernel_c;
    svt_ext_all_sad_calculation_8x8_16x16      = svt_ext_all_sad_calculation_8x8_16x16_c;
    svt_ext_eight_sad_calculation_32x32_64x64  = svt_ext_eight_sad_calculation_32x32_64x64_c;
    svt_ext_sad_calculation_8x8_16x16          = svt_ext_sad_calculation_8x8_16x16_c;
    svt_ext_sad_calculation_32x32_64x64        = svt_ext_sad_calculation_32x32_64x64_c;
    svt_initialize_buffer_32bits               = svt_initialize_buffer_32bits_c;
    svt_nxm_sad_kernel                         = svt_nxm_sad_kernel_helper_c;
    svt_nxm_sad_kernel_sub_sampled             = svt_nxm_sad_kernel_helper_c;
    downsample_2d                              = svt_aom_downsample_2d_c;
    svt_compute_mean_8x8                       = svt_compute_mean_c;                /* aom_dsp_rtcd.c:515-518 */
    svt_compute_mean_square_values_8x8         = svt_compute_mean_squared_values_c;
    svt_compute_sub_mean_8x8                   = svt_compute_sub_mean_8x8_c;
    svt_compute_interm_var_four8x8             = svt_compute_interm_var_four8x8_c;
    svt_aom_un_pack2d_16_bit_src_mul4          = svt_enc_msb_un_pack2_d; /* common_dsp_rtcd.c:518, C entry */
#ifdef REF_HAVE_AVX2
    if (simd && __builtin_cpu_supports("avx2")) {
        svt_sad_loop_kernel                       = svt_sad_loop_kernel_avx2_intrin;
        svt_ext_all_sad_calculation_8x8_16x16     = svt_ext_all_sad_calculation_8x8_16x16_avx2;
        svt_ext_eight_sad_calculation_32x32_64x64 = svt_ext_eight_sad_calculation_32x32_64x64_avx2;
        svt_ext_sad_calculation_8x8_16x16         = svt_ext_sad_calculation_8x8_16x16_avx2_intrin;
        svt_nxm_sad_kernel                        = svt_nxm_sad_kernel_helper_avx2;
        svt_nxm_sad_kernel_sub_sampled            = svt_nxm_sad_kernel_helper_avx2;
        downsample_2d                             = svt_aom_downsample_2d_avx2; /* aom_dsp_rtcd.c:507 */
        svt_compute_mean_8x8                      = svt_compute_mean8x8_avx2_intrin;
        svt_compute_mean_square_values_8x8        = svt_compute_mean_of_squared_values8x8_sse2_intrin;
        svt_compute_sub_mean_8x8                  = svt_compute_sub_mean8x8_sse2_intrin;
        svt_compute_interm_var_four8x8            = svt_compute_interm_var_four8x8_avx2_intrin;
        g_simd                                    = 1;
        return 1;
    }
#endif
    g_simd = 0;
    return 0;
}

/* ---- planes ---- */
typedef struct RefPlane {
    uint8_t *buffer_y;
    uint32_t stride_y, org_x, org_y, width, height;
} RefPlane;
typedef struct RefPyramid {
    RefPlane full, quarter, sixteenth;
    uint64_t picture_number;
} RefPyramid;

static void to_desc(const RefPlane *pl, EbPictureBufferDesc *d) {
    memset(d, 0, sizeof(*d));
    d->buffer_y  = pl->buffer_y;
    d->stride_y  = (uint16_t)pl->stride_y;
    d->org_x     = (uint16_t)pl->org_x;
    d->org_y     = (uint16_t)pl->org_y;
    d->width     = (uint16_t)pl->width;
    d->height    = (uint16_t)pl->height;
    d->max_width = (uint16_t)pl->width;
    d->max_height = (uint16_t)pl->height;
}

static int plane_alloc(RefPlane *pl, uint32_t w, uint32_t h, uint32_t pad) {
    pl->width = w; pl->height = h; pl->org_x = pl->org_y = pad;
    pl->stride_y = w + 2 * pad;
    pl->buffer_y = (uint8_t *)calloc((size_t)pl->stride_y * (h + 2 * pad), 1);
    return pl->buffer_y ? 0 : -1;
}

void ref_pyramid_free(RefPyramid *pyr) {
    free(pyr->full.buffer_y); free(pyr->quarter.buffer_y); free(pyr->sixteenth.buffer_y);
    memset(pyr, 0, sizeof(*pyr));
}

/* Luma part of svt_aom_pad_input_pictures + svt_aom_downsample_filtering_input_picture, run
 * through the reference functions. Same geometry as orc_pyramid_build (pads 72/32/16). */
int ref_pyramid_build(RefPyramid *pyr, const uint8_t *luma, uint32_t stride, uint32_t width, uint32_t height,
                      int hme_level1, uint64_t picture_number) {
    if (g_simd < 0) ref_set_simd(0);
    const int reuse = pyr->full.buffer_y && pyr->quarter.buffer_y && pyr->sixteenth.buffer_y && pyr->full.width == width &&
        pyr->full.height == height; /* a recycled pa-reference object: refill in place */
    if (!reuse) {
        memset(pyr, 0, sizeof(*pyr));
        if (plane_alloc(&pyr->full, width, height, 72) || plane_alloc(&pyr->quarter, width >> 1, height >> 1, 32) ||
            plane_alloc(&pyr->sixteenth, width >> 2, height >> 2, 16)) {
            ref_pyramid_free(pyr);
            return -1;
        }
    }
    pyr->picture_number = picture_number;
    RefPlane *f = &pyr->full;
    for (uint32_t y = 0; y < height; ++y)
        memcpy(f->buffer_y + (size_t)(f->org_y + y) * f->stride_y + f->org_x, luma + (size_t)y * stride, width);
    svt_aom_generate_padding(f->buffer_y, f->stride_y, f->width, f->height, f->org_x, f->org_y);
    EbPictureBufferDesc df, dq, ds;
    to_desc(&pyr->full, &df); to_desc(&pyr->quarter, &dq); to_desc(&pyr->sixteenth, &ds);
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    if (!pcs) return -1;
    pcs->enable_hme_flag = 1;
    pcs->enable_hme_level0_flag = 1;
    pcs->enable_hme_level1_flag = hme_level1 ? 1 : 0;
    svt_aom_downsample_filtering_input_picture(pcs, &df, &dq, &ds);
    free(pcs);
    return 0;
}

/* ---- parameters ---- */
#include "ref_ctx_map.h"

/* Runs the reference's own svt_aom_sig_deriv_me + svt_aom_get_enable_me_8x8 + input-resolution
 * classification on a minimal scs/pcs and reports the MeContext it produced. */
int ref_derive_params(const SvtMeDeriveInput *in, SvtMeParams *out) {
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    MeContext               *ctx = (MeContext *)calloc(1, sizeof(*ctx));
    if (!scs || !pcs || !ctx) return -1;
    memset(out, 0, sizeof(*out));
    EbInputResolution res;
    svt_aom_derive_input_resolution(&res, in->width * in->height);
    scs->input_resolution = res;
    scs->static_config.pred_structure = in->rtc_tune ? SVT_AV1_PRED_LOW_DELAY_B : SVT_AV1_PRED_RANDOM_ACCESS;
    scs->static_config.qp = in->qp;
    scs->frame_rate = in->frame_rate_q16;
    scs->mrp_ctrls.safe_limit_nref = in->safe_limit_nref;
    scs->mrp_ctrls.safe_limit_zz_th = in->safe_limit_zz_th;
    pcs->scs = scs;
    pcs->enc_mode = (EncMode)in->enc_mode;
    pcs->sc_class1 = in->sc_class1;
    pcs->hierarchical_levels = in->hierarchical_levels;
    pcs->temporal_layer_index = in->temporal_layer_index;
    pcs->use_best_me_unipred_cand_only = in->use_best_me_unipred_cand_only;
    /* HME level flags: enc_mode_config.c:1798-1809 (first-pass derivation), open-loop values */
    pcs->enable_hme_flag = 1;
    pcs->enable_hme_level0_flag = 1;
    pcs->enable_hme_level1_flag = 1;
    pcs->enable_hme_level2_flag = (in->sc_class1 || in->enc_mode <= ENC_M6) ? 1 : 0;
    svt_aom_sig_deriv_me(scs, pcs, ctx);
    ctx_to_params(ctx, out);
    out->enable_me_8x8 = svt_aom_get_enable_me_8x8((EncMode)in->enc_mode, in->rtc_tune, res);
    out->enable_me_16x16 = svt_aom_get_enable_me_16x16((EncMode)in->enc_mode);
    out->input_resolution_le_480p = res <= INPUT_SIZE_480p_RANGE;
    uint8_t mr, mc;
    svt_aom_get_max_allocated_me_refs(in->ref_count_used_list0, in->ref_count_used_list1, &mr, &mc);
    out->max_refs = mr; out->max_cand = mc; out->max_l0 = in->ref_count_used_list0;
    free(scs); free(pcs); free(ctx);
    return 0;
}

/* Temporal-filter ME parameters from the reference's own svt_aom_sig_deriv_me_tf (enc_mode_config.c:829), followed by what
 * produce_temporally_filtered_pic does per block before it calls svt_aom_motion_estimation_b64 (temporal_filtering.c:296,
 * 315,3046-3074): me_type, FULL_SAD HME, one list / one reference, set_hme_search_params_mctf(ctx, 0). */
void svt_aom_sig_deriv_me_tf(PictureParentControlSet *pcs, MeContext *me_ctx);
int ref_derive_params_tf(const SvtMeTfDeriveInput *in, SvtMeParams *out) {
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    MeContext               *ctx = (MeContext *)calloc(1, sizeof(*ctx));
    if (!scs || !pcs || !ctx) return -1;
    memset(out, 0, sizeof(*out));
    EbInputResolution res;
    svt_aom_derive_input_resolution(&res, in->width * in->height);
    scs->input_resolution = res;
    scs->static_config.qp = in->qp;
    pcs->scs = scs;
    pcs->enc_mode = (EncMode)in->enc_mode;
    pcs->tf_ctrls.hme_me_level = in->hme_me_level;
    pcs->tf_ctrls.qp_opt = in->qp_opt;
    pcs->tf_ctrls.me_exit_th = in->me_exit_th;
    /* enc_mode_config.c:1810-1834 */
    pcs->tf_enable_hme_flag = pcs->tf_enable_hme_level0_flag = 1;
    pcs->tf_enable_hme_level1_flag = in->hme_me_level <= 2;
    pcs->tf_enable_hme_level2_flag = in->hme_me_level == 0;
    svt_aom_sig_deriv_me_tf(pcs, ctx);
    ctx->me_type = ME_MCTF;
    ctx->hme_search_method = FULL_SAD_SEARCH;
    ctx->hme_l0_sa = ctx->hme_l0_sa_default_tf; /* set_hme_search_params_mctf(ctx, 0) */
    ctx->tf_me_exit_th = pcs->tf_ctrls.me_exit_th;
    ctx_to_params(ctx, out);
    out->num_of_list_to_search = 1;
    out->num_of_ref_pic_to_search[0] = 1;
    out->temporal_layer_index = in->temporal_layer_index;
    out->is_ref = in->is_ref;
    out->max_number_of_pus_per_sb = 85;
    out->enable_me_8x8 = out->enable_me_16x16 = 1;
    out->max_refs = out->max_cand = out->max_l0 = 1;
    out->input_resolution_le_480p = res <= INPUT_SIZE_480p_RANGE;
    free(scs); free(pcs); free(ctx);
    return 0;
}

/* ---- whole pictures through svt_aom_motion_estimation_b64 ----
 * PicShell: the picture-level objects one ME picture needs (scs / pcs / MeSbResults shells pointing straight
 * into the caller's SvtMeOut arrays; outputs the caller does not want go to scratch). Shared read-only by the
 * worker threads: svt_aom_motion_estimation_b64 writes only per-b64 entries (motion_estimation.c:2711,2902,
 * 3017-3076,3216-3217). ThreadCtx: what an ME thread owns (MeContext, me_process.c:75-88). */
typedef struct PicShell {
    const SvtMeParams *p;
    const RefPyramid  *cur;
    const RefPyramid *const *refs; /* [2][4] flattened */
    SvtMeOut          *out;
    uint32_t           bw, bh, nb, n_pu;
    SequenceControlSet      *scs;
    PictureParentControlSet *pcs;
    MotionEstimationData    *med;
    MeSbResults             *res;
    MeSbResults            **resp;
    B64Geom                 *geom;
    void                    *scratch[11];
} PicShell;

typedef struct ThreadCtx {
    MeContext          *ctx;
    EbPictureBufferDesc dcur, dref[2][4][3];
    const PicShell     *bound;
} ThreadCtx;

static void shell_free(PicShell *s) {
    free(s->scs); free(s->pcs); free(s->med); free(s->res); free(s->resp); free(s->geom);
    for (int i = 0; i < 11; ++i) free(s->scratch[i]);
    memset(s, 0, sizeof(*s));
}

static int shell_init(PicShell *s, const SvtMeParams *p, const RefPyramid *cur, const RefPyramid *const *refs, SvtMeOut *out) {
    memset(s, 0, sizeof(*s));
    s->p = p; s->cur = cur; s->refs = refs; s->out = out;
    const uint32_t W = cur->full.width, H = cur->full.height;
    s->bw = (W + 63) / 64; s->bh = (H + 63) / 64; s->nb = s->bw * s->bh;
    s->n_pu = p->enable_me_16x16 ? (p->enable_me_8x8 ? p->max_number_of_pus_per_sb : 21u) : 5u;
    const uint32_t nb = s->nb, n_pu = s->n_pu;
    s->scs  = (SequenceControlSet *)calloc(1, sizeof(*s->scs));
    s->pcs  = (PictureParentControlSet *)calloc(1, sizeof(*s->pcs));
    s->med  = (MotionEstimationData *)calloc(1, sizeof(*s->med));
    s->res  = (MeSbResults *)calloc(nb, sizeof(*s->res));
    s->resp = (MeSbResults **)calloc(nb, sizeof(*s->resp));
    s->geom = (B64Geom *)calloc(nb, sizeof(*s->geom));
    if (!s->scs || !s->pcs || !s->med || !s->res || !s->resp || !s->geom) { shell_free(s); return -1; }
    int k = 0, bad = 0;
#define OUTPTR(field, bytes) ((out->field) ? (void *)(out->field) : (s->scratch[k] = calloc((bytes), 1), bad |= !s->scratch[k], s->scratch[k++]))
    uint8_t  *tot  = (uint8_t *)OUTPTR(total_me_candidate_index, (size_t)nb * n_pu);
    uint32_t *mvs  = (uint32_t *)OUTPTR(me_mv_array, (size_t)nb * n_pu * p->max_refs * 4);
    uint8_t  *cand = (uint8_t *)OUTPTR(me_candidate_array, (size_t)nb * n_pu * p->max_cand);
    PictureParentControlSet *pcs = s->pcs;
    pcs->me_64x64_distortion         = (uint32_t *)OUTPTR(me_64x64_distortion, (size_t)nb * 4);
    pcs->me_32x32_distortion         = (uint32_t *)OUTPTR(me_32x32_distortion, (size_t)nb * 4);
    pcs->me_16x16_distortion         = (uint32_t *)OUTPTR(me_16x16_distortion, (size_t)nb * 4);
    pcs->me_8x8_distortion           = (uint32_t *)OUTPTR(me_8x8_distortion, (size_t)nb * 4);
    pcs->me_8x8_cost_variance        = (uint32_t *)OUTPTR(me_8x8_cost_variance, (size_t)nb * 4);
    pcs->rc_me_distortion            = (uint32_t *)OUTPTR(rc_me_distortion, (size_t)nb * 4);
    pcs->stationary_block_present_sb = (uint8_t *)OUTPTR(stationary_block_present_sb, nb);
    pcs->rc_me_allow_gm              = (uint8_t *)OUTPTR(rc_me_allow_gm, nb);
#undef OUTPTR
    if (bad) { shell_free(s); return -1; }
    for (uint32_t b = 0; b < nb; ++b) {
        s->res[b].total_me_candidate_index = tot + (size_t)b * n_pu;
        s->res[b].me_mv_array              = (MvCandidate *)(mvs + (size_t)b * n_pu * p->max_refs);
        s->res[b].me_candidate_array       = (MeCandidate *)(cand + (size_t)b * n_pu * p->max_cand);
        s->resp[b]                         = &s->res[b];
        uint32_t ox = (b % s->bw) * 64, oy = (b / s->bw) * 64;
        s->geom[b].org_x = (uint16_t)ox; s->geom[b].org_y = (uint16_t)oy;
        s->geom[b].width = (uint8_t)((W - ox) < 64 ? W - ox : 64);
        s->geom[b].height = (uint8_t)((H - oy) < 64 ? H - oy : 64);
    }
    s->med->me_results = s->resp; s->med->max_cand = p->max_cand; s->med->max_refs = p->max_refs; s->med->max_l0 = p->max_l0;
    s->scs->input_resolution = p->input_resolution_le_480p ? INPUT_SIZE_480p_RANGE : INPUT_SIZE_1080p_RANGE;
    s->scs->mrp_ctrls.only_l_bwd = p->only_l_bwd;
    s->scs->static_config.restricted_motion_vector = 0;
    pcs->scs = s->scs; pcs->pa_me_data = s->med; pcs->b64_geom = s->geom;
    pcs->picture_number = p->picture_number;
    pcs->aligned_width = (uint16_t)W; pcs->aligned_height = (uint16_t)H;
    pcs->enable_me_8x8 = p->enable_me_8x8; pcs->enable_me_16x16 = p->enable_me_16x16;
    pcs->max_number_of_pus_per_sb = p->max_number_of_pus_per_sb;
    pcs->hierarchical_levels = p->hierarchical_levels; pcs->temporal_layer_index = p->temporal_layer_index;
    pcs->similar_brightness_refs = p->similar_brightness_refs;
    pcs->gm_ctrls.enabled = p->gm_enabled; pcs->gm_ctrls.use_distance_based_active_th = p->gm_use_distance_based_active_th;
    return 0;
}

/* What me_process.c:219-262 does when an ME thread picks up a picture: per-picture scalars and the
 * me_ds_ref_array pointers go into the thread's MeContext. */
static void thread_bind(ThreadCtx *t, const PicShell *s) {
    if (t->bound == s) return;
    const SvtMeParams *p = s->p;
    memset(t->ctx, 0, sizeof(*t->ctx));
    params_to_ctx(p, t->ctx);
    to_desc(&s->cur->full, &t->dcur);
    for (int li = 0; li < p->num_of_list_to_search; ++li)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            const RefPyramid *r = s->refs[li * 4 + ri];
            to_desc(&r->full, &t->dref[li][ri][0]);
            to_desc(&r->quarter, &t->dref[li][ri][1]);
            to_desc(&r->sixteenth, &t->dref[li][ri][2]);
            t->ctx->me_ds_ref_array[li][ri].picture_ptr           = &t->dref[li][ri][0];
            t->ctx->me_ds_ref_array[li][ri].quarter_picture_ptr   = &t->dref[li][ri][1];
            t->ctx->me_ds_ref_array[li][ri].sixteenth_picture_ptr = &t->dref[li][ri][2];
            t->ctx->me_ds_ref_array[li][ri].picture_number        = p->ref_picture_number[li][ri];
        }
    t->bound = s;
}

/* The b64 loop of me_process.c:174-289 over b64 rows [row0, row1), with its pointer set-up (:189-213). */
static void run_rows(ThreadCtx *t, const PicShell *s, uint32_t row0, uint32_t row1) {
    thread_bind(t, s);
    MeContext *ctx = t->ctx;
    SvtMeOut  *out = s->out;
    const RefPyramid *cur = s->cur;
    const uint32_t bw = s->bw;
    if (row1 > s->bh) row1 = s->bh;
    for (uint32_t by = row0; by < row1; ++by)
        for (uint32_t bx = 0; bx < bw; ++bx) {
            const uint32_t b = bx + by * bw, ox = bx * 64, oy = by * 64;
            const RefPlane *f = &cur->full, *q = &cur->quarter, *x = &cur->sixteenth;
            ctx->b64_src_ptr    = f->buffer_y + (size_t)(f->org_y + oy) * f->stride_y + f->org_x + ox;
            ctx->b64_src_stride = f->stride_y;
            ctx->quarter_b64_buffer = q->buffer_y + (size_t)(q->org_y + (oy >> 1)) * q->stride_y + q->org_x + (ox >> 1);
            ctx->quarter_b64_buffer_stride = q->stride_y;
            ctx->sixteenth_b64_buffer = x->buffer_y + (size_t)(x->org_y + (oy >> 2)) * x->stride_y + x->org_x + (ox >> 2);
            ctx->sixteenth_b64_buffer_stride = x->stride_y;
            ctx->me_type = s->p->me_type == 1 ? ME_MCTF : ME_OPEN_LOOP;
            /* not part of the reference: clear the tables of refs the block will not search so the
             * deep-state dump is comparable (they are never read before being initialised). */
            if (out->p_sb_best_sad) memset(ctx->p_sb_best_sad, 0, sizeof(ctx->p_sb_best_sad));
            svt_aom_motion_estimation_b64(s->pcs, b, ox, oy, ctx, &t->dcur);
            if (out->p_sb_best_sad) memcpy(out->p_sb_best_sad + (size_t)b * 2 * 4 * 85, ctx->p_sb_best_sad, sizeof(ctx->p_sb_best_sad));
            if (out->p_sb_best_mv) memcpy(out->p_sb_best_mv + (size_t)b * 2 * 4 * 85, ctx->p_sb_best_mv, sizeof(ctx->p_sb_best_mv));
            if (out->hme_sc || out->hme_sad || out->do_ref)
                for (int li = 0; li < 2; ++li)
                    for (int ri = 0; ri < 4; ++ri) {
                        const size_t k = ((size_t)b * 2 + li) * 4 + ri;
                        if (out->hme_sc) { out->hme_sc[2 * k] = ctx->search_results[li][ri].hme_sc_x; out->hme_sc[2 * k + 1] = ctx->search_results[li][ri].hme_sc_y; }
                        if (out->hme_sad) out->hme_sad[k] = ctx->search_results[li][ri].hme_sad;
                        if (out->do_ref) out->do_ref[k] = ctx->search_results[li][ri].do_ref;
                    }
        }
}

/* Single thread, raster b64 order, rows [row0,row1) of b64s. */
int ref_me_picture_rows(const SvtMeParams *p, const RefPyramid *cur, const RefPyramid *refs[2][4], SvtMeOut *out,
                        uint32_t row0, uint32_t row1) {
    if (g_simd < 0) ref_set_simd(0);
    PicShell  s;
    ThreadCtx t;
    memset(&t, 0, sizeof(t));
    t.ctx = (MeContext *)calloc(1, sizeof(*t.ctx));
    if (!t.ctx || shell_init(&s, p, cur, (const RefPyramid *const *)refs, out)) { free(t.ctx); return -1; }
    run_rows(&t, &s, row0, row1);
    shell_free(&s);
    free(t.ctx);
    return 0;
}

int ref_me_picture(const SvtMeParams *p, const RefPyramid *cur, const RefPyramid *refs[2][4], SvtMeOut *out) {
    return ref_me_picture_rows(p, cur, refs, out, 0, 0xffffffffu);
}

/* ---- all-cores form (bench.py's reference arm): the reference's own threading scheme — ME threads pulling
 * (picture, segment) tasks from one queue (me_process.c:97-124; segments enc_handle.c:443-461, here a segment =
 * seg_rows rows of b64s of one picture) — without the encoder around it. Every thread owns a MeContext like the
 * reference's ME threads do; results go straight into each job's SvtMeOut arrays. */
typedef struct RefMtJob {
    SvtMeParams       params;
    const RefPyramid *cur;
    const RefPyramid *refs[2][4];
    SvtMeOut          out;
} RefMtJob;

typedef struct MeMt {
    PicShell  *shells;
    ThreadCtx *threads;
    uint32_t   segs_per_pic, seg_rows;
} MeMt;

static void me_mt_item(void *vctx, uint32_t item, uint32_t thread) {
    MeMt *m = (MeMt *)vctx;
    const uint32_t pic = item / m->segs_per_pic, seg = item % m->segs_per_pic;
    run_rows(&m->threads[thread], &m->shells[pic], seg * m->seg_rows, (seg + 1) * m->seg_rows);
}

int ref_me_pictures_mt(RefMtJob *jobs, uint32_t n, uint32_t threads, uint32_t seg_rows) {
    if (g_simd < 0) ref_set_simd(0);
    if (!jobs || !n) return -1;
    if (threads < 1) threads = 1;
    if (seg_rows < 1) seg_rows = 1;
    MeMt m;
    m.shells  = (PicShell *)calloc(n, sizeof(PicShell));
    m.threads = (ThreadCtx *)calloc(threads, sizeof(ThreadCtx));
    int rc = (!m.shells || !m.threads) ? -1 : 0;
    uint32_t bh = 0;
    for (uint32_t i = 0; i < n && !rc; ++i) {
        rc = shell_init(&m.shells[i], &jobs[i].params, jobs[i].cur, &jobs[i].refs[0][0], &jobs[i].out);
        if (!rc) {
            if (i && m.shells[i].bh != bh) rc = -1; /* one picture size per call */
            bh = m.shells[i].bh;
        }
    }
    for (uint32_t t = 0; t < threads && !rc; ++t)
        if (!(m.threads[t].ctx = (MeContext *)calloc(1, sizeof(MeContext)))) rc = -1;
    if (!rc) {
        m.seg_rows     = seg_rows;
        m.segs_per_pic = (bh + seg_rows - 1) / seg_rows;
        rc = mt_run(n * m.segs_per_pic, threads, me_mt_item, &m);
    }
    if (m.shells) for (uint32_t i = 0; i < n; ++i) shell_free(&m.shells[i]);
    if (m.threads) for (uint32_t t = 0; t < threads; ++t) free(m.threads[t].ctx);
    free(m.shells); free(m.threads);
    return rc;
}

/* Pyramids of n pictures on the pool (one task per picture). Planes that are already allocated with the right
 * size are refilled in place, like the encoder's recycled pa-reference objects. */
typedef struct PyrMt {
    RefPyramid *pyr;
    const uint8_t *const *luma;
    const uint64_t *nums;
    uint32_t stride, width, height;
    int hme_level1, failed;
} PyrMt;

static void pyr_mt_item(void *vctx, uint32_t item, uint32_t thread) {
    (void)thread;
    PyrMt *m = (PyrMt *)vctx;
    if (ref_pyramid_build(&m->pyr[item], m->luma[item], m->stride, m->width, m->height, m->hme_level1, m->nums[item]))
        __atomic_store_n(&m->failed, 1, __ATOMIC_RELAXED);
}

int ref_pyramids_build_mt(RefPyramid *pyr, const uint8_t *const *luma, uint32_t stride, uint32_t width, uint32_t height,
                          int hme_level1, const uint64_t *picture_numbers, uint32_t n, uint32_t threads) {
    if (g_simd < 0) ref_set_simd(0);
    PyrMt m = {pyr, luma, picture_numbers, stride, width, height, hme_level1, 0};
    int rc = mt_run(n, threads, pyr_mt_item, &m);
    return (rc || m.failed) ? -1 : 0;
}
unsigned ref_sizeof_mt_job(void) { return (unsigned)sizeof(RefMtJob); }

int ref_simd_level(void) { return g_simd; }

/* 10-bit source -> the 8-bit plane ME reads: svt_aom_un_pack2d (Codec/pic_operators.c:228-244) with no 2-bit plane. */
void ref_unpack_msb(const uint16_t *in16, uint32_t in_stride, uint8_t *out8, uint32_t out8_stride, uint32_t width,
                    uint32_t height) {
    if (g_simd < 0) ref_set_simd(0);
    svt_aom_un_pack2d((uint16_t *)in16, in_stride, out8, out8_stride, NULL, 0, width, height);
}
unsigned ref_sizeof_params(void) { return (unsigned)sizeof(SvtMeParams); }
unsigned ref_sizeof_derive_input(void) { return (unsigned)sizeof(SvtMeDeriveInput); }
unsigned ref_sizeof_out(void) { return (unsigned)sizeof(SvtMeOut); }

/* ---- picture statistics: svt_aom_gathering_picture_statistics (pic_analysis_process.c:1560) on a built pyramid ----
 * The function takes the sequence and picture control sets; the shells below hold the fields it reads
 * (scs: calc_hist, calculate_variance, block_mean_calc_prec, the region counts, static_config.scene_change_detection /
 * enable_adaptive_quantization / variance_octile; pcs: b64_total_count, b64_geom[].org_x/org_y) and the arrays it
 * writes (variance, picture_histogram, average_intensity_per_region, avg_luma, pic_avg_variance). */
int ref_picture_statistics(const RefPyramid *pyr, const SvtMeStatsParams *prm, SvtMeStatsOut *out) {
    if (g_simd < 0) ref_set_simd(0);
    if (prm->regions_per_width < 1 || prm->regions_per_width > MAX_NUMBER_OF_REGIONS_IN_WIDTH || prm->regions_per_height < 1 ||
        prm->regions_per_height > MAX_NUMBER_OF_REGIONS_IN_HEIGHT)
        return -1;
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    if (!scs || !pcs) { free(scs); free(pcs); return -1; }
    const uint32_t cols = (pyr->full.width + 63) / 64, rows = (pyr->full.height + 63) / 64, n_b64 = cols * rows;
    const uint32_t n_var = prm->all_block_sizes ? 85 : 1;
    scs->calc_hist                                      = prm->calc_hist;
    scs->calculate_variance                             = prm->calculate_variance;
    scs->block_mean_calc_prec                           = prm->block_mean_prec_sub ? BLOCK_MEAN_PREC_SUB : BLOCK_MEAN_PREC_FULL;
    scs->picture_analysis_number_of_regions_per_width   = prm->regions_per_width;
    scs->picture_analysis_number_of_regions_per_height  = prm->regions_per_height;
    scs->static_config.scene_change_detection           = prm->scene_change_detection;
    scs->static_config.enable_adaptive_quantization     = prm->all_block_sizes ? 1 : 0;
    scs->static_config.variance_octile                  = 0;
    pcs->b64_total_count = (uint16_t)n_b64;
    pcs->b64_geom        = (B64Geom *)calloc(n_b64, sizeof(B64Geom));
    pcs->variance        = (uint16_t **)calloc(n_b64, sizeof(uint16_t *));
    uint16_t *var_rows   = (uint16_t *)calloc((size_t)n_b64 * n_var, sizeof(uint16_t));
    uint32_t *hist       = (uint32_t *)calloc(4 * 4 * 256, sizeof(uint32_t));
    uint32_t **hist_y[MAX_NUMBER_OF_REGIONS_IN_WIDTH];
    uint32_t  *hist_p[MAX_NUMBER_OF_REGIONS_IN_WIDTH][MAX_NUMBER_OF_REGIONS_IN_HEIGHT];
    int rc = -1;
    if (pcs->b64_geom && pcs->variance && var_rows && hist) {
        for (uint32_t b = 0; b < n_b64; ++b) {
            pcs->b64_geom[b].org_x = (uint16_t)((b % cols) * 64);
            pcs->b64_geom[b].org_y = (uint16_t)((b / cols) * 64);
            pcs->variance[b]       = var_rows + (size_t)b * n_var;
        }
        for (int x = 0; x < MAX_NUMBER_OF_REGIONS_IN_WIDTH; ++x) {
            for (int y = 0; y < MAX_NUMBER_OF_REGIONS_IN_HEIGHT; ++y) hist_p[x][y] = hist + ((size_t)x * 4 + y) * 256;
            hist_y[x] = hist_p[x];
        }
        pcs->picture_histogram = hist_y;
        EbPictureBufferDesc df, ds;
        to_desc(&pyr->full, &df); to_desc(&pyr->sixteenth, &ds);
        svt_aom_gathering_picture_statistics(scs, pcs, &df, &ds);
        if (prm->calculate_variance && out->variance) memcpy(out->variance, var_rows, (size_t)n_b64 * n_var * sizeof(uint16_t));
        if (out->pic_avg_variance) *out->pic_avg_variance = pcs->pic_avg_variance;
        if (prm->calc_hist) {
            if (out->picture_histogram) memcpy(out->picture_histogram, hist, 4 * 4 * 256 * sizeof(uint32_t));
            if (out->average_intensity_per_region)
                for (int x = 0; x < 4; ++x)
                    for (int y = 0; y < 4; ++y) out->average_intensity_per_region[x * 4 + y] = pcs->average_intensity_per_region[x][y];
        }
        if (out->avg_luma) *out->avg_luma = pcs->avg_luma;
        rc = 0;
    }
    free(hist); free(var_rows); free(pcs->variance); free(pcs->b64_geom); free(pcs); free(scs);
    return rc;
}

/* ---- dynamic-GOP detector HME: dg_detector_hme_level0 (pd_process.c:491) with ONE segment covering the picture ----
 * The function takes its pictures from ppcs->pa_ref_pic_wrapper / ppcs->dg_detector->ref_pic->pa_ref_pic_wrapper and adds
 * into ppcs->dg_detector->metrics under a mutex; the three thread primitives it calls are no-ops here (one thread). */
EbErrorType svt_block_on_mutex(EbHandle mutex_handle) { (void)mutex_handle; return EB_ErrorNone; }
EbErrorType svt_release_mutex(EbHandle mutex_handle) { (void)mutex_handle; return EB_ErrorNone; }
EbErrorType svt_post_semaphore(EbHandle semaphore_handle) { (void)semaphore_handle; return EB_ErrorNone; }
void dg_detector_hme_level0(struct PictureParentControlSet *ppcs, uint32_t seg_idx); /* pd_process.c:491 */

int ref_dg_detector_hme(const RefPyramid *src, const RefPyramid *ref, uint32_t input_resolution, SvtMeDgMetrics *out) {
    if (g_simd < 0) ref_set_simd(0);
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs)), *rpcs = (PictureParentControlSet *)calloc(1, sizeof(*rpcs));
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    if (!pcs || !rpcs || !scs) { free(pcs); free(rpcs); free(scs); return -1; }
    EbPictureBufferDesc  ds, dr;
    EbPaReferenceObject  pa_s, pa_r;
    EbObjectWrapper      w_s, w_r;
    DGDetectorSeg        dg;
    memset(&pa_s, 0, sizeof(pa_s)); memset(&pa_r, 0, sizeof(pa_r)); memset(&w_s, 0, sizeof(w_s)); memset(&w_r, 0, sizeof(w_r));
    memset(&dg, 0, sizeof(dg));
    to_desc(&src->sixteenth, &ds); to_desc(&ref->sixteenth, &dr);
    pa_s.sixteenth_downsampled_picture_ptr = &ds;  pa_r.sixteenth_downsampled_picture_ptr = &dr;
    w_s.object_ptr = &pa_s;  w_r.object_ptr = &pa_r;
    pcs->pa_ref_pic_wrapper = &w_s;  rpcs->pa_ref_pic_wrapper = &w_r;
    dg.ref_pic = rpcs;
    pcs->dg_detector = &dg;
    pcs->scs = scs;
    scs->b64_size = 64;
    pcs->input_resolution = (EbInputResolution)input_resolution;
    pcs->aligned_width = (uint16_t)src->full.width;  pcs->aligned_height = (uint16_t)src->full.height;
    pcs->me_segments_column_count = 1;  pcs->me_segments_row_count = 1;
    dg_detector_hme_level0(pcs, 0);
    out->tot_dist = dg.metrics.tot_dist;  out->tot_cplx = dg.metrics.tot_cplx;  out->tot_active = dg.metrics.tot_active;
    out->sum_in_vectors = dg.metrics.sum_in_vectors;
    const int done = dg.metrics.seg_completed == 1;
    free(pcs); free(rpcs); free(scs);
    return done ? 0 : -1;
}

/*
 * svtme_b200.h — C ABI of the B200-native open-loop motion-estimation path.
 *
 * Drop-in boundary for ONE stage of SVT-AV1 (reference: Patman86/SVT-AV1-Mod-by-Patman):
 * the pre-ME pyramid (1/4, 1/16 pictures), pre-HME, HME level 0/1/2 and the full-pel integer
 * search that fills the per-64x64 best-MV/SAD tables of the 85 square PUs.
 *
 * Two groups of entry points:
 *   (1) production ABI  svtme_*         — picture granular; replaces the body of the PAME branch of
 *       svt_aom_motion_estimation_kernel (Source/Lib/Codec/me_process.c:126-313) and
 *       svt_aom_downsample_filtering_input_picture (Source/Lib/Codec/pic_analysis_process.c:1922).
 *   (2) kernel-test ABI *_b200          — the RTCD signatures of Source/Lib/Codec/aom_dsp_rtcd.h
 *       (lines 776, 838-854) as synchronous host wrappers (H2D, launch, D2H) so that SadTest /
 *       PictureOperatorTest shaped harnesses can be pointed at the CUDA kernels unchanged.
 *
 * Plain C: pointers + sizes only, no C++/torch types. All functions return 0 on success and a
 * negative SVTME_ERR_* otherwise (kernel-test wrappers that mirror `void` RTCD functions abort()
 * on a CUDA failure — there is NO CPU fallback anywhere in this library).
 */
#ifndef SVTME_B200_H
#define SVTME_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVTME_OK 0
#define SVTME_ERR_INVALID -1   /* bad argument / unsupported parameter combination */
#define SVTME_ERR_NOMEM -2     /* host or device allocation failed (EB_ErrorInsufficientResources) */
#define SVTME_ERR_CUDA -3      /* CUDA runtime error; svtme_last_error() has the text */
#define SVTME_ERR_NOTFOUND -4  /* picture number not resident */

#define SVTME_MAX_LISTS 2      /* MAX_NUM_OF_REF_PIC_LIST  definitions.h:2349 */
#define SVTME_MAX_REFS 4       /* MAX_REF_IDX / REF_LIST_MAX_DEPTH definitions.h:2364 */
#define SVTME_PU_COUNT 85      /* SQUARE_PU_COUNT: 1 + 4 + 16 + 64 */
#define SVTME_MAX_SAD_VALUE (128u * 128u * 255u) /* MAX_SAD_VALUE motion_estimation.h:85 */

typedef struct SvtMeSearchArea {
    uint16_t width, height;    /* SearchArea me_context.h:319-322 */
} SvtMeSearchArea;

typedef struct SvtMeSearchAreaMinMax {
    SvtMeSearchArea sa_min, sa_max; /* SearchAreaMinMax me_context.h:323-326 */
} SvtMeSearchAreaMinMax;

/*
 * Scalar ME parameters of one picture. Field names and meaning follow MeContext
 * (Source/Lib/Codec/me_context.h:366-509) as filled by svt_aom_sig_deriv_me
 * (enc_mode_config.c:684-823) plus the per-picture fields set in me_process.c:219-227 and the
 * PictureParentControlSet / SequenceControlSet scalars motion_estimation.c reads.
 * svtme_derive_params() fills it from (preset, resolution, qp, ...) exactly like the reference.
 */
typedef struct SvtMeParams {
    /* search areas */
    SvtMeSearchAreaMinMax me_sa;          /* full-pel min/max search area */
    SvtMeSearchAreaMinMax hme_l0_sa;      /* total HME level-0 area (all quadrants) */
    SvtMeSearchArea hme_l1_sa, hme_l2_sa; /* refinement windows */
    uint16_t num_hme_sa_w, num_hme_sa_h;  /* always 2x2 (enc_mode_config.c:144-145) */
    /* flags */
    uint8_t enable_hme_flag, enable_hme_level0_flag, enable_hme_level1_flag, enable_hme_level2_flag;
    uint8_t hme_search_method, me_search_method; /* 0 SUB_SAD_SEARCH, 1 FULL_SAD_SEARCH (definitions.h:2079-2080) */
    uint8_t reduce_hme_l0_sr_th_min, reduce_hme_l0_sr_th_max;
    /* PreHmeCtrls me_context.h:334-339 */
    uint8_t prehme_enable, prehme_skip_search_line, prehme_l1_early_exit, _pad0;
    SvtMeSearchAreaMinMax prehme_sa_cfg[2];
    /* MeHmeRefPruneCtrls me_context.h:280-290 */
    uint8_t enable_me_hme_ref_pruning, _pad1;
    uint16_t prune_ref_if_hme_sad_dev_bigger_than_th;
    uint16_t prune_ref_if_me_sad_dev_bigger_than_th;
    uint16_t zz_sad_pct;
    uint32_t zz_sad_th;
    uint32_t phme_sad_th;
    uint16_t phme_sad_pct;
    /* MeSrCtrls me_context.h:292-305 */
    uint8_t enable_me_sr_adjustment, distance_based_hme_resizing;
    uint16_t reduce_me_sr_based_on_mv_length_th;
    uint16_t stationary_hme_sad_abs_th;
    uint16_t stationary_me_sr_divisor;
    uint16_t reduce_me_sr_based_on_hme_sad_abs_th;
    uint16_t me_sr_divisor_for_low_hme_sad;
    /* Me8x8VarCtrls me_context.h:310-318 */
    uint8_t me_8x8_var_enabled, _pad2;
    uint32_t me_sr_div4_th, me_sr_div2_th, me_sr_mult2_th;
    /* MvBasedSearchAdj me_context.h:355-363 */
    uint8_t mv_sa_adj_enabled, mv_sa_adj_nearest_ref_only;
    uint16_t mv_sa_adj_mv_size_th, mv_sa_adj_sa_multiplier;
    /* misc MeContext scalars */
    int32_t prune_me_candidates_th;
    uint8_t use_best_unipred_cand_only;
    /* me_ctx->me_type (me_context.h:44-51,459): 0 = the open-loop ME of the PAME tasks (ME_OPEN_LOOP), 1 = ME_MCTF, the
     * temporal-filter ME of TASK_TFME (temporal_filtering.c:3075): no reference pruning, unscaled picture distance in the
     * full-pel area (motion_estimation.c:1302), early return after HME when search_results[0][0].hme_sad < tf_me_exit_th
     * (:3179-3183), no candidate arrays / distortions: its consumers read p_sb_best_sad / p_sb_best_mv and the HME
     * results (the "deep" SvtMeOut arrays). */
    uint8_t me_type;
    uint16_t tf_me_exit_th;               /* me_ctx->tf_me_exit_th = pcs->tf_ctrls.me_exit_th (temporal_filtering.c:3064) */
    uint32_t me_early_exit_th, me_safe_limit_zz_th, prev_me_stage_based_exit_th;
    /* per-picture, me_process.c:219-227 */
    uint8_t num_of_list_to_search;
    uint8_t num_of_ref_pic_to_search[2];
    uint8_t temporal_layer_index;
    uint8_t is_ref;
    /* pcs / scs scalars read by motion_estimation.c */
    uint8_t hierarchical_levels;        /* pcs->hierarchical_levels        :2496 */
    uint8_t similar_brightness_refs;    /* pcs->similar_brightness_refs    :2497 */
    uint8_t enable_me_8x8;              /* pcs->enable_me_8x8              :2785 */
    uint8_t enable_me_16x16;            /* pcs->enable_me_16x16 */
    uint8_t max_number_of_pus_per_sb;   /* pcs->max_number_of_pus_per_sb (85) */
    uint8_t only_l_bwd;                 /* scs->mrp_ctrls.only_l_bwd       :2852 */
    uint8_t gm_enabled;                 /* pcs->gm_ctrls.enabled           :3219 */
    uint8_t gm_use_distance_based_active_th;
    uint8_t input_resolution_le_480p;   /* scs->input_resolution <= INPUT_SIZE_480p_RANGE */
    uint8_t max_cand, max_refs, max_l0; /* pcs->pa_me_data->max_* (pd_process.c:3510-3512) */
    uint8_t _pad4[3];
    uint64_t picture_number;                                 /* pcs->picture_number */
    uint64_t ref_picture_number[SVTME_MAX_LISTS][SVTME_MAX_REFS]; /* me_ds_ref_array[][].picture_number */
} SvtMeParams;

/* Encoder-level inputs of svt_aom_sig_deriv_me + the flag derivations around it. */
typedef struct SvtMeDeriveInput {
    int32_t enc_mode;            /* preset: -1 (MR) .. 13 */
    uint32_t width, height;      /* luma size -> EbInputResolution (sequence_control_set.c:113-131) */
    uint32_t qp;                 /* scs->static_config.qp (default 35) */
    uint8_t rtc_tune;            /* pred_structure == SVT_AV1_PRED_LOW_DELAY_B */
    uint8_t sc_class1;           /* screen-content class 1 */
    uint8_t hierarchical_levels; /* pcs->hierarchical_levels */
    uint8_t temporal_layer_index;
    uint8_t is_ref;
    uint8_t slice_type_b;        /* 1: B_SLICE (two lists), 0: P_SLICE */
    uint8_t ref_list0_count_try, ref_list1_count_try; /* pd_process.c:3666-3677 */
    uint32_t frame_rate_q16;     /* scs->frame_rate (Q16) — low-frame-rate scaling :346-350 */
    uint8_t safe_limit_nref;     /* scs->mrp_ctrls.safe_limit_nref */
    uint8_t only_l_bwd;
    uint8_t use_best_me_unipred_cand_only;
    uint8_t similar_brightness_refs;
    uint32_t safe_limit_zz_th;
    uint8_t ref_count_used_list0, ref_count_used_list1; /* MeSbResults allocation pcs.c:90-95 */
    uint8_t gm_level_enabled, gm_use_distance_based_active_th;
} SvtMeDeriveInput;

/* Encoder-level inputs of the temporal-filter ME parameters (svt_aom_sig_deriv_me_tf, enc_mode_config.c:829-875;
 * tf_set_me_hme_params_oq :601-681; TF HME flags :1810-1834; set_hme_search_params_mctf level 0 and the FULL_SAD HME
 * of create_me_context_and_picture_control, temporal_filtering.c:2730-2749,315). */
typedef struct SvtMeTfDeriveInput {
    int32_t enc_mode;
    uint32_t width, height;      /* luma size -> EbInputResolution */
    uint32_t qp;                 /* scs->static_config.qp (used when qp_opt) */
    uint8_t hme_me_level;        /* pcs->tf_ctrls.hme_me_level, 0..4 */
    uint8_t qp_opt;              /* pcs->tf_ctrls.qp_opt */
    uint16_t me_exit_th;         /* pcs->tf_ctrls.me_exit_th */
    uint8_t temporal_layer_index, is_ref; /* of the centre picture (temporal_filtering.c:3049-3052) */
    uint8_t _pad[2];
} SvtMeTfDeriveInput;
int svtme_derive_params_tf(const SvtMeTfDeriveInput *in, SvtMeParams *out);

/* Host-side mirror of svt_aom_sig_deriv_me (enc_mode_config.c:684-823), set_hme_search_params
 * (:141-213), set_me_search_params (:218-353), the HME level flags (:1798-1809) and
 * svt_aom_get_enable_me_8x8 (:76-96). Pure host code. */
int svtme_derive_params(const SvtMeDeriveInput *in, SvtMeParams *out);

/*
 * Per-picture outputs, host memory, caller-owned. n_b64 = ceil(aligned_w/64)*ceil(aligned_h/64),
 * n_pu = exported PU count (85 / 21 / 5, see svtme_exported_pu_count()). Layouts are those of
 * MeSbResults (me_sb_results.h:29-50; allocation pcs.c:97-119), one record per b64 back to back,
 * and of the PictureParentControlSet per-b64 arrays (motion_estimation.c:3065-3076,3216-3217).
 * Any pointer may be NULL (that output is skipped).
 */
typedef struct SvtMeOut {
    uint8_t *total_me_candidate_index; /* [n_b64][n_pu] */
    uint32_t *me_mv_array;             /* [n_b64][n_pu*max_refs]  MvCandidate.as_int */
    uint8_t *me_candidate_array;       /* [n_b64][n_pu*max_cand]  MeCandidate bitfield byte */
    uint32_t *me_64x64_distortion;     /* [n_b64] */
    uint32_t *me_32x32_distortion;
    uint32_t *me_16x16_distortion;
    uint32_t *me_8x8_distortion;
    uint32_t *me_8x8_cost_variance;
    uint32_t *rc_me_distortion;
    uint8_t *stationary_block_present_sb;
    uint8_t *rc_me_allow_gm;
    /* optional deep state for parity checks (MeContext.p_sb_best_sad/mv, search_results) */
    uint32_t *p_sb_best_sad;           /* [n_b64][2][4][85], entries of refs not searched = 0 */
    uint32_t *p_sb_best_mv;            /* [n_b64][2][4][85] */
    int16_t *hme_sc;                   /* [n_b64][2][4][2]  (x,y) search centres */
    uint64_t *hme_sad;                 /* [n_b64][2][4]     search_results[].hme_sad at exit */
    uint8_t *do_ref;                   /* [n_b64][2][4] */
} SvtMeOut;

typedef struct SvtMeHandle SvtMeHandle;

/* One handle per GPU (one process per GPU in multi-GPU runs). width/height: luma size of the
 * source pictures; capacity: number of pyramids kept resident (pa-ref ring). */
int svtme_open(int device, uint32_t width, uint32_t height, uint32_t capacity, SvtMeHandle **out);
void svtme_close(SvtMeHandle *h);
const char *svtme_last_error(void);

uint32_t svtme_b64_count(const SvtMeHandle *h);
uint32_t svtme_exported_pu_count(const SvtMeParams *p);

/* Upload one 8-bit luma plane (host memory) and build pad + 1/4 + 1/16 on the device.
 * Replaces svt_aom_pad_input_pictures (pic_analysis_process.c:1981) for luma and
 * svt_aom_downsample_filtering_input_picture (:1922). Asynchronous on the handle's stream. */
int svtme_put_picture(SvtMeHandle *h, uint64_t pic_num, const uint8_t *luma, uint32_t stride);
/* Same, source already in device memory (tightly described by stride). */
int svtme_put_picture_device(SvtMeHandle *h, uint64_t pic_num, const void *d_luma, uint32_t stride);

/* Batched forms: n pictures, two kernel launches in total (decimation, padding). For device input
 * with 16-byte aligned rows the decimation kernel reads the caller's buffers directly. The encoder
 * side is the same loop as above run over a look-ahead window of pictures. */
int svtme_put_pictures(SvtMeHandle *h, uint32_t n, const uint64_t *picture_numbers, const uint8_t *const *luma,
                       uint32_t stride);
int svtme_put_pictures_device(SvtMeHandle *h, uint32_t n, const uint64_t *picture_numbers, const void *const *d_luma,
                              uint32_t stride);
/* 10-bit sources: 16-bit samples, stride in SAMPLES. The ME planes get (uint8_t)(sample >> 2), i.e. the 8-bit buffer
 * svt_aom_un_pack2d -> svt_enc_msb_un_pack2_d (Codec/pic_operators.c:228-244, C_DEFAULT/pack_unpack_c.c:202-218) fills
 * and that input_padded_pic->buffer_y aliases for 10-bit encodes; the conversion is fused into the decimation kernel. */
int svtme_put_pictures_10bit(SvtMeHandle *h, uint32_t n, const uint64_t *picture_numbers, const uint16_t *const *luma16,
                             uint32_t stride_samples);
int svtme_put_pictures_10bit_device(SvtMeHandle *h, uint32_t n, const uint64_t *picture_numbers, const void *const *d_luma16,
                                    uint32_t stride_samples);
/* Copy the padded planes back (level 0 full, 1 quarter, 2 sixteenth) incl. the reference's pads
 * (72/32/16): the CPU stages that still read them (picture statistics, TF) need them. */
int svtme_get_plane(SvtMeHandle *h, uint64_t pic_num, int level, uint8_t *dst, uint32_t dst_stride,
                    uint32_t pad);
int svtme_release_picture(SvtMeHandle *h, uint64_t pic_num);
/* Device pointers of a resident pyramid level (for P2P / NCCL exchange between GPUs). */
int svtme_plane_device_ptr(SvtMeHandle *h, uint64_t pic_num, int level, void **ptr, size_t *bytes);
/* Declare a picture resident whose planes were written externally (NCCL recv) */
int svtme_adopt_picture(SvtMeHandle *h, uint64_t pic_num);
/* Ordering for pyramids that another GPU writes into / reads from ring slots on the CALLER's stream (NCCL send / recv /
 * all-gather over NVLink), without host synchronisation. No reference counterpart: the reference shares pa-reference
 * objects by pointer in host memory (me_process.c:248-262). begin: the pictures get ring slots if they have none and
 * `cuda_stream` (a cudaStream_t) is made to wait for the pyramid builds enqueued so far and for every launch in flight
 * that still reads one of those slots. end: everything enqueued on `cuda_stream` so far completes before any later
 * launch, upload or build of this handle starts. */
int svtme_external_write_begin(SvtMeHandle *h, uint32_t n, const uint64_t *picture_numbers, void *cuda_stream);
int svtme_external_write_end(SvtMeHandle *h, void *cuda_stream);
/* One process driving several GPUs (the reference is one process with a pool of ME threads, enc_handle.c:768-805): copy
 * the pyramid of pic_num from the handle that built it into a ring slot of a handle on another device
 * (cudaMemcpyPeerAsync on the destination's upload stream, ordered after the source's builds and the destination's
 * readers of the slot). Both handles must have been opened for the same picture size. */
int svtme_share_picture(SvtMeHandle *src, SvtMeHandle *dst, uint64_t pic_num);

/* ---- picture statistics of a resident picture --------------------------------------------
 * Replaces svt_aom_gathering_picture_statistics (Codec/pic_analysis_process.c:1560-1578), which the picture-analysis
 * kernel calls right after the pyramid build (:2137) on the same two buffers: the padded luma and the 1/16 plane.
 *   variance  compute_picture_spatial_statistics / compute_block_mean_compute_variance (:307-1378, :1530-1552): per b64
 *             the means and means of squares of the 64 8x8 blocks (svt_compute_interm_var_four8x8, every other row, for
 *             BLOCK_MEAN_PREC_SUB — what enc_handle.c:4257 sets — or svt_compute_mean_8x8 / _square_values_8x8 for
 *             BLOCK_MEAN_PREC_FULL), averaged up to 16x16 / 32x32 / 64x64 with >> 2, variance = (E[x^2] - E[x]^2) >> 16
 *             as uint16_t in raster-scan order (64x64 = 0, 32x32 = 1..4, 16x16 = 5..20, 8x8 = 21..84); an incomplete b64 at
 *             the right / bottom edge reads the padding like the reference does. pic_avg_variance = mean of entry 0.
 *   histogram sub_sample_luma_generate_pixel_intensity_histogram_bins (:1462-1528) on the 1/16 plane: per region 256
 *             bins that start at 1, every decim-th sample of every decim-th row (decim = 1 with scene-change detection
 *             on, else 4), bins scaled by 16 * decim^2, average intensity per region, avg_luma over the picture. */
typedef struct SvtMeStatsParams {
    uint8_t calc_hist;              /* scs->calc_hist */
    uint8_t calculate_variance;     /* scs->calculate_variance (off: pic_avg_variance = 0, variance untouched) */
    uint8_t block_mean_prec_sub;    /* scs->block_mean_calc_prec == BLOCK_MEAN_PREC_SUB (definitions.h:2403-2404) */
    uint8_t all_block_sizes;        /* enable_adaptive_quantization == 1 || variance_octile (:1110): all 85 entries are
                                     * written; otherwise only entry 0 (64x64) */
    uint8_t scene_change_detection; /* static_config.scene_change_detection (:1494) */
    uint8_t _pad[3];
    uint32_t regions_per_width, regions_per_height; /* scs->picture_analysis_number_of_regions_per_width / height, 1..4 */
} SvtMeStatsParams;

typedef struct SvtMeStatsOut {          /* host arrays of the caller; layouts of PictureParentControlSet (pcs.h) */
    uint16_t *variance;                 /* [b64_count][all_block_sizes ? 85 : 1]: pcs->variance (allocation pcs.c:1306-1313) */
    uint16_t *pic_avg_variance;         /* 1 */
    uint32_t *picture_histogram;        /* [4][4][256], index [region x][region y][bin]: pcs->picture_histogram */
    uint64_t *average_intensity_per_region; /* [4][4]: pcs->average_intensity_per_region (the value is a uint8_t) */
    uint64_t *avg_luma;                 /* 1: pcs->avg_luma (INVALID_LUMA = 256 when calc_hist is off, :1563) */
} SvtMeStatsOut;
/* Synchronous: the arrays are filled on return. The picture must be resident (svtme_put_picture*). */
int svtme_picture_statistics(SvtMeHandle *h, uint64_t pic_num, const SvtMeStatsParams *params, SvtMeStatsOut *out);
/* n pictures, one launch sequence, one download (the picture-analysis kernel of the reference handles one picture per
 * call; a look-ahead window's worth is the batch here). */
int svtme_pictures_statistics(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, const SvtMeStatsParams *params, SvtMeStatsOut *outs);
/* Profiling aid: device milliseconds of the kernels and of the result download of the last svtme_picture_statistics, svtme_pictures_statistics or
 * svtme_dg_detector_hme call. */
int svtme_last_statistics_ms(SvtMeHandle *h, float *kernels_ms, float *download_ms);
uint32_t svtme_sizeof_stats_params(void);

/* ---- dynamic-GOP detector HME between two resident pictures ---------------------------------
 * Replaces dg_detector_hme_level0 (Codec/pd_process.c:491-589) over ALL segments of the picture (the reference runs it
 * per ME segment from the ME kernel's TASK_DG_DETECTOR_HME branch and adds into the same metrics under a mutex): per
 * b64 one early_hme_b64 (:389-487) = full-SAD svt_sad_loop_kernel of the 16x16 block of the source's 1/16 picture over a
 * window of 16 x 16 (<= 360p), 64 x 64 (<= 480p) or 128 x 128 positions of `ref_pic`'s 1/16 plane, clipped like HME
 * level 0; metrics as in DGDetectorMetrics (pcs.h:696-702, without seg_completed). input_resolution = the
 * EbInputResolution value of ppcs->input_resolution (0 = 240p range ... 6 = 8K range, definitions.h:2086-2094). */
typedef struct SvtMeDgMetrics {
    uint64_t tot_dist;       /* sum of the best SADs */
    uint32_t tot_cplx;       /* b64 whose best SAD exceeds 16 * 16 * 30 */
    uint32_t tot_active;     /* b64 with a non-zero vector */
    int32_t sum_in_vectors;  /* vectors pointing towards (+) / away from (-) the picture centre, per component */
} SvtMeDgMetrics;
int svtme_dg_detector_hme(SvtMeHandle *h, uint64_t src_pic, uint64_t ref_pic, uint32_t input_resolution, SvtMeDgMetrics *out);

/* Open-loop ME of one picture against refs[list][idx] (picture numbers previously put).
 * Replaces the b64 loop of me_process.c:174-289 calling svt_aom_motion_estimation_b64
 * (motion_estimation.c:3146). Synchronous: results are in `out` on return. */
int svtme_run_picture(SvtMeHandle *h, const SvtMeParams *params, uint64_t cur,
                      const uint64_t refs[SVTME_MAX_LISTS][SVTME_MAX_REFS], SvtMeOut *out);

/* Batched form: n independent pictures, one launch sequence, one D2H. */
typedef struct SvtMeJob {
    SvtMeParams params;
    uint64_t cur;
    uint64_t refs[SVTME_MAX_LISTS][SVTME_MAX_REFS];
    SvtMeOut out;
} SvtMeJob;
int svtme_run_pictures(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n);
/* Device-only part of the batched form (no D2H); svtme_fetch_results completes it. */
int svtme_launch_pictures(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n);
int svtme_fetch_results(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n);
/* Queued (pipelined) form of the batched call, for a look-ahead window processed mini-GOP by mini-GOP:
 * svtme_submit_pictures enqueues the batch and its result download and returns; several batches may be
 * in flight. svtme_collect_results blocks until the OLDEST submitted batch is complete and fills its
 * SvtMeOut arrays (jobs/n must describe that batch). Host pictures given to svtme_put_pictures travel on
 * an upload stream of their own, so with   put(k+1); submit(k+1); collect(k)   the upload of batch k+1,
 * the pipeline of batch k and the download of batch k-1 overlap. No reference counterpart: the
 * reference's ME threads consume pictures one at a time from the picture-decision FIFO
 * (me_process.c:97-124); this is that FIFO at batch granularity. */
int svtme_submit_pictures(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n);
int svtme_collect_results(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n);
uint32_t svtme_pending_batches(SvtMeHandle *h);
/* Page-lock caller memory in place (cudaHostRegister / cudaHostUnregister). When the three big SvtMeOut tables of a job
 * (total_me_candidate_index, me_mv_array, me_candidate_array — the MeSbResults arrays of pcs.c:97-119) are page-locked,
 * the queued form downloads them straight into the caller's arrays: no staging copy, no host-side scatter. Pageable arrays
 * keep working (they are filled from a pinned mirror by svtme_collect_results). */
int svtme_pin_host_memory(void *ptr, size_t bytes);
int svtme_unpin_host_memory(void *ptr);
/* Timeline aid for the queued form (measurement only): mark an origin on the main stream; after each
 * svtme_collect_results, ms3 = {pipeline start, pipeline end, download end} of that batch in ms after the mark. */
int svtme_debug_mark(SvtMeHandle *h);
int svtme_debug_last_span(SvtMeHandle *h, float *ms3);
int svtme_sync(SvtMeHandle *h);
/* CUDA-event time (ms) of the ME kernel launches of the last svtme_launch_pictures call, and
 * the number of kernels this library has launched so far. */
float svtme_last_kernel_ms(SvtMeHandle *h);
uint64_t svtme_launch_count(void);
/* Run the handle's work on a caller-owned CUDA stream (cudaStream_t passed as void*); NULL = private stream. */
int svtme_set_stream(SvtMeHandle *h, void *cuda_stream);
/* Algorithmic SAD byte-ops (search positions x block pixels compared) executed by the stage kernel in
 * the last svtme_launch_pictures call: the numerator of the integer-SAD roofline. Synchronises. */
int svtme_last_work_ops(SvtMeHandle *h, uint64_t *ops);

/* Profiling aid (no reference counterpart): device milliseconds of each launch slot of the stage
 * pipeline in the last launch. Slots: 0 state clear, 1 zz-SAD, 2 pre-HME, 3 HME L0, 4 HME L1,
 * 5 HME L2 + search centre + HME pruning, 6 integer search (check_00_center, probe, full-pel),
 * 7 ME pruning + candidates + statistics. */
#define SVTME_PHASES 8
int svtme_last_phase_ms(SvtMeHandle *h, float *ms, uint32_t n);
/* Algorithmic SAD byte-ops (positions x block pixels compared) per launch slot, same slots. */
int svtme_last_phase_work(SvtMeHandle *h, uint64_t *ops, uint32_t n);
/* Profiling aid (no reference counterpart): a slot holds its search launches (the kernels that compute SADs) AND the
 * one-thread-per-b64 bookkeeping launches between them. With search timing on, every later launch also brackets the
 * search launches of each slot with their own events; svtme_last_search_ms returns their device milliseconds per slot
 * (the duration of the search kernels alone; 0 for a slot without one). Off by default (32 more events per launch). */
int svtme_set_search_timing(SvtMeHandle *h, int on);
int svtme_last_search_ms(SvtMeHandle *h, float *ms, uint32_t n);
/* Measured issue rate (Ginst/s, whole GPU) of VABSDIFF4.U8.ACC, 4 byte-SADs per instruction: the
 * integer-SAD roofline denominator (not part of MEASURED_PEAKS.json). */
double svtme_measure_vabsdiff4_ginst_per_s(void);
/* sizeof() of the ABI structs as compiled into the library (layout check for FFI bindings). */
uint32_t svtme_sizeof_params(void);
uint32_t svtme_sizeof_out(void);
uint32_t svtme_sizeof_job(void);
uint32_t svtme_sizeof_derive_input(void);
uint32_t svtme_sizeof_tf_derive_input(void);

/* ---------------- kernel-test ABI: RTCD signatures (aom_dsp_rtcd.h) ---------------- */

/* svt_sad_loop_kernel  aom_dsp_rtcd.h:776 / compute_sad_c.c:58 */
void svt_sad_loop_kernel_b200(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                              uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                              int16_t *x_search_center, int16_t *y_search_center,
                              uint32_t src_stride_raw, uint8_t skip_search_line,
                              int16_t search_area_width, int16_t search_area_height);
/* svt_ext_all_sad_calculation_8x8_16x16  aom_dsp_rtcd.h:850 / motion_estimation.c:337 */
void svt_ext_all_sad_calculation_8x8_16x16_b200(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                                uint32_t ref_stride, uint32_t mv,
                                                uint32_t *p_best_sad_8x8, uint32_t *p_best_sad_16x16,
                                                uint32_t *p_best_mv8x8, uint32_t *p_best_mv16x16,
                                                uint32_t p_eight_sad16x16[16][8],
                                                uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad);
/* svt_ext_eight_sad_calculation_32x32_64x64  aom_dsp_rtcd.h:851 / motion_estimation.c:369 */
void svt_ext_eight_sad_calculation_32x32_64x64_b200(uint32_t p_sad16x16[16][8],
                                                    uint32_t *p_best_sad_32x32,
                                                    uint32_t *p_best_sad_64x64,
                                                    uint32_t *p_best_mv32x32,
                                                    uint32_t *p_best_mv64x64, uint32_t mv,
                                                    uint32_t p_sad32x32[4][8]);
/* svt_ext_sad_calculation_8x8_16x16  aom_dsp_rtcd.h:839 / motion_estimation.c:98 */
void svt_ext_sad_calculation_8x8_16x16_b200(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                            uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                            uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                            uint32_t *p_best_mv16x16, uint32_t mv,
                                            uint32_t *p_sad16x16, uint32_t *p_sad8x8, uint8_t sub_sad);
/* svt_ext_sad_calculation_32x32_64x64  aom_dsp_rtcd.h:845 / motion_estimation.c:172 */
void svt_ext_sad_calculation_32x32_64x64_b200(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                              uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                              uint32_t *p_best_mv64x64, uint32_t mv,
                                              uint32_t *p_sad32x32);
/* svt_initialize_buffer_32bits  aom_dsp_rtcd.h:852 / me_sad_calculation.c:14 */
void svt_initialize_buffer_32bits_b200(uint32_t *pointer, uint32_t count128, uint32_t count32,
                                       uint32_t value);
/* svt_nxm_sad_kernel  aom_dsp_rtcd.h:854 / compute_sad_c.c:209 */
uint32_t svt_nxm_sad_kernel_b200(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                 uint32_t ref_stride, uint32_t height, uint32_t width);
/* downsample_2d  aom_dsp_rtcd.h:838 / pic_analysis_process.c:131 */
void svt_aom_downsample_2d_b200(uint8_t *input_samples, uint32_t input_stride,
                                uint32_t input_area_width, uint32_t input_area_height,
                                uint8_t *decim_samples, uint32_t decim_stride, uint32_t decim_step);
/* svt_aom_generate_padding  pic_operators.c:338 */
void svt_aom_generate_padding_b200(uint8_t *src_pic, uint32_t src_stride, uint32_t original_src_width,
                                   uint32_t original_src_height, uint32_t padding_width,
                                   uint32_t padding_height);
/* Whole full-pel search of one 64x64 block (open_loop_me_fullpel_search_sblock,
 * motion_estimation.c:781): src = 64x64 block, ref points at the top-left search position.
 * best tables (85 entries, search order) are read-modify-write like p_sb_best_sad/mv. */
void svt_open_loop_me_fullpel_search_sblock_b200(const uint8_t *src, uint32_t src_stride,
                                                 const uint8_t *ref, uint32_t ref_stride,
                                                 int16_t x_search_area_origin,
                                                 int16_t y_search_area_origin,
                                                 uint32_t search_area_width,
                                                 uint32_t search_area_height, uint8_t sub_sad,
                                                 uint32_t *p_sb_best_sad, uint32_t *p_sb_best_mv);

/* Batched form of the above for meaningful timing: n searches of 64x64 blocks in ONE launch, everything in
 * DEVICE memory. Item i searches the block at d_src + src_offset against the window whose top-left search position is
 * d_ref + ref_offset (i.e. MV (x_search_area_origin, y_search_area_origin)), search_area_width x search_area_height
 * positions; d_best_sad / d_best_mv hold n x 85 entries (search order), read-modify-write like p_sb_best_sad / p_sb_best_mv.
 * The reference rows must be readable from 16 bytes before to 8 rows + 256 bytes past the window (padded planes are).
 * Asynchronous on cuda_stream. Returns 0 or a negative SVTME_ERR_*. */
typedef struct SvtMeFullpelItem {
    uint32_t src_offset, ref_offset;              /* bytes from d_src / d_ref */
    int16_t x_search_area_origin, y_search_area_origin;
    uint16_t search_area_width, search_area_height;
} SvtMeFullpelItem;
int svt_open_loop_me_fullpel_search_sblock_b200_batch(void *cuda_stream, uint32_t n, const uint8_t *d_src, uint32_t src_stride,
                                                      const uint8_t *d_ref, uint32_t ref_stride, const SvtMeFullpelItem *d_items,
                                                      uint8_t sub_sad, uint32_t *d_best_sad, uint32_t *d_best_mv);

/* ---------------- allocation-free batched svt_sad_loop_kernel ----------------
 * NO COUNTERPART IN THE REFERENCE; added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). A variant of
 * svt_sad_loop_kernel_b200 above: n_searches independent searches of one geometry, search i using the block at
 * d_src + i * src_step and the window whose top-left search position is d_ref + i * ref_step. All pointers are
 * DEVICE pointers owned by the caller; d_best_sad / d_x_search_center / d_y_search_center are arrays of n_searches
 * elements with the RTCD meaning per element (best_sad = 0xffffff and x / y untouched when no position beat the
 * initial best). The reference rows must be readable from 16 bytes before to 4 rows + 256 bytes past the window,
 * like the wrapper's own staging copy. Asynchronous on `cuda_stream` (a cudaStream_t; NULL = the default stream).
 * The per-search result keys, the kernel's only scratch memory, live in a static device array of
 * SVTME_NOALLOC_MAX_SEARCHES entries (64 KB per GPU, held until the process exits): more searches per call are
 * rejected with SVTME_ERR_INVALID. No cudaMalloc / cudaFree / allocating library call is made. The array is shared
 * by all calls on a device: calls that overlap on the GPU race (serialise them on one stream); nothing in it has
 * to be cleared between calls. svtme_noalloc_setup(device) does the once-per-device set-up (device address of the
 * array, kernel attributes) ahead of time; otherwise the first call on a device does it. */
#define SVTME_NOALLOC_MAX_SEARCHES 8192
int svtme_noalloc_setup(int device);
int svt_sad_loop_kernel_b200_noalloc(void *cuda_stream, uint32_t n_searches, const uint8_t *d_src, uint32_t src_stride,
                                     size_t src_step, const uint8_t *d_ref, uint32_t ref_stride, size_t ref_step,
                                     uint32_t block_height, uint32_t block_width, uint32_t src_stride_raw,
                                     uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height,
                                     uint64_t *d_best_sad, int16_t *d_x_search_center, int16_t *d_y_search_center);

#ifdef __cplusplus
}
#endif
#endif /* SVTME_B200_H */

/*
 * svtme_dump.c — dump / replay hooks for the REAL reference encoder. TEST INFRASTRUCTURE ONLY (see ../me_oracle.h).
 *
 * Not part of the product and not part of the reference: oracle/encoder_dump/build_encoder.sh copies the reference
 * tree to a scratch directory, adds this file to its Codec library and inserts two calls around
 * svt_aom_motion_estimation_b64 in the PAME branch of svt_aom_motion_estimation_kernel
 * (Source/Lib/Codec/me_process.c:266). The patched tree and its binaries live outside the repository
 * (oracle/_ref/, git-ignored); no reference source is copied into the repository.
 *
 *   SVTME_DUMP_DIR=<dir>   per ME picture <dir>/me_%06llu.bin: a header with what the encoder handed to ME
 *                          (SvtMeParams rebuilt from MeContext + pcs/scs, the encoder-level SvtMeDeriveInput, the
 *                          reference picture numbers) followed by one record per b64 with the MeSbResults arrays
 *                          (me_sb_results.h:29-50) and the per-b64 pcs arrays (motion_estimation.c:3065-3076,
 *                          3216-3217); with SVTME_DUMP_DEEP=1 also p_sb_best_sad / p_sb_best_mv / search_results.
 *                          <dir>/luma_%06llu.bin: the 8-bit source luma ME read (input_padded_pic), for the picture
 *                          and for each of its references.
 *   SVTME_REPLAY_DIR=<dir> after the reference's own ME of a b64, overwrite its results with the record of
 *                          <dir>/me_%06llu.bin (same layout, e.g. written from the GPU outputs): the rest of the
 *                          encoder then consumes THOSE results, and the bitstream shows whether they are the same.
 * Records are written with pwrite at fixed offsets: any number of ME threads / segments.
 */
#define _GNU_SOURCE
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "pcs.h"
#include "sequence_control_set.h"
#include "me_context.h"
#include "motion_estimation.h"
#include "svtme_b200.h"
#include "ref_ctx_map.h"

#define DUMP_MAGIC "SVTMEDMP"
#define DUMP_HEADER_BYTES 4096

typedef struct DumpHeader {
    char     magic[8];
    uint32_t version, header_bytes, record_bytes, width, height, b64_cols, b64_rows, n_pu, max_refs, max_cand, max_l0, deep;
    uint32_t input_resolution, me_segments, decode_order, slice_type;
    uint64_t picture_number;
    uint64_t ref_picture_number[2][4];
    SvtMeDeriveInput din;
    SvtMeParams      params;
} DumpHeader;

static const char *g_dump_dir, *g_replay_dir;
static int g_deep, g_init;

static void hook_init(void) {
    if (g_init) return;
    g_dump_dir   = getenv("SVTME_DUMP_DIR");
    g_replay_dir = getenv("SVTME_REPLAY_DIR");
    g_deep       = getenv("SVTME_DUMP_DEEP") != NULL;
    g_init       = 1;
}

static size_t up8(size_t v) { return (v + 7) & ~(size_t)7; }

static size_t record_bytes(uint32_t n_pu, uint32_t max_refs, uint32_t max_cand, int deep) {
    size_t r = up8(n_pu) + up8((size_t)n_pu * max_refs * 4) + up8((size_t)n_pu * max_cand) + up8(6 * 4 + 2);
    if (deep) r += 2 * (size_t)2 * 4 * 85 * 4 + up8(2 * 4 * 2 * 2) + 2 * 4 * 8 + up8(2 * 4);
    return r;
}

static void fill_header(DumpHeader *h, PictureParentControlSet *pcs, MeContext *ctx) {
    SequenceControlSet *scs = pcs->scs;
    memset(h, 0, sizeof(*h));
    memcpy(h->magic, DUMP_MAGIC, 8);
    h->version      = 1;
    h->header_bytes = DUMP_HEADER_BYTES;
    h->width        = pcs->aligned_width;
    h->height       = pcs->aligned_height;
    h->b64_cols     = (pcs->aligned_width + 63) / 64;
    h->b64_rows     = (pcs->aligned_height + 63) / 64;
    h->n_pu         = pcs->enable_me_16x16 ? (pcs->enable_me_8x8 ? pcs->max_number_of_pus_per_sb : 21u) : 5u; /* pcs.c:107-112 */
    h->max_refs = pcs->pa_me_data->max_refs; h->max_cand = pcs->pa_me_data->max_cand; h->max_l0 = pcs->pa_me_data->max_l0;
    h->deep         = (uint32_t)g_deep;
    h->record_bytes = (uint32_t)record_bytes(h->n_pu, h->max_refs, h->max_cand, g_deep);
    h->input_resolution = scs->input_resolution;
    h->me_segments  = pcs->me_segments_total_count;
    h->decode_order = (uint32_t)pcs->decode_order;
    h->slice_type   = pcs->slice_type;
    h->picture_number = pcs->picture_number;
    SvtMeParams *p = &h->params;
    ctx_to_params(ctx, p);
    p->num_of_list_to_search       = ctx->num_of_list_to_search;
    p->num_of_ref_pic_to_search[0] = ctx->num_of_ref_pic_to_search[0];
    p->num_of_ref_pic_to_search[1] = ctx->num_of_list_to_search > 1 ? ctx->num_of_ref_pic_to_search[1] : 0;
    p->temporal_layer_index = ctx->temporal_layer_index;
    p->is_ref               = ctx->is_ref;
    p->hierarchical_levels  = pcs->hierarchical_levels;
    p->similar_brightness_refs = pcs->similar_brightness_refs;
    p->enable_me_8x8 = pcs->enable_me_8x8; p->enable_me_16x16 = pcs->enable_me_16x16;
    p->max_number_of_pus_per_sb = pcs->max_number_of_pus_per_sb;
    p->only_l_bwd = scs->mrp_ctrls.only_l_bwd;
    p->gm_enabled = pcs->gm_ctrls.enabled;
    p->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    p->input_resolution_le_480p = scs->input_resolution <= INPUT_SIZE_480p_RANGE;
    p->max_cand = pcs->pa_me_data->max_cand; p->max_refs = pcs->pa_me_data->max_refs; p->max_l0 = pcs->pa_me_data->max_l0;
    p->picture_number = pcs->picture_number;
    for (int li = 0; li < p->num_of_list_to_search; ++li)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri)
            h->ref_picture_number[li][ri] = p->ref_picture_number[li][ri] = ctx->me_ds_ref_array[li][ri].picture_number;
    SvtMeDeriveInput *d = &h->din;
    d->enc_mode = pcs->enc_mode;
    d->width = scs->max_input_luma_width; d->height = scs->max_input_luma_height;
    d->qp = scs->static_config.qp;
    d->rtc_tune = scs->static_config.pred_structure == SVT_AV1_PRED_LOW_DELAY_B;
    d->sc_class1 = pcs->sc_class1;
    d->hierarchical_levels = pcs->hierarchical_levels;
    d->temporal_layer_index = pcs->temporal_layer_index;
    d->is_ref = pcs->is_ref;
    d->slice_type_b = pcs->slice_type == B_SLICE;
    d->ref_list0_count_try = pcs->ref_list0_count_try;
    d->ref_list1_count_try = pcs->slice_type == B_SLICE ? pcs->ref_list1_count_try : 0;
    d->frame_rate_q16 = scs->frame_rate;
    d->safe_limit_nref = scs->mrp_ctrls.safe_limit_nref;
    d->only_l_bwd = scs->mrp_ctrls.only_l_bwd;
    d->use_best_me_unipred_cand_only = pcs->use_best_me_unipred_cand_only;
    d->similar_brightness_refs = pcs->similar_brightness_refs;
    d->safe_limit_zz_th = scs->mrp_ctrls.safe_limit_zz_th;
    d->ref_count_used_list0 = pcs->pa_me_data->max_l0;
    d->ref_count_used_list1 = (uint8_t)(pcs->pa_me_data->max_refs - pcs->pa_me_data->max_l0);
    d->gm_level_enabled = pcs->gm_ctrls.enabled;
    d->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
}

static void write_luma(const char *dir, uint64_t pic, const EbPictureBufferDesc *pl, uint32_t W, uint32_t H) {
    char path[1024];
    snprintf(path, sizeof(path), "%s/luma_%06llu.bin", dir, (unsigned long long)pic);
    if (access(path, F_OK) == 0) return;
    char tmp[1100];
    snprintf(tmp, sizeof(tmp), "%s.%d.tmp", path, (int)getpid());
    FILE *f = fopen(tmp, "wb");
    if (!f) return;
    for (uint32_t y = 0; y < H; ++y) fwrite(pl->buffer_y + (size_t)(pl->org_y + y) * pl->stride_y + pl->org_x, 1, W, f);
    fclose(f);
    rename(tmp, path);
}

/* Before svt_aom_motion_estimation_b64 (dump mode with deep state only): clear the best tables, so that entries of
 * references the block does not search are comparable. They are never read before being initialised. */
void svtme_hook_pre_b64(MeContext *ctx) {
    hook_init();
    if (g_dump_dir && g_deep) memset(ctx->p_sb_best_sad, 0, sizeof(ctx->p_sb_best_sad));
}

void svtme_hook_post_b64(PictureParentControlSet *pcs, uint32_t b64, MeContext *ctx, EbPictureBufferDesc *input_padded_pic) {
    hook_init();
    if (!g_dump_dir && !g_replay_dir) return;
    DumpHeader h;
    fill_header(&h, pcs, ctx);
    const uint32_t n_pu = h.n_pu, max_refs = h.max_refs, max_cand = h.max_cand;
    MeSbResults *r = pcs->pa_me_data->me_results[b64];
    char path[1024];
    if (g_dump_dir) {
        snprintf(path, sizeof(path), "%s/me_%06llu.bin", g_dump_dir, (unsigned long long)pcs->picture_number);
        const int fd = open(path, O_CREAT | O_WRONLY, 0644);
        if (fd < 0) { perror(path); abort(); }
        uint8_t *rec = (uint8_t *)calloc(1, h.record_bytes);
        size_t o = 0;
        memcpy(rec + o, r->total_me_candidate_index, n_pu); o += up8(n_pu);
        memcpy(rec + o, r->me_mv_array, (size_t)n_pu * max_refs * 4); o += up8((size_t)n_pu * max_refs * 4);
        memcpy(rec + o, r->me_candidate_array, (size_t)n_pu * max_cand); o += up8((size_t)n_pu * max_cand);
        uint32_t six[6] = {pcs->me_64x64_distortion[b64], pcs->me_32x32_distortion[b64], pcs->me_16x16_distortion[b64],
                           pcs->me_8x8_distortion[b64], pcs->me_8x8_cost_variance[b64], pcs->rc_me_distortion[b64]};
        memcpy(rec + o, six, 24);
        rec[o + 24] = pcs->stationary_block_present_sb[b64];
        rec[o + 25] = pcs->rc_me_allow_gm[b64];
        o += up8(26);
        if (g_deep) {
            memcpy(rec + o, ctx->p_sb_best_sad, 2 * 4 * 85 * 4); o += 2 * 4 * 85 * 4;
            memcpy(rec + o, ctx->p_sb_best_mv, 2 * 4 * 85 * 4); o += 2 * 4 * 85 * 4;
            int16_t  sc[2][4][2];
            uint64_t hs[2][4];
            uint8_t  dr[2][4];
            for (int li = 0; li < 2; ++li)
                for (int ri = 0; ri < 4; ++ri) {
                    sc[li][ri][0] = ctx->search_results[li][ri].hme_sc_x; sc[li][ri][1] = ctx->search_results[li][ri].hme_sc_y;
                    hs[li][ri] = ctx->search_results[li][ri].hme_sad; dr[li][ri] = ctx->search_results[li][ri].do_ref;
                }
            memcpy(rec + o, sc, sizeof(sc)); o += up8(sizeof(sc));
            memcpy(rec + o, hs, sizeof(hs)); o += sizeof(hs);
            memcpy(rec + o, dr, sizeof(dr)); o += up8(sizeof(dr));
        }
        uint8_t hdr[DUMP_HEADER_BYTES] = {0};
        memcpy(hdr, &h, sizeof(h));
        if (pwrite(fd, hdr, sizeof(hdr), 0) != (ssize_t)sizeof(hdr) ||
            pwrite(fd, rec, h.record_bytes, (off_t)DUMP_HEADER_BYTES + (off_t)b64 * h.record_bytes) != (ssize_t)h.record_bytes) {
            perror("svtme_dump pwrite"); abort();
        }
        close(fd);
        free(rec);
        if (b64 == 0) {
            write_luma(g_dump_dir, pcs->picture_number, input_padded_pic, h.width, h.height);
            for (int li = 0; li < h.params.num_of_list_to_search; ++li)
                for (int ri = 0; ri < h.params.num_of_ref_pic_to_search[li]; ++ri)
                    write_luma(g_dump_dir, h.ref_picture_number[li][ri], ctx->me_ds_ref_array[li][ri].picture_ptr, h.width, h.height);
        }
    }
    if (g_replay_dir) {
        snprintf(path, sizeof(path), "%s/me_%06llu.bin", g_replay_dir, (unsigned long long)pcs->picture_number);
        const int fd = open(path, O_RDONLY);
        if (fd < 0) { perror(path); abort(); }
        uint8_t hdr[DUMP_HEADER_BYTES];
        DumpHeader fh;
        if (pread(fd, hdr, sizeof(hdr), 0) != (ssize_t)sizeof(hdr)) { perror("svtme replay header"); abort(); }
        memcpy(&fh, hdr, sizeof(fh));
        if (memcmp(fh.magic, DUMP_MAGIC, 8) || fh.n_pu != n_pu || fh.max_refs != max_refs || fh.max_cand != max_cand ||
            fh.width != h.width || fh.height != h.height) {
            fprintf(stderr, "svtme replay: %s does not describe this picture\n", path); abort();
        }
        uint8_t *rec = (uint8_t *)malloc(fh.record_bytes);
        if (pread(fd, rec, fh.record_bytes, (off_t)DUMP_HEADER_BYTES + (off_t)b64 * fh.record_bytes) != (ssize_t)fh.record_bytes) {
            perror("svtme replay record"); abort();
        }
        close(fd);
        size_t o = 0;
        memcpy(r->total_me_candidate_index, rec + o, n_pu); o += up8(n_pu);
        memcpy(r->me_mv_array, rec + o, (size_t)n_pu * max_refs * 4); o += up8((size_t)n_pu * max_refs * 4);
        memcpy(r->me_candidate_array, rec + o, (size_t)n_pu * max_cand); o += up8((size_t)n_pu * max_cand);
        uint32_t six[6];
        memcpy(six, rec + o, 24);
        pcs->me_64x64_distortion[b64] = six[0]; pcs->me_32x32_distortion[b64] = six[1]; pcs->me_16x16_distortion[b64] = six[2];
        pcs->me_8x8_distortion[b64] = six[3]; pcs->me_8x8_cost_variance[b64] = six[4]; pcs->rc_me_distortion[b64] = six[5];
        pcs->stationary_block_present_sb[b64] = rec[o + 24];
        pcs->rc_me_allow_gm[b64]              = rec[o + 25];
        free(rec);
    }
}

/* ---- picture statistics and dynamic-GOP detector: what the real encoder computed, for tests/golden/encdump_aux_*.npz ----
 * svtme_hook_post_statistics: called right after svt_aom_gathering_picture_statistics (pic_analysis_process.c:2137).
 *   <dir>/stats_%06llu.bin = 16 uint32 {magic 'SVST', width, height, n_b64, n_var, calc_hist, calculate_variance,
 *   block_mean_prec_sub, all_block_sizes, scene_change_detection, regions_w, regions_h, pic_avg_variance, 0, 0, 0},
 *   uint64 avg_luma, uint64 average_intensity[4][4], uint32 histogram[4][4][256], uint16 variance[n_b64][n_var].
 * svtme_hook_post_dg: called inside dg_detector_hme_level0 (pd_process.c:584-586) when the last segment is done, under the
 *   detector's mutex, right before the frame is signalled done:
 *   <dir>/dg_%06llu_%06llu.bin (source, reference) = uint64 tot_dist, uint32 tot_cplx, tot_active, int32 sum_in_vectors,
 *   uint32 input_resolution. */
void svtme_hook_post_statistics(SequenceControlSet *scs, PictureParentControlSet *pcs) {
    hook_init();
    if (!g_dump_dir) return;
    char path[600];
    snprintf(path, sizeof(path), "%s/stats_%06llu.bin", g_dump_dir, (unsigned long long)pcs->picture_number);
    FILE *f = fopen(path, "wb");
    if (!f) { perror("svtme_dump stats"); abort(); }
    const uint32_t all = scs->static_config.enable_adaptive_quantization == 1 || scs->static_config.variance_octile;
    const uint32_t n_var = all ? 85 : 1;
    const uint32_t hd[16] = {0x54535653u, pcs->aligned_width, pcs->aligned_height, pcs->b64_total_count, n_var, scs->calc_hist,
                             scs->calculate_variance, scs->block_mean_calc_prec == BLOCK_MEAN_PREC_SUB, all,
                             scs->static_config.scene_change_detection != 0, scs->picture_analysis_number_of_regions_per_width,
                             scs->picture_analysis_number_of_regions_per_height, pcs->pic_avg_variance, 0, 0, 0};
    fwrite(hd, sizeof(hd), 1, f);
    const uint64_t luma = pcs->avg_luma;
    fwrite(&luma, 8, 1, f);
    uint64_t inten[16];
    static uint32_t hist[16 * 256];
    memset(inten, 0, sizeof(inten));
    memset(hist, 0, sizeof(hist));
    if (scs->calc_hist)
        for (uint32_t x = 0; x < scs->picture_analysis_number_of_regions_per_width; ++x)
            for (uint32_t y = 0; y < scs->picture_analysis_number_of_regions_per_height; ++y) {
                inten[x * 4 + y] = pcs->average_intensity_per_region[x][y];
                memcpy(hist + (x * 4 + y) * 256, pcs->picture_histogram[x][y], 256 * sizeof(uint32_t));
            }
    fwrite(inten, sizeof(inten), 1, f);
    fwrite(hist, sizeof(hist), 1, f);
    if (scs->calculate_variance)
        for (uint32_t b = 0; b < pcs->b64_total_count; ++b) fwrite(pcs->variance[b], sizeof(uint16_t), n_var, f);
    fclose(f);
}

void svtme_hook_post_dg(PictureParentControlSet *pcs) {
    hook_init();
    if (!g_dump_dir || !pcs->dg_detector || !pcs->dg_detector->ref_pic) return;
    const DGDetectorMetrics *m = &pcs->dg_detector->metrics;
    if (m->seg_completed != (uint16_t)(pcs->me_segments_column_count * pcs->me_segments_row_count)) return;
    char path[600];
    snprintf(path, sizeof(path), "%s/dg_%06llu_%06llu.bin", g_dump_dir, (unsigned long long)pcs->picture_number,
             (unsigned long long)pcs->dg_detector->ref_pic->picture_number);
    FILE *f = fopen(path, "wb");
    if (!f) { perror("svtme_dump dg"); abort(); }
    const uint64_t dist = m->tot_dist;
    const uint32_t rest[4] = {m->tot_cplx, m->tot_active, (uint32_t)m->sum_in_vectors, (uint32_t)pcs->input_resolution};
    fwrite(&dist, 8, 1, f);
    fwrite(rest, sizeof(rest), 1, f);
    fclose(f);
}

/* ---- temporal-filter ME (ME_MCTF): what the real encoder hands to svt_aom_motion_estimation_b64 from
 * produce_temporally_filtered_pic (temporal_filtering.c:3046-3080) and what it gets back ----
 * svtme_hook_post_tf_b64: called right after that call. <dir>/tf_%06llu_%06llu.bin (centre picture, neighbouring frame):
 *   a 4096-byte header {magic 'SVTMETF1', uint32 width, height, b64_cols, b64_rows, record_bytes, 0; uint64 centre, frame;
 *   SvtMeTfDeriveInput as the encoder's pcs / scs had it; SvtMeParams rebuilt from the MeContext}, then one 696-byte record
 *   per block: p_sb_best_sad[0][0][85], p_sb_best_mv[0][0][85], hme_sc_x, hme_sc_y (int16), 4 pad bytes, hme_sad (uint64).
 *   <dir>/tfluma_%06llu_%06llu_{c,f}.bin: the two luma planes the call read (a neighbouring frame that was itself filtered
 *   earlier is no longer the source picture; the fixture generator keeps only pairs of untouched frames). */
typedef struct TfDumpHeader {
    char               magic[8];
    uint32_t           width, height, b64_cols, b64_rows, record_bytes, reserved;
    uint64_t           centre, frame;
    SvtMeTfDeriveInput tin;
    SvtMeParams        params;
} TfDumpHeader;
#define TF_RECORD_BYTES (85 * 4 * 2 + 8 + 8)

static void write_plane(const char *path, const EbPictureBufferDesc *pl, uint32_t W, uint32_t H) {
    if (access(path, F_OK) == 0) return;
    FILE *f = fopen(path, "wb");
    if (!f) return;
    for (uint32_t y = 0; y < H; ++y) fwrite(pl->buffer_y + (size_t)(pl->org_y + y) * pl->stride_y + pl->org_x, 1, W, f);
    fclose(f);
}

void svtme_hook_post_tf_b64(PictureParentControlSet *centre, uint32_t b64, MeContext *ctx, EbPictureBufferDesc *central) {
    hook_init();
    if (!g_dump_dir) return;
    SequenceControlSet *scs = centre->scs;
    const uint64_t      frame = ctx->me_ds_ref_array[0][0].picture_number;
    TfDumpHeader        h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, "SVTMETF1", 8);
    h.width = centre->aligned_width; h.height = centre->aligned_height;
    h.b64_cols = (centre->aligned_width + 63) / 64; h.b64_rows = (centre->aligned_height + 63) / 64;
    h.record_bytes = TF_RECORD_BYTES;
    h.centre = centre->picture_number; h.frame = frame;
    h.tin.enc_mode = centre->enc_mode;
    h.tin.width = scs->max_input_luma_width; h.tin.height = scs->max_input_luma_height;
    h.tin.qp = scs->static_config.qp;
    h.tin.hme_me_level = centre->tf_ctrls.hme_me_level; h.tin.qp_opt = centre->tf_ctrls.qp_opt;
    h.tin.me_exit_th = centre->tf_ctrls.me_exit_th;
    h.tin.temporal_layer_index = centre->temporal_layer_index; h.tin.is_ref = centre->is_ref;
    SvtMeParams *p = &h.params;
    ctx_to_params(ctx, p);
    p->num_of_list_to_search = ctx->num_of_list_to_search;
    p->num_of_ref_pic_to_search[0] = ctx->num_of_ref_pic_to_search[0];
    p->num_of_ref_pic_to_search[1] = 0;
    p->temporal_layer_index = ctx->temporal_layer_index;
    p->is_ref = ctx->is_ref;
    p->picture_number = centre->picture_number;
    p->ref_picture_number[0][0] = frame;
    char path[1024];
    snprintf(path, sizeof(path), "%s/tf_%06llu_%06llu.bin", g_dump_dir, (unsigned long long)h.centre, (unsigned long long)frame);
    const int fd = open(path, O_CREAT | O_WRONLY, 0644);
    if (fd < 0) { perror(path); abort(); }
    uint8_t rec[TF_RECORD_BYTES];
    memset(rec, 0, sizeof(rec));
    memcpy(rec, ctx->p_sb_best_sad[0][0], 85 * 4);
    memcpy(rec + 340, ctx->p_sb_best_mv[0][0], 85 * 4);
    const int16_t sc[2] = {ctx->search_results[0][0].hme_sc_x, ctx->search_results[0][0].hme_sc_y};
    memcpy(rec + 680, sc, 4);
    const uint64_t hs = ctx->search_results[0][0].hme_sad;
    memcpy(rec + 688, &hs, 8);
    uint8_t hdr[DUMP_HEADER_BYTES] = {0};
    memcpy(hdr, &h, sizeof(h));
    if (pwrite(fd, hdr, sizeof(hdr), 0) != (ssize_t)sizeof(hdr) ||
        pwrite(fd, rec, sizeof(rec), (off_t)DUMP_HEADER_BYTES + (off_t)b64 * TF_RECORD_BYTES) != (ssize_t)sizeof(rec)) {
        perror("svtme_dump tf pwrite"); abort();
    }
    close(fd);
    if (b64 == 0) {
        snprintf(path, sizeof(path), "%s/tfluma_%06llu_%06llu_c.bin", g_dump_dir, (unsigned long long)h.centre, (unsigned long long)frame);
        write_plane(path, central, h.width, h.height);
        snprintf(path, sizeof(path), "%s/tfluma_%06llu_%06llu_f.bin", g_dump_dir, (unsigned long long)h.centre, (unsigned long long)frame);
        write_plane(path, ctx->me_ds_ref_array[0][0].picture_ptr, h.width, h.height);
    }
}

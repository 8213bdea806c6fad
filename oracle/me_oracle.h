/*
 * me_oracle.h — CPU restatement (plain C) of the reference's open-loop ME path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build, load or
 * call it, and only as the checker / reported CPU baseline. The product (svtme_b200) never links
 * or calls this code and has no CPU fallback.
 *
 * Parity status: PINNED. Every function here is checked bit-for-bit against the reference's own
 * C sources compiled from /root/reference (oracle/Makefile -> oracle/_ref/libsvtref.so) by
 * tests/test_oracle_vs_ref.py, and against golden vectors generated from that build
 * (tests/golden/, generator tests/golden/make_golden.py). The reference stores no golden vectors
 * of its own for this path: its tests are C-vs-SIMD differential tests.
 *
 * All file:line citations are into /root/reference/Source/Lib/.
 */
#ifndef ME_ORACLE_H
#define ME_ORACLE_H

#include <stdint.h>
#include "svtme_b200.h" /* SvtMeParams / SvtMeOut type definitions only */

#ifdef __cplusplus
extern "C" {
#endif

/* One padded 8-bit plane: EbPictureBufferDesc subset (Codec/pic_buffer_desc.h). */
typedef struct OrcPlane {
    uint8_t *buffer_y;
    uint32_t stride_y;
    uint32_t org_x, org_y;
    uint32_t width, height;
} OrcPlane;

/* EbPaReferenceObject subset (Codec/reference_object.h:73-89): full, quarter, sixteenth. */
typedef struct OrcPyramid {
    OrcPlane full, quarter, sixteenth;
    uint64_t picture_number;
} OrcPyramid;

/* ---- kernels (RTCD signatures) ---- */
void orc_sad_loop_kernel(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                         uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                         int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                         uint8_t skip_search_line, int16_t search_area_width,
                         int16_t search_area_height);
uint32_t orc_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                            uint32_t ref_stride, uint32_t height, uint32_t width);
void orc_ext_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                       uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                       uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                       uint32_t *p_best_mv16x16, uint32_t mv, uint32_t *p_sad16x16,
                                       uint32_t *p_sad8x8, uint8_t sub_sad);
void orc_ext_sad_calculation_32x32_64x64(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                         uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                         uint32_t *p_best_mv64x64, uint32_t mv, uint32_t *p_sad32x32);
void orc_ext_all_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                           uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                           uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                           uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                           uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad);
void orc_ext_eight_sad_calculation_32x32_64x64(uint32_t p_sad16x16[16][8], uint32_t *p_best_sad_32x32,
                                               uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                               uint32_t *p_best_mv64x64, uint32_t mv,
                                               uint32_t p_sad32x32[4][8]);
void orc_initialize_buffer_32bits(uint32_t *pointer, uint32_t count128, uint32_t count32,
                                  uint32_t value);
void orc_downsample_2d(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                       uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                       uint32_t decim_step);
void orc_generate_padding(uint8_t *src_pic, uint32_t src_stride, uint32_t original_src_width,
                          uint32_t original_src_height, uint32_t padding_width,
                          uint32_t padding_height);
/* open_loop_me_fullpel_search_sblock over caller-owned 85-entry tables (search order). */
void orc_fullpel_search_sblock(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                               uint32_t ref_stride, int16_t x_search_area_origin,
                               int16_t y_search_area_origin, uint32_t search_area_width,
                               uint32_t search_area_height, uint8_t sub_sad, uint32_t *p_sb_best_sad,
                               uint32_t *p_sb_best_mv);

/* ---- pyramid ---- */
/* Allocate the three planes with the reference geometry (pads 72/32/16, stride = w + 2*pad),
 * copy the w x h luma in, pad, decimate, pad. hme_level1 selects the 2-step (1) or direct
 * step-4 (0) sixteenth path (pic_analysis_process.c:1949-1969). Returns 0, -1 on alloc failure. */
int orc_pyramid_build(OrcPyramid *pyr, const uint8_t *luma, uint32_t stride, uint32_t width,
                      uint32_t height, int hme_level1, uint64_t picture_number);
void orc_pyramid_free(OrcPyramid *pyr);

/* ---- stage ---- */
/* svt_aom_motion_estimation_b64 (motion_estimation.c:3146) for every b64 of the picture, in
 * raster order, single thread. refs[list][idx] must be valid for idx < num_of_ref_pic_to_search. */
int orc_me_picture(const SvtMeParams *p, const OrcPyramid *cur,
                   const OrcPyramid *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS], SvtMeOut *out);
/* Same, b64 rows [row0,row1) only (for multi-threaded CPU baselines). */
int orc_me_picture_rows(const SvtMeParams *p, const OrcPyramid *cur,
                        const OrcPyramid *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS], SvtMeOut *out,
                        uint32_t row0, uint32_t row1);

/* ---- picture statistics (me_oracle_stats.c) ---- */
/* svt_aom_gathering_picture_statistics (pic_analysis_process.c:1560) on a built pyramid: variance of the padded luma,
 * histogram / intensities of the 1/16 plane. `out` arrays as in svtme_picture_statistics. */
int orc_picture_statistics(const OrcPyramid *pyr, const SvtMeStatsParams *prm, SvtMeStatsOut *out);
/* dg_detector_hme_level0 (pd_process.c:491) over all segments: src's 16x16 blocks searched in ref's 1/16 plane. */
int orc_dg_detector_hme(const OrcPyramid *src, const OrcPyramid *ref, uint32_t input_resolution, SvtMeDgMetrics *out);

#ifdef __cplusplus
}
#endif
#endif

/*
 * me_oracle_kernels.c — CPU restatement of the reference's ME kernels and pyramid (plain C).
 * TEST INFRASTRUCTURE ONLY (see me_oracle.h). Citations: /root/reference/Source/Lib/.
 */
#include <stdlib.h>
#include <string.h>
#include "me_oracle.h"

static inline uint32_t absdiff_u8(uint8_t a, uint8_t b) { return a > b ? (uint32_t)(a - b) : (uint32_t)(b - a); }

/* Sum of |src - ref| over rows x cols. svt_fast_loop_nxm_sad_kernel, C_DEFAULT/compute_sad_c.c:20-37;
 * svt_nxm_sad_kernel_helper_c :209-212. */
uint32_t orc_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                            uint32_t ref_stride, uint32_t height, uint32_t width) {
    uint32_t acc = 0;
    for (uint32_t r = 0; r < height; ++r)
        for (uint32_t c = 0; c < width; ++c)
            acc += absdiff_u8(src[r * src_stride + c], ref[r * ref_stride + c]);
    return acc;
}

/* Exhaustive search, first strict minimum in (y, x) raster order; best starts at 0xffffff and
 * x/y stay untouched if nothing beats it. skip_search_line drops EVEN search rows, only for
 * 16-wide blocks of height <= 16. C_DEFAULT/compute_sad_c.c:58-101. */
void orc_sad_loop_kernel(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                         uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                         int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                         uint8_t skip_search_line, int16_t search_area_width,
                         int16_t search_area_height) {
    const int skip_even = (block_width == 16 && block_height <= 16 && skip_search_line);
    uint64_t best = 0xffffff;
    for (int16_t sy = 0; sy < search_area_height; ++sy) {
        const uint8_t *row0 = ref + (size_t)sy * src_stride_raw;
        if (skip_even && (sy & 1) == 0)
            continue;
        for (int16_t sx = 0; sx < search_area_width; ++sx) {
            uint32_t sad = orc_nxm_sad_kernel(src, src_stride, row0 + sx, ref_stride, block_height, block_width);
            if (sad < best) {
                best             = sad;
                *x_search_center = sx;
                *y_search_center = sy;
            }
        }
    }
    *best_sad = best;
}

/* 8 wide, `rows` rows. svt_aom_compute8x4_sad_kernel_c / compute8x8_sad_kernel_c,
 * Codec/motion_estimation.c:41-93. */
static uint32_t sad_8xn(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                        int rows) {
    return orc_nxm_sad_kernel(src, src_stride, ref, ref_stride, (uint32_t)rows, 8);
}

/* The four 8x8 SADs of one 16x16 block at one position; sub_sad = rows 0,2,4,6 doubled.
 * Codec/motion_estimation.c:105-135 (single) and :224-330 (eight). */
static void four_8x8_sads(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                          int sub_sad, uint32_t out[4]) {
    for (int k = 0; k < 4; ++k) {
        const uint8_t *s = src + (k >> 1) * 8 * src_stride + (k & 1) * 8;
        const uint8_t *r = ref + (k >> 1) * 8 * ref_stride + (k & 1) * 8;
        out[k] = sub_sad ? (sad_8xn(s, 2 * src_stride, r, 2 * ref_stride, 4) << 1)
                         : sad_8xn(s, src_stride, r, ref_stride, 8);
    }
}

/* Codec/motion_estimation.c:98-165 */
void orc_ext_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                       uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                       uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                       uint32_t *p_best_mv16x16, uint32_t mv, uint32_t *p_sad16x16,
                                       uint32_t *p_sad8x8, uint8_t sub_sad) {
    four_8x8_sads(src, src_stride, ref, ref_stride, sub_sad, p_sad8x8);
    uint32_t sum = 0;
    for (int k = 0; k < 4; ++k) {
        if (p_sad8x8[k] < p_best_sad_8x8[k]) {
            p_best_sad_8x8[k] = p_sad8x8[k];
            p_best_mv8x8[k]   = mv;
        }
        sum += p_sad8x8[k];
    }
    if (sum < p_best_sad_16x16[0]) {
        p_best_sad_16x16[0] = sum;
        p_best_mv16x16[0]   = mv;
    }
    *p_sad16x16 = sum;
}

/* Codec/motion_estimation.c:172-205 */
void orc_ext_sad_calculation_32x32_64x64(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                         uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                         uint32_t *p_best_mv64x64, uint32_t mv, uint32_t *p_sad32x32) {
    uint32_t total = 0;
    for (int q = 0; q < 4; ++q) {
        uint32_t s = p_sad16x16[4 * q] + p_sad16x16[4 * q + 1] + p_sad16x16[4 * q + 2] + p_sad16x16[4 * q + 3];
        p_sad32x32[q] = s;
        if (s < p_best_sad_32x32[q]) {
            p_best_sad_32x32[q] = s;
            p_best_mv32x32[q]   = mv;
        }
        total += s;
    }
    if (total < p_best_sad_64x64[0]) {
        p_best_sad_64x64[0] = total;
        p_best_mv64x64[0]   = mv;
    }
}

static inline uint32_t mv_plus_x(uint32_t mv, int k) {
    /* _MVXT/_MVYT unpack, x += k, repack. Codec/motion_estimation.c:231-235 */
    int16_t x = (int16_t)(mv & 0xffff);
    int16_t y = (int16_t)(mv >> 16);
    x         = (int16_t)(x + k);
    return ((uint32_t)(uint16_t)y << 16) | (uint16_t)x;
}

/* 16 blocks in raster order, each for 8 consecutive x positions; block -> z-order slot.
 * Codec/motion_estimation.c:210-362. p_eight_sad8x8 is unused by the reference too. */
void orc_ext_all_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                           uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                           uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                           uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                           uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad) {
    (void)p_eight_sad8x8;
    for (int by = 0; by < 4; ++by) {
        for (int bx = 0; bx < 4; ++bx) {
            const int slot = (by >> 1) * 8 + (bx >> 1) * 4 + (by & 1) * 2 + (bx & 1);
            const uint8_t *s = src + 16 * by * src_stride + 16 * bx;
            const uint8_t *r = ref + 16 * by * ref_stride + 16 * bx;
            for (int k = 0; k < 8; ++k) {
                uint32_t sad8[4];
                four_8x8_sads(s, src_stride, r + k, ref_stride, sub_sad, sad8);
                uint32_t sum = 0;
                for (int q = 0; q < 4; ++q) {
                    if (sad8[q] < p_best_sad_8x8[4 * slot + q]) {
                        p_best_sad_8x8[4 * slot + q] = sad8[q];
                        p_best_mv8x8[4 * slot + q]   = mv_plus_x(mv, k);
                    }
                    sum += sad8[q];
                }
                p_eight_sad16x16[slot][k] = sum;
                if (sum < p_best_sad_16x16[slot]) {
                    p_best_sad_16x16[slot] = sum;
                    p_best_mv16x16[slot]   = mv_plus_x(mv, k);
                }
            }
        }
    }
}

/* Codec/motion_estimation.c:369-425 */
void orc_ext_eight_sad_calculation_32x32_64x64(uint32_t p_sad16x16[16][8], uint32_t *p_best_sad_32x32,
                                               uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                               uint32_t *p_best_mv64x64, uint32_t mv,
                                               uint32_t p_sad32x32[4][8]) {
    for (int k = 0; k < 8; ++k) {
        uint32_t total = 0;
        for (int q = 0; q < 4; ++q) {
            uint32_t s = p_sad16x16[4 * q][k] + p_sad16x16[4 * q + 1][k] + p_sad16x16[4 * q + 2][k] +
                p_sad16x16[4 * q + 3][k];
            p_sad32x32[q][k] = s;
            if (s < p_best_sad_32x32[q]) {
                p_best_sad_32x32[q] = s;
                p_best_mv32x32[q]   = mv_plus_x(mv, k);
            }
            total += s;
        }
        if (total < p_best_sad_64x64[0]) {
            p_best_sad_64x64[0] = total;
            p_best_mv64x64[0]   = mv_plus_x(mv, k);
        }
    }
}

/* Codec/me_sad_calculation.c:14-17 */
void orc_initialize_buffer_32bits(uint32_t *pointer, uint32_t count128, uint32_t count32, uint32_t value) {
    for (uint32_t i = 0; i < count128 * 4 + count32; ++i) pointer[i] = value;
}

/* open_loop_me_fullpel_search_sblock + the two per-position helpers,
 * Codec/motion_estimation.c:429-474 (8 positions), :476-776 (1 position), :781-817 (loop nest).
 * Table layout (search order): [0]=64x64, [1..4]=32x32, [5..20]=16x16 z-order, [21..84]=8x8. */
void orc_fullpel_search_sblock(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                               uint32_t ref_stride, int16_t x_search_area_origin,
                               int16_t y_search_area_origin, uint32_t search_area_width,
                               uint32_t search_area_height, uint8_t sub_sad, uint32_t *p_sb_best_sad,
                               uint32_t *p_sb_best_mv) {
    uint32_t *bs64 = p_sb_best_sad, *bs32 = p_sb_best_sad + 1, *bs16 = p_sb_best_sad + 5, *bs8 = p_sb_best_sad + 21;
    uint32_t *bm64 = p_sb_best_mv, *bm32 = p_sb_best_mv + 1, *bm16 = p_sb_best_mv + 5, *bm8 = p_sb_best_mv + 21;
    const uint32_t w8 = search_area_width & ~7u;
    uint32_t eight16[16][8], eight32[4][8];
    for (uint32_t sy = 0; sy < search_area_height; ++sy) {
        for (uint32_t sx = 0; sx < w8; sx += 8) {
            const uint8_t *r  = ref + (size_t)sy * ref_stride + sx;
            uint32_t       mv = ((uint32_t)(uint16_t)((int32_t)sy + y_search_area_origin) << 16) |
                (uint16_t)((int32_t)sx + x_search_area_origin);
            orc_ext_all_sad_calculation_8x8_16x16((uint8_t *)src, src_stride, (uint8_t *)r, ref_stride, mv, bs8,
                                                  bs16, bm8, bm16, eight16, NULL, sub_sad);
            orc_ext_eight_sad_calculation_32x32_64x64(eight16, bs32, bs64, bm32, bm64, mv, eight32);
        }
        for (uint32_t sx = w8; sx < search_area_width; ++sx) {
            const uint8_t *r  = ref + (size_t)sy * ref_stride + sx;
            uint32_t       mv = ((uint32_t)(uint16_t)((int32_t)sy + y_search_area_origin) << 16) |
                (uint16_t)((int32_t)sx + x_search_area_origin);
            uint32_t sad16[16], sad8[64], sad32[4];
            for (int by = 0; by < 4; ++by)
                for (int bx = 0; bx < 4; ++bx) {
                    const int slot = (by >> 1) * 8 + (bx >> 1) * 4 + (by & 1) * 2 + (bx & 1);
                    orc_ext_sad_calculation_8x8_16x16((uint8_t *)src + 16 * by * src_stride + 16 * bx, src_stride,
                                                      (uint8_t *)r + 16 * by * ref_stride + 16 * bx, ref_stride,
                                                      &bs8[4 * slot], &bs16[slot], &bm8[4 * slot], &bm16[slot], mv,
                                                      &sad16[slot], &sad8[4 * slot], sub_sad);
                }
            orc_ext_sad_calculation_32x32_64x64(sad16, bs32, bs64, bm32, bm64, mv, sad32);
        }
    }
}

/* 2x2 box at rows {v-1,v}, cols {h-1,h} with v,h starting at step/2 and advancing by step.
 * Codec/pic_analysis_process.c:131-161 */
void orc_downsample_2d(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                       uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                       uint32_t decim_step) {
    const uint32_t half = decim_step >> 1;
    uint32_t       oy   = 0;
    for (uint32_t v = half; v < input_area_height; v += decim_step, ++oy) {
        const uint8_t *lo = input_samples + (size_t)v * input_stride;
        const uint8_t *hi = lo - input_stride;
        uint32_t       ox = 0;
        for (uint32_t h = half; h < input_area_width; h += decim_step, ++ox)
            decim_samples[(size_t)oy * decim_stride + ox] =
                (uint8_t)(((uint32_t)hi[h - 1] + hi[h] + lo[h - 1] + lo[h] + 2) >> 2);
    }
}

/* Replicate edges: horizontal first (per interior row), then whole-stride rows up and down.
 * Codec/pic_operators.c:338-382 */
void orc_generate_padding(uint8_t *src_pic, uint32_t src_stride, uint32_t original_src_width,
                          uint32_t original_src_height, uint32_t padding_width,
                          uint32_t padding_height) {
    uint8_t *row = src_pic + padding_width + (size_t)padding_height * src_stride;
    for (uint32_t y = 0; y < original_src_height; ++y, row += src_stride) {
        memset(row - padding_width, row[0], padding_width);
        memset(row + original_src_width, row[original_src_width - 1], padding_width);
    }
    uint8_t *top = src_pic + (size_t)padding_height * src_stride;
    uint8_t *bot = src_pic + (size_t)(padding_height + original_src_height - 1) * src_stride;
    for (uint32_t k = 1; k <= padding_height; ++k) {
        memcpy(top - (size_t)k * src_stride, top, src_stride);
        memcpy(bot + (size_t)k * src_stride, bot, src_stride);
    }
}

static int plane_alloc(OrcPlane *pl, uint32_t w, uint32_t h, uint32_t pad) {
    pl->width    = w;
    pl->height   = h;
    pl->org_x    = pad;
    pl->org_y    = pad;
    pl->stride_y = w + 2 * pad; /* pic_buffer_desc.c:59-60 */
    pl->buffer_y = (uint8_t *)calloc((size_t)pl->stride_y * (h + 2 * pad), 1);
    return pl->buffer_y ? 0 : -1;
}

/* svt_aom_pad_input_pictures (luma) + svt_aom_downsample_filtering_input_picture,
 * Codec/pic_analysis_process.c:1922-1986. Pads: source 72 (enc_handle.c:4082-4094),
 * quarter 32, sixteenth 16 (enc_handle.c:1292-1311). Destination offset uses org_x for both
 * axes like the reference (:1934-1935) — equal because org_x == org_y. */
int orc_pyramid_build(OrcPyramid *pyr, const uint8_t *luma, uint32_t stride, uint32_t width,
                      uint32_t height, int hme_level1, uint64_t picture_number) {
    memset(pyr, 0, sizeof(*pyr));
    pyr->picture_number = picture_number;
    if (plane_alloc(&pyr->full, width, height, 72) || plane_alloc(&pyr->quarter, width >> 1, height >> 1, 32) ||
        plane_alloc(&pyr->sixteenth, width >> 2, height >> 2, 16)) {
        orc_pyramid_free(pyr);
        return -1;
    }
    OrcPlane *f = &pyr->full, *q = &pyr->quarter, *s = &pyr->sixteenth;
    for (uint32_t y = 0; y < height; ++y)
        memcpy(f->buffer_y + (size_t)(f->org_y + y) * f->stride_y + f->org_x, luma + (size_t)y * stride, width);
    orc_generate_padding(f->buffer_y, f->stride_y, f->width, f->height, f->org_x, f->org_y);
    uint8_t *f0 = f->buffer_y + f->org_x + (size_t)f->org_y * f->stride_y;
    uint8_t *q0 = q->buffer_y + q->org_x + (size_t)q->org_x * q->stride_y;
    uint8_t *s0 = s->buffer_y + s->org_x + (size_t)s->org_x * s->stride_y;
    orc_downsample_2d(f0, f->stride_y, f->width, f->height, q0, q->stride_y, 2);
    orc_generate_padding(q->buffer_y, q->stride_y, q->width, q->height, q->org_x, q->org_y);
    if (hme_level1)
        orc_downsample_2d(q0, q->stride_y, q->width, q->height, s0, s->stride_y, 2);
    else
        orc_downsample_2d(f0, f->stride_y, f->width, f->height, s0, s->stride_y, 4);
    orc_generate_padding(s->buffer_y, s->stride_y, s->width, s->height, s->org_x, s->org_y);
    return 0;
}

void orc_pyramid_free(OrcPyramid *pyr) {
    free(pyr->full.buffer_y);
    free(pyr->quarter.buffer_y);
    free(pyr->sixteenth.buffer_y);
    memset(pyr, 0, sizeof(*pyr));
}

/* 10-bit source -> the 8-bit plane the open-loop ME reads. Follows svt_enc_msb_un_pack2_d,
 * Source/Lib/C_DEFAULT/pack_unpack_c.c:202-218 (called through svt_aom_un_pack2d, Codec/pic_operators.c:228-244):
 * out = (uint8_t)(sample >> 2). */
void orc_unpack_msb(const uint16_t *in16, uint32_t in_stride, uint8_t *out8, uint32_t out8_stride, uint32_t width,
                    uint32_t height) {
    for (uint32_t j = 0; j < height; ++j)
        for (uint32_t k = 0; k < width; ++k) out8[k + (size_t)j * out8_stride] = (uint8_t)(in16[k + (size_t)j * in_stride] >> 2);
}

/*
 * me_oracle_stats.c — CPU restatement of the picture statistics the picture-analysis kernel gathers right after the
 * pyramid build (svt_aom_gathering_picture_statistics, Codec/pic_analysis_process.c:1560-1578). TEST INFRASTRUCTURE
 * ONLY (see me_oracle.h). Pinned against the reference's own function compiled from /root/reference
 * (oracle/ref_harness.c: ref_picture_statistics) by tests/test_stats.py and against tests/golden/stats_*.npz.
 * All file:line citations are into /root/reference/Source/Lib/Codec/pic_analysis_process.c.
 */
#include <string.h>
#include "me_oracle.h"

#define VARIANCE_PRECISION 16 /* definitions.h */

/* svt_compute_mean_c / svt_compute_mean_squared_values_c on an 8x8 block (:197-232): every row */
static void mean8x8_full(const uint8_t *p, uint32_t stride, uint64_t *mean, uint64_t *msq) {
    uint64_t s = 0, q = 0;
    for (int y = 0; y < 8; ++y, p += stride)
        for (int x = 0; x < 8; ++x) {
            s += p[x];
            q += (uint64_t)p[x] * p[x];
        }
    *mean = (s << (VARIANCE_PRECISION >> 1)) / 64;
    *msq = (q << VARIANCE_PRECISION) / 64;
}

/* svt_compute_sub_mean_8x8_c / svt_aom_compute_sub_mean_squared_values_c (:234-272): rows 0, 2, 4, 6 */
static void mean8x8_sub(const uint8_t *p, uint32_t stride, uint64_t *mean, uint64_t *msq) {
    uint64_t s = 0, q = 0;
    for (int y = 0; y < 8; y += 2)
        for (int x = 0; x < 8; ++x) {
            const uint8_t v = p[(size_t)y * stride + x];
            s += v;
            q += (uint64_t)v * v;
        }
    *mean = s << 3;
    *msq = q << 11;
}

static uint16_t var_of(uint64_t msq, uint64_t mean) { return (uint16_t)((msq - mean * mean) >> VARIANCE_PRECISION); }

/* compute_block_mean_compute_variance (:307-1378) for the b64 whose top-left sample is `p` */
static void b64_variance(const uint8_t *p, uint32_t stride, int prec_sub, int all_sizes, uint16_t *out) {
    uint64_t m8[64], q8[64], m16[16], q16[16], m32[4], q32[4];
    for (int by = 0; by < 8; ++by)
        for (int bx = 0; bx < 8; ++bx) {
            const uint8_t *b = p + (size_t)(8 * by) * stride + 8 * bx;
            if (prec_sub)
                mean8x8_sub(b, stride, &m8[by * 8 + bx], &q8[by * 8 + bx]);
            else
                mean8x8_full(b, stride, &m8[by * 8 + bx], &q8[by * 8 + bx]);
        }
    for (int y = 0; y < 4; ++y)
        for (int x = 0; x < 4; ++x) {
            const int k = (2 * y) * 8 + 2 * x;
            m16[y * 4 + x] = (m8[k] + m8[k + 1] + m8[k + 8] + m8[k + 9]) >> 2;
            q16[y * 4 + x] = (q8[k] + q8[k + 1] + q8[k + 8] + q8[k + 9]) >> 2;
        }
    for (int y = 0; y < 2; ++y)
        for (int x = 0; x < 2; ++x) {
            const int k = (2 * y) * 4 + 2 * x;
            m32[y * 2 + x] = (m16[k] + m16[k + 1] + m16[k + 4] + m16[k + 5]) >> 2;
            q32[y * 2 + x] = (q16[k] + q16[k + 1] + q16[k + 4] + q16[k + 5]) >> 2;
        }
    const uint64_t m64 = (m32[0] + m32[1] + m32[2] + m32[3]) >> 2, q64 = (q32[0] + q32[1] + q32[2] + q32[3]) >> 2;
    if (all_sizes) { /* (:1110) raster-scan indices: ME_TIER_ZERO_PU_* of me_context.h:54-140 */
        for (int k = 0; k < 64; ++k) out[21 + k] = var_of(q8[k], m8[k]);
        for (int k = 0; k < 16; ++k) out[5 + k] = var_of(q16[k], m16[k]);
        for (int k = 0; k < 4; ++k) out[1 + k] = var_of(q32[k], m32[k]);
    }
    out[0] = var_of(q64, m64);
}

int orc_picture_statistics(const OrcPyramid *pyr, const SvtMeStatsParams *prm, SvtMeStatsOut *out) {
    /* svt_aom_gathering_picture_statistics :1560 */
    uint64_t avg_luma = 256; /* INVALID_LUMA */
    if (prm->calc_hist) {
        /* sub_sample_luma_generate_pixel_intensity_histogram_bins :1462-1528 on the 1/16 plane */
        const OrcPlane *pl = &pyr->sixteenth;
        const uint32_t nw = prm->regions_per_width, nh = prm->regions_per_height;
        if (nw < 1 || nw > 4 || nh < 1 || nh > 4) return -1;
        const uint32_t rw = pl->width / nw, rh = pl->height / nh;
        const uint32_t decim = prm->scene_change_detection ? 1 : 4;
        uint64_t total = 0;
        for (uint32_t ix = 0; ix < nw; ++ix)
            for (uint32_t iy = 0; iy < nh; ++iy) {
                uint32_t *hist = out->picture_histogram + ((size_t)ix * 4 + iy) * 256;
                for (int b = 0; b < 256; ++b) hist[b] = 1; /* "Initialize bins to 1" :1481 */
                const uint32_t w = rw + (ix == nw - 1 ? pl->width - nw * rw : 0);
                const uint32_t h = rh + (iy == nh - 1 ? pl->height - nh * rh : 0);
                const uint8_t *p = pl->buffer_y + (size_t)(pl->org_y + iy * rh) * pl->stride_y + pl->org_x + ix * rw;
                uint64_t sum = 0;
                for (uint32_t y = 0; y < h; y += decim) /* calculate_histogram :166-184 */
                    for (uint32_t x = 0; x < w; x += decim) {
                        const uint8_t v = p[(size_t)y * pl->stride_y + x];
                        ++hist[v];
                        sum += v;
                    }
                sum *= decim * decim;
                out->average_intensity_per_region[ix * 4 + iy] = (uint8_t)((sum + ((w * h) >> 1)) / (w * h));
                total += sum;
                for (int b = 0; b < 256; ++b) hist[b] = hist[b] * 4 * 4 * decim * decim;
            }
        avg_luma = total / ((uint64_t)pl->width * pl->height);
    }
    *out->avg_luma = avg_luma;
    if (prm->calculate_variance) {
        /* compute_picture_spatial_statistics :1530-1552 */
        const OrcPlane *f = &pyr->full;
        const uint32_t cols = (f->width + 63) / 64, rows = (f->height + 63) / 64, n = prm->all_block_sizes ? 85 : 1;
        uint64_t tot = 0;
        for (uint32_t b = 0; b < cols * rows; ++b) {
            const uint8_t *p = f->buffer_y + (size_t)(f->org_y + (b / cols) * 64) * f->stride_y + f->org_x + (b % cols) * 64;
            uint16_t v[85];
            b64_variance(p, f->stride_y, prm->block_mean_prec_sub, prm->all_block_sizes, v);
            memcpy(out->variance + (size_t)b * n, v, n * sizeof(uint16_t));
            tot += v[0];
        }
        *out->pic_avg_variance = (uint16_t)(tot / (cols * rows));
    } else
        *out->pic_avg_variance = 0;
    return 0;
}

/* ---- dynamic-GOP detector HME: dg_detector_hme_level0 over all segments (pd_process.c:491-589) ----
 * early_hme_b64 (:389-487): the clipping of hme_level_0 with the pad limits org - 1 = 15 of the reference's 1/16 planes,
 * one full-SAD svt_sad_loop_kernel, vector x 4. */
int orc_dg_detector_hme(const OrcPyramid *src, const OrcPyramid *ref, uint32_t input_resolution, SvtMeDgMetrics *out) {
    const OrcPlane *sp = &src->sixteenth, *rp = &ref->sixteenth;
    const int16_t S = input_resolution <= 1 ? 16 : (input_resolution <= 2 ? 64 : 128); /* (:497-498) */
    const uint32_t cols = (src->full.width + 63) / 64, rows = (src->full.height + 63) / 64;
    memset(out, 0, sizeof(*out));
    uint64_t best = (uint64_t)~0;
    int16_t cx = 0, cy = 0; /* sr_center lives across the b64 loop (:501) */
    for (uint32_t by = 0; by < rows; ++by)
        for (uint32_t bx = 0; bx < cols; ++bx) {
            const int16_t org_x = (int16_t)(bx * 64) >> 2, org_y = (int16_t)(by * 64) >> 2;
            int16_t sa_w = (int16_t)((S + 7) & ~0x07), sa_h = S;
            const int16_t pad_w = (int16_t)rp->org_x - 1, pad_h = (int16_t)rp->org_y - 1;
            const int16_t W = (int16_t)rp->width, H = (int16_t)rp->height;
            int16_t ox = -(int16_t)(sa_w >> 1), oy = -(int16_t)(sa_h >> 1);
            if ((org_x + ox) < -pad_w) {
                ox = -pad_w - org_x;
                sa_w = sa_w - (-pad_w - (org_x + ox));
            }
            if ((org_x + ox) > W - 1) ox = ox - ((org_x + ox) - (W - 1));
            if ((org_x + ox + sa_w) > W) sa_w = (int16_t)((1 > sa_w - ((org_x + ox + sa_w) - W)) ? 1 : sa_w - ((org_x + ox + sa_w) - W));
            sa_w = (sa_w < 8) ? sa_w : sa_w & ~0x07;
            if ((org_y + oy) < -pad_h) {
                oy = -pad_h - org_y;
                sa_h = sa_h - (-pad_h - (org_y + oy));
            }
            if ((org_y + oy) > H - 1) oy = oy - ((org_y + oy) - (H - 1));
            if (org_y + oy + sa_h > H) sa_h = (int16_t)((1 > sa_h - ((org_y + oy + sa_h) - H)) ? 1 : sa_h - ((org_y + oy + sa_h) - H));
            uint8_t *s = sp->buffer_y + (size_t)(sp->org_y + org_y) * sp->stride_y + sp->org_x + org_x;
            uint8_t *r = rp->buffer_y + (size_t)((int)rp->org_y + org_y + oy) * rp->stride_y + (int)rp->org_x + org_x + ox;
            orc_sad_loop_kernel(s, sp->stride_y, r, rp->stride_y, 16, 16, &best, &cx, &cy, rp->stride_y, 0, sa_w, sa_h);
            cx = (int16_t)((int16_t)(cx + ox) * 4);
            cy = (int16_t)((int16_t)(cy + oy) * 4);
            out->tot_dist += best;
            out->tot_cplx += best > 16 * 16 * 30;
            out->tot_active += (cx != 0 || cy != 0);
            if (by < rows / 2)
                out->sum_in_vectors += cy > 0 ? -1 : (cy < 0 ? 1 : 0);
            else if (by > rows / 2)
                out->sum_in_vectors += cy > 0 ? 1 : (cy < 0 ? -1 : 0);
            if (bx < cols / 2)
                out->sum_in_vectors += cx > 0 ? -1 : (cx < 0 ? 1 : 0);
            else if (bx > cols / 2)
                out->sum_in_vectors += cx > 0 ? 1 : (cx < 0 ? -1 : 0);
        }
    return 0;
}

/*
 * me_oracle_stage.c — CPU restatement (plain C) of svt_aom_motion_estimation_b64 and everything
 * it calls for ME_OPEN_LOOP. TEST INFRASTRUCTURE ONLY (see me_oracle.h).
 * Citations: /root/reference/Source/Lib/Codec/motion_estimation.c unless another file is named.
 *
 * Out of scope here (same as the product): ME_MCTF (temporal filtering), restricted_motion_vector
 * tiles, super-res / resize scaled references.
 */
#include <stdlib.h>
#include <string.h>
#include "me_oracle.h"
#include "mt_pool.h"

#define NL SVTME_MAX_LISTS
#define NR SVTME_MAX_REFS
#define MAXU32 0xFFFFFFFFu
#define IABS(v) ((v) < 0 ? -(v) : (v))
#define IMIN(a, b) ((a) < (b) ? (a) : (b))
#define IMAX(a, b) ((a) > (b) ? (a) : (b))

typedef struct PreHme {
    uint16_t sa_w, sa_h;
    int16_t  col, row;
    uint64_t sad;
    uint8_t  valid;
} PreHme;

typedef struct RefResult {
    int16_t  hme_sc_x, hme_sc_y;
    uint64_t hme_sad;
    uint8_t  do_ref;
} RefResult;

/* Per-b64 working state: the MeContext members (me_context.h:366-509) the open-loop path touches. */
typedef struct B64State {
    const SvtMeParams *p;
    const OrcPyramid  *cur;
    const OrcPyramid *(*refs)[NR];
    uint32_t org_x, org_y, b64_w, b64_h;
    int16_t  pic_w, pic_h;
    const uint8_t *src, *qsrc, *ssrc; /* b64_src_ptr, quarter_b64_buffer, sixteenth_b64_buffer */
    uint32_t src_stride, qstride, sstride;
    SvtMeSearchAreaMinMax hme_l0_sa;
    int16_t  cx[3][NL][NR][2][2], cy[3][NL][NR][2][2]; /* [level][list][ref][sr_w][sr_h] */
    uint64_t lsad[3][NL][NR][2][2];
    RefResult sr[NL][NR];
    uint32_t  div[NL][NR];
    uint32_t  zz[NL][NR];
    PreHme    ph[NL][NR][2];
    uint8_t   performed_phme[NL][NR][2];
    uint32_t  best_sad[NL][NR][SVTME_PU_COUNT], best_mv[NL][NR][SVTME_PU_COUNT];
    uint32_t  me_distortion[SVTME_PU_COUNT];
} B64State;

/* svt_aom_get_scaled_picture_distance :1239-1243 */
static uint16_t scaled_dist(uint16_t dist) { return (uint16_t)(((dist * 5) / 8) + ((dist % 8) ? 1 : 0)); }

/* get_me_reference's *dist :1232-1235 */
static uint16_t ref_dist(const B64State *s, int li, int ri) {
    int64_t d = (int64_t)s->p->picture_number - (int64_t)s->p->ref_picture_number[li][ri];
    return (uint16_t)(int16_t)IABS(d);
}

/* init_me_hme_data :3080-3140 (everything is zero-initialised per b64 here, see header note) */
static void init_state(B64State *s) {
    memset(s->cx, 0, sizeof(s->cx));
    memset(s->cy, 0, sizeof(s->cy));
    memset(s->lsad, 0, sizeof(s->lsad));
    memset(s->best_mv, 0, sizeof(s->best_mv));
    memset(s->best_sad, 0, sizeof(s->best_sad));
    memset(s->ph, 0, sizeof(s->ph));
    memset(s->performed_phme, 0, sizeof(s->performed_phme));
    for (int li = 0; li < NL; ++li)
        for (int ri = 0; ri < NR; ++ri) {
            s->sr[li][ri].hme_sc_x = 0;
            s->sr[li][ri].hme_sc_y = 0;
            s->sr[li][ri].do_ref   = 1;
            s->sr[li][ri].hme_sad  = MAXU32;
            s->div[li][ri]         = 1;
            s->zz[li][ri]          = MAXU32;
        }
    s->hme_l0_sa = s->p->hme_l0_sa;
}

/* get_zz_sad :1737-1759 — 64 x (h/2) SAD on every other row, doubled. */
static uint32_t zz_sad_of(const B64State *s, const OrcPlane *ref) {
    const uint8_t *r = ref->buffer_y + (int16_t)ref->org_x + (int16_t)s->org_x +
        (size_t)((int16_t)ref->org_y + (int16_t)s->org_y) * ref->stride_y;
    return orc_nxm_sad_kernel(s->src, s->src_stride << 1, r, ref->stride_y << 1, s->b64_h >> 1, s->b64_w) << 1;
}

/* init_zz_sad :2452-2507 */
static void init_zz_sad(B64State *s) {
    const SvtMeParams *p    = s->p;
    uint32_t           best = MAXU32;
    for (int li = 0; li < p->num_of_list_to_search; ++li)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri)
            if (p->temporal_layer_index > 0 || li == 0) {
                uint32_t z = zz_sad_of(s, &s->refs[li][ri]->full);
                z          = (z * 64 * 64) / (s->b64_w * s->b64_h);
                s->zz[li][ri] = z;
                best          = IMIN(best, z);
            }
    if (p->temporal_layer_index > 0 && best < p->zz_sad_th)
        for (int li = 0; li < p->num_of_list_to_search; ++li)
            for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ++ri)
                if ((uint32_t)((s->zz[li][ri] - best) * 100) > (uint32_t)(p->zz_sad_pct * best))
                    s->sr[li][ri].do_ref = 0;
    if (p->me_safe_limit_zz_th) {
        int limit = p->hierarchical_levels > 0 && p->num_of_list_to_search == 2 &&
            p->temporal_layer_index >= p->hierarchical_levels && p->similar_brightness_refs &&
            s->zz[0][0] < p->me_safe_limit_zz_th && s->zz[1][0] < p->me_safe_limit_zz_th;
        if (limit)
            for (int li = 0; li < p->num_of_list_to_search; ++li)
                for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ++ri) s->sr[li][ri].do_ref = 0;
    }
}

/* One svt_sad_loop_kernel call on a decimated plane with the HME row sub-sampling convention
 * (:891-913): SUB_SAD -> doubled strides, half height, result x2. */
static void hme_sad_loop(const B64State *s, const uint8_t *blk, uint32_t blk_stride, const OrcPlane *ref,
                         int32_t region_x, int32_t region_y, uint32_t bw, uint32_t bh, uint8_t skip_line,
                         int16_t sa_w, int16_t sa_h, uint64_t *best, int16_t *bx, int16_t *by) {
    const int full = s->p->hme_search_method == 1; /* FULL_SAD_SEARCH definitions.h:2080 */
    uint8_t  *r    = ref->buffer_y + (uint32_t)(region_x + region_y * (int32_t)ref->stride_y);
    orc_sad_loop_kernel((uint8_t *)blk, full ? blk_stride : blk_stride * 2, r,
                        full ? ref->stride_y : ref->stride_y * 2, full ? bh : bh >> 1, bw, best, bx, by,
                        ref->stride_y, skip_line, sa_w, sa_h);
    if (!full)
        *best *= 2;
}

/* prehme_core :1641-1735 */
static void prehme_core(B64State *s, int16_t org_x, int16_t org_y, uint32_t bw, uint32_t bh,
                        const OrcPlane *ref, PreHme *d) {
    int16_t pad_w = (int16_t)ref->org_x - 1, pad_h = (int16_t)ref->org_y - 1;
    int16_t sa_w = (int16_t)d->sa_w, sa_h = (int16_t)d->sa_h;
    int16_t ox = -(int16_t)(sa_w >> 1), oy = -(int16_t)(sa_h >> 1);
    int16_t W = (int16_t)ref->width, H = (int16_t)ref->height;

    ox   = ((org_x + ox) < -pad_w) ? -pad_w - org_x : ox;
    sa_w = ((org_x + ox) < -pad_w) ? sa_w - (-pad_w - (org_x + ox)) : sa_w;
    ox   = ((org_x + ox) > W - 1) ? ox - ((org_x + ox) - (W - 1)) : ox;
    sa_w = ((org_x + ox + sa_w) > W) ? IMAX(1, sa_w - ((org_x + ox + sa_w) - W)) : sa_w;
    oy   = ((org_y + oy) < -pad_h) ? -pad_h - org_y : oy;
    sa_h = ((org_y + oy) < -pad_h) ? sa_h - (-pad_h - (org_y + oy)) : sa_h;
    oy   = ((org_y + oy) > H - 1) ? oy - ((org_y + oy) - (H - 1)) : oy;
    sa_h = (org_y + oy + sa_h > H) ? IMAX(1, sa_h - ((org_y + oy + sa_h) - H)) : sa_h;

    int16_t x_tl = ((int16_t)ref->org_x + org_x) + ox;
    int16_t y_tl = ((int16_t)ref->org_y + org_y) + oy;
    hme_sad_loop(s, s->ssrc, s->sstride, ref, x_tl, y_tl, bw, bh, s->p->prehme_skip_search_line, sa_w, sa_h,
                 &d->sad, &d->col, &d->row);
    d->col   = (int16_t)((int16_t)(d->col + ox) * 4);
    d->row   = (int16_t)((int16_t)(d->row + oy) * 4);
    d->valid = 1;
}

/* check_prehme_early_exit :1763-1788 */
static int prehme_early_exit(B64State *s, int li, int ri, int sr) {
    const SvtMeParams *p = s->p;
    PreHme            *d = &s->ph[li][ri][sr];
    if (p->me_early_exit_th && s->zz[li][ri] < p->me_early_exit_th) {
        d->col = d->row = 0;
        d->sad          = 0;
        d->valid        = 1;
        return 1;
    }
    if (p->prehme_l1_early_exit) {
        const PreHme *l0 = &s->ph[0][ri][sr];
        if (li == 1 && l0->valid && ((l0->sad < (32 * 32)) || ((IABS(l0->col) < 16) && (IABS(l0->row) < 16)))) {
            d->col   = (int16_t)-l0->col;
            d->row   = (int16_t)-l0->row;
            d->sad   = l0->sad;
            d->valid = 1;
            return 1;
        }
    }
    return 0;
}

/* prehme_b64 :1791-1866 */
static void prehme_b64(B64State *s) {
    const SvtMeParams *p        = s->p;
    uint32_t           best_sad = MAXU32;
    for (int li = 0; li < p->num_of_list_to_search; ++li) {
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            if (p->temporal_layer_index > 0 || li == 0) {
                uint32_t factor = scaled_dist(ref_dist(s, li, ri));
                for (int sr = 0; sr < 2; ++sr) {
                    if (prehme_early_exit(s, li, ri, sr))
                        continue;
                    PreHme *d = &s->ph[li][ri][sr];
                    if (!s->sr[li][ri].do_ref) {
                        d->col = d->row = 0;
                        d->sad          = MAXU32;
                        continue;
                    }
                    d->sa_w = (uint16_t)IMIN(p->prehme_sa_cfg[sr].sa_min.width * factor,
                                             (uint32_t)p->prehme_sa_cfg[sr].sa_max.width);
                    d->sa_h = (uint16_t)IMIN(p->prehme_sa_cfg[sr].sa_min.height * factor,
                                             (uint32_t)p->prehme_sa_cfg[sr].sa_max.height);
                    prehme_core(s, (int16_t)s->org_x >> 2, (int16_t)s->org_y >> 2, s->b64_w >> 2, s->b64_h >> 2,
                                &s->refs[li][ri]->sixteenth, d);
                    s->performed_phme[li][ri][sr] = 1;
                }
                uint32_t m = (uint32_t)IMIN(s->ph[li][ri][0].sad, s->ph[li][ri][1].sad);
                best_sad   = IMIN(best_sad, m);
            } else {
                for (int sr = 0; sr < 2; ++sr) {
                    s->ph[1][ri][sr].col = (int16_t)-s->ph[0][ri][sr].col;
                    s->ph[1][ri][sr].row = (int16_t)-s->ph[0][ri][sr].row;
                    s->ph[1][ri][sr].sad = s->ph[0][ri][sr].sad;
                }
            }
        }
    }
    if (p->temporal_layer_index > 0 && best_sad < p->phme_sad_th)
        for (int li = 0; li < p->num_of_list_to_search; ++li)
            for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
                if (!s->sr[li][ri].do_ref)
                    continue;
                uint32_t m = (uint32_t)IMIN(s->ph[li][ri][0].sad, s->ph[li][ri][1].sad);
                if ((uint32_t)((m - best_sad) * 100) > (uint32_t)(p->phme_sad_pct * best_sad))
                    s->sr[li][ri].do_ref = 0;
            }
}

/* get_hme_l0_search_area :1870-1935 (modifies s->hme_l0_sa like the reference; caller restores) */
static void hme_l0_area(B64State *s, int li, int ri, uint16_t dist, int16_t *sa_w, int16_t *sa_h) {
    const SvtMeParams *p = s->p;
    if (p->enable_me_sr_adjustment && p->distance_based_hme_resizing) {
        uint8_t is_hor = 1, is_ver = 1, is_still = 0;
        if (p->reduce_hme_l0_sr_th_min && p->reduce_hme_l0_sr_th_max && (li || ri)) {
            int16_t mvx = s->cx[0][0][0][0][0], mvy = s->cy[0][0][0][0][0];
            is_ver   = (IABS(mvx) < p->reduce_hme_l0_sr_th_min) && (IABS(mvy) > p->reduce_hme_l0_sr_th_max);
            is_hor   = (IABS(mvx) > p->reduce_hme_l0_sr_th_max) && (IABS(mvy) < p->reduce_hme_l0_sr_th_min);
            is_still = (IABS(mvx) < (p->reduce_hme_l0_sr_th_min * 3)) && (IABS(mvy) < (p->reduce_hme_l0_sr_th_min * 3));
        }
        uint8_t xo = is_hor ? 1 : 2, yo = is_ver ? 1 : 2;
        if (p->enable_me_sr_adjustment == 2 && is_still)
            xo = yo = 4;
        s->hme_l0_sa.sa_min.width  = (uint16_t)(s->hme_l0_sa.sa_min.width / (xo + ri));
        s->hme_l0_sa.sa_min.height = (uint16_t)(s->hme_l0_sa.sa_min.height / (yo + ri));
        s->hme_l0_sa.sa_max.width  = (uint16_t)(s->hme_l0_sa.sa_max.width / (xo + ri));
        s->hme_l0_sa.sa_max.height = (uint16_t)(s->hme_l0_sa.sa_max.height / (yo + ri));
    }
    int32_t factor = scaled_dist(dist);
    int16_t w      = (int16_t)(s->hme_l0_sa.sa_min.width / p->num_hme_sa_w);
    w              = (int16_t)IMIN((((w * factor) + 15) & ~0x0F),
                      (((s->hme_l0_sa.sa_max.width / p->num_hme_sa_w) + 15) & ~0x0F));
    int16_t h      = (int16_t)(s->hme_l0_sa.sa_min.height / p->num_hme_sa_h);
    h              = (int16_t)IMIN((h * factor), s->hme_l0_sa.sa_max.height / p->num_hme_sa_h);
    *sa_w          = w;
    *sa_h          = h;
}

/* Shared tail of hme_level_0/1/2 (:820-1113): clip the window (with the no-op left/top size
 * corrections kept), search, add the origin, scale the vector. */
static void hme_level_search(B64State *s, int level, const uint8_t *blk, uint32_t blk_stride,
                             const OrcPlane *ref, int16_t org_x, int16_t org_y, uint32_t bw, uint32_t bh,
                             int16_t sa_w, int16_t sa_h, int16_t sa_ox, int16_t sa_oy, int16_t pad_w,
                             int16_t pad_h, int scale, uint64_t *best, int16_t *scx, int16_t *scy) {
    int16_t W = (int16_t)ref->width, H = (int16_t)ref->height;
    (void)level;
    if ((org_x + sa_ox) < -pad_w) {
        sa_ox = -pad_w - org_x;
        sa_w  = sa_w - (-pad_w - (org_x + sa_ox));
    }
    if ((org_x + sa_ox) > W - 1)
        sa_ox = sa_ox - ((org_x + sa_ox) - (W - 1));
    if ((org_x + sa_ox + sa_w) > W)
        sa_w = IMAX(1, sa_w - ((org_x + sa_ox + sa_w) - W));
    sa_w = (sa_w < 8) ? sa_w : sa_w & ~0x07;
    if ((org_y + sa_oy) < -pad_h) {
        sa_oy = -pad_h - org_y;
        sa_h  = sa_h - (-pad_h - (org_y + sa_oy));
    }
    if ((org_y + sa_oy) > H - 1)
        sa_oy = sa_oy - ((org_y + sa_oy) - (H - 1));
    if (org_y + sa_oy + sa_h > H)
        sa_h = IMAX(1, sa_h - ((org_y + sa_oy + sa_h) - H));
    int16_t x_tl = ((int16_t)ref->org_x + org_x) + sa_ox;
    int16_t y_tl = ((int16_t)ref->org_y + org_y) + sa_oy;
    hme_sad_loop(s, blk, blk_stride, ref, x_tl, y_tl, bw, bh, 0, sa_w, sa_h, best, scx, scy);
    *scx = (int16_t)((int16_t)(*scx + sa_ox) * scale);
    *scy = (int16_t)((int16_t)(*scy + sa_oy) * scale);
}

static void set_quadrants(B64State *s, int lvl, int li, int ri, int16_t x, int16_t y, uint64_t sad) {
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
            s->cx[lvl][li][ri][a][b]   = x;
            s->cy[lvl][li][ri][a][b]   = y;
            s->lsad[lvl][li][ri][a][b] = sad;
        }
}

/* hme_level0_b64 :1976-2106 (+ get_worst_quadrant :1939-1971) */
static void hme_level0_b64(B64State *s) {
    const SvtMeParams          *p    = s->p;
    const SvtMeSearchAreaMinMax base = s->hme_l0_sa;
    for (int li = 0; li < p->num_of_list_to_search; ++li) {
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            if (p->me_early_exit_th && s->zz[li][ri] < (p->me_early_exit_th >> 2)) {
                set_quadrants(s, 0, li, ri, 0, 0, 0);
                continue;
            }
            if (p->prev_me_stage_based_exit_th) {
                int sr = s->ph[li][ri][0].sad <= s->ph[li][ri][1].sad ? 0 : 1;
                if (s->performed_phme[li][ri][sr] &&
                    s->ph[li][ri][sr].sad < (p->prev_me_stage_based_exit_th >> 4)) {
                    set_quadrants(s, 0, li, ri, s->ph[li][ri][sr].col, s->ph[li][ri][sr].row, s->ph[li][ri][sr].sad);
                    continue;
                }
            }
            if (!s->sr[li][ri].do_ref) {
                set_quadrants(s, 0, li, ri, 0, 0, MAXU32);
                continue;
            }
            if (!(p->temporal_layer_index > 0 || li == 0))
                continue;
            const OrcPlane *ref = &s->refs[li][ri]->sixteenth;
            int16_t         sa_w = 0, sa_h = 0;
            hme_l0_area(s, li, ri, ref_dist(s, li, ri), &sa_w, &sa_h);
            for (int sh = 0; sh < p->num_hme_sa_h; ++sh)
                for (int sw = 0; sw < p->num_hme_sa_w; ++sw) {
                    /* hme_level_0 :836-846 */
                    int16_t w  = (int16_t)((sa_w + 7) & ~0x07);
                    int16_t ox = (int16_t)(-(int16_t)((w * p->num_hme_sa_w) >> 1) + (int16_t)(w * sw));
                    int16_t oy = (int16_t)(-(int16_t)((sa_h * p->num_hme_sa_h) >> 1) + (int16_t)(sa_h * sh));
                    hme_level_search(s, 0, s->ssrc, s->sstride, ref, (int16_t)s->org_x >> 2, (int16_t)s->org_y >> 2,
                                     s->b64_w >> 2, s->b64_h >> 2, w, sa_h, ox, oy, (int16_t)ref->org_x - 1,
                                     (int16_t)ref->org_y - 1, 4, &s->lsad[0][li][ri][sw][sh],
                                     &s->cx[0][li][ri][sw][sh], &s->cy[0][li][ri][sw][sh]);
                }
            if (p->enable_me_sr_adjustment && p->distance_based_hme_resizing)
                s->hme_l0_sa = base;
            if (p->prehme_enable) {
                uint64_t(*q)[2] = s->lsad[0][li][ri];
                uint8_t  bw = 0, bh = 0;
                uint64_t mx = 0;
                if (q[0][0] > mx) { mx = q[0][0]; bw = 0; bh = 0; }
                if (q[1][0] > mx) { mx = q[1][0]; bw = 1; bh = 0; }
                if (q[0][1] > mx) { mx = q[0][1]; bw = 0; bh = 1; }
                if (q[1][1] > mx) { bw = 1; bh = 1; }
                int sr = s->ph[li][ri][0].sad <= s->ph[li][ri][1].sad ? 0 : 1;
                if (s->ph[li][ri][sr].sad < q[bw][bh]) {
                    q[bw][bh]                 = s->ph[li][ri][sr].sad;
                    s->cx[0][li][ri][bw][bh] = s->ph[li][ri][sr].col;
                    s->cy[0][li][ri][bw][bh] = s->ph[li][ri][sr].row;
                }
            }
        }
    }
}

/* hme_level1_b64 :2111-2192 */
static void hme_level1_b64(B64State *s) {
    const SvtMeParams *p = s->p;
    for (int li = 0; li < p->num_of_list_to_search; ++li)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            if (!(p->temporal_layer_index > 0 || li == 0))
                continue;
            if (p->me_early_exit_th && s->zz[li][ri] < (p->me_early_exit_th >> 2)) {
                set_quadrants(s, 1, li, ri, 0, 0, 0);
                continue;
            }
            if (!s->sr[li][ri].do_ref) {
                set_quadrants(s, 1, li, ri, 0, 0, MAXU32);
                continue;
            }
            const OrcPlane *ref = &s->refs[li][ri]->quarter;
            for (int sh = 0; sh < p->num_hme_sa_h; ++sh)
                for (int sw = 0; sw < p->num_hme_sa_w; ++sw) {
                    if (p->prev_me_stage_based_exit_th &&
                        s->lsad[0][li][ri][sw][sh] < (p->prev_me_stage_based_exit_th >> 5)) {
                        s->cx[1][li][ri][sw][sh]   = s->cx[0][li][ri][sw][sh];
                        s->cy[1][li][ri][sw][sh]   = s->cy[0][li][ri][sw][sh];
                        s->lsad[1][li][ri][sw][sh] = s->lsad[0][li][ri][sw][sh];
                        continue;
                    }
                    /* hme_level_1 :940-947 */
                    int16_t w  = (int16_t)(((int16_t)p->hme_l1_sa.width + 7) & ~0x07);
                    int16_t h  = (int16_t)p->hme_l1_sa.height;
                    int16_t ox = (int16_t)(-(w >> 1) + (s->cx[0][li][ri][sw][sh] >> 1));
                    int16_t oy = (int16_t)(-(h >> 1) + (s->cy[0][li][ri][sw][sh] >> 1));
                    hme_level_search(s, 1, s->qsrc, s->qstride, ref, (int16_t)s->org_x >> 1, (int16_t)s->org_y >> 1,
                                     s->b64_w >> 1, s->b64_h >> 1, w, h, ox, oy, (int16_t)ref->org_x - 1,
                                     (int16_t)ref->org_y - 1, 2, &s->lsad[1][li][ri][sw][sh],
                                     &s->cx[1][li][ri][sw][sh], &s->cy[1][li][ri][sw][sh]);
                }
        }
}

/* hme_level2_b64 :2197-2247 — note: no do_ref / early-exit test at this level in the reference. */
static void hme_level2_b64(B64State *s) {
    const SvtMeParams *p = s->p;
    for (int li = 0; li < p->num_of_list_to_search; ++li)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            if (!(p->temporal_layer_index > 0 || li == 0))
                continue;
            const OrcPlane *ref = &s->refs[li][ri]->full;
            for (int sh = 0; sh < p->num_hme_sa_h; ++sh)
                for (int sw = 0; sw < p->num_hme_sa_w; ++sw) {
                    if (p->prev_me_stage_based_exit_th &&
                        s->lsad[1][li][ri][sw][sh] < (p->prev_me_stage_based_exit_th >> 2)) {
                        s->cx[2][li][ri][sw][sh]   = s->cx[1][li][ri][sw][sh];
                        s->cy[2][li][ri][sw][sh]   = s->cy[1][li][ri][sw][sh];
                        s->lsad[2][li][ri][sw][sh] = s->lsad[1][li][ri][sw][sh];
                        continue;
                    }
                    /* hme_level_2 :1041-1047 */
                    int16_t w  = (int16_t)(((int16_t)p->hme_l2_sa.width + 7) & ~0x07);
                    int16_t h  = (int16_t)p->hme_l2_sa.height;
                    int16_t ox = (int16_t)(-(w >> 1) + s->cx[1][li][ri][sw][sh]);
                    int16_t oy = (int16_t)(-(h >> 1) + s->cy[1][li][ri][sw][sh]);
                    hme_level_search(s, 2, s->src, s->src_stride, ref, (int16_t)s->org_x, (int16_t)s->org_y, s->b64_w,
                                     s->b64_h, w, h, ox, oy, 63, 63, 1, &s->lsad[2][li][ri][sw][sh],
                                     &s->cx[2][li][ri][sw][sh], &s->cy[2][li][ri][sw][sh]);
                }
        }
}

/* set_final_seach_centre_sb :2252-2450. hme_mv_sad / centres deliberately persist across the
 * (list, ref) loop: base-layer list-1 refs inherit the previous ref's SAD (quirk #6). */
static void set_final_search_centre(B64State *s) {
    const SvtMeParams *p = s->p;
    int16_t  hx = 0, hy = 0, cx = 0, cy = 0;
    uint64_t hsad = 0;
    for (int li = 0; li < p->num_of_list_to_search; ++li)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            if (p->temporal_layer_index > 0 || li == 0) {
                if (p->enable_hme_flag) {
                    int lvl = -1;
                    /* the three blocks run in this order; a later matching block overwrites */
                    for (int pass = 0; pass < 3; ++pass) {
                        int on = pass == 0 ? (p->enable_hme_level0_flag && !p->enable_hme_level1_flag && !p->enable_hme_level2_flag)
                            : pass == 1   ? (p->enable_hme_level1_flag && !p->enable_hme_level2_flag)
                                          : p->enable_hme_level2_flag;
                        if (!on)
                            continue;
                        lvl  = pass;
                        hx   = s->cx[lvl][li][ri][0][0];
                        hy   = s->cy[lvl][li][ri][0][0];
                        hsad = s->lsad[lvl][li][ri][0][0];
                        uint32_t w = 1, h = 0;
                        while (h < p->num_hme_sa_h) {
                            while (w < p->num_hme_sa_w) {
                                if (s->lsad[lvl][li][ri][w][h] < hsad) {
                                    hx   = s->cx[lvl][li][ri][w][h];
                                    hy   = s->cy[lvl][li][ri][w][h];
                                    hsad = s->lsad[lvl][li][ri][w][h];
                                }
                                w++;
                            }
                            w = 0;
                            h++;
                        }
                    }
                    cx = hx;
                    cy = hy;
                }
            } else {
                cx = 0;
                cy = 0;
            }
            s->sr[li][ri].hme_sc_x = cx;
            s->sr[li][ri].hme_sc_y = cy;
            s->sr[li][ri].hme_sad  = hsad;
        }
}

/* hme_prune_ref_and_adjust_sr :2547-2588 — loops cover all 2x4 slots incl. unused ones. */
static void hme_prune_ref_and_adjust_sr(B64State *s) {
    const SvtMeParams *p  = s->p;
    uint16_t           th = p->prune_ref_if_hme_sad_dev_bigger_than_th;
    if (p->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = (uint64_t)~0;
        for (int li = 0; li < NL; ++li)
            for (int ri = 0; ri < NR; ++ri)
                if (s->sr[li][ri].hme_sad < best)
                    best = s->sr[li][ri].hme_sad;
        for (int li = 0; li < NL; ++li)
            for (int ri = 1; ri < NR; ++ri)
                if ((s->sr[li][ri].hme_sad - best) * 100 > (th * best))
                    s->sr[li][ri].do_ref = 0;
    }
    if (p->enable_me_sr_adjustment)
        for (int li = 0; li < NL; ++li)
            for (int ri = 0; ri < NR; ++ri) {
                const RefResult *r = &s->sr[li][ri];
                if (IABS(r->hme_sc_x) <= p->reduce_me_sr_based_on_mv_length_th &&
                    IABS(r->hme_sc_y) <= p->reduce_me_sr_based_on_mv_length_th &&
                    r->hme_sad < p->stationary_hme_sad_abs_th)
                    s->div[li][ri] = p->stationary_me_sr_divisor;
                else if (r->hme_sad < p->reduce_me_sr_based_on_hme_sad_abs_th)
                    s->div[li][ri] = p->me_sr_divisor_for_low_hme_sad;
            }
}

/* 64 x (h/2) sub-sampled SAD at offset (dx,dy) of the full-res ref, doubled by the caller. */
static uint32_t sad64_at(const B64State *s, const OrcPlane *ref, int32_t idx) {
    return orc_nxm_sad_kernel(s->src, s->src_stride << 1, ref->buffer_y + (uint32_t)idx, ref->stride_y << 1,
                              s->b64_h >> 1, s->b64_w);
}

/* check_00_center :1139-1206 */
static uint32_t check_00_center(const B64State *s, const OrcPlane *ref, int16_t *xc, int16_t *yc, uint32_t zz) {
    int16_t  org_x = (int16_t)s->org_x, org_y = (int16_t)s->org_y;
    int16_t  W = (int16_t)ref->width, H = (int16_t)ref->height;
    uint32_t zero_sad;
    uint32_t idx = (uint32_t)((int16_t)ref->org_x + org_x + ((int16_t)ref->org_y + org_y) * (int32_t)ref->stride_y);
    if (s->p->me_early_exit_th)
        zero_sad = zz;
    else
        zero_sad = sad64_at(s, ref, (int32_t)idx);
    zero_sad <<= 1;
    *xc = ((org_x + *xc) < -63) ? -63 - org_x : *xc;
    *xc = ((org_x + *xc) > W - 1) ? *xc - ((org_x + *xc) - (W - 1)) : *xc;
    *yc = ((org_y + *yc) < -63) ? -63 - org_y : *yc;
    *yc = ((org_y + *yc) > H - 1) ? *yc - ((org_y + *yc) - (H - 1)) : *yc;
    uint64_t zero_cost = (uint64_t)(zero_sad << 8); /* COST_PRECISION 8, 32-bit shift as in the reference */
    idx                = (uint32_t)((int16_t)(ref->org_x + org_x) + *xc +
                     ((int16_t)(ref->org_y + org_y) + *yc) * (int32_t)ref->stride_y);
    uint32_t hme_sad = sad64_at(s, ref, (int32_t)idx) << 1;
    uint64_t hme_cost = (uint64_t)(hme_sad << 8);
    uint64_t best     = IMIN(zero_cost, hme_cost);
    *xc               = (best == zero_cost) ? 0 : *xc;
    *yc               = (best == zero_cost) ? 0 : *yc;
    return hme_sad;
}

/* integer_search_b64 :1249-1586 (restricted_motion_vector == 0 only) */
static void integer_search_b64(B64State *s) {
    const SvtMeParams *p = s->p;
    const int16_t org_x = (int16_t)s->org_x, org_y = (int16_t)s->org_y;
    const uint8_t sub_sad = p->me_search_method == 0; /* SUB_SAD_SEARCH definitions.h:2079 */
    for (int li = 0; li < p->num_of_list_to_search; ++li) {
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            const OrcPlane *ref = &s->refs[li][ri]->full;
            if (s->sr[li][ri].do_ref == 0)
                continue;
            int16_t  xc = s->sr[li][ri].hme_sc_x, yc = s->sr[li][ri].hme_sc_y;
            int16_t  sa_w = (int16_t)p->me_sa.sa_min.width, sa_h = (int16_t)p->me_sa.sa_min.height;
            uint16_t dist = ref_dist(s, li, ri);
            if (p->me_type != 1) /* :1302: the temporal-filter ME (ME_MCTF) uses the picture distance as it is */
                dist = scaled_dist(dist);
            sa_w          = (int16_t)IMIN((sa_w * dist), p->me_sa.sa_max.width);
            sa_h          = (int16_t)IMIN((sa_h * dist), p->me_sa.sa_max.height);
            if (p->mv_sa_adj_enabled && (!p->mv_sa_adj_nearest_ref_only || ri == 0)) {
                if (IABS(xc) > p->mv_sa_adj_mv_size_th)
                    sa_w = (int16_t)(sa_w * p->mv_sa_adj_sa_multiplier);
                if (IABS(yc) > p->mv_sa_adj_mv_size_th)
                    sa_h = (int16_t)(sa_h * p->mv_sa_adj_sa_multiplier);
            }
            sa_w = (int16_t)((IMAX(1, (int32_t)(sa_w / s->div[li][ri])) + 7) & ~0x07);
            sa_h = (int16_t)IMAX(3, (int32_t)(sa_h / s->div[li][ri]));
            const int16_t h_before = sa_h, w_before = sa_w;
            uint64_t      best_hme_sad = (uint64_t)~0;
            if (p->me_early_exit_th) {
                if (s->zz[li][ri] < (p->me_early_exit_th / 6)) {
                    sa_w = 1;
                    sa_h = 1;
                }
            } else {
                uint8_t accurate = 1;
                if ((xc != 0 || yc != 0) && p->is_ref) {
                    best_hme_sad = check_00_center(s, ref, &xc, &yc, s->zz[li][ri]);
                    if (xc == 0 && yc == 0)
                        accurate = 0;
                }
                if (p->enable_me_sr_adjustment == 2) {
                    if ((accurate && (best_hme_sad < (24 * 24))) || (p->is_ref && s->sr[li][ri].hme_sad < (24 * 24)))
                        sa_h = sa_h / 2;
                    if ((li || ri) && s->best_sad[0][0][0] < 5000 && sa_h == h_before && sa_w == w_before) {
                        sa_h = sa_h >> 1;
                        sa_w = sa_w >> 1;
                    }
                }
            }
            uint32_t *bsad = s->best_sad[li][ri], *bmv = s->best_mv[li][ri];
            orc_initialize_buffer_32bits(bsad, 21, 1, SVTME_MAX_SAD_VALUE);
            if (p->me_8x8_var_enabled && (sa_w * sa_h > 24)) {
                /* one-position probe; its result stays in the tables (:1393-1441) */
                const uint8_t *r = ref->buffer_y + (uint32_t)(((int16_t)(ref->org_x + s->org_x) + xc) +
                                                              ((int16_t)(ref->org_y + s->org_y) + yc) * (int32_t)ref->stride_y);
                orc_fullpel_search_sblock(s->src, s->src_stride, r, ref->stride_y, xc, yc, 1, 1, sub_sad, bsad, bmv);
                const uint32_t mean = bsad[0] / 64;
                uint32_t       ssq  = 0;
                for (int i = 0; i < 64; ++i) {
                    const int32_t d = (int32_t)bsad[21 + i] - (int32_t)mean;
                    ssq += (uint32_t)(d * d);
                }
                uint32_t var = ssq / 64;
                if (var > p->me_sr_mult2_th) {
                    sa_w = (int16_t)((IMAX(1, sa_w * 3 / 2) + 7) & ~0x7);
                    sa_h = (int16_t)IMAX(1, sa_h * 3 / 2);
                }
                if (var < p->me_sr_div4_th) {
                    sa_w = (int16_t)((IMAX(1, sa_w >> 2) + 7) & ~0x7);
                    sa_h = (int16_t)IMAX(1, sa_h >> 2);
                    sa_h = (int16_t)IMAX(3, sa_h);
                } else if (var < p->me_sr_div2_th) {
                    sa_w = (int16_t)((IMIN(sa_w, sa_w >> 1) + 7) & ~0x7);
                    sa_h = (int16_t)IMIN(sa_h, sa_h >> 1);
                    sa_h = (int16_t)IMAX(3, sa_h);
                }
            }
            int16_t ox = (int16_t)(xc - (sa_w >> 1)), oy = (int16_t)(yc - (sa_h >> 1));
            ox   = ((org_x + ox) < -63) ? -63 - org_x : ox;
            sa_w = ((org_x + ox) < -63) ? sa_w - (-63 - (org_x + ox)) : sa_w;
            ox   = ((org_x + ox) > s->pic_w - 1) ? ox - ((org_x + ox) - (s->pic_w - 1)) : ox;
            sa_w = ((org_x + ox + sa_w) > s->pic_w) ? IMAX(1, sa_w - ((org_x + ox + sa_w) - s->pic_w)) : sa_w;
            sa_w = (sa_w < 8) ? sa_w : sa_w & ~0x07;
            oy   = ((org_y + oy) < -63) ? -63 - org_y : oy;
            sa_h = ((org_y + oy) < -63) ? sa_h - (-63 - (org_y + oy)) : sa_h;
            oy   = ((org_y + oy) > s->pic_h - 1) ? oy - ((org_y + oy) - (s->pic_h - 1)) : oy;
            sa_h = (org_y + oy + sa_h > s->pic_h) ? IMAX(1, sa_h - ((org_y + oy + sa_h) - s->pic_h)) : sa_h;
            const uint8_t *r = ref->buffer_y + (uint32_t)(((int16_t)(ref->org_x + s->org_x) + ox) +
                                                          ((int16_t)(ref->org_y + s->org_y) + oy) * (int32_t)ref->stride_y);
            orc_fullpel_search_sblock(s->src, s->src_stride, r, ref->stride_y, ox, oy, (uint32_t)sa_w, (uint32_t)sa_h,
                                      sub_sad, bsad, bmv);
        }
    }
}

static const uint8_t k_tab8x8[64] = {0,  1,  4,  5,  16, 17, 20, 21, 2,  3,  6,  7,  18, 19, 22, 23,
                                     8,  9,  12, 13, 24, 25, 28, 29, 10, 11, 14, 15, 26, 27, 30, 31,
                                     32, 33, 36, 37, 48, 49, 52, 53, 34, 35, 38, 39, 50, 51, 54, 55,
                                     40, 41, 44, 45, 56, 57, 60, 61, 42, 43, 46, 47, 58, 59, 62, 63};

/* me_prune_ref :1592-1635 */
static void me_prune_ref(B64State *s) {
    const SvtMeParams *p = s->p;
    for (int li = 0; li < p->num_of_list_to_search; ++li)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
            s->sr[li][ri].hme_sad = 0;
            if (s->sr[li][ri].do_ref == 0) {
                s->sr[li][ri].hme_sad = (uint64_t)SVTME_MAX_SAD_VALUE * 64;
                continue;
            }
            for (int i = 0; i < 64; ++i) s->sr[li][ri].hme_sad += s->best_sad[li][ri][21 + k_tab8x8[i]];
        }
    uint16_t th = p->prune_ref_if_me_sad_dev_bigger_than_th;
    if (p->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = (uint64_t)~0;
        for (int li = 0; li < NL; ++li)
            for (int ri = 0; ri < NR; ++ri)
                if (s->sr[li][ri].hme_sad < best)
                    best = s->sr[li][ri].hme_sad;
        for (int li = 0; li < NL; ++li)
            for (int ri = 1; ri < NR; ++ri)
                if ((s->sr[li][ri].hme_sad - best) * 100 > (th * best))
                    s->sr[li][ri].do_ref = 0;
    }
}

/* z_to_raster :2590-2601 — search-order PU index -> exported raster PU index */
static const uint8_t k_z_to_raster[85] = {
    0,  1,  2,  3,  4,  5,  6,  9,  10, 7,  8,  11, 12, 13, 14, 17, 18, 15, 16, 19, 20, 21, 22, 29, 30, 23, 24, 31, 32,
    37, 38, 45, 46, 39, 40, 47, 48, 25, 26, 33, 34, 27, 28, 35, 36, 41, 42, 49, 50, 43, 44, 51, 52, 53, 54, 61, 62, 55,
    56, 63, 64, 69, 70, 77, 78, 71, 72, 79, 80, 57, 58, 65, 66, 59, 60, 67, 68, 73, 74, 81, 82, 75, 76, 83, 84};

/* MeCandidate bitfield byte (me_sb_results.h:29-35): direction:2 ref_idx_l0:2 ref_idx_l1:2
 * ref0_list:1 ref1_list:1, LSB first; the 1-bit fields keep bit 0 of what is assigned (24 -> 0). */
static uint8_t cand_byte(unsigned dir, unsigned i0, unsigned i1, unsigned l0, unsigned l1) {
    return (uint8_t)((dir & 3) | ((i0 & 3) << 2) | ((i1 & 3) << 4) | ((l0 & 1) << 6) | ((l1 & 1) << 7));
}

static uint32_t exported_pus(const SvtMeParams *p) {
    return p->enable_me_16x16 ? (p->enable_me_8x8 ? p->max_number_of_pus_per_sb : 21u) : 5u;
}

typedef struct B64Out {
    uint8_t  *total; /* [n_pu] */
    uint32_t *mv;    /* [n_pu*max_refs] */
    uint8_t  *cand;  /* [n_pu*max_cand] */
} B64Out;

static int use_pu(const SvtMeParams *p, uint32_t n_idx) {
    return p->enable_me_16x16 ? (p->enable_me_8x8 || n_idx < 21) : n_idx < 5;
}

/* construct_me_candidate_array_single_ref :2716-2766 */
static void construct_single_ref(B64State *s, B64Out *o) {
    const SvtMeParams *p = s->p;
    uint8_t do_ref = s->sr[0][0].do_ref;
    memset(o->total, 1, exported_pus(p));
    for (uint32_t n = 0; n < p->max_number_of_pus_per_sb; ++n) {
        const uint8_t pu      = k_z_to_raster[n];
        s->me_distortion[pu]  = s->best_sad[0][0][n];
        if (do_ref == 0)
            continue;
        if (use_pu(p, n)) {
            o->cand[pu * p->max_cand] = cand_byte(0, 0, 0, 0, 0);
            o->mv[pu * p->max_refs]   = s->best_mv[0][0][n];
        }
    }
}

/* construct_me_candidate_array_mrp_off :2602-2715 */
static void construct_mrp_off(B64State *s, B64Out *o, uint32_t nlists) {
    const SvtMeParams *p = s->p;
    uint8_t org[2];
    org[0] = s->sr[0][0].do_ref;
    org[1] = (nlists == 1) ? 0 : s->sr[1][0].do_ref;
    if (nlists < 2 || !s->sr[1][0].do_ref)
        nlists = 1;
    const uint32_t th = (org[0] && org[1]) ? (uint32_t)p->prune_me_candidates_th : 0;
    memset(o->total, 1, exported_pus(p));
    for (uint32_t n = 0; n < p->max_number_of_pus_per_sb; ++n) {
        const uint8_t pu   = k_z_to_raster[n];
        uint8_t       off  = 0;
        const int     use  = use_pu(p, n);
        uint8_t      *cand = use ? &o->cand[pu * p->max_cand] : NULL;
        uint8_t       dr[2] = {org[0], org[1]};
        const uint32_t best = (org[0] && org[1]) ? IMIN(s->best_sad[0][0][n], s->best_sad[1][0][n])
            : org[0]                             ? s->best_sad[0][0][n]
                                                 : s->best_sad[1][0][n];
        s->me_distortion[pu] = best;
        int8_t min_list      = -1;
        if (p->use_best_unipred_cand_only && dr[0] && dr[1])
            min_list = s->best_sad[0][0][n] < s->best_sad[1][0][n] ? 0 : 1;
        for (int li = 0; (uint32_t)li < nlists && (use || off == 0); ++li) {
            if (dr[li] == 0)
                continue;
            if (th > 0) {
                uint32_t d = (s->best_sad[li][0][n] - best) * 100;
                if (d > (best * th)) {
                    dr[li] = 0;
                    continue;
                }
            }
            if (min_list != -1 && min_list != li) {
                if (use)
                    o->mv[pu * p->max_refs + (li ? p->max_l0 : 0)] = s->best_mv[li][0][n];
                continue;
            }
            if (use) {
                cand[off] = cand_byte((unsigned)li, 0, 0, li == 0 ? (unsigned)li : 24, li == 1 ? (unsigned)li : 24);
                o->mv[pu * p->max_refs + (li ? p->max_l0 : 0)] = s->best_mv[li][0][n];
            }
            off++;
        }
        if (dr[0] && dr[1] && use) {
            cand[off]    = cand_byte(2, 0, 0, 0, 1);
            o->total[pu] = (uint8_t)(off + 1);
        }
    }
}

/* construct_me_candidate_array :2768-2905 */
static void construct_general(B64State *s, B64Out *o, uint32_t nlists) {
    const SvtMeParams *p = s->p;
    for (uint32_t n = 0; n < p->max_number_of_pus_per_sb; ++n) {
        const uint8_t pu   = (n > 4) ? k_z_to_raster[n] : (uint8_t)n;
        uint8_t       off  = 0;
        const int     use  = use_pu(p, n);
        uint8_t      *cand = use ? &o->cand[pu * p->max_cand] : NULL;
        uint8_t       dr[NL][NR];
        const uint32_t th   = (uint32_t)p->prune_me_candidates_th;
        uint32_t       best = MAXU32;
        memset(dr, 0, sizeof(dr));
        for (uint32_t li = 0; li < nlists; ++li)
            for (uint32_t ri = 0; ri < p->num_of_ref_pic_to_search[li]; ++ri) {
                dr[li][ri] = s->sr[li][ri].do_ref;
                if (dr[li][ri] == 0)
                    continue;
                best = s->best_sad[li][ri][n] < best ? s->best_sad[li][ri][n] : best;
            }
        s->me_distortion[pu] = best;
        for (uint32_t li = 0; li < nlists && (use || off == 0); ++li)
            for (uint32_t ri = 0; ri < p->num_of_ref_pic_to_search[li] && (use || off == 0); ++ri) {
                if (dr[li][ri] == 0)
                    continue;
                if (th > 0) {
                    uint32_t d = (s->best_sad[li][ri][n] - best) * 100;
                    if (d > (best * th)) {
                        dr[li][ri] = 0;
                        continue;
                    }
                }
                if (use) {
                    cand[off] = cand_byte(li, ri, ri, li == 0 ? li : 24, li == 1 ? li : 24);
                    o->mv[pu * p->max_refs + (li ? p->max_l0 : 0) + ri] = s->best_mv[li][ri][n];
                }
                off++;
            }
        if (nlists == 2 && use) {
            for (uint32_t a = 0; a < p->num_of_ref_pic_to_search[0]; ++a)
                for (uint32_t b = 0; b < p->num_of_ref_pic_to_search[1]; ++b) {
                    if (p->only_l_bwd && (a > 0 || b > 0))
                        continue;
                    if (dr[0][a] && dr[1][b])
                        cand[off++] = cand_byte(2, a, b, 0, 1);
                }
            if (!p->only_l_bwd) {
                for (uint32_t a = 1; a < p->num_of_ref_pic_to_search[0]; ++a)
                    if (dr[0][0] && dr[0][a])
                        cand[off++] = cand_byte(2, 0, a, 0, 0);
                if (p->num_of_ref_pic_to_search[1] == 3 && dr[1][0] && dr[1][2])
                    cand[off++] = cand_byte(2, 0, 2, 1, 1);
            }
        }
        if (use)
            o->total[pu] = off;
    }
}

/* compute_distortion :3034-3077 */
static void compute_distortion(const B64State *s, uint32_t b64, SvtMeOut *out) {
    const uint32_t *d  = s->me_distortion;
    uint32_t        d64 = d[0], d32 = 0, d16 = 0, d8 = 0;
    for (int i = 0; i < 4; ++i) d32 += d[1 + i];
    for (int i = 0; i < 16; ++i) d16 += d[5 + i];
    for (int i = 0; i < 64; ++i) d8 += d[21 + i];
    uint64_t mean = d8 / 64, ssq = 0;
    for (int i = 0; i < 64; ++i) {
        const int64_t diff = (int64_t)d[21 + i] - (int64_t)mean;
        ssq += (uint64_t)(diff * diff);
    }
    const uint32_t pix = s->b64_w * s->b64_h;
    if (out->me_8x8_cost_variance) out->me_8x8_cost_variance[b64] = (uint32_t)(ssq / 64);
    if (out->rc_me_distortion) out->rc_me_distortion[b64] = s->p->input_resolution_le_480p ? d8 : d16;
    if (out->me_64x64_distortion) out->me_64x64_distortion[b64] = (d64 * 4096u) / pix;
    if (out->me_32x32_distortion) out->me_32x32_distortion[b64] = (d32 * 4096u) / pix;
    if (out->me_16x16_distortion) out->me_16x16_distortion[b64] = (d16 * 4096u) / pix;
    if (out->me_8x8_distortion) out->me_8x8_distortion[b64] = (d8 * 4096u) / pix;
}

static const uint8_t k_8x8_to_16x16[64] = {5,  5,  6,  6,  7,  7,  8,  8,  5,  5,  6,  6,  7,  7,  8,  8,
                                           9,  9,  10, 10, 11, 11, 12, 12, 9,  9,  10, 10, 11, 11, 12, 12,
                                           13, 13, 14, 14, 15, 15, 16, 16, 13, 13, 14, 14, 15, 15, 16, 16,
                                           17, 17, 18, 18, 19, 19, 20, 20, 17, 17, 18, 18, 19, 19, 20, 20};
static const uint8_t k_16x16_to_32x32[16] = {1, 1, 2, 2, 1, 1, 2, 2, 3, 3, 4, 4, 3, 3, 4, 4};

/* perform_gm_detection :2908-3031. Keeps the reference's index mixing: n_idx addresses the
 * candidate array as a raster PU index and the best-MV table as a search-order index. */
static void gm_detection(const B64State *s, const B64Out *o, uint32_t b64, SvtMeOut *out) {
    const SvtMeParams *p = s->p;
    uint64_t cnt[NL][NR][2][2];
    uint64_t stationary = 0, tot = 0;
    memset(cnt, 0, sizeof(cnt));
    const int small = p->input_resolution_le_480p;
    const unsigned count = small ? 64 : 16;
    for (unsigned i = 0; i < count; ++i) {
        uint8_t n = (uint8_t)((small ? 21 : 5) + i);
        if (small) {
            if (!p->enable_me_8x8) {
                if (n >= 21) n = k_8x8_to_16x16[n - 21];
                if (!p->enable_me_16x16 && n >= 5) n = k_16x16_to_32x32[n - 5];
            }
        } else if (!p->enable_me_16x16 && n >= 5)
            n = k_16x16_to_32x32[n - 5];
        const uint8_t c   = o->cand[n * p->max_cand];
        const unsigned dir = c & 3, i0 = (c >> 2) & 3, i1 = (c >> 4) & 3, l0 = (c >> 6) & 1, l1 = (c >> 7) & 1;
        const uint32_t li = (dir == 0 || dir == 2) ? l0 : l1;
        const uint32_t ri = (dir == 0 || dir == 2) ? i0 : i1;
        const uint64_t a = p->picture_number, b = p->ref_picture_number[li][ri];
        int th;
        if (small) {
            uint16_t dist = (uint16_t)IABS((int16_t)(IMAX(a, b) - IMIN(a, b)));
            th            = p->gm_use_distance_based_active_th ? IMAX(dist >> 1, 4) : 4;
        } else {
            uint16_t dist = (uint16_t)IABS((int16_t)(a - b));
            th            = p->gm_use_distance_based_active_th ? IMAX(dist * 16, 32) : 32;
        }
        const uint32_t mv = s->best_mv[li][ri][n];
        int mx = (int16_t)(mv & 0xffff) << 2, my = (int16_t)(mv >> 16) << 2;
        if (mx < -th) cnt[li][ri][0][0]++;
        else if (mx > th) cnt[li][ri][0][1]++;
        if (my < -th) cnt[li][ri][1][0]++;
        else if (my > th) cnt[li][ri][1][1]++;
        const int sth = small ? 0 : 4;
        if (abs(mx) <= sth && abs(my) <= sth) stationary++;
        tot++;
    }
    if (stationary > ((tot * 5) / 100) && out->stationary_block_present_sb)
        out->stationary_block_present_sb[b64] = 1;
    for (int l = 0; l < NL; ++l)
        for (int r = 0; r < NR; ++r)
            for (int c = 0; c < 2; ++c)
                for (int k = 0; k < 2; ++k)
                    if (cnt[l][r][c][k] > (tot / 2) && out->rc_me_allow_gm)
                        out->rc_me_allow_gm[b64] = 1;
}

/* svt_aom_motion_estimation_b64 :3146-3223 for one block */
static void me_b64(B64State *s, uint32_t b64, SvtMeOut *out, B64Out *o) {
    const SvtMeParams *p = s->p;
    const int mctf = p->me_type == 1; /* ME_MCTF: the temporal-filter ME (temporal_filtering.c:3075) */
    const uint8_t prune_ref = p->enable_hme_flag && !mctf; /* :3173 */
    init_state(s);
    /* hme_b64 :2511-2545 */
    if (p->me_early_exit_th || p->me_safe_limit_zz_th)
        init_zz_sad(s);
    if (p->prehme_enable)
        prehme_b64(s);
    if (p->enable_hme_flag) {
        if (p->enable_hme_level0_flag) hme_level0_b64(s);
        if (p->enable_hme_level1_flag) hme_level1_b64(s);
        if (p->enable_hme_level2_flag) hme_level2_b64(s);
    }
    set_final_search_centre(s);
    if (mctf && s->sr[0][0].hme_sad < p->tf_me_exit_th) /* :3179-3183: the caller sees tf_use_pred_64x64_only_th = ~0 */
        return;
    if (prune_ref)
        hme_prune_ref_and_adjust_sr(s);
    integer_search_b64(s);
    if (prune_ref && p->enable_me_hme_ref_pruning)
        me_prune_ref(s);
    if (mctf) /* :3196: no candidate arrays, distortions or GM detection for the temporal filter */
        return;
    if (p->num_of_ref_pic_to_search[0] == 1 && p->num_of_ref_pic_to_search[1] == 0)
        construct_single_ref(s, o);
    else if (p->num_of_ref_pic_to_search[0] == 1 && p->num_of_ref_pic_to_search[1] == 1)
        construct_mrp_off(s, o, p->num_of_list_to_search);
    else
        construct_general(s, o, p->num_of_list_to_search);
    compute_distortion(s, b64, out);
    if (out->stationary_block_present_sb) out->stationary_block_present_sb[b64] = 0;
    if (out->rc_me_allow_gm) out->rc_me_allow_gm[b64] = 0;
    if (p->gm_enabled)
        gm_detection(s, o, b64, out);
}

int orc_me_picture_rows(const SvtMeParams *p, const OrcPyramid *cur,
                        const OrcPyramid *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS], SvtMeOut *out,
                        uint32_t row0, uint32_t row1) {
    const uint32_t W = cur->full.width, H = cur->full.height;
    if ((W & 7) || (H & 7) || p->max_number_of_pus_per_sb != 85)
        return -1;
    const uint32_t bw = (W + 63) / 64, bh = (H + 63) / 64;
    const uint32_t n_pu = exported_pus(p);
    B64State *s = (B64State *)calloc(1, sizeof(B64State));
    uint8_t  *tmp_total = (uint8_t *)calloc(n_pu, 1);
    uint32_t *tmp_mv    = (uint32_t *)calloc((size_t)n_pu * p->max_refs, 4);
    uint8_t  *tmp_cand  = (uint8_t *)calloc((size_t)n_pu * p->max_cand, 1);
    if (!s || !tmp_total || !tmp_mv || !tmp_cand) {
        free(s); free(tmp_total); free(tmp_mv); free(tmp_cand);
        return -1;
    }
    s->p = p;
    s->cur = cur;
    s->refs = (const OrcPyramid *(*)[NR])refs;
    s->pic_w = (int16_t)W;
    s->pic_h = (int16_t)H;
    if (row1 > bh) row1 = bh;
    for (uint32_t by = row0; by < row1; ++by)
        for (uint32_t bx = 0; bx < bw; ++bx) {
            const uint32_t b64 = bx + by * bw;
            s->org_x = bx * 64;
            s->org_y = by * 64;
            s->b64_w = (W - s->org_x) < 64 ? W - s->org_x : 64;
            s->b64_h = (H - s->org_y) < 64 ? H - s->org_y : 64;
            /* me_process.c:189-213 */
            const OrcPlane *f = &cur->full, *q = &cur->quarter, *x = &cur->sixteenth;
            s->src        = f->buffer_y + (size_t)(f->org_y + s->org_y) * f->stride_y + f->org_x + s->org_x;
            s->src_stride = f->stride_y;
            s->qsrc       = q->buffer_y + (size_t)(q->org_y + (s->org_y >> 1)) * q->stride_y + q->org_x + (s->org_x >> 1);
            s->qstride    = q->stride_y;
            s->ssrc       = x->buffer_y + (size_t)(x->org_y + (s->org_y >> 2)) * x->stride_y + x->org_x + (s->org_x >> 2);
            s->sstride    = x->stride_y;
            B64Out o;
            o.total = out->total_me_candidate_index ? out->total_me_candidate_index + (size_t)b64 * n_pu : tmp_total;
            o.mv    = out->me_mv_array ? out->me_mv_array + (size_t)b64 * n_pu * p->max_refs : tmp_mv;
            o.cand  = out->me_candidate_array ? out->me_candidate_array + (size_t)b64 * n_pu * p->max_cand : tmp_cand;
            memset(o.total, 0, n_pu);
            memset(o.mv, 0, (size_t)n_pu * p->max_refs * 4);
            memset(o.cand, 0, (size_t)n_pu * p->max_cand);
            me_b64(s, b64, out, &o);
            if (out->p_sb_best_sad) memcpy(out->p_sb_best_sad + (size_t)b64 * NL * NR * 85, s->best_sad, sizeof(s->best_sad));
            if (out->p_sb_best_mv) memcpy(out->p_sb_best_mv + (size_t)b64 * NL * NR * 85, s->best_mv, sizeof(s->best_mv));
            for (int li = 0; li < NL; ++li)
                for (int ri = 0; ri < NR; ++ri) {
                    const size_t k = ((size_t)b64 * NL + li) * NR + ri;
                    if (out->hme_sc) {
                        out->hme_sc[2 * k]     = s->sr[li][ri].hme_sc_x;
                        out->hme_sc[2 * k + 1] = s->sr[li][ri].hme_sc_y;
                    }
                    if (out->hme_sad) out->hme_sad[k] = s->sr[li][ri].hme_sad;
                    if (out->do_ref) out->do_ref[k] = s->sr[li][ri].do_ref;
                }
        }
    free(s); free(tmp_total); free(tmp_mv); free(tmp_cand);
    return 0;
}

int orc_me_picture(const SvtMeParams *p, const OrcPyramid *cur,
                   const OrcPyramid *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS], SvtMeOut *out) {
    return orc_me_picture_rows(p, cur, refs, out, 0, 0xffffffffu);
}

/* ---- all-cores form for bench.py's CPU arm when oracle/_ref is absent ("port"): (picture, b64-row segment)
 * tasks on a pthread pool, the reference's ME-thread scheme (me_process.c:97-124, enc_handle.c:443-461,768-805). */
typedef struct OrcMtJob {
    SvtMeParams       params;
    const OrcPyramid *cur;
    const OrcPyramid *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS];
    SvtMeOut          out;
} OrcMtJob;
typedef struct OrcMeMt {
    OrcMtJob *jobs;
    uint32_t  segs_per_pic, seg_rows;
    int       failed;
} OrcMeMt;
static void orc_me_mt_item(void *vctx, uint32_t item, uint32_t thread) {
    (void)thread;
    OrcMeMt *m = (OrcMeMt *)vctx;
    OrcMtJob *j = &m->jobs[item / m->segs_per_pic];
    const uint32_t seg = item % m->segs_per_pic;
    if (orc_me_picture_rows(&j->params, j->cur, j->refs, &j->out, seg * m->seg_rows, (seg + 1) * m->seg_rows))
        __atomic_store_n(&m->failed, 1, __ATOMIC_RELAXED);
}
int orc_me_pictures_mt(OrcMtJob *jobs, uint32_t n, uint32_t threads, uint32_t seg_rows) {
    if (!jobs || !n) return -1;
    if (seg_rows < 1) seg_rows = 1;
    const uint32_t bh = (jobs[0].cur->full.height + 63) / 64;
    OrcMeMt m = {jobs, (bh + seg_rows - 1) / seg_rows, seg_rows, 0};
    const int rc = mt_run(n * m.segs_per_pic, threads, orc_me_mt_item, &m);
    return (rc || m.failed) ? -1 : 0;
}
typedef struct OrcPyrMt {
    OrcPyramid *pyr;
    const uint8_t *const *luma;
    const uint64_t *nums;
    uint32_t stride, width, height;
    int hme_level1, failed;
} OrcPyrMt;
static void orc_pyr_mt_item(void *vctx, uint32_t item, uint32_t thread) {
    (void)thread;
    OrcPyrMt *m = (OrcPyrMt *)vctx;
    orc_pyramid_free(&m->pyr[item]);
    if (orc_pyramid_build(&m->pyr[item], m->luma[item], m->stride, m->width, m->height, m->hme_level1, m->nums[item]))
        __atomic_store_n(&m->failed, 1, __ATOMIC_RELAXED);
}
int orc_pyramids_build_mt(OrcPyramid *pyr, const uint8_t *const *luma, uint32_t stride, uint32_t width, uint32_t height,
                          int hme_level1, const uint64_t *picture_numbers, uint32_t n, uint32_t threads) {
    OrcPyrMt m = {pyr, luma, picture_numbers, stride, width, height, hme_level1, 0};
    const int rc = mt_run(n, threads, orc_pyr_mt_item, &m);
    return (rc || m.failed) ? -1 : 0;
}
unsigned orc_sizeof_mt_job(void) { return (unsigned)sizeof(OrcMtJob); }

unsigned orc_sizeof_params(void) { return (unsigned)sizeof(SvtMeParams); }
unsigned orc_sizeof_out(void) { return (unsigned)sizeof(SvtMeOut); }

// svtme_scalar.cuh — per-b64 state and the SCALAR control flow of svt_aom_motion_estimation_b64
// (Source/Lib/Codec/motion_estimation.c:3146-3223; citations below are into that file): search-area
// derivation, edge clipping in int16, early exits, reference pruning, candidate construction,
// statistics. Every function works on a StageState through a pointer, so the same code runs on
// a state block in global memory (one thread per b64, "scalar kernels") or in shared memory.
#pragma once
#include "svtme_search.cuh"

#define NL SVTME_NL
#define NR SVTME_NR
#define MAXU32 0xFFFFFFFFu
#define MAX_JOBS SVTME_MAX_JOBS

struct PreHme {
    uint16_t sa_w, sa_h;
    int16_t col, row;
    unsigned long long sad;
    uint8_t valid;
};

struct RefResult {
    int16_t hme_sc_x, hme_sc_y;
    unsigned long long hme_sad;
    uint8_t do_ref;
};

struct __align__(8) JobMeta { // what the post step needs to know about an HJob (8 bytes, fetched with one load)
    int8_t li, ri, a, b; // a,b: search region / quadrant indices
    int16_t ox, oy;      // search-area origin that was applied
};

struct StageState {
    SvtMeSearchAreaMinMax hme_l0_sa;
    int org_x, org_y, b64_w, b64_h, pic_w, pic_h;
    int16_t cx[3][NL][NR][2][2], cy[3][NL][NR][2][2];
    unsigned long long lsad[3][NL][NR][2][2];
    RefResult sr[NL][NR];
    uint32_t div[NL][NR];
    uint32_t zz[NL][NR];
    PreHme ph[NL][NR][2];
    uint8_t performed_phme[NL][NR][2];
    uint32_t best_sad[NL][NR][SVTME_PU_COUNT], best_mv[NL][NR][SVTME_PU_COUNT];
    uint32_t me_distortion[SVTME_PU_COUNT];
    HJob jobs[MAX_JOBS];
    JobMeta meta[MAX_JOBS];
    int n_jobs;
    int16_t fp_parked[NL][NR][4]; // integer search parked for the big-arena kernel: search origin (x, y), area (w, h)
    uint8_t fp_deferred[NL][NR];
    uint8_t job_idx[NL][NR][4]; // index of the job last posted for (list, ref, 2*b + a): a hint, verified against meta[] by its reader
    // integer-search hand-over between plan steps
    int fp_do, fp_sa_w, fp_sa_h, fp_ox, fp_oy, fp_probe, fp_phase;
    int16_t fp_xc, fp_yc;
    int c00_needed;
    int tf_exit; // ME_MCTF: search_results[0][0].hme_sad < tf_me_exit_th, the function returns after HME (:3179-3183)
};

// The slice of the state integer_search_b64 works on for ONE reference (int_kernel keeps it in
// shared memory). Member names match StageState so that the plan_int_* steps compile for both.
struct IntState {
    int org_x, org_y, b64_w, b64_h, pic_w, pic_h;
    RefResult sr[NL][NR];
    uint32_t div[NL][NR];
    uint32_t zz[NL][NR];
    uint32_t best_sad[1][1][1]; // p_sb_best_sad of (list 0, ref 0), 64x64 PU (:1345)
    FpTables tab;
    HJob jobs[2];
    JobMeta meta[2];
    int n_jobs;
    uint8_t job_idx[NL][NR][4];
    int fp_do, fp_sa_w, fp_sa_h, fp_ox, fp_oy, fp_probe, fp_phase;
    int16_t fp_xc, fp_yc;
    int c00_needed;
    uint32_t fp_var;
};

// The part of the state block the scalar steps work on (everything but the PU tables) is ~3 KB at a 9 KB stride per
// b64: to a thread walking it, every access is a new line that half of the time comes from DRAM. The step starts by
// asking L2 for all of those lines at once.
#ifndef SVTME_SCALAR_PREFETCH
#define SVTME_SCALAR_PREFETCH 0 // measured: 1080p 3.236 (on) vs 3.212 ms (off), 4K 67.3 vs 66.8 — the extra requests cost more than they hide
#endif
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_hot_state(const StageState *s) {
#if SVTME_SCALAR_PREFETCH
    const char *b = reinterpret_cast<const char *>(s);
    constexpr int head = (int)offsetof(StageState, best_sad), tail = (int)offsetof(StageState, jobs), end = (int)sizeof(StageState);
#pragma unroll
    for (int o = 0; o < head; o += 128) prefetch_l2(b + o);
    prefetch_l2(b + head - 1);
#pragma unroll
    for (int o = tail; o < end; o += 128) prefetch_l2(b + o);
    prefetch_l2(b + end - 1);
#endif
}

__device__ __forceinline__ int iabs(int v) { return v < 0 ? -v : v; }
__device__ __forceinline__ uint16_t scaled_dist(uint16_t dist) { // svt_aom_get_scaled_picture_distance :1239
    return (uint16_t)(((dist * 5) / 8) + ((dist % 8) ? 1 : 0));
}
__device__ __forceinline__ uint16_t ref_dist(const SvtMeParams &p, int li, int ri) { // get_me_reference :1232
    long long d = (long long)p.picture_number - (long long)p.ref_picture_number[li][ri];
    return (uint16_t)(int16_t)(d < 0 ? -d : d);
}

// `nj`: the caller's running job count. It lives in a register and is stored to s->n_jobs once, when the plan step is
// complete: s->n_jobs++ here was a global load that had to wait for the previous job's store, one round trip per job.
template <class S>
__device__ void add_job(S *s, int &nj, const DevPlane &pl, int x_tl, int y_tl, int sa_w, int sa_h, int skip, int li,
                        int ri, int a, int b, int16_t ox, int16_t oy) {
    const int n = nj++;
    HJob &j = s->jobs[n];
    // (x_tl, y_tl) already include the plane origin of the REFERENCE geometry; rebase on ours
    j.ref = pl.base + (size_t)y_tl * pl.stride + x_tl;
    j.stride = pl.stride;
    j.sa_w = sa_w;
    j.sa_h = sa_h;
    j.skip_even = skip;
    j.key = HME_KEY_INIT;
    s->meta[n] = JobMeta{(int8_t)li, (int8_t)ri, (int8_t)a, (int8_t)b, ox, oy};
    s->job_idx[li][ri][(2 * b + a) & 3] = (uint8_t)n;
}

// Window clipping shared by hme_level_0/1/2 (:820-1113), with the reference's no-op left/top size
// corrections kept. pad_w/pad_h are the LOGICAL limits (org-1 of the reference's planes, 63 at L2).
__device__ void clip_hme(int16_t org_x, int16_t org_y, int16_t W, int16_t H, int16_t pad_w, int16_t pad_h,
                         int16_t &sa_w, int16_t &sa_h, int16_t &ox, int16_t &oy) {
    if ((org_x + ox) < -pad_w) {
        ox = -pad_w - org_x;
        sa_w = sa_w - (-pad_w - (org_x + ox));
    }
    if ((org_x + ox) > W - 1) ox = ox - ((org_x + ox) - (W - 1));
    if ((org_x + ox + sa_w) > W) sa_w = max(1, sa_w - ((org_x + ox + sa_w) - W));
    sa_w = (sa_w < 8) ? sa_w : sa_w & ~0x07;
    if ((org_y + oy) < -pad_h) {
        oy = -pad_h - org_y;
        sa_h = sa_h - (-pad_h - (org_y + oy));
    }
    if ((org_y + oy) > H - 1) oy = oy - ((org_y + oy) - (H - 1));
    if (org_y + oy + sa_h > H) sa_h = max(1, sa_h - ((org_y + oy + sa_h) - H));
}

__device__ void set_quadrants(StageState *s, int lvl, int li, int ri, int16_t x, int16_t y, unsigned long long sad) {
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
            s->cx[lvl][li][ri][a][b] = x;
            s->cy[lvl][li][ri][a][b] = y;
            s->lsad[lvl][li][ri][a][b] = sad;
        }
}

// ---------------------------------------------------------------- plan / post steps: a TEAM of 8 lanes per b64
// The scalar bookkeeping between two search rounds (post: scatter the round's results into the state; plan: decide
// and describe the next round's searches) loops over the NL x NR = 8 (list, reference) slots. One thread per b64 walked
// those loops as one long chain of dependent global accesses, 32 b64 = 32 different cache lines per warp instruction.
// Here lane t of a team owns slot (list, ref) = (t >> 2, t & 3): the chains are 5-8x shorter, the 8 lanes of a team
// touch neighbouring fields of one state block, and the few values that cross slots (minima, the running search
// centre) move by shuffles. Jobs keep the order of the reference's loops: a lane counts the jobs its slot will post,
// an exclusive prefix sum over the team gives its first index, then it writes them.
struct Team {
    unsigned mask; // the 8 lanes of this team inside the warp
    int tl;        // lane inside the team = slot index li * NR + ri
    int li, ri;
};
__device__ __forceinline__ Team team_of() {
    const int lane = threadIdx.x & 31;
    Team t;
    t.tl = lane & 7;
    t.mask = 0xffu << (lane & 24);
    t.li = t.tl >> 2;
    t.ri = t.tl & 3;
    return t;
}
__device__ __forceinline__ uint32_t team_min(const Team &t, uint32_t v) {
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(t.mask, v, o, 8));
    return v;
}
__device__ __forceinline__ unsigned long long team_min(const Team &t, unsigned long long v) {
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        const unsigned long long u = __shfl_xor_sync(t.mask, v, o, 8);
        v = u < v ? u : v;
    }
    return v;
}
// first job index of this lane's slot and the team's total, from the per-lane job counts
__device__ __forceinline__ int team_job_base(const Team &t, int cnt, int &total) {
    int v = cnt;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const int u = __shfl_up_sync(t.mask, v, o, 8);
        if (t.tl >= o) v += u;
    }
    total = __shfl_sync(t.mask, v, 7, 8);
    return v - cnt;
}
__device__ __forceinline__ bool slot_in_range(const SvtMeParams &p, const Team &t) {
    return t.li < p.num_of_list_to_search && t.ri < p.num_of_ref_pic_to_search[t.li];
}
// The job the last plan step posted for (this slot, quadrant / region q), or -1: add_job left its index in job_idx, the
// entry is verified against meta[] (a stale index of an earlier round points at another slot's job or beyond nj).
__device__ __forceinline__ int slot_job(const StageState *s, const Team &t, int q, int nj, JobMeta &m) {
    const int k = s->job_idx[t.li][t.ri][q];
    if (k >= nj) return -1;
    m = s->meta[k];
    return (m.li == t.li && m.ri == t.ri && ((2 * m.b + m.a) & 3) == q) ? k : -1;
}

// init_me_hme_data (:3080-3140) on a state block the host has zero-filled: only the non-zero fields
__device__ __forceinline__ void team_init_state(StageState *s, const SvtMeParams &p, const Team &t) {
    s->sr[t.li][t.ri].do_ref = 1;
    s->sr[t.li][t.ri].hme_sad = MAXU32;
    s->div[t.li][t.ri] = 1;
    s->zz[t.li][t.ri] = MAXU32;
    if (t.tl == 0) s->hme_l0_sa = p.hme_l0_sa;
}

// zz-SAD jobs: one 1x1 search per reference at MV (0,0) (get_zz_sad :1737-1759). Returns the number of jobs.
__device__ __forceinline__ int team_plan_zz(StageState *s, const DevJob &J, const Team &t, int org_x, int org_y) {
    const SvtMeParams &p = J.prm;
    const bool mine = slot_in_range(p, t) && (p.temporal_layer_index > 0 || t.li == 0);
    int total, nj = team_job_base(t, mine ? 1 : 0, total);
    if (mine) {
        const DevPlane &pl = J.ref[t.li][t.ri].lvl[0];
        add_job(s, nj, pl, pl.org_x + org_x, pl.org_y + org_y, 1, 1, 0, t.li, t.ri, 0, 0, 0, 0);
    }
    return total;
}

// init_zz_sad :2452-2507
__device__ __forceinline__ void team_post_zz(StageState *s, const SvtMeParams &p, const Team &t, int nj) {
    uint32_t zcur = s->zz[t.li][t.ri], zjob = MAXU32;
    JobMeta m;
    const int k = slot_job(s, t, 0, nj, m);
    if (k >= 0) {
        uint32_t z = (uint32_t)(s->jobs[k].key >> 32) << 1;
        z = (z * 64 * 64) / (uint32_t)(s->b64_w * s->b64_h);
        s->zz[t.li][t.ri] = z;
        zcur = zjob = z;
    }
    const uint32_t best = team_min(t, zjob);
    const bool prunable = slot_in_range(p, t) && t.ri >= 1;
    if (p.temporal_layer_index > 0 && best < p.zz_sad_th && prunable)
        if ((uint32_t)((zcur - best) * 100) > (uint32_t)(p.zz_sad_pct * best)) s->sr[t.li][t.ri].do_ref = 0;
    const uint32_t zz00 = __shfl_sync(t.mask, zcur, 0, 8), zz10 = __shfl_sync(t.mask, zcur, 4, 8);
    if (p.me_safe_limit_zz_th) {
        const bool limit = p.hierarchical_levels > 0 && p.num_of_list_to_search == 2 &&
            p.temporal_layer_index >= p.hierarchical_levels && p.similar_brightness_refs && zz00 < p.me_safe_limit_zz_th &&
            zz10 < p.me_safe_limit_zz_th;
        if (limit && prunable) s->sr[t.li][t.ri].do_ref = 0;
    }
}

// check_prehme_early_exit :1763-1788
__device__ __forceinline__ bool prehme_early_exit(StageState *s, const SvtMeParams &p, int li, int ri, int sr) {
    PreHme &d = s->ph[li][ri][sr];
    if (p.me_early_exit_th && s->zz[li][ri] < p.me_early_exit_th) {
        d.col = d.row = 0;
        d.sad = 0;
        d.valid = 1;
        return true;
    }
    if (p.prehme_l1_early_exit) {
        const PreHme &l0 = s->ph[0][ri][sr];
        if (li == 1 && l0.valid && ((l0.sad < (32 * 32)) || ((iabs(l0.col) < 16) && (iabs(l0.row) < 16)))) {
            d.col = (int16_t)-l0.col;
            d.row = (int16_t)-l0.row;
            d.sad = l0.sad;
            d.valid = 1;
            return true;
        }
    }
    return false;
}

// prehme_b64 (:1791-1866): decide per (ref, region) and emit jobs (prehme_core :1641). do_list[l]: list l is planned in
// this step (both lists are one round unless list 1 waits for list 0's result, prehme_l1_early_exit). Returns the job count.
__device__ __forceinline__ int team_plan_prehme(StageState *s, const DevJob &J, const Team &t, bool do_list0, bool do_list1) {
    const SvtMeParams &p = J.prm;
    const int li = t.li, ri = t.ri;
    const bool mine = slot_in_range(p, t) && (li == 0 ? do_list0 : do_list1) && (p.temporal_layer_index > 0 || li == 0);
    unsigned emit = 0; // bit sr: region sr gets a search
    if (mine)
        for (int sr = 0; sr < 2; ++sr) {
            if (prehme_early_exit(s, p, li, ri, sr)) continue;
            PreHme &d = s->ph[li][ri][sr];
            if (!s->sr[li][ri].do_ref) {
                d.col = d.row = 0;
                d.sad = MAXU32;
                continue;
            }
            emit |= 1u << sr;
        }
    int total, nj = team_job_base(t, __popc(emit), total);
    if (emit) {
        const uint32_t factor = scaled_dist(ref_dist(p, li, ri));
        const DevPlane &pl = J.ref[li][ri].lvl[2];
        const int16_t org_x = (int16_t)s->org_x >> 2, org_y = (int16_t)s->org_y >> 2;
        const int16_t pad_w = 15, pad_h = 15; // sixteenth org (16) - 1
        const int16_t W = (int16_t)pl.width, H = (int16_t)pl.height;
        for (int sr = 0; sr < 2; ++sr) {
            if (!(emit >> sr & 1)) continue;
            PreHme &d = s->ph[li][ri][sr];
            const uint16_t dw = (uint16_t)min(p.prehme_sa_cfg[sr].sa_min.width * factor, (uint32_t)p.prehme_sa_cfg[sr].sa_max.width);
            const uint16_t dh = (uint16_t)min(p.prehme_sa_cfg[sr].sa_min.height * factor, (uint32_t)p.prehme_sa_cfg[sr].sa_max.height);
            d.sa_w = dw;
            d.sa_h = dh;
            int16_t sa_w = (int16_t)dw, sa_h = (int16_t)dh;
            int16_t ox = -(int16_t)(sa_w >> 1), oy = -(int16_t)(sa_h >> 1);
            ox = ((org_x + ox) < -pad_w) ? -pad_w - org_x : ox;
            sa_w = ((org_x + ox) < -pad_w) ? sa_w - (-pad_w - (org_x + ox)) : sa_w;
            ox = ((org_x + ox) > W - 1) ? ox - ((org_x + ox) - (W - 1)) : ox;
            sa_w = ((org_x + ox + sa_w) > W) ? max(1, sa_w - ((org_x + ox + sa_w) - W)) : sa_w;
            oy = ((org_y + oy) < -pad_h) ? -pad_h - org_y : oy;
            sa_h = ((org_y + oy) < -pad_h) ? sa_h - (-pad_h - (org_y + oy)) : sa_h;
            oy = ((org_y + oy) > H - 1) ? oy - ((org_y + oy) - (H - 1)) : oy;
            sa_h = (org_y + oy + sa_h > H) ? max(1, sa_h - ((org_y + oy + sa_h) - H)) : sa_h;
            // svt_sad_loop_kernel skips the even search rows only for a block that is 16 wide (compute_sad_c.c:74-79): the
            // 1/16 block of an incomplete b64 at the right picture edge is narrower and searches every row
            const int skip = (p.prehme_skip_search_line && (s->b64_w >> 2) == 16) ? 1 : 0;
            add_job(s, nj, pl, pl.org_x + org_x + ox, pl.org_y + org_y + oy, sa_w, sa_h, skip, li, ri, sr, 0, ox, oy);
        }
    }
    return total;
}

__device__ __forceinline__ void team_post_prehme_jobs(StageState *s, const SvtMeParams &p, const Team &t, int nj) {
    const bool full = p.hme_search_method == 1;
    for (int q = 0; q < 2; ++q) {
        JobMeta m;
        const int k = slot_job(s, t, q, nj, m);
        if (k < 0) continue;
        const unsigned long long key = s->jobs[k].key;
        PreHme &d = s->ph[t.li][t.ri][q];
        d.sad = (key >> 32) * (full ? 1 : 2);
        d.col = (int16_t)((int16_t)((int16_t)(key & 0xffff) + m.ox) * 4);
        d.row = (int16_t)((int16_t)((int16_t)((key >> 16) & 0xffff) + m.oy) * 4);
        d.valid = 1;
        s->performed_phme[t.li][t.ri][q] = 1;
    }
}

// tail of prehme_b64 :1841-1866 (runs after the team has synchronised on the post step above)
__device__ __forceinline__ void team_post_prehme_final(StageState *s, const SvtMeParams &p, const Team &t) {
    const bool inr = slot_in_range(p, t), tlok = p.temporal_layer_index > 0 || t.li == 0;
    uint32_t mval = MAXU32;
    if (inr) {
        if (tlok)
            mval = (uint32_t)min(s->ph[t.li][t.ri][0].sad, s->ph[t.li][t.ri][1].sad);
        else
            for (int sr = 0; sr < 2; ++sr) {
                s->ph[1][t.ri][sr].col = (int16_t)-s->ph[0][t.ri][sr].col;
                s->ph[1][t.ri][sr].row = (int16_t)-s->ph[0][t.ri][sr].row;
                s->ph[1][t.ri][sr].sad = s->ph[0][t.ri][sr].sad;
            }
    }
    const uint32_t best_sad = team_min(t, (inr && tlok) ? mval : MAXU32);
    if (p.temporal_layer_index > 0 && best_sad < p.phme_sad_th && inr && t.ri >= 1 && s->sr[t.li][t.ri].do_ref)
        if ((uint32_t)((mval - best_sad) * 100) > (uint32_t)(p.phme_sad_pct * best_sad)) s->sr[t.li][t.ri].do_ref = 0;
}

// get_hme_l0_search_area :1870-1935 (works on a local copy of hme_l0_sa == modify + restore)
__device__ __forceinline__ void hme_l0_area(const StageState *s, const SvtMeParams &p, int li, int ri, uint16_t dist, int16_t *sa_w,
                                            int16_t *sa_h) {
    SvtMeSearchAreaMinMax a = s->hme_l0_sa;
    if (p.enable_me_sr_adjustment && p.distance_based_hme_resizing) {
        uint8_t is_hor = 1, is_ver = 1, is_still = 0;
        if (p.reduce_hme_l0_sr_th_min && p.reduce_hme_l0_sr_th_max && (li || ri)) {
            int16_t mvx = s->cx[0][0][0][0][0], mvy = s->cy[0][0][0][0][0];
            is_ver = (iabs(mvx) < p.reduce_hme_l0_sr_th_min) && (iabs(mvy) > p.reduce_hme_l0_sr_th_max);
            is_hor = (iabs(mvx) > p.reduce_hme_l0_sr_th_max) && (iabs(mvy) < p.reduce_hme_l0_sr_th_min);
            is_still = (iabs(mvx) < (p.reduce_hme_l0_sr_th_min * 3)) && (iabs(mvy) < (p.reduce_hme_l0_sr_th_min * 3));
        }
        uint8_t xo = is_hor ? 1 : 2, yo = is_ver ? 1 : 2;
        if (p.enable_me_sr_adjustment == 2 && is_still) xo = yo = 4;
        a.sa_min.width = (uint16_t)(a.sa_min.width / (xo + ri));
        a.sa_min.height = (uint16_t)(a.sa_min.height / (yo + ri));
        a.sa_max.width = (uint16_t)(a.sa_max.width / (xo + ri));
        a.sa_max.height = (uint16_t)(a.sa_max.height / (yo + ri));
    }
    const int factor = scaled_dist(dist);
    int16_t w = (int16_t)(a.sa_min.width / p.num_hme_sa_w);
    w = (int16_t)min((((w * factor) + 15) & ~0x0F), (((a.sa_max.width / p.num_hme_sa_w) + 15) & ~0x0F));
    int16_t h = (int16_t)(a.sa_min.height / p.num_hme_sa_h);
    h = (int16_t)min((h * factor), a.sa_max.height / p.num_hme_sa_h);
    *sa_w = w;
    *sa_h = h;
}

// hme_level0_b64 :1976-2106. first_only: plan (list0, ref0) only; skip_first: everything else. Returns the job count.
__device__ __forceinline__ int team_plan_l0(StageState *s, const DevJob &J, const Team &t, bool first_only, bool skip_first) {
    const SvtMeParams &p = J.prm;
    const int li = t.li, ri = t.ri;
    const bool first = t.tl == 0;
    bool search = false;
    if (slot_in_range(p, t) && !((first_only && !first) || (skip_first && first))) {
        bool done = false;
        if (p.me_early_exit_th && s->zz[li][ri] < (p.me_early_exit_th >> 2)) {
            set_quadrants(s, 0, li, ri, 0, 0, 0);
            done = true;
        }
        if (!done && p.prev_me_stage_based_exit_th) {
            const int sr = s->ph[li][ri][0].sad <= s->ph[li][ri][1].sad ? 0 : 1;
            if (s->performed_phme[li][ri][sr] && s->ph[li][ri][sr].sad < (p.prev_me_stage_based_exit_th >> 4)) {
                set_quadrants(s, 0, li, ri, s->ph[li][ri][sr].col, s->ph[li][ri][sr].row, s->ph[li][ri][sr].sad);
                done = true;
            }
        }
        if (!done && !s->sr[li][ri].do_ref) {
            set_quadrants(s, 0, li, ri, 0, 0, MAXU32);
            done = true;
        }
        search = !done && (p.temporal_layer_index > 0 || li == 0);
    }
    int total, nj = team_job_base(t, search ? p.num_hme_sa_h * p.num_hme_sa_w : 0, total);
    if (search) {
        const DevPlane &pl = J.ref[li][ri].lvl[2];
        int16_t sa_w0 = 0, sa_h0 = 0;
        hme_l0_area(s, p, li, ri, ref_dist(p, li, ri), &sa_w0, &sa_h0);
        const int16_t org_x = (int16_t)s->org_x >> 2, org_y = (int16_t)s->org_y >> 2;
        for (int sh = 0; sh < p.num_hme_sa_h; ++sh)
            for (int sw = 0; sw < p.num_hme_sa_w; ++sw) {
                int16_t w = (int16_t)((sa_w0 + 7) & ~0x07), h = sa_h0;
                int16_t ox = (int16_t)(-(int16_t)((w * p.num_hme_sa_w) >> 1) + (int16_t)(w * sw));
                int16_t oy = (int16_t)(-(int16_t)((sa_h0 * p.num_hme_sa_h) >> 1) + (int16_t)(sa_h0 * sh));
                clip_hme(org_x, org_y, (int16_t)pl.width, (int16_t)pl.height, 15, 15, w, h, ox, oy);
                add_job(s, nj, pl, pl.org_x + org_x + ox, pl.org_y + org_y + oy, w, h, 0, li, ri, sw, sh, ox, oy);
            }
    }
    return total;
}

// results of an HME level into lsad / cx / cy of this slot; returns true when the slot had searches in the round
__device__ __forceinline__ bool team_post_level_jobs(StageState *s, const SvtMeParams &p, const Team &t, int nj, int lvl, int scale) {
    const bool full = p.hme_search_method == 1;
    bool any = false;
    JobMeta mm[4];
    int kk[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) kk[q] = slot_job(s, t, q, nj, mm[q]);
    unsigned long long key[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) key[q] = kk[q] >= 0 ? s->jobs[kk[q]].key : 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (kk[q] < 0) continue;
        const JobMeta &m = mm[q];
        const unsigned long long k = key[q];
        s->lsad[lvl][m.li][m.ri][m.a][m.b] = (k >> 32) * (full ? 1 : 2);
        s->cx[lvl][m.li][m.ri][m.a][m.b] = (int16_t)((int16_t)((int16_t)(k & 0xffff) + m.ox) * scale);
        s->cy[lvl][m.li][m.ri][m.a][m.b] = (int16_t)((int16_t)((int16_t)((k >> 16) & 0xffff) + m.oy) * scale);
        any = true;
    }
    return any;
}

// worst-quadrant replacement by the pre-HME result (:2073-2101, get_worst_quadrant :1939-1971), for a slot that searched
__device__ __forceinline__ void team_post_l0_prehme(StageState *s, const SvtMeParams &p, const Team &t) {
    if (!p.prehme_enable) return;
    const int li = t.li, ri = t.ri;
    unsigned long long(*q)[2] = s->lsad[0][li][ri];
    int bw = 0, bh = 0;
    unsigned long long mx = 0;
    if (q[0][0] > mx) { mx = q[0][0]; bw = 0; bh = 0; }
    if (q[1][0] > mx) { mx = q[1][0]; bw = 1; bh = 0; }
    if (q[0][1] > mx) { mx = q[0][1]; bw = 0; bh = 1; }
    if (q[1][1] > mx) { bw = 1; bh = 1; }
    const int sr = s->ph[li][ri][0].sad <= s->ph[li][ri][1].sad ? 0 : 1;
    if (s->ph[li][ri][sr].sad < q[bw][bh]) {
        q[bw][bh] = s->ph[li][ri][sr].sad;
        s->cx[0][li][ri][bw][bh] = s->ph[li][ri][sr].col;
        s->cy[0][li][ri][bw][bh] = s->ph[li][ri][sr].row;
    }
}

// hme_level1_b64 :2111-2192 (level 1) and hme_level2_b64 :2197-2247 (level 2: no do_ref / early-exit test in the reference)
__device__ __forceinline__ int team_plan_refine(StageState *s, const DevJob &J, const Team &t, int lvl) {
    const SvtMeParams &p = J.prm;
    const int li = t.li, ri = t.ri, prev = lvl - 1;
    unsigned emit = 0; // bit 2 * sh + sw
    if (slot_in_range(p, t) && (p.temporal_layer_index > 0 || li == 0)) {
        bool done = false;
        if (lvl == 1) {
            if (p.me_early_exit_th && s->zz[li][ri] < (p.me_early_exit_th >> 2)) {
                set_quadrants(s, 1, li, ri, 0, 0, 0);
                done = true;
            } else if (!s->sr[li][ri].do_ref) {
                set_quadrants(s, 1, li, ri, 0, 0, MAXU32);
                done = true;
            }
        }
        if (!done)
            for (int sh = 0; sh < p.num_hme_sa_h; ++sh)
                for (int sw = 0; sw < p.num_hme_sa_w; ++sw) {
                    if (p.prev_me_stage_based_exit_th && s->lsad[prev][li][ri][sw][sh] < (p.prev_me_stage_based_exit_th >> (lvl == 1 ? 5 : 2))) {
                        s->cx[lvl][li][ri][sw][sh] = s->cx[prev][li][ri][sw][sh];
                        s->cy[lvl][li][ri][sw][sh] = s->cy[prev][li][ri][sw][sh];
                        s->lsad[lvl][li][ri][sw][sh] = s->lsad[prev][li][ri][sw][sh];
                        continue;
                    }
                    emit |= 1u << (2 * sh + sw);
                }
    }
    int total, nj = team_job_base(t, __popc(emit), total);
    if (emit) {
        const DevPlane &pl = J.ref[li][ri].lvl[lvl == 1 ? 1 : 0];
        const int16_t org_x = lvl == 1 ? (int16_t)s->org_x >> 1 : (int16_t)s->org_x, org_y = lvl == 1 ? (int16_t)s->org_y >> 1 : (int16_t)s->org_y;
        const int16_t pad = lvl == 1 ? 31 : 63;
        const SvtMeSearchArea &sa = lvl == 1 ? p.hme_l1_sa : p.hme_l2_sa;
        for (int sh = 0; sh < p.num_hme_sa_h; ++sh)
            for (int sw = 0; sw < p.num_hme_sa_w; ++sw) {
                if (!(emit >> (2 * sh + sw) & 1)) continue;
                int16_t w = (int16_t)(((int16_t)sa.width + 7) & ~0x07), h = (int16_t)sa.height;
                const int16_t pcx = s->cx[prev][li][ri][sw][sh], pcy = s->cy[prev][li][ri][sw][sh];
                int16_t ox = (int16_t)(-(w >> 1) + (lvl == 1 ? (pcx >> 1) : pcx));
                int16_t oy = (int16_t)(-(h >> 1) + (lvl == 1 ? (pcy >> 1) : pcy));
                clip_hme(org_x, org_y, (int16_t)pl.width, (int16_t)pl.height, pad, pad, w, h, ox, oy);
                add_job(s, nj, pl, pl.org_x + org_x + ox, pl.org_y + org_y + oy, w, h, 0, li, ri, sw, sh, ox, oy);
            }
    }
    return total;
}

// set_final_seach_centre_sb :2252-2450. The reference walks the slots in loop order with running values (hsad and the
// centres persist from one slot to the next when a slot does not set them, e.g. list 1 of a base-layer picture keeps the
// last list-0 SAD): every lane computes its own candidate, then all lanes replay the walk over the 8 slots with shuffles
// and keep the values the walk has at their own slot. Returns this slot's (hme_sc_x, hme_sc_y, hme_sad) as stored, or the
// values in memory for a slot outside the searched range.
__device__ __forceinline__ void team_final_search_centre(StageState *s, const SvtMeParams &p, const Team &t, int16_t &out_x, int16_t &out_y,
                                                         unsigned long long &out_sad) {
    const int li = t.li, ri = t.ri;
    const bool inr = slot_in_range(p, t);
    const int pass = p.enable_hme_level2_flag ? 2 : (p.enable_hme_level1_flag ? 1 : (p.enable_hme_level0_flag ? 0 : -1));
    const bool sets = inr && (p.temporal_layer_index > 0 || li == 0) && p.enable_hme_flag && pass >= 0;
    int16_t ox = 0, oy = 0;
    unsigned long long osad = 0;
    if (sets) {
        ox = s->cx[pass][li][ri][0][0];
        oy = s->cy[pass][li][ri][0][0];
        osad = s->lsad[pass][li][ri][0][0];
        uint32_t w = 1, h = 0;
        while (h < p.num_hme_sa_h) {
            while (w < p.num_hme_sa_w) {
                if (s->lsad[pass][li][ri][w][h] < osad) {
                    ox = s->cx[pass][li][ri][w][h];
                    oy = s->cy[pass][li][ri][w][h];
                    osad = s->lsad[pass][li][ri][w][h];
                }
                w++;
            }
            w = 0;
            h++;
        }
    }
    int16_t hx = 0, hy = 0, cx = 0, cy = 0, my_x = 0, my_y = 0;
    unsigned long long hsad = 0, my_sad = 0;
    const int oxy = ((int)(uint16_t)oy << 16) | (uint16_t)ox;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int r_xy = __shfl_sync(t.mask, oxy, r, 8);
        const unsigned long long r_sad = __shfl_sync(t.mask, osad, r, 8);
        const int r_flags = __shfl_sync(t.mask, (inr ? 1 : 0) | (sets ? 2 : 0), r, 8);
        if (r_flags & 1) {
            const bool r_tlok = p.temporal_layer_index > 0 || (r >> 2) == 0;
            if (r_tlok) {
                if (p.enable_hme_flag) {
                    if (r_flags & 2) {
                        hx = (int16_t)(r_xy & 0xffff);
                        hy = (int16_t)(r_xy >> 16);
                        hsad = r_sad;
                    }
                    cx = hx;
                    cy = hy;
                }
            } else {
                cx = 0;
                cy = 0;
            }
            if (r == t.tl) {
                my_x = cx;
                my_y = cy;
                my_sad = hsad;
            }
        }
    }
    if (inr) {
        s->sr[li][ri].hme_sc_x = my_x;
        s->sr[li][ri].hme_sc_y = my_y;
        s->sr[li][ri].hme_sad = my_sad;
        out_x = my_x; out_y = my_y; out_sad = my_sad;
    } else {
        out_x = s->sr[li][ri].hme_sc_x;
        out_y = s->sr[li][ri].hme_sc_y;
        out_sad = s->sr[li][ri].hme_sad;
    }
}

// hme_prune_ref_and_adjust_sr :2547-2588 — covers all 2x4 slots incl. unused ones. (x, y, sad): this slot's search result.
__device__ __forceinline__ void team_hme_prune_and_adjust(StageState *s, const SvtMeParams &p, const Team &t, int16_t x, int16_t y,
                                                          unsigned long long sad) {
    const uint16_t th = p.prune_ref_if_hme_sad_dev_bigger_than_th;
    if (p.enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        const unsigned long long best = team_min(t, sad);
        if (t.ri >= 1 && (sad - best) * 100 > ((unsigned long long)th * best)) s->sr[t.li][t.ri].do_ref = 0;
    }
    if (p.enable_me_sr_adjustment) {
        if (iabs(x) <= p.reduce_me_sr_based_on_mv_length_th && iabs(y) <= p.reduce_me_sr_based_on_mv_length_th &&
            sad < p.stationary_hme_sad_abs_th)
            s->div[t.li][t.ri] = p.stationary_me_sr_divisor;
        else if (sad < p.reduce_me_sr_based_on_hme_sad_abs_th)
            s->div[t.li][t.ri] = p.me_sr_divisor_for_low_hme_sad;
    }
}

// integer_search_b64 :1249-1586, one reference, split at the points where a SAD search is needed.
// Step A: search-area size before the centre test; emits the two check_00_center SADs if needed.
template <class S>
__device__ void plan_int_a(S *s, const DevJob &J, int li, int ri) {
    const SvtMeParams &p = J.prm;
    s->n_jobs = 0;
    s->c00_needed = 0;
    s->fp_do = s->sr[li][ri].do_ref != 0;
    if (!s->fp_do) return;
    int16_t xc = s->sr[li][ri].hme_sc_x, yc = s->sr[li][ri].hme_sc_y;
    if (!p.me_early_exit_th && (xc != 0 || yc != 0) && p.is_ref) {
        // check_00_center :1139-1206: clip the centre, then SAD at (0,0) and at the centre
        const DevPlane &pl = J.ref[li][ri].lvl[0];
        const int16_t org_x = (int16_t)s->org_x, org_y = (int16_t)s->org_y;
        const int16_t W = (int16_t)pl.width, H = (int16_t)pl.height;
        xc = ((org_x + xc) < -63) ? -63 - org_x : xc;
        xc = ((org_x + xc) > W - 1) ? xc - ((org_x + xc) - (W - 1)) : xc;
        yc = ((org_y + yc) < -63) ? -63 - org_y : yc;
        yc = ((org_y + yc) > H - 1) ? yc - ((org_y + yc) - (H - 1)) : yc;
        int nj = 0;
        add_job(s, nj, pl, pl.org_x + org_x, pl.org_y + org_y, 1, 1, 0, li, ri, 0, 0, 0, 0);
        add_job(s, nj, pl, pl.org_x + org_x + xc, pl.org_y + org_y + yc, 1, 1, 0, li, ri, 1, 0, 0, 0);
        s->n_jobs = nj;
        s->c00_needed = 1;
    }
    s->fp_xc = xc;
    s->fp_yc = yc;
}

// Step B: finish the size logic, initialise the tables, decide on the 8x8-variance probe.
template <class S>
__device__ void plan_int_b(S *s, const DevJob &J, int li, int ri) {
    const SvtMeParams &p = J.prm;
    int16_t xc = s->fp_xc, yc = s->fp_yc;
    int16_t sa_w = (int16_t)p.me_sa.sa_min.width, sa_h = (int16_t)p.me_sa.sa_min.height;
    // (:1302: ME_MCTF, the temporal-filter ME, uses the picture distance as it is)
    const uint16_t dist = p.me_type == 1 ? ref_dist(p, li, ri) : scaled_dist(ref_dist(p, li, ri));
    sa_w = (int16_t)min((int)(sa_w * dist), (int)p.me_sa.sa_max.width);
    sa_h = (int16_t)min((int)(sa_h * dist), (int)p.me_sa.sa_max.height);
    // NB: the reference applies the mv-based scaling to the UNCLIPPED hme centre (:1309-1314)
    if (p.mv_sa_adj_enabled && (!p.mv_sa_adj_nearest_ref_only || ri == 0)) {
        if (iabs(s->sr[li][ri].hme_sc_x) > p.mv_sa_adj_mv_size_th) sa_w = (int16_t)(sa_w * p.mv_sa_adj_sa_multiplier);
        if (iabs(s->sr[li][ri].hme_sc_y) > p.mv_sa_adj_mv_size_th) sa_h = (int16_t)(sa_h * p.mv_sa_adj_sa_multiplier);
    }
    sa_w = (int16_t)((max(1u, (uint32_t)sa_w / s->div[li][ri]) + 7) & ~0x07u);
    sa_h = (int16_t)max(3u, (uint32_t)sa_h / s->div[li][ri]);
    const int16_t h_before = sa_h, w_before = sa_w;
    unsigned long long best_hme_sad = ~0ULL;
    if (p.me_early_exit_th) {
        if (s->zz[li][ri] < (p.me_early_exit_th / 6)) {
            sa_w = 1;
            sa_h = 1;
        }
    } else {
        uint8_t accurate = 1;
        if (s->c00_needed) {
            const uint32_t zero_sad = (uint32_t)(s->jobs[0].key >> 32) << 1;
            const uint32_t hme_sad = (uint32_t)(s->jobs[1].key >> 32) << 1;
            const unsigned long long zero_cost = (uint32_t)(zero_sad << 8), hme_cost = (uint32_t)(hme_sad << 8);
            const unsigned long long best = min(zero_cost, hme_cost);
            xc = (best == zero_cost) ? 0 : xc;
            yc = (best == zero_cost) ? 0 : yc;
            best_hme_sad = hme_sad;
            if (xc == 0 && yc == 0) accurate = 0;
        }
        if (p.enable_me_sr_adjustment == 2) {
            if ((accurate && (best_hme_sad < (24 * 24))) || (p.is_ref && s->sr[li][ri].hme_sad < (24 * 24))) sa_h = sa_h / 2;
            if ((li || ri) && s->best_sad[0][0][0] < 5000 && sa_h == h_before && sa_w == w_before) {
                sa_h = sa_h >> 1;
                sa_w = sa_w >> 1;
            }
        }
    }
    s->fp_xc = xc;
    s->fp_yc = yc;
    s->fp_sa_w = sa_w;
    s->fp_sa_h = sa_h;
    s->fp_probe = p.me_8x8_var_enabled && (sa_w * sa_h > 24);
}

// Variance of the 64 best 8x8 SADs after the one-point probe (:1416-1425), by the 32 lanes of one warp:
// mean = SAD64x64 / 64, sum of squared differences in uint32 arithmetic like the reference (wraps mod 2^32,
// so the order of the additions does not matter), / 64. Every lane returns the result.
template <class S>
__device__ __forceinline__ uint32_t probe_variance(const S *s) {
    const int lane = threadIdx.x & 31;
    const uint32_t mean = (uint32_t)(s->tab.key[0] >> 32) / 64;
    const int32_t d0 = (int32_t)(uint32_t)(s->tab.key[21 + lane] >> 32) - (int32_t)mean;
    const int32_t d1 = (int32_t)(uint32_t)(s->tab.key[21 + 32 + lane] >> 32) - (int32_t)mean;
    uint32_t ssq = (uint32_t)(d0 * d0) + (uint32_t)(d1 * d1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
    return ssq / 64;
}

// Step C: after the probe — 8x8-variance scaling (:1416-1441) and picture clipping (:1470-1561).
template <class S>
__device__ void plan_int_c(S *s, const DevJob &J) {
    const SvtMeParams &p = J.prm;
    int16_t sa_w = (int16_t)s->fp_sa_w, sa_h = (int16_t)s->fp_sa_h;
    const int16_t xc = s->fp_xc, yc = s->fp_yc;
    if (s->fp_probe) {
        const uint32_t var = s->fp_var; // variance of the 64 probe 8x8 SADs, computed lane-parallel by the caller (probe_variance)
        if (var > p.me_sr_mult2_th) {
            sa_w = (int16_t)((max(1, sa_w * 3 / 2) + 7) & ~0x7);
            sa_h = (int16_t)max(1, sa_h * 3 / 2);
        }
        if (var < p.me_sr_div4_th) {
            sa_w = (int16_t)((max(1, sa_w >> 2) + 7) & ~0x7);
            sa_h = (int16_t)max(1, sa_h >> 2);
            sa_h = (int16_t)max(3, (int)sa_h);
        } else if (var < p.me_sr_div2_th) {
            sa_w = (int16_t)((min((int)sa_w, sa_w >> 1) + 7) & ~0x7);
            sa_h = (int16_t)min((int)sa_h, sa_h >> 1);
            sa_h = (int16_t)max(3, (int)sa_h);
        }
    }
    const int16_t org_x = (int16_t)s->org_x, org_y = (int16_t)s->org_y;
    const int16_t W = (int16_t)s->pic_w, H = (int16_t)s->pic_h;
    int16_t ox = (int16_t)(xc - (sa_w >> 1)), oy = (int16_t)(yc - (sa_h >> 1));
    ox = ((org_x + ox) < -63) ? -63 - org_x : ox;
    sa_w = ((org_x + ox) < -63) ? sa_w - (-63 - (org_x + ox)) : sa_w;
    ox = ((org_x + ox) > W - 1) ? ox - ((org_x + ox) - (W - 1)) : ox;
    sa_w = ((org_x + ox + sa_w) > W) ? max(1, sa_w - ((org_x + ox + sa_w) - W)) : sa_w;
    sa_w = (sa_w < 8) ? sa_w : sa_w & ~0x07;
    oy = ((org_y + oy) < -63) ? -63 - org_y : oy;
    sa_h = ((org_y + oy) < -63) ? sa_h - (-63 - (org_y + oy)) : sa_h;
    oy = ((org_y + oy) > H - 1) ? oy - ((org_y + oy) - (H - 1)) : oy;
    sa_h = (org_y + oy + sa_h > H) ? max(1, sa_h - ((org_y + oy + sa_h) - H)) : sa_h;
    s->fp_sa_w = sa_w;
    s->fp_sa_h = sa_h;
    s->fp_ox = ox;
    s->fp_oy = oy;
}

__constant__ uint8_t k_tab8x8[64] = {0,  1,  4,  5,  16, 17, 20, 21, 2,  3,  6,  7,  18, 19, 22, 23,
                                     8,  9,  12, 13, 24, 25, 28, 29, 10, 11, 14, 15, 26, 27, 30, 31,
                                     32, 33, 36, 37, 48, 49, 52, 53, 34, 35, 38, 39, 50, 51, 54, 55,
                                     40, 41, 44, 45, 56, 57, 60, 61, 42, 43, 46, 47, 58, 59, 62, 63};
__constant__ uint8_t k_z_to_raster[85] = {
    0,  1,  2,  3,  4,  5,  6,  9,  10, 7,  8,  11, 12, 13, 14, 17, 18, 15, 16, 19, 20, 21, 22, 29, 30, 23, 24, 31, 32,
    37, 38, 45, 46, 39, 40, 47, 48, 25, 26, 33, 34, 27, 28, 35, 36, 41, 42, 49, 50, 43, 44, 51, 52, 53, 54, 61, 62, 55,
    56, 63, 64, 69, 70, 77, 78, 71, 72, 79, 80, 57, 58, 65, 66, 59, 60, 67, 68, 73, 74, 81, 82, 75, 76, 83, 84};
__constant__ uint8_t k_8x8_to_16x16[64] = {5,  5,  6,  6,  7,  7,  8,  8,  5,  5,  6,  6,  7,  7,  8,  8,
                                           9,  9,  10, 10, 11, 11, 12, 12, 9,  9,  10, 10, 11, 11, 12, 12,
                                           13, 13, 14, 14, 15, 15, 16, 16, 13, 13, 14, 14, 15, 15, 16, 16,
                                           17, 17, 18, 18, 19, 19, 20, 20, 17, 17, 18, 18, 19, 19, 20, 20};
__constant__ uint8_t k_16x16_to_32x32[16] = {1, 1, 2, 2, 1, 1, 2, 2, 3, 3, 4, 4, 3, 3, 4, 4};

// me_prune_ref :1592-1635 — called by ALL threads of the CTA (4 warps): one warp per reference sums its 64
// best 8x8 SADs (the reference walks them in tab8x8 order; a sum does not care), thread 0 prunes.
__device__ void me_prune_ref(StageState *s, const SvtMeParams &p) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int slot = warp; slot < NL * NR; slot += nw) {
        const int li = slot / NR, ri = slot % NR;
        if (li >= p.num_of_list_to_search || ri >= (int)p.num_of_ref_pic_to_search[li]) continue;
        if (s->sr[li][ri].do_ref == 0) {
            if (lane == 0) s->sr[li][ri].hme_sad = (unsigned long long)SVTME_MAX_SAD_VALUE * 64;
            continue;
        }
        unsigned long long v = (unsigned long long)s->best_sad[li][ri][21 + lane] + s->best_sad[li][ri][21 + 32 + lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s->sr[li][ri].hme_sad = v;
    }
    __syncthreads();
    const uint16_t th = p.prune_ref_if_me_sad_dev_bigger_than_th;
    if (threadIdx.x == 0 && p.enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        unsigned long long best = ~0ULL;
        for (int li = 0; li < NL; ++li)
            for (int ri = 0; ri < NR; ++ri)
                if (s->sr[li][ri].hme_sad < best) best = s->sr[li][ri].hme_sad;
        for (int li = 0; li < NL; ++li)
            for (int ri = 1; ri < NR; ++ri)
                if ((s->sr[li][ri].hme_sad - best) * 100 > ((unsigned long long)th * best)) s->sr[li][ri].do_ref = 0;
    }
    __syncthreads();
}

__device__ __forceinline__ uint8_t cand_byte(unsigned dir, unsigned i0, unsigned i1, unsigned l0, unsigned l1) {
    return (uint8_t)((dir & 3) | ((i0 & 3) << 2) | ((i1 & 3) << 4) | ((l0 & 1) << 6) | ((l1 & 1) << 7));
}
__device__ __forceinline__ bool use_pu(const SvtMeParams &p, uint32_t n) {
    return p.enable_me_16x16 ? (p.enable_me_8x8 || n < 21) : n < 5;
}

// construct_me_candidate_array{,_mrp_off,_single_ref} :2602-2905 — one thread per search-order PU n.
__device__ void construct_candidates(StageState *s, const SvtMeParams &p, uint32_t n, uint8_t *total, uint32_t *mvo,
                                     uint8_t *cando) {
    const uint8_t n0 = p.num_of_ref_pic_to_search[0], n1 = p.num_of_ref_pic_to_search[1];
    uint32_t nlists = p.num_of_list_to_search;
    const bool use = use_pu(p, n);
    if (n0 == 1 && n1 == 0) { // _single_ref :2716
        const uint8_t pu = k_z_to_raster[n];
        if (use) total[pu] = 1;
        s->me_distortion[pu] = s->best_sad[0][0][n];
        if (s->sr[0][0].do_ref == 0) return;
        if (use) {
            cando[pu * p.max_cand] = cand_byte(0, 0, 0, 0, 0);
            mvo[pu * p.max_refs] = s->best_mv[0][0][n];
        }
        return;
    }
    if (n0 == 1 && n1 == 1) { // _mrp_off :2602
        const uint8_t pu = k_z_to_raster[n];
        uint8_t org[2];
        org[0] = s->sr[0][0].do_ref;
        org[1] = (nlists == 1) ? 0 : s->sr[1][0].do_ref;
        if (nlists < 2 || !s->sr[1][0].do_ref) nlists = 1;
        const uint32_t th = (org[0] && org[1]) ? (uint32_t)p.prune_me_candidates_th : 0;
        if (use) total[pu] = 1;
        uint8_t off = 0;
        uint8_t *cand = use ? &cando[pu * p.max_cand] : nullptr;
        uint8_t dr[2] = {org[0], org[1]};
        const uint32_t best = (org[0] && org[1]) ? min(s->best_sad[0][0][n], s->best_sad[1][0][n])
            : org[0]                             ? s->best_sad[0][0][n]
                                                 : s->best_sad[1][0][n];
        s->me_distortion[pu] = best;
        int min_list = -1;
        if (p.use_best_unipred_cand_only && dr[0] && dr[1]) min_list = s->best_sad[0][0][n] < s->best_sad[1][0][n] ? 0 : 1;
        for (int li = 0; (uint32_t)li < nlists && (use || off == 0); ++li) {
            if (dr[li] == 0) continue;
            if (th > 0) {
                uint32_t d = (s->best_sad[li][0][n] - best) * 100;
                if (d > (best * th)) {
                    dr[li] = 0;
                    continue;
                }
            }
            if (min_list != -1 && min_list != li) {
                if (use) mvo[pu * p.max_refs + (li ? p.max_l0 : 0)] = s->best_mv[li][0][n];
                continue;
            }
            if (use) {
                cand[off] = cand_byte((unsigned)li, 0, 0, li == 0 ? (unsigned)li : 24, li == 1 ? (unsigned)li : 24);
                mvo[pu * p.max_refs + (li ? p.max_l0 : 0)] = s->best_mv[li][0][n];
            }
            off++;
        }
        if (dr[0] && dr[1] && use) {
            cand[off] = cand_byte(2, 0, 0, 0, 1);
            total[pu] = (uint8_t)(off + 1);
        }
        return;
    }
    // general :2768
    const uint8_t pu = (n > 4) ? k_z_to_raster[n] : (uint8_t)n;
    uint8_t off = 0;
    uint8_t *cand = use ? &cando[pu * p.max_cand] : nullptr;
    uint8_t dr[NL][NR];
    const uint32_t th = (uint32_t)p.prune_me_candidates_th;
    uint32_t best = MAXU32;
    for (int li = 0; li < NL; ++li)
        for (int ri = 0; ri < NR; ++ri) dr[li][ri] = 0;
    for (uint32_t li = 0; li < nlists; ++li)
        for (uint32_t ri = 0; ri < p.num_of_ref_pic_to_search[li]; ++ri) {
            dr[li][ri] = s->sr[li][ri].do_ref;
            if (dr[li][ri] == 0) continue;
            best = s->best_sad[li][ri][n] < best ? s->best_sad[li][ri][n] : best;
        }
    s->me_distortion[pu] = best;
    for (uint32_t li = 0; li < nlists && (use || off == 0); ++li)
        for (uint32_t ri = 0; ri < p.num_of_ref_pic_to_search[li] && (use || off == 0); ++ri) {
            if (dr[li][ri] == 0) continue;
            if (th > 0) {
                uint32_t d = (s->best_sad[li][ri][n] - best) * 100;
                if (d > (best * th)) {
                    dr[li][ri] = 0;
                    continue;
                }
            }
            if (use) {
                cand[off] = cand_byte(li, ri, ri, li == 0 ? li : 24, li == 1 ? li : 24);
                mvo[pu * p.max_refs + (li ? p.max_l0 : 0) + ri] = s->best_mv[li][ri][n];
            }
            off++;
        }
    if (nlists == 2 && use) {
        for (uint32_t a = 0; a < n0; ++a)
            for (uint32_t b = 0; b < n1; ++b) {
                if (p.only_l_bwd && (a > 0 || b > 0)) continue;
                if (dr[0][a] && dr[1][b]) cand[off++] = cand_byte(2, a, b, 0, 1);
            }
        if (!p.only_l_bwd) {
            for (uint32_t a = 1; a < n0; ++a)
                if (dr[0][0] && dr[0][a]) cand[off++] = cand_byte(2, 0, a, 0, 0);
            if (n1 == 3 && dr[1][0] && dr[1][2]) cand[off++] = cand_byte(2, 0, 2, 1, 1);
        }
    }
    if (use) total[pu] = off;
}

// compute_distortion :3034-3077 + perform_gm_detection :2908-3031 (thread 0)
// compute_distortion :3034-3077 + perform_gm_detection :2908-3031 — called by the 32 lanes of ONE warp: the
// sums over the 85 PU distortions are lane-parallel, lane 0 writes (and runs the rare GM histogram).
__device__ void finish_b64(StageState *s, const DevJob &J, uint32_t b64, const uint8_t *cando) {
    const SvtMeParams &p = J.prm;
    const uint32_t *d = s->me_distortion;
    const int lane = threadIdx.x & 31;
    const uint32_t d64 = d[0];
    uint32_t d32 = lane < 4 ? d[1 + lane] : 0u, d16 = lane < 16 ? d[5 + lane] : 0u;
    const uint32_t e0 = d[21 + lane], e1 = d[21 + 32 + lane];
    uint32_t d8 = e0 + e1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        d32 += __shfl_xor_sync(0xffffffffu, d32, o);
        d16 += __shfl_xor_sync(0xffffffffu, d16, o);
        d8 += __shfl_xor_sync(0xffffffffu, d8, o);
    }
    const long long mean = d8 / 64;
    const long long f0 = (long long)e0 - mean, f1 = (long long)e1 - mean;
    unsigned long long ssq = (unsigned long long)(f0 * f0) + (unsigned long long)(f1 * f1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
    if (lane != 0) return;
    const uint32_t pix = (uint32_t)(s->b64_w * s->b64_h);
    J.var8[b64] = (uint32_t)(ssq / 64);
    J.rc_dist[b64] = p.input_resolution_le_480p ? d8 : d16;
    J.dist64[b64] = (d64 * 4096u) / pix;
    J.dist32[b64] = (d32 * 4096u) / pix;
    J.dist16[b64] = (d16 * 4096u) / pix;
    J.dist8[b64] = (d8 * 4096u) / pix;
    uint8_t stationary_flag = 0, gm_flag = 0;
    if (p.gm_enabled) {
        unsigned long long cnt[NL][NR][2][2];
        unsigned long long stationary = 0, tot = 0;
        for (int a = 0; a < NL; ++a)
            for (int b = 0; b < NR; ++b)
                for (int c = 0; c < 2; ++c) cnt[a][b][c][0] = cnt[a][b][c][1] = 0;
        const bool small = p.input_resolution_le_480p;
        const unsigned count = small ? 64 : 16;
        for (unsigned i = 0; i < count; ++i) {
            uint8_t n = (uint8_t)((small ? 21 : 5) + i);
            if (small) {
                if (!p.enable_me_8x8) {
                    if (n >= 21) n = k_8x8_to_16x16[n - 21];
                    if (!p.enable_me_16x16 && n >= 5) n = k_16x16_to_32x32[n - 5];
                }
            } else if (!p.enable_me_16x16 && n >= 5)
                n = k_16x16_to_32x32[n - 5];
            const uint8_t c = cando[n * p.max_cand];
            const unsigned dir = c & 3, i0 = (c >> 2) & 3, i1 = (c >> 4) & 3, l0 = (c >> 6) & 1, l1 = (c >> 7) & 1;
            const uint32_t li = (dir == 0 || dir == 2) ? l0 : l1, ri = (dir == 0 || dir == 2) ? i0 : i1;
            const unsigned long long a = p.picture_number, b = p.ref_picture_number[li][ri];
            int th;
            if (small) {
                uint16_t dist = (uint16_t)iabs((int16_t)(max(a, b) - min(a, b)));
                th = p.gm_use_distance_based_active_th ? max(dist >> 1, 4) : 4;
            } else {
                uint16_t dist = (uint16_t)iabs((int16_t)(a - b));
                th = p.gm_use_distance_based_active_th ? max(dist * 16, 32) : 32;
            }
            const uint32_t mv = s->best_mv[li][ri][n];
            const int mx = (int16_t)(mv & 0xffff) << 2, my = (int16_t)(mv >> 16) << 2;
            if (mx < -th) cnt[li][ri][0][0]++;
            else if (mx > th) cnt[li][ri][0][1]++;
            if (my < -th) cnt[li][ri][1][0]++;
            else if (my > th) cnt[li][ri][1][1]++;
            const int sth = small ? 0 : 4;
            if (iabs(mx) <= sth && iabs(my) <= sth) stationary++;
            tot++;
        }
        if (stationary > ((tot * 5) / 100)) stationary_flag = 1;
        for (int l = 0; l < NL; ++l)
            for (int r = 0; r < NR; ++r)
                for (int c = 0; c < 2; ++c)
                    for (int k = 0; k < 2; ++k)
                        if (cnt[l][r][c][k] > (tot / 2)) gm_flag = 1;
    }
    J.stationary[b64] = stationary_flag;
    J.allow_gm[b64] = gm_flag;
}


// svtme_stage.cu — the open-loop ME stage as a PIPELINE OF WIDE KERNELS over all 64x64 blocks (b64)
// of all pictures of a batch. Device twin of svt_aom_motion_estimation_b64
// (Source/Lib/Codec/motion_estimation.c:3146-3223) called for every b64 by me_process.c:174-289.
//
// The reference's per-b64 flow alternates scalar decisions (search areas, clipping, pruning) with
// SAD searches. Run inside ONE CTA per b64 the scalar steps serialise on one thread while the rest
// of the CTA waits at a barrier (measured: ~45 % of the kernel time). Here the two kinds of work are
// separate launches over the whole batch:
//   scalar_kernel<STEP>  one THREAD per b64: plan / post steps on a state block in global memory
//   engine_kernel        one CTA per b64: executes the round of searches the last plan step posted
//                        (zz-SAD, pre-HME, HME L0 / L1 / L2) through hme_search_multi
//   int_kernel           one CTA per (b64, reference): check_00_center, 8x8-variance probe and the
//                        85-PU full-pel search (integer_search_b64 :1249-1586)
//   finish_kernel        one CTA per b64: me_prune_ref, candidate arrays, statistics
// Dependencies between launches are exactly the reference's statement order; b64s never interact.
#include <cstdlib>
#include <mutex>
#include "svtme_scalar.cuh"

struct WideArgs {
    const DevJob *jobs;
    StageState *st;
    int b64_per_job, n_b64_total;
};

size_t svtme_stage_state_bytes() { return sizeof(StageState); }
int svtme_fp_direct_max() { return SVTME_FP_DIRECT_MAX; }

// global -> shared copy of a whole struct with cp.async (all chunks of the CTA in flight at once: ONE round trip, where the
// register copy loop paid one per unrolled group of iterations). The caller waits (cp_async_wait_all) and synchronises.
template <class T>
__device__ __forceinline__ void async_copy_struct(T *dst, const T *src) {
    static_assert(sizeof(T) % 8 == 0 && alignof(T) >= 8, "copied as 8-byte chunks");
    const uint8_t *g = reinterpret_cast<const uint8_t *>(src);
    uint8_t *d = reinterpret_cast<uint8_t *>(dst);
    for (int i = threadIdx.x; i < (int)(sizeof(T) / 8); i += blockDim.x) cp_async8(d + 8 * i, g + 8 * i);
}
__device__ __forceinline__ void copy_job_to_smem(DevJob *dst, const DevJob *src) {
    async_copy_struct(dst, src);
    cp_async_wait_all(); // (a __syncthreads of the caller publishes it)
}

__host__ __device__ __forceinline__ bool l0_dependent(const SvtMeParams &p) { // :1882-1895
    return p.enable_me_sr_adjustment && p.distance_based_hme_resizing && p.reduce_hme_l0_sr_th_min &&
        p.reduce_hme_l0_sr_th_max;
}

// ---------------------------------------------------------------- scalar steps (a team of 8 lanes per b64, svtme_scalar.cuh)
enum { S_INIT = 0, S_ZZ, S_PREHME1, S_PREHME2, S_L0A, S_L0B, S_L1, S_L2, S_N };
#define SCALAR_TEAM 8

template <int STEP>
__device__ __forceinline__ void scalar_step(const WideArgs &a, int gid, const Team &t) {
    const DevJob &J = a.jobs[gid / a.b64_per_job];
    const uint32_t b64 = (uint32_t)(gid % a.b64_per_job);
    const SvtMeParams &p = J.prm;
    StageState *s = &a.st[gid];
    const bool hme_l0 = p.enable_hme_flag && p.enable_hme_level0_flag;
    const bool hme_l1 = p.enable_hme_flag && p.enable_hme_level1_flag;
    const bool hme_l2 = p.enable_hme_flag && p.enable_hme_level2_flag;
    const bool zz = p.me_early_exit_th || p.me_safe_limit_zz_th;
    // nj: the jobs of the round that just ran (every lane reads it before lane 0 replaces it at the end of the step);
    // total: the jobs this step posts for the next round. The team synchronises wherever a lane reads what another wrote.
    const int nj = STEP == S_INIT ? 0 : s->n_jobs;
    __syncwarp(t.mask);
    int total = 0;
    if (STEP == S_INIT) { // the state block was zero-filled by the host (best tables start at 0)
        const int org_x = (int)(b64 % J.b64_cols) * 64, org_y = (int)(b64 / J.b64_cols) * 64;
        if (t.tl == 0) {
            s->org_x = org_x;
            s->org_y = org_y;
            s->pic_w = J.width;
            s->pic_h = J.height;
            s->b64_w = min(64, J.width - org_x);
            s->b64_h = min(64, J.height - org_y);
        }
        team_init_state(s, p, t);
        if (zz) total = team_plan_zz(s, J, t, org_x, org_y);
    } else if (STEP == S_ZZ) {
        if (zz) team_post_zz(s, p, t, nj);
        __syncwarp(t.mask);
        // list 1 depends on list 0 only through prehme_l1_early_exit (:1763-1788): otherwise both lists are one round
        if (p.prehme_enable) total = team_plan_prehme(s, J, t, true, p.num_of_list_to_search > 1 && !p.prehme_l1_early_exit);
    } else if (STEP == S_PREHME1) {
        if (p.prehme_enable) team_post_prehme_jobs(s, p, t, nj);
        __syncwarp(t.mask);
        if (p.prehme_enable && p.num_of_list_to_search > 1 && p.prehme_l1_early_exit) total = team_plan_prehme(s, J, t, false, true);
    } else if (STEP == S_PREHME2) {
        if (p.prehme_enable) {
            team_post_prehme_jobs(s, p, t, nj);
            __syncwarp(t.mask);
            team_post_prehme_final(s, p, t);
        }
        __syncwarp(t.mask);
        if (hme_l0) total = team_plan_l0(s, J, t, l0_dependent(p), false);
    } else if (STEP == S_L0A) {
        if (hme_l0 && team_post_level_jobs(s, p, t, nj, 0, 4)) team_post_l0_prehme(s, p, t);
        __syncwarp(t.mask);
        if (hme_l0 && l0_dependent(p)) total = team_plan_l0(s, J, t, false, true);
    } else if (STEP == S_L0B) {
        if (hme_l0 && l0_dependent(p) && team_post_level_jobs(s, p, t, nj, 0, 4)) team_post_l0_prehme(s, p, t);
        __syncwarp(t.mask);
        if (hme_l1) total = team_plan_refine(s, J, t, 1);
    } else if (STEP == S_L1) {
        if (hme_l1) team_post_level_jobs(s, p, t, nj, 1, 2);
        __syncwarp(t.mask);
        if (hme_l2) total = team_plan_refine(s, J, t, 2);
    } else if (STEP == S_L2) {
        if (hme_l2) team_post_level_jobs(s, p, t, nj, 2, 1);
        __syncwarp(t.mask);
        int16_t x, y;
        unsigned long long sad;
        team_final_search_centre(s, p, t, x, y, sad);
        if (p.me_type == 1) { // ME_MCTF (:3173-3183): no pruning / area adjustment; early return when the HME SAD is small enough
            if (t.tl == 0) s->tf_exit = sad < p.tf_me_exit_th;
        } else if (p.enable_hme_flag)
            team_hme_prune_and_adjust(s, p, t, x, y, sad);
    }
    __syncwarp(t.mask);
    if (t.tl == 0) s->n_jobs = total;
    __syncwarp(t.mask);
}

// Steps [FIRST, LAST] back to back in one launch: used where no search round separates them in this batch
// (e.g. the two pre-HME post steps when list 1 does not wait for list 0).
template <int FIRST, int LAST>
__global__ void __launch_bounds__(128) scalar_kernel(WideArgs a) {
    const int gid = (int)((blockIdx.x * blockDim.x + threadIdx.x) / SCALAR_TEAM);
    if (gid >= a.n_b64_total) return; // (whole teams leave together)
    const Team t = team_of();
    scalar_step<FIRST>(a, gid, t);
    if (LAST != FIRST) scalar_step<LAST>(a, gid, t);
}

// ---------------------------------------------------------------- HME engine launch (one CTA per b64)
struct EngSmem {
    HJob jobs[MAX_JOBS];
    HPlan plan;
    DevPlane cur; // the level of the current picture this round reads its source block from
    __align__(16) uint8_t src[64 * 64];
    __align__(16) uint8_t arena[SVTME_HME_ARENA_BYTES];
};

// level: 0 = 64x64 block, every other row (zz-SAD); 1/2/3 = HME level 0/1/2 block
__global__ void __launch_bounds__(SVTME_HME_THREADS, SVTME_HME_CTAS_PER_SM) engine_kernel(WideArgs a, int level, int wslot) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    EngSmem *sm = reinterpret_cast<EngSmem *>(smem_raw);
    const int gid = blockIdx.x, tid = threadIdx.x;
    StageState *s = &a.st[gid];
    const int n = s->n_jobs;
    if (n == 0) return;
    const DevJob &J = a.jobs[gid / a.b64_per_job];
    const bool full = J.prm.hme_search_method == 1;
    const int b64_w = s->b64_w, b64_h = s->b64_h, org_x = s->org_x, org_y = s->org_y;
    const int sh = level == 1 ? 2 : (level == 2 ? 1 : 0); // down-scaling of the level
    const DevPlane &pl = J.cur.lvl[level == 1 ? 2 : (level == 2 ? 1 : 0)];
    const int side = 64 >> sh; // source block is side x side
    {   // source block -> shared memory, 16-byte chunks (b64_src_ptr / quarter / sixteenth buffers, me_process.c:189-213)
        const int cpr = side >> 4;
        for (int i = tid; i < side * cpr; i += SVTME_HME_THREADS) {
            const int r = i / cpr, c = i - r * cpr;
            *reinterpret_cast<uint4 *>(&sm->src[r * side + c * 16]) =
                *reinterpret_cast<const uint4 *>(pl.at((org_x >> sh) + c * 16, (org_y >> sh) + r));
        }
    }
    for (int i = tid; i < n * (int)(sizeof(HJob) / 4); i += SVTME_HME_THREADS)
        reinterpret_cast<uint32_t *>(sm->jobs)[i] = reinterpret_cast<const uint32_t *>(s->jobs)[i];
    __syncthreads();
    const bool sub = level == 0 || !full;
    const HBlk blk = make_hblk(sm->src, sub ? 2 * side : side, b64_w >> sh, sub ? (b64_h >> (sh + 1)) : (b64_h >> sh), sub ? 2 : 1);
    hme_search_multi(blk, sm->jobs, n, &sm->plan, sm->arena);
    unsigned long long w = 0;
    if (tid < n) {
        s->jobs[tid].key = sm->jobs[tid].key;
        const int rows = sm->jobs[tid].skip_even ? sm->jobs[tid].sa_h / 2 : sm->jobs[tid].sa_h;
        w = (unsigned long long)sm->jobs[tid].sa_w * rows * blk.bw * blk.nrows;
    }
    if (tid < 32 && J.work_ops) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if (tid == 0 && w) atomicAdd(J.work_ops + wslot, w);
    }
}

// ---------------------------------------------------------------- HME level 1 / 2 refinement, small windows
// hme_level_1 / hme_level_2 (:923-1113) refine each quadrant centre inside a window of a few dozen
// positions (8x3 at every preset but M0) with a LARGE block (32x32 of the quarter picture, 64x64 of the
// full one), up to 32 windows per b64. Here ONE WARP owns one window, no CTA barrier anywhere:
//   staging : the rows the window touches go to the warp's own shared-memory tile as RAW aligned 16-byte
//             chunks (LDG.128 -> STS.128, every reference row read once); the window's misalignment is
//             applied by the readers. (Reading the rows straight through L1 was measured 4x slower:
//             every load instruction touched 8 cache lines.)
//   search  : lane = (rq = lane >> 2: block rows rq, rq+8, ...; c = lane & 3: byte phase), positions
//             x = c and c + 4 of an 8-wide x-tile, REFINE_YT consecutive y per pass. Per (block row, y)
//             a lane reads BWW+2 words of the staged row, shifts them to its byte phase and feeds
//             2*BWW VABSDIFF4; the source row (16-byte aligned in the plane) is loaded once per block
//             row. Row pitch + a 16-byte skew every 8 rows keep the 8 row groups on distinct banks.
//   argmin  : row-group partial sums meet in a 3-step xor butterfly; the winner is the minimum over
//             keys sad << 32 | y << 16 | x (first strict minimum in raster order, compute_sad_c.c:73-95).
// Any sa_w / sa_h works (tiles of 8 x REFINE_YT positions); the host selects this kernel only when
// every picture of the batch has small refinement windows.
#define REFINE_YT 3
template <int BWW> // block width in 32-bit words: 8 = HME level 1, 16 = level 2
struct RefineCfg {
    static constexpr int NCH = BWW == 8 ? 4 : 6;                  // 16-byte chunks staged per row: 15 + 8 + 4*BWW + 3 bytes
    static constexpr int PITCH = BWW == 8 ? 80 : 112;             // bytes
    static constexpr int ROWS = (BWW == 8 ? 32 : 64) + REFINE_YT; // >= (nrows-1)*row_step + REFINE_YT for both row steps
    static constexpr int BYTES = ROWS * PITCH + 32;               // + skew + slack for the (shifted-out) last word
};
__device__ __forceinline__ int refine_rowoff(int R, int pitch) { return R * pitch + ((R >> 3) & 1) * 16; }

template <int BWW, bool FULLW>
__device__ __forceinline__ void refine_rows(const uint8_t *src0, int src_pitch, const uint8_t *tile, int boff, int rs, int nrows, int kw,
                                            int rq, int ny, uint32_t (&acc)[REFINE_YT][2]) {
    constexpr int PITCH = RefineCfg<BWW>::PITCH;
    const uint32_t shb = (uint32_t)(boff & 3) * 8;
    const uint8_t *col = tile + (boff & ~3);
    for (int r = rq; r < nrows; r += 8) {
        uint32_t sw[BWW];
        const uint4 *sp = reinterpret_cast<const uint4 *>(src0 + r * src_pitch); // shared memory: the four lanes of a row group broadcast
#pragma unroll
        for (int k = 0; k < BWW / 4; ++k) {
            const uint4 v = sp[k];
            sw[4 * k] = v.x; sw[4 * k + 1] = v.y; sw[4 * k + 2] = v.z; sw[4 * k + 3] = v.w;
        }
#pragma unroll
        for (int ty = 0; ty < REFINE_YT; ++ty) {
            if (ty < ny) {
                const uint32_t *rp = reinterpret_cast<const uint32_t *>(col + refine_rowoff(ty + r * rs, PITCH));
                uint32_t t[BWW + 2];
#pragma unroll
                for (int k = 0; k < BWW + 2; ++k) t[k] = rp[k];
#pragma unroll
                for (int k = 0; k < BWW + 1; ++k) t[k] = __funnelshift_r(t[k], t[k + 1], shb);
#pragma unroll
                for (int k = 0; k < BWW; ++k) {
                    if (FULLW || k < kw) {
                        acc[ty][0] = vsad4(sw[k], t[k], acc[ty][0]);
                        acc[ty][1] = vsad4(sw[k], t[k + 1], acc[ty][1]);
                    }
                }
            }
        }
    }
}

#ifndef REFINE_CTAS_PER_SM
#define REFINE_CTAS_PER_SM 8
#endif
struct RefineJob { // the fields of an HJob the kernel reads, staged once per CTA
    const uint8_t *ref;
    int stride, sa_w, sa_h, skip_even;
};
template <int BWW>
struct RefineSmem {
    RefineJob jobs[MAX_JOBS];
    __align__(16) uint8_t src[(BWW == 8 ? 32 : 64) * BWW * 4]; // the compared rows of the source block, pitch BWW * 4 bytes
    __align__(16) uint8_t tile[4][RefineCfg<BWW>::BYTES];
};
template <int BWW>
__global__ void __launch_bounds__(128, REFINE_CTAS_PER_SM) refine_kernel(WideArgs a, int sh, int wslot) {
    using Cfg = RefineCfg<BWW>;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    RefineSmem<BWW> *sm = reinterpret_cast<RefineSmem<BWW> *>(smem_raw);
    const int gid = blockIdx.x, tid = threadIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    StageState *s = &a.st[gid];
    const int n = s->n_jobs;
    if (n == 0) return;
    uint8_t *tile = sm->tile[warp];
    const DevJob &J = a.jobs[gid / a.b64_per_job];
    const bool full_sad = J.prm.hme_search_method == 1;
    const int rs = full_sad ? 1 : 2;
    const DevPlane &cp = J.cur.lvl[sh]; // sh = 1: quarter picture (HME level 1), 0: full picture (level 2)
    const int bw = s->b64_w >> sh, kw = bw >> 2;
    const int nrows = full_sad ? (s->b64_h >> sh) : (s->b64_h >> (sh + 1));
    {   // once per CTA: the job descriptors and the compared rows of the source block go to shared memory (the windows of a
        // b64 — up to 32 — all use the same block; reading its rows through L1 per window kept l1tex 80 % busy)
        if (tid < n) {
            const HJob &hj = s->jobs[tid];
            sm->jobs[tid] = RefineJob{hj.ref, hj.stride, hj.sa_w, hj.sa_h, hj.skip_even};
        }
        const uint8_t *g0 = cp.at(s->org_x >> sh, s->org_y >> sh); // 16-byte aligned: block origins are multiples of 32 >> sh ... of 64 >> sh
        const size_t gp = (size_t)cp.stride * rs;
        for (int u = tid; u < nrows * (BWW / 4); u += 128) {
            const int r = u / (BWW / 4), ch = u - r * (BWW / 4);
            cp_async16(sm->src + r * (BWW * 4) + 16 * ch, g0 + r * gp + 16 * ch);
        }
        cp_async_wait_all();
    }
    __syncthreads();
    const int c = lane & 3, rq = lane >> 2;
    unsigned long long work = 0;
    for (int j = warp; j < n; j += 4) {
        const RefineJob jb = sm->jobs[j];
        const uint8_t *ref = jb.ref;
        const int stride = jb.stride, sa_w = jb.sa_w, sa_h = jb.sa_h, skip_even = jb.skip_even;
        if (sa_w <= 0 || sa_h <= 0) continue;
        unsigned long long best = ~0ULL;
        for (int y0 = 0; y0 < sa_h; y0 += REFINE_YT) {
            const int ny = min(REFINE_YT, sa_h - y0);
            const int rows = (nrows - 1) * rs + ny;
            for (int x0 = 0; x0 < sa_w; x0 += 8) {
                const uintptr_t a0 = reinterpret_cast<uintptr_t>(ref + (size_t)y0 * stride + x0);
                const uint8_t *g = reinterpret_cast<const uint8_t *>(a0 & ~(uintptr_t)15);
                __syncwarp(); // the previous tile's readers are done
                for (int u = lane; u < rows * Cfg::NCH; u += 32) { // raw aligned 16-byte chunks, asynchronously (LDGSTS)
                    const int R = u / Cfg::NCH, ch = u - R * Cfg::NCH;
                    cp_async16(tile + refine_rowoff(R, Cfg::PITCH) + 16 * ch, g + (size_t)R * stride + 16 * ch);
                }
                cp_async_wait_all();
                __syncwarp();
                uint32_t acc[REFINE_YT][2];
#pragma unroll
                for (int ty = 0; ty < REFINE_YT; ++ty) acc[ty][0] = acc[ty][1] = 0;
                const int boff = (int)(a0 & 15) + c;
                if (kw == BWW)
                    refine_rows<BWW, true>(sm->src, BWW * 4, tile, boff, rs, nrows, kw, rq, ny, acc);
                else
                    refine_rows<BWW, false>(sm->src, BWW * 4, tile, boff, rs, nrows, kw, rq, ny, acc);
#pragma unroll
                for (int ty = 0; ty < REFINE_YT; ++ty)
#pragma unroll
                    for (int xh = 0; xh < 2; ++xh) {
                        uint32_t v = acc[ty][xh];
                        v += __shfl_xor_sync(0xffffffffu, v, 4);
                        v += __shfl_xor_sync(0xffffffffu, v, 8);
                        v += __shfl_xor_sync(0xffffffffu, v, 16);
                        const int y = y0 + ty, x = x0 + c + 4 * xh;
                        const bool ok = ty < ny && x < sa_w && !(skip_even && !(y & 1));
                        const unsigned long long k = ((unsigned long long)v << 32) | ((unsigned)y << 16) | (unsigned)x;
                        best = (ok && k < best) ? k : best;
                    }
            }
        }
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other < best ? other : best;
        }
        if (lane == 0) {
            if (best < HME_KEY_INIT) s->jobs[j].key = best;
            work += (unsigned long long)sa_w * (skip_even ? sa_h / 2 : sa_h) * bw * nrows;
        }
    }
    if (lane == 0 && work && J.work_ops) atomicAdd(J.work_ops + wslot, work);
}

// ---------------------------------------------------------------- pre-HME in strip mode
// The two pre-HME windows of a b64 (prehme_core :1641) are a tall 8-wide column and a flat, wide row
// around the block; the b64s of one column (one row) of the picture therefore search almost the same
// pixels of the sixteenth plane. strip_kernel gives ONE CTA a line segment of up to STRIP_SEG b64s for
// one (list, reference, region): it stages the union of their windows once and runs all their search
// items as one flattened list (thousands of items per barrier instead of a few hundred, staging
// amortised over the segment). Regions with index 0 are grouped along columns, index 1 along rows —
// a choice that only affects speed: any set of windows is handled, the segment is split until the
// union fits the arena, and a single window larger than the arena takes the tiled path.
#define STRIP_SEG 32
#ifndef STRIP_STREAM
#define STRIP_STREAM 1            // tall pre-HME windows as streamed columns (strip_columns); 0: tiled items only
#endif
#ifndef STRIP_STREAM_MIN_H
#define STRIP_STREAM_MIN_H 32     // shorter windows stay tiled: a column spends 7 extra steps on ramp-up / ramp-down
#endif
struct StripSmem {
    HJob jobs[STRIP_SEG];
    HPlan plan;                 // plan.geo[j]: window j inside the strip
    StripJob sj[STRIP_SEG];
    int gidx[STRIP_SEG], jidx[STRIP_SEG], sx[STRIP_SEG], sy[STRIP_SEG];
    HJob strip_job;             // the strip itself, as the one staging job
    HGeo strip_geo;
    int n, next, item_total, fits, col_total;
    __align__(16) uint8_t src[STRIP_SEG][8 * 16]; // the 8 compared rows (every other one) of each 16x16 sixteenth block
    __align__(16) uint8_t arena[SVTME_HME_ARENA_BYTES];
};

__device__ __forceinline__ int warp_min(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int warp_max(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// MODE 1 (HME level 0) has many short tasks (near references search 16x4 .. 16x12 positions): its
// instantiation trades registers and arena for residency (6 CTAs per SM, L0_STRIP_ARENA bytes of arena)
// so that the per-task latency chain (state -> job -> source blocks -> plan -> staging) overlaps.
#ifndef L0_STRIP_ARENA
#define L0_STRIP_ARENA (16 * 1024)
#endif
#ifndef L0_STRIP_CTAS
#define L0_STRIP_CTAS 6
#endif
template <int MODE>
#ifndef STRIP0_CTAS
#define STRIP0_CTAS 5 // measured 4 -> 5: 1080p 1.168 -> 1.147 ms, 4K 14.80 -> 14.49 ms (96 registers: the tiled items spill a little)
#endif
__global__ void __launch_bounds__(SVTME_HME_THREADS, MODE == 0 ? STRIP0_CTAS : L0_STRIP_CTAS)
    strip_kernel(WideArgs a, int cols, int rows, int wslot, int arena_limit, int stream_min_h) {
    constexpr int mode = MODE;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    StripSmem *sm = reinterpret_cast<StripSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int segs_v = (rows + STRIP_SEG - 1) / STRIP_SEG, segs_h = (cols + STRIP_SEG - 1) / STRIP_SEG;
    // mode 0: pre-HME (region 0 along columns, region 1 along rows); mode 1: HME level 0, the four quadrant
    // windows of a (list, reference) each along rows (hme_level_0 searches around (0,0): the windows of a
    // row of b64s are the same rectangle shifted by 16 pixels)
    const int n_col_tasks = mode == 0 ? cols * segs_v : 0, per_slot = n_col_tasks + (mode == 0 ? 1 : 4) * rows * segs_h;
    // Task order: kind-major, then (list, reference), then picture. CTAs resident on an SM at the same time
    // then run the same kind of window (same unrolled item body): the bodies are 6-16 KB each and mixed
    // kinds thrashed the instruction cache (ncu: "no instruction" was the top stall reason).
    const int n_row_kinds = mode == 0 ? 1 : 4, row_tasks = rows * segs_h;
    const int n_pics = (int)gridDim.x / (per_slot * NL * NR);
    int t = blockIdx.x, region, bx0, by0, len, step, slot, pic;
    const int all_col = n_col_tasks * NL * NR * n_pics;
    if (t < all_col) {
        const int task = t % n_col_tasks;
        t /= n_col_tasks;
        pic = t % n_pics; slot = t / n_pics;
        region = 0; bx0 = task / segs_v; by0 = (task % segs_v) * STRIP_SEG; len = min(STRIP_SEG, rows - by0); step = cols;
    } else if (mode == 0) {
        t -= all_col;
        const int u = t % row_tasks;
        t /= row_tasks;
        pic = t % n_pics; slot = t / n_pics;
        region = 1;
        by0 = u / segs_h; bx0 = (u % segs_h) * STRIP_SEG; len = min(STRIP_SEG, cols - bx0); step = 1;
    } else { // HME level 0: picture-major (window sizes differ per reference: mixing them balances the tail)
        const int u = t % row_tasks;
        t /= row_tasks;
        region = t % n_row_kinds; t /= n_row_kinds;
        slot = t % (NL * NR); pic = t / (NL * NR);
        by0 = u / segs_h; bx0 = (u % segs_h) * STRIP_SEG; len = min(STRIP_SEG, cols - bx0); step = 1;
    }
    const int li = slot / NR, ri = slot % NR;
    const DevJob &J = a.jobs[pic];
    if (li >= J.prm.num_of_list_to_search || ri >= (int)J.prm.num_of_ref_pic_to_search[li]) return;
    const int g0 = pic * a.b64_per_job + by0 * cols + bx0;
    // ---- the segment's jobs of this (list, reference, region), compacted (warp 0, lane <-> b64)
    if (warp == 0) {
        int found = -1;
        StageState *s = nullptr;
        if (lane < len) {
            s = &a.st[g0 + lane * step];
            // the plan step left the job's index in job_idx (a scan of meta[] here cost up to 20 dependent
            // global loads with three warps waiting at the barrier below: 19 % of the HME level 0 launch)
            const int nj = s->n_jobs, k = s->job_idx[li][ri][region];
            if (k < nj) {
                const JobMeta m = s->meta[k];
                if (m.li == li && m.ri == ri && (mode == 0 ? m.a : m.b * 2 + m.a) == region) found = k;
            }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, found >= 0);
        const int pos = __popc(bal & ((1u << lane) - 1u));
        if (found >= 0) {
            sm->jobs[pos] = s->jobs[found];
            sm->gidx[pos] = g0 + lane * step;
            sm->jidx[pos] = found;
            sm->sj[pos] = StripJob{sm->src[pos], s->b64_w >> 2, s->b64_h >> 3};
            sm->sx[pos] = s->org_x >> 2;
            sm->sy[pos] = s->org_y >> 2;
        }
        if (lane == 0) sm->n = __popc(bal);
    }
    __syncthreads();
    const int n = sm->n;
    if (n == 0) return;
    {   // sixteenth source blocks (me_process.c:205-213): the sub-sampled SAD reads every other row, only those are kept
        const DevPlane &cur = J.cur.lvl[2];
        for (int i = tid; i < n * 8; i += SVTME_HME_THREADS) {
            const int k = i >> 3, r = i & 7;
            *reinterpret_cast<uint4 *>(&sm->src[k][r * 16]) = *reinterpret_cast<const uint4 *>(cur.at(sm->sx[k], sm->sy[k] + 2 * r));
        }
    }
    const DevPlane &rp = J.ref[li][ri].lvl[2];
    const int rs = 2, src_pitch = 16; // SUB_SAD_SEARCH (the host selects strip mode only then); compact source rows
    int ra = 0;
    while (ra < n) {
        if (warp == 0) {
            const HJob jb = sm->jobs[min(lane, n - 1)];
            const int nrows_j = max(sm->sj[min(lane, n - 1)].nrows, 1);
            const uint32_t off = (uint32_t)(jb.ref - rp.base); // (a plane is far below 4 GB: 32-bit division)
            const int y_tl = (int)(off / (uint32_t)rp.stride), x_tl = (int)(off - (uint32_t)y_tl * (uint32_t)rp.stride);
            const bool live = jb.sa_w > 0 && jb.sa_h > 0;
            // tall windows (width a multiple of 8) of full-width blocks are STREAMED (strip_columns: the 8 x H pre-HME windows,
            // the HME level 0 quadrants of far references); everything else runs as tiled items
            const bool stream_ok = STRIP_STREAM && live && jb.sa_w >= 8 && (jb.sa_w & 7) == 0 && jb.sa_h >= stream_min_h &&
                sm->sj[min(lane, n - 1)].bw == 16;
            const int G = (jb.sa_w <= 8 || stream_ok) ? 2 : (jb.sa_w <= 16 ? 4 : 8);
            const int ny = (jb.sa_h + rs - 1) / rs;
            // y-tile: 2 rows for the widest items, 8 for tall narrow windows, and for 16-wide windows 2 where a
            // 4-row tile would be half empty (ny = 2, 6, 10: the 16x4 / 16x12 / 16x20 windows of HME level 0)
            const int TY = G == 8 ? 2 : ((G == 2 && ny >= 16) ? 8 : ((G == 4 && (ny & 3) != 0 && (ny & 3) <= 2 && ny <= 12) ? 2 : 4));
            const int n_yt = (ny + TY - 1) / TY;
            int rb = n, x0, y0, xs, xo, xt_n, dy, rows_s, rows_real, words, bytes, cpr, m;
            bool stream;
            for (;;) {
                const bool in = lane >= ra && lane < rb;
                x0 = warp_min(in ? x_tl : 0x7fffffff);
                y0 = warp_min(in ? y_tl : 0x7fffffff);
                const int dx = x_tl - x0;
                dy = y_tl - y0;
                // tile origin: a multiple of the tile width for the narrow tiles; 16 bytes (what the widest loads need) for
                // the 32-wide ones, so that windows 16 pixels apart (neighbouring b64) all start on a tile boundary
                xs = dx & ~(min(4 * G, 16) - 1);
                xo = dx - xs;
                stream = stream_ok && xo == 0; // (windows of one b64 column start at the same x: always)
                xt_n = (xo + jb.sa_w + 4 * G - 1) / (4 * G);
                // rows the job reads: whole y-tiles of 8-row sub-blocks (items), or the last position's last row (columns)
                rows_s = warp_max(in ? (stream ? dy + jb.sa_h + 8 * rs : dy + n_yt * TY * rs + 7 * rs) : 0); // (+ the prefetched row)
                rows_real = warp_max(in ? dy + jb.sa_h + (nrows_j - 1) * rs : 1);
                words = warp_max(in ? (xs >> 2) + xt_n * G + 5 : 0);
                cpr = (words + 3) >> 2;
                m = cpr | 1;
                bytes = (rows_s * m + (rows_s >> 3) + 1) * 16;
                if (bytes <= arena_limit || rb - ra == 1) break;
                rb = ra + (rb - ra + 1) / 2;
            }
            const bool in = lane >= ra && lane < rb;
            // flat three-row windows of full blocks: one item takes all three y-positions (hme_item_wide3)
            const bool wide3 = mode == 0 && G == 8 && jb.sa_h == 3 && !jb.skip_even && sm->sj[min(lane, n - 1)].bw == 16 &&
                sm->sj[min(lane, n - 1)].nrows == 8;
            const int n_items = (in && live && !stream) ? (wide3 ? xt_n * 4 : rs * n_yt * xt_n * 4) : 0;
            // Columns: (row parity, byte phase) = 8 per streamed job, one per thread. The jobs that fill whole passes of the
            // CTA run their columns in one piece; the columns of the r jobs left over are cut into NT / (8 r) y-chunks each,
            // so that the last pass is short instead of a full column walked by a few lanes (17 b64 rows at 1080p: 16 jobs
            // in pass one, the 17th as 16 x 8 short chunks) — the other warps would wait for it at the barrier.
            const unsigned sbal = __ballot_sync(0xffffffffu, in && stream);
            const int s_idx = __popc(sbal & ((1u << lane) - 1u)), s_tot = __popc(sbal), per_pass = SVTME_HME_THREADS / 8;
            const int s_full = (s_tot / per_pass) * per_pass;
            int chunks = 1;
            if (in && stream && jb.sa_w == 8 && s_idx >= s_full) chunks = max(1, min(per_pass / (s_tot - s_full), (jb.sa_h >> 1) / 8));
            const int n_cols = (in && stream) ? 8 * (jb.sa_w >> 3) * chunks : 0; // (wider windows give enough units by themselves)
            int ps = n_items, pc = n_cols;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, ps, o), vc = __shfl_up_sync(0xffffffffu, pc, o);
                if (lane >= o) { ps += v; pc += vc; }
            }
            const int total = __shfl_sync(0xffffffffu, ps, 31), total_c = __shfl_sync(0xffffffffu, pc, 31);
            if (in) {
                HGeo &g = sm->plan.geo[lane];
                g.arena_off = (uint32_t)xs;
                g.sad_off = 0;
                g.copy_bytes = bytes;
                g.m = m;
                g.sa_w = jb.sa_w;
                g.bh = jb.sa_h;
                g.x0 = g.y0 = 0;
                g.xt_n = xt_n;
                g.n_items = n_items;
                g.G = G;
                g.skip_even = jb.skip_even;
                g.variant = wide3 ? 5 : (G == 2 ? (TY == 8 ? 3 : 0) : (G == 4 ? (TY == 2 ? 4 : 1) : 2));
                g.mg_xt = magic_div((uint32_t)max(xt_n, 1));
                g.mg_wb = magic_div((uint32_t)max(jb.sa_w, 1));
                g.item_base = ps - n_items;
                g.row_off = dy;
                g.x_shift = xo;
                g.n_cols = n_cols;
                g.col_base = pc - n_cols;
                g.col_chunks = chunks;
            }
            if (lane == 0) {
                sm->strip_job.ref = rp.base + (size_t)y0 * rp.stride + x0;
                sm->strip_job.stride = rp.stride;
                HGeo &g = sm->strip_geo;
                g.arena_off = 0;
                g.sad_off = 0;
                g.ref_off = 0;
                g.m = m;
                g.cpr = cpr;
                g.mg_cpr = magic_div((uint32_t)cpr);
                g.rows = rows_s;
                g.rows_real = rows_real;
                g.stage_base = 0;
                g.stage_units = rows_s * cpr;
                sm->next = rb;
                sm->item_total = total;
                sm->col_total = total_c;
                sm->fits = bytes <= arena_limit;
            }
        }
        __syncthreads();
        if (sm->fits) {
            if (sm->item_total > 0 || sm->col_total > 0) {
                const HGroup SG = HGroup{0, 1, 0, sm->strip_geo.stage_units, 0, 0};
                hme_group_stage(&sm->strip_job, &sm->strip_geo, SG, sm->arena);
                __syncthreads();
                if (sm->col_total > 0) strip_columns(sm->sj, src_pitch, sm->jobs, sm->plan.geo, ra, sm->col_total, sm->arena);
                if (sm->item_total > 0) strip_group_items(sm->sj, src_pitch, rs, sm->jobs, sm->plan.geo, ra, sm->item_total, sm->arena);
            }
        } else { // one window larger than the arena: tiled like any big HME job
            const HBlk blk = make_hblk(sm->sj[ra].src, src_pitch, sm->sj[ra].bw, sm->sj[ra].nrows, rs);
            hme_search_tiled(blk, sm->jobs, ra, &sm->plan, sm->arena, arena_limit);
        }
        __syncthreads();
        ra = sm->next;
    }
    unsigned long long w = 0;
    if (tid < n) {
        a.st[sm->gidx[tid]].jobs[sm->jidx[tid]].key = sm->jobs[tid].key;
        const int rws = sm->jobs[tid].skip_even ? sm->jobs[tid].sa_h / 2 : sm->jobs[tid].sa_h;
        w = (unsigned long long)sm->jobs[tid].sa_w * rws * sm->sj[tid].bw * sm->sj[tid].nrows;
    }
    if (tid < 32 && J.work_ops) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if (tid == 0 && w) atomicAdd(J.work_ops + wslot, w);
    }
}

// ---------------------------------------------------------------- rounds of single-position SADs (zz-SAD)
// No staging, no shared memory: one warp per job reads source and reference rows straight from
// global memory (get_zz_sad :1737-1759 -> svt_nxm_sad_kernel on every other row of the 64x64 block).
__global__ void __launch_bounds__(128) points_kernel(WideArgs a) {
    const int gid = blockIdx.x;
    StageState *s = &a.st[gid];
    const int n = s->n_jobs;
    if (n == 0) return;
    const DevJob &J = a.jobs[gid / a.b64_per_job];
    const DevPlane &pl = J.cur.lvl[0];
    const HBlk blk = make_hblk(pl.at(s->org_x, s->org_y), 2 * pl.stride, s->b64_w, s->b64_h >> 1, 2);
    hme_points(blk, s->jobs, n);
    if (threadIdx.x == 0 && J.work_ops) atomicAdd(J.work_ops + 1, (unsigned long long)n * blk.bw * blk.nrows);
}

// ---------------------------------------------------------------- integer search (one CTA per b64 x reference)
// SMALL: every search area of the batch is small enough for the raw-staged full-pel path (decided on
// the host from the parameters; a larger area would still be searched correctly, only slower), so
// the CTA needs only a 10 KB arena: more CTAs per SM.
// Outcome of the planning the lean kernel does up front for every reference of its b64 (plan_int_a / _b / _c, one lane per
// reference): 16 bytes per reference, the budget that keeps 14 CTAs per SM.
struct IntPlanRec {
    int16_t fp_sa_w, fp_sa_h, fp_ox, fp_oy, fp_xc, fp_yc;
    uint8_t fp_do, fp_probe, fp_phase, _pad;
};
template <bool SMALL>
struct IntSmem {
    IntState st;
    DevJob job;
    IntPlanRec rec[SMALL ? NL * NR : 1];
    int n_c00;
    __align__(16) uint8_t src64[64 * 64];
    __align__(16) uint8_t arena[SMALL ? SVTME_FP_SMALL_ARENA_BYTES : SVTME_ARENA_BYTES];
};

// What plan_int_a / _b / _c read and write for ONE reference when a lane plans it on its own: the slot's HME results are
// read in place from the state block (global memory), the two check_00_center jobs and the plan live with the lane.
struct IntLaneView {
    int org_x, org_y, b64_w, b64_h, pic_w, pic_h;
    const RefResult (*sr)[NR];
    const uint32_t (*div)[NR];
    const uint32_t (*zz)[NR];
    const uint32_t (*best_sad)[NR][SVTME_PU_COUNT];
    HJob *jobs;
    JobMeta *meta;
    uint8_t (*job_idx)[NR][4];
    int n_jobs;
    int fp_do, fp_sa_w, fp_sa_h, fp_ox, fp_oy, fp_probe, fp_phase;
    int16_t fp_xc, fp_yc;
    int c00_needed;
    uint32_t fp_var;
};

#define T0(stmt)                        \
    do {                                \
        if (threadIdx.x == 0) { stmt; } \
        __syncthreads();                \
    } while (0)

// The (list, ref_idx) slots a launch covers: only those some picture of the batch uses. Slot 0 =
// (list 0, ref 0) runs in a launch of its own when the other references' search areas may depend on
// its result (:1345-1350, enable_me_sr_adjustment == 2); otherwise all slots are one launch.
enum { INT_ALL = 0, INT_DEFER = 1, INT_RESUME = 2 };
struct IntSlots {
    int n;
    int8_t slot[NL * NR];
};
// integer_search_b64 for ONE (list, reference) of the CTA's b64; the source block and the job descriptor are
// already in shared memory. All threads call with identical arguments.
template <bool SMALL>
__device__ __forceinline__ void int_one_ref(IntSmem<SMALL> *sm, StageState *g, int li, int ri, int mode, unsigned long long &work) {
    const int tid = threadIdx.x;
    IntState *s = &sm->st;
    const DevJob &J = sm->job;
    const SvtMeParams &p = J.prm;
    __syncthreads(); // the previous reference of this CTA is done with the state and the tables
    if (tid == 0) {
        s->org_x = g->org_x; s->org_y = g->org_y;
        s->b64_w = g->b64_w; s->b64_h = g->b64_h;
        s->pic_w = g->pic_w; s->pic_h = g->pic_h;
        s->sr[li][ri] = g->sr[li][ri];
        s->div[li][ri] = g->div[li][ri];
        s->zz[li][ri] = g->zz[li][ri];
        s->best_sad[0][0][0] = g->best_sad[0][0][0];
        if (mode == INT_RESUME) {
            s->fp_ox = g->fp_parked[li][ri][0]; s->fp_oy = g->fp_parked[li][ri][1];
            s->fp_sa_w = g->fp_parked[li][ri][2]; s->fp_sa_h = g->fp_parked[li][ri][3];
            s->fp_phase = 2;
            s->n_jobs = 0;
        }
    }
    const int org_x = g->org_x, org_y = g->org_y;
    // svt_initialize_buffer_32bits(p_sb_best_sad, 21, 1, MAX_SAD_VALUE) :1368
    for (int pu = tid; pu < SVTME_PU_COUNT; pu += (int)blockDim.x) {
        // resumed job: the parked entries rank before every search position (order 0), like the probe's did
        s->tab.key[pu] = (unsigned long long)(mode == INT_RESUME ? g->best_sad[li][ri][pu] : SVTME_MAX_SAD_VALUE) << 32;
        s->tab.mv[pu] = mode == INT_RESUME ? g->best_mv[li][ri][pu] : 0;
    }
    __syncthreads();
    // The three scalar steps run back to back on thread 0 whenever no search separates them
    // (no check_00_center, no 8x8-variance probe): one serial section instead of three.
    if (mode != INT_RESUME)
    T0(plan_int_a(s, J, li, ri); s->fp_phase = 0;
       if (!s->n_jobs) {
           plan_int_b(s, J, li, ri);
           s->fp_phase = 1;
           if (!s->fp_probe) {
               plan_int_c(s, J);
               s->fp_phase = 2;
           }
       });
    const bool sub_sad = p.me_search_method == 0;
    const DevPlane &pl = J.ref[li][ri].lvl[0];
    if (s->fp_phase == 0) { // check_00_center: two single-position SADs
        const HBlk blk_zz = make_hblk(sm->src64, 128, s->b64_w, s->b64_h >> 1, 2);
        hme_points(blk_zz, s->jobs, s->n_jobs);
        work += (unsigned long long)s->n_jobs * blk_zz.bw * blk_zz.nrows;
        T0(plan_int_b(s, J, li, ri); s->fp_phase = 1;
           if (!s->fp_probe) {
               plan_int_c(s, J);
               s->fp_phase = 2;
           });
    }
    if (s->fp_phase == 1) {
        const int xc = s->fp_xc, yc = s->fp_yc;
        fullpel_probe(sm->src64, pl.at(org_x + xc, org_y + yc), pl.stride, xc, yc, sub_sad, 1, &s->tab);
        work += 64ull * (sub_sad ? 32 : 64);
        if (tid < 32) {
            const uint32_t var = probe_variance(s);
            if (tid == 0) {
                s->fp_var = var;
                plan_int_c(s, J);
            }
        }
        __syncthreads();
    }
    const bool park = mode == INT_DEFER && s->fp_sa_w * s->fp_sa_h > SVTME_FP_DIRECT_MAX;
    if (!park) {
        const int ox = s->fp_ox, oy = s->fp_oy;
        fullpel_search(sm->src64, pl.at(org_x + ox, org_y + oy), pl.stride, ox, oy, s->fp_sa_w, s->fp_sa_h, sub_sad, 2,
                       &s->tab, sm->arena, (int)sizeof(sm->arena));
        work += (unsigned long long)s->fp_sa_w * s->fp_sa_h * 64ull * (sub_sad ? 32 : 64);
    } else if (tid == 0) {
        g->fp_parked[li][ri][0] = (int16_t)s->fp_ox; g->fp_parked[li][ri][1] = (int16_t)s->fp_oy;
        g->fp_parked[li][ri][2] = (int16_t)s->fp_sa_w; g->fp_parked[li][ri][3] = (int16_t)s->fp_sa_h;
        g->fp_deferred[li][ri] = 1;
    }
    for (int pu = tid; pu < SVTME_PU_COUNT; pu += (int)blockDim.x) {
        g->best_sad[li][ri][pu] = (uint32_t)(s->tab.key[pu] >> 32);
        g->best_mv[li][ri][pu] = s->tab.mv[pu];
    }
}

// The lean kernel's planning, hoisted: lane k of warp 0 plans the k-th reference of the launch's slot list (plan_int_a),
// the check_00_center SADs of ALL references that need them run as one round of points (two jobs per reference, compacted
// at the start of the arena, which no search uses yet), then the lanes finish (plan_int_b, and plan_int_c where no probe
// follows). Before: every reference in turn copied its slot state to shared memory and planned on thread 0 with the
// other 63 threads at a barrier — a global round trip, ~250 single-lane instructions and three barriers per reference.
template <bool SMALL>
__device__ __forceinline__ void int_plan_all(IntSmem<SMALL> *sm, const StageState *g, const IntSlots &sl, unsigned todo, unsigned long long &work) {
    const int tid = threadIdx.x, lane = tid & 31;
    const DevJob &J = sm->job;
    HJob *c00_jobs = reinterpret_cast<HJob *>(sm->arena);               // [2 * NL * NR]
    JobMeta *junk_meta = reinterpret_cast<JobMeta *>(sm->arena + 1024); // add_job's bookkeeping nobody reads here
    uint8_t(*junk_idx)[NR][4] = reinterpret_cast<uint8_t(*)[NR][4]>(sm->arena + 1024 + 512);
    IntLaneView v;
    HJob mine[2];
    bool c00 = false, planned = false;
    int li = 0, ri = 0, base = 0;
    __syncthreads(); // the job descriptor and the source block were staged by all threads (cp.async, each waited for its own)
    if (tid < 32) {
        planned = lane < sl.n && (todo >> lane & 1);
        if (planned) {
            li = sl.slot[lane] / NR;
            ri = sl.slot[lane] % NR;
            v.org_x = g->org_x; v.org_y = g->org_y;
            v.b64_w = g->b64_w; v.b64_h = g->b64_h;
            v.pic_w = g->pic_w; v.pic_h = g->pic_h;
            v.sr = g->sr; v.div = g->div; v.zz = g->zz; v.best_sad = g->best_sad;
            v.jobs = mine;
            v.meta = junk_meta + 2 * lane;
            v.job_idx = junk_idx;
            v.fp_sa_w = v.fp_sa_h = v.fp_ox = v.fp_oy = v.fp_probe = 0;
            v.fp_xc = v.fp_yc = 0;
            v.fp_var = 0;
            plan_int_a(&v, J, li, ri);
            v.fp_phase = 0;
            c00 = v.n_jobs == 2;
        }
        const unsigned cmask = __ballot_sync(0xffffffffu, c00);
        base = 2 * __popc(cmask & ((1u << lane) - 1u));
        if (c00) {
            c00_jobs[base] = mine[0];
            c00_jobs[base + 1] = mine[1];
        }
        if (lane == 0) sm->n_c00 = 2 * __popc(cmask);
    }
    __syncthreads();
    const int n_c00 = sm->n_c00;
    if (n_c00) { // check_00_center: two single-position SADs per reference
        const HBlk blk_zz = make_hblk(sm->src64, 128, g->b64_w, g->b64_h >> 1, 2);
        hme_points(blk_zz, c00_jobs, n_c00); // (ends with a barrier)
        work += (unsigned long long)n_c00 * blk_zz.bw * blk_zz.nrows;
    }
    if (planned) {
        if (c00) v.jobs = c00_jobs + base;
        plan_int_b(&v, J, li, ri);
        v.fp_phase = 1;
        if (!v.fp_probe) {
            plan_int_c(&v, J);
            v.fp_phase = 2;
        }
        IntPlanRec &r = sm->rec[lane];
        r.fp_sa_w = (int16_t)v.fp_sa_w; r.fp_sa_h = (int16_t)v.fp_sa_h;
        r.fp_ox = (int16_t)v.fp_ox; r.fp_oy = (int16_t)v.fp_oy;
        r.fp_xc = v.fp_xc; r.fp_yc = v.fp_yc;
        r.fp_do = (uint8_t)v.fp_do; r.fp_probe = (uint8_t)v.fp_probe; r.fp_phase = (uint8_t)v.fp_phase;
    }
    __syncthreads();
}

// integer_search_b64 for ONE (list, reference) whose plan is in sm->rec[k]: tables, the 8x8-variance probe, the search
// (or parking it for the big-arena kernel), the write-back. All threads call with identical arguments.
template <bool SMALL>
__device__ __forceinline__ void int_one_ref_planned(IntSmem<SMALL> *sm, StageState *g, int k, int li, int ri, int mode, unsigned long long &work) {
    const int tid = threadIdx.x;
    IntState *s = &sm->st;
    const DevJob &J = sm->job;
    const SvtMeParams &p = J.prm;
    __syncthreads(); // the previous reference of this CTA is done with the state and the tables
    if (tid == 0) {
        const IntPlanRec &r = sm->rec[k];
        s->org_x = g->org_x; s->org_y = g->org_y;
        s->pic_w = g->pic_w; s->pic_h = g->pic_h;
        s->fp_do = r.fp_do; s->fp_probe = r.fp_probe; s->fp_phase = r.fp_phase;
        s->fp_sa_w = r.fp_sa_w; s->fp_sa_h = r.fp_sa_h; s->fp_ox = r.fp_ox; s->fp_oy = r.fp_oy;
        s->fp_xc = r.fp_xc; s->fp_yc = r.fp_yc;
    }
    const int org_x = g->org_x, org_y = g->org_y;
    // svt_initialize_buffer_32bits(p_sb_best_sad, 21, 1, MAX_SAD_VALUE) :1368
    for (int pu = tid; pu < SVTME_PU_COUNT; pu += (int)blockDim.x) {
        s->tab.key[pu] = (unsigned long long)SVTME_MAX_SAD_VALUE << 32;
        s->tab.mv[pu] = 0;
    }
    __syncthreads();
    const bool sub_sad = p.me_search_method == 0;
    const DevPlane &pl = J.ref[li][ri].lvl[0];
    if (s->fp_phase == 1) {
        const int xc = s->fp_xc, yc = s->fp_yc;
        fullpel_probe(sm->src64, pl.at(org_x + xc, org_y + yc), pl.stride, xc, yc, sub_sad, 1, &s->tab);
        work += 64ull * (sub_sad ? 32 : 64);
        if (tid < 32) {
            const uint32_t var = probe_variance(s);
            if (tid == 0) {
                s->fp_var = var;
                plan_int_c(s, J);
            }
        }
        __syncthreads();
    }
    const bool park = mode == INT_DEFER && s->fp_sa_w * s->fp_sa_h > SVTME_FP_DIRECT_MAX;
    if (!park) {
        const int ox = s->fp_ox, oy = s->fp_oy;
        fullpel_search(sm->src64, pl.at(org_x + ox, org_y + oy), pl.stride, ox, oy, s->fp_sa_w, s->fp_sa_h, sub_sad, 2,
                       &s->tab, sm->arena, (int)sizeof(sm->arena));
        work += (unsigned long long)s->fp_sa_w * s->fp_sa_h * 64ull * (sub_sad ? 32 : 64);
    } else if (tid == 0) {
        g->fp_parked[li][ri][0] = (int16_t)s->fp_ox; g->fp_parked[li][ri][1] = (int16_t)s->fp_oy;
        g->fp_parked[li][ri][2] = (int16_t)s->fp_sa_w; g->fp_parked[li][ri][3] = (int16_t)s->fp_sa_h;
        g->fp_deferred[li][ri] = 1;
    }
    for (int pu = tid; pu < SVTME_PU_COUNT; pu += (int)blockDim.x) {
        g->best_sad[li][ri][pu] = (uint32_t)(s->tab.key[pu] >> 32);
        g->best_mv[li][ri][pu] = s->tab.mv[pu];
    }
}

// One CTA per b64: job descriptor and source block go to shared memory once, then the references of the
// launch's slot list run one after the other (slot order = the reference's loop order, so a reference that
// depends on (list 0, ref 0)'s result, :1345-1350, finds it in the state block).
template <bool SMALL>
__global__ void __launch_bounds__(SMALL ? SVTME_INT_SMALL_THREADS : SVTME_THREADS, SMALL ? SVTME_INT_SMALL_CTAS_PER_SM : SVTME_CTAS_PER_SM)
    int_kernel(WideArgs a, IntSlots sl, int mode) {
    // mode INT_ALL: the whole function. A batch whose areas CAN be large (presets <= 6) still has mostly small
    // ones (stationary blocks divide the area by 8): the lean kernel runs first in mode INT_DEFER — plan, probe,
    // and the search if the area turned out small, else it parks tables + area in the state block — and the
    // big-arena kernel follows in mode INT_RESUME for the parked jobs only.
    extern __shared__ __align__(16) uint8_t smem_raw[];
    IntSmem<SMALL> *sm = reinterpret_cast<IntSmem<SMALL> *>(smem_raw);
    const int gid = blockIdx.x, tid = threadIdx.x;
    const DevJob &Jg = a.jobs[gid / a.b64_per_job];
    StageState *g = &a.st[gid];
    // which slots of the list have work: lane k looks at slot k (one round trip to the state block instead of a serial
    // scan by every thread — 30 % of the stall samples of the resume launch at 4K, where most b64 have nothing parked)
    unsigned todo = 0; // bit per slot of the list
    {
        const int k = tid & 31;
        bool mine = false;
        if (k < sl.n) {
            const int li = sl.slot[k] / NR, ri = sl.slot[k] % NR;
            mine = !g->tf_exit && li < Jg.prm.num_of_list_to_search && ri < Jg.prm.num_of_ref_pic_to_search[li] && g->sr[li][ri].do_ref &&
                (mode != INT_RESUME || g->fp_deferred[li][ri]); // (pruned reference: its tables keep the zero fill, :1287-1290)
        }
        todo = __ballot_sync(0xffffffffu, mine);
    }
    if (!todo) return;
    copy_job_to_smem(&sm->job, &Jg);
    {
        const DevPlane &f = Jg.cur.lvl[0];
        const int org_x = g->org_x, org_y = g->org_y;
        for (int i = tid; i < 256; i += (int)blockDim.x) { // 64 rows x 4 chunks of 16 bytes, asynchronously (LDGSTS)
            const int r = i >> 2, c4 = i & 3;
            cp_async16(&sm->src64[r * 64 + c4 * 16], f.at(org_x + c4 * 16, org_y + r));
        }
        cp_async_wait_all(); // (the first __syncthreads of int_one_ref publishes the block to the CTA)
    }
    unsigned long long work = 0;
    if (SMALL && mode != INT_RESUME) { // (the lean kernel: every reference planned up front, then one after the other)
        int_plan_all<SMALL>(sm, g, sl, todo, work);
        for (int k = 0; k < sl.n; ++k)
            if (todo >> k & 1) int_one_ref_planned<SMALL>(sm, g, k, sl.slot[k] / NR, sl.slot[k] % NR, mode, work);
    } else
        for (int k = 0; k < sl.n; ++k)
            if (todo >> k & 1) int_one_ref<SMALL>(sm, g, sl.slot[k] / NR, sl.slot[k] % NR, mode, work);
    if (tid == 0 && Jg.work_ops) atomicAdd(Jg.work_ops + 6, work);
}

// ---------------------------------------------------------------- pruning, candidates, statistics (one CTA per b64)
struct FinSmem {
    StageState st;
    DevJob job;
};

__global__ void __launch_bounds__(128) finish_kernel(WideArgs a) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    FinSmem *sm = reinterpret_cast<FinSmem *>(smem_raw);
    const int gid = blockIdx.x, tid = threadIdx.x;
    const uint32_t b64 = (uint32_t)(gid % a.b64_per_job);
    // job descriptor and whole state block -> shared memory (cp.async: one round trip), then the reference's scalar tail
    // runs on the copy
    async_copy_struct(&sm->job, &a.jobs[gid / a.b64_per_job]);
    async_copy_struct(&sm->st, &a.st[gid]);
    cp_async_wait_all();
    __syncthreads();
    StageState *s = &sm->st;
    const DevJob &J = sm->job;
    const SvtMeParams &p = J.prm;
    // ---- pruning, candidates, statistics :3192-3221 (none of it for ME_MCTF, :3173,3196: the temporal filter reads
    // p_sb_best_sad / p_sb_best_mv and the HME results, i.e. the deep state below)
    if (p.me_type != 1) {
        if (p.enable_hme_flag && p.enable_me_hme_ref_pruning) me_prune_ref(s, p); // (CTA-uniform condition; syncs inside)
        uint8_t *total = J.total + (size_t)b64 * J.n_pu;
        uint32_t *mvo = J.mv + (size_t)b64 * J.n_pu * p.max_refs;
        uint8_t *cando = J.cand + (size_t)b64 * J.n_pu * p.max_cand;
        if (tid < p.max_number_of_pus_per_sb) construct_candidates(s, p, (uint32_t)tid, total, mvo, cando);
        __threadfence_block();
        __syncthreads();
        if (tid < 32) finish_b64(s, J, b64, cando);
    }
    // deep state for parity checks
    if (J.best_sad)
        for (int i = tid; i < NL * NR * SVTME_PU_COUNT; i += blockDim.x) {
            J.best_sad[(size_t)b64 * NL * NR * SVTME_PU_COUNT + i] = (&s->best_sad[0][0][0])[i];
            J.best_mv[(size_t)b64 * NL * NR * SVTME_PU_COUNT + i] = (&s->best_mv[0][0][0])[i];
        }
    if (J.hme_sc && tid < NL * NR) {
        const int li = tid / NR, ri = tid % NR;
        const size_t k = (size_t)b64 * NL * NR + tid;
        J.hme_sc[2 * k] = s->sr[li][ri].hme_sc_x;
        J.hme_sc[2 * k + 1] = s->sr[li][ri].hme_sc_y;
        J.hme_sad[k] = s->sr[li][ri].hme_sad;
        J.do_ref[k] = s->sr[li][ri].do_ref;
    }
}

// ---------------------------------------------------------------- host: the launch sequence
// Arena bytes strip_kernel may fill; lowered only by tests (SVTME_STRIP_ARENA_LIMIT in the environment,
// read once) to exercise the segment-splitting and tiled paths with ordinary pictures.
static int strip_arena_limit_from_env() {
    const char *e = getenv("SVTME_STRIP_ARENA_LIMIT");
    const int v = e ? atoi(e) : 0;
    return (v >= 1024 && v <= SVTME_HME_ARENA_BYTES) ? v : SVTME_HME_ARENA_BYTES;
}
static const int g_strip_arena_limit = strip_arena_limit_from_env();
// Minimum window height for streamed columns; SVTME_STRIP_STREAM_MIN_H in the environment (read once) lets tests and
// A/B timings move it (a huge value turns streaming off: every window runs as tiled items).
static int strip_stream_min_h_from_env() {
    const char *e = getenv("SVTME_STRIP_STREAM_MIN_H");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : STRIP_STREAM_MIN_H;
}
static const int g_strip_stream_min_h = strip_stream_min_h_from_env();
// HME level 0: streaming its quadrant windows (16 x 4 ... 48 x 48, 2 to 24 positions per column and parity) was measured
// SLOWER than the tiled items at every threshold (1080p 0.60 -> 0.65-0.66 ms, 4K 11.7 -> 12.2 ms: a column of n positions
// spends n + 7 steps, and the narrow tile geometry streaming needs makes the 16 KB strips shorter): off unless
// SVTME_L0_STREAM_MIN_H says otherwise.
static int l0_stream_min_h_from_env() {
    const char *e = getenv("SVTME_L0_STREAM_MIN_H");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : (1 << 30);
}
static const int g_l0_stream_min_h = l0_stream_min_h_from_env();

// Launch slots (one event after each; names in SVTME_PHASE_NAMES of svtme_common.cuh).
SvtmeStageFlags svtme_stage_flags(const SvtMeParams &p) {
    SvtmeStageFlags f;
    f.small_areas = false; // filled in by the caller (svtme_api.cu)
    f.zz = p.me_early_exit_th || p.me_safe_limit_zz_th;
    f.prehme = p.prehme_enable != 0;
    f.prehme_l1_round = f.prehme && p.num_of_list_to_search > 1 && p.prehme_l1_early_exit;
    f.hme_l0 = p.enable_hme_flag && p.enable_hme_level0_flag;
    f.hme_l0_round_b = f.hme_l0 && l0_dependent(p);
    f.hme_l1 = p.enable_hme_flag && p.enable_hme_level1_flag;
    f.hme_l2 = p.enable_hme_flag && p.enable_hme_level2_flag;
    f.int_dependent = p.enable_me_sr_adjustment == 2; // plan_int_b reads p_sb_best_sad[0][0][0] (:1345-1350)
    f.max_refs[0] = (int)p.num_of_ref_pic_to_search[0];
    f.max_refs[1] = p.num_of_list_to_search > 1 ? (int)p.num_of_ref_pic_to_search[1] : 0;
    // refinement windows of at most 64 positions (8x3 everywhere but M0) take the warp-per-window kernel
    f.refine_small = (!f.hme_l1 || ((p.hme_l1_sa.width + 7u) & ~7u) * p.hme_l1_sa.height <= 64u) &&
        (!f.hme_l2 || ((p.hme_l2_sa.width + 7u) & ~7u) * p.hme_l2_sa.height <= 64u);
    return f;
}

// cudaFuncSetAttribute acts on the CURRENT device: once per device, under a lock (a process may open handles on several
// GPUs and launch from several ME threads).
struct DgAcc;
__global__ void dg_hme_kernel(DevPlane src16, DevPlane ref16, int cols, int rows, int sa, DgAcc *acc); // (defined with the DG-detector section below)
void svtme_configure_stage_kernels() {
    static std::mutex mu;
    static bool done[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
    std::lock_guard<std::mutex> lk(mu);
    if (done[dev]) return;
    cudaFuncSetAttribute(engine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EngSmem));
    cudaFuncSetAttribute(dg_hme_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EngSmem));
    cudaFuncSetAttribute(strip_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(StripSmem));
    cudaFuncSetAttribute(strip_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(StripSmem));
    cudaFuncSetAttribute(int_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IntSmem<false>));
    cudaFuncSetAttribute(int_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IntSmem<true>));
    cudaFuncSetAttribute(finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FinSmem));
    done[dev] = true;
}

void svtme_launch_me_stage(const DevJob *d_jobs, int n_jobs, int b64_cols, int b64_rows, void *d_states, cudaStream_t st,
                           cudaEvent_t *ev, const SvtmeStageFlags &f, cudaEvent_t *sev, uint32_t *sev_mask) {
    const int b64_per_job = b64_cols * b64_rows;
    svtme_configure_stage_kernels();
    WideArgs a;
    a.jobs = d_jobs;
    a.st = reinterpret_cast<StageState *>(d_states);
    a.b64_per_job = b64_per_job;
    a.n_b64_total = n_jobs * b64_per_job;
    const int N = a.n_b64_total, sg = (N * SCALAR_TEAM + 127) / 128;
    int slot = 0;
    auto mark = [&]() {
        if (ev) cudaEventRecord(ev[slot], st);
        ++slot;
    };
    // optional: the SEARCH launches of a slot bracketed by their own events (sev[(2 * slot + round) * 2 + {0, 1}]), so that
    // a kernel's duration can be read without the scalar bookkeeping launches that share its slot
    uint32_t smask = 0;
    auto search_begin = [&](int round) {
        if (sev) cudaEventRecord(sev[(2 * slot + round) * 2], st);
    };
    auto search_end = [&](int round) {
        if (sev) cudaEventRecord(sev[(2 * slot + round) * 2 + 1], st), smask |= 1u << (2 * slot + round);
    };
    svtme_launch_zero_fill(d_states, ((size_t)N * sizeof(StageState) + 15) & ~(size_t)15, st); // (the allocation is rounded up likewise)
    mark(); // 0: state clear
    int launches = 0;
    // scalar steps always run (they also carry the bookkeeping of disabled phases); engine launches
    // that no picture of the batch needs are skipped
    scalar_kernel<S_INIT, S_INIT><<<sg, 128, 0, st>>>(a);
    if (f.zz) {
        search_begin(0);
        points_kernel<<<N, 128, 0, st>>>(a), ++launches;
        search_end(0);
    }
    scalar_kernel<S_ZZ, S_ZZ><<<sg, 128, 0, st>>>(a);
    mark(); // 1: zz-SAD
    // pre-HME rounds: strip mode when every picture uses the sub-sampled SAD, else one CTA per b64
    const int strip_tasks = n_jobs * NL * NR *
        (b64_cols * ((b64_rows + STRIP_SEG - 1) / STRIP_SEG) + b64_rows * ((b64_cols + STRIP_SEG - 1) / STRIP_SEG));
    auto prehme_round = [&](int round) {
        search_begin(round);
        if (f.prehme_strip)
            strip_kernel<0><<<strip_tasks, SVTME_HME_THREADS, sizeof(StripSmem), st>>>(a, b64_cols, b64_rows, 2, g_strip_arena_limit, g_strip_stream_min_h);
        else
            engine_kernel<<<N, SVTME_HME_THREADS, sizeof(EngSmem), st>>>(a, 1, 2);
        ++launches;
        search_end(round);
    };
    if (f.prehme) prehme_round(0);
    int fused = 0;
    if (f.prehme_l1_round) {
        scalar_kernel<S_PREHME1, S_PREHME1><<<sg, 128, 0, st>>>(a);
        prehme_round(1);
        scalar_kernel<S_PREHME2, S_PREHME2><<<sg, 128, 0, st>>>(a);
    } else
        scalar_kernel<S_PREHME1, S_PREHME2><<<sg, 128, 0, st>>>(a), ++fused;
    mark(); // 2: pre-HME
    const int l0_tasks = n_jobs * NL * NR * 4 * b64_rows * ((b64_cols + STRIP_SEG - 1) / STRIP_SEG);
    auto l0_round = [&](int round) {
        search_begin(round);
        if (f.prehme_strip && f.l0_strip)
            {
            // the arena is the last member of StripSmem: a smaller arena is a smaller dynamic allocation
            const int lim = g_strip_arena_limit < L0_STRIP_ARENA ? g_strip_arena_limit : L0_STRIP_ARENA;
            strip_kernel<1><<<l0_tasks, SVTME_HME_THREADS, sizeof(StripSmem) - SVTME_HME_ARENA_BYTES + L0_STRIP_ARENA, st>>>(
                a, b64_cols, b64_rows, 3, lim, g_l0_stream_min_h);
        }
        else
            engine_kernel<<<N, SVTME_HME_THREADS, sizeof(EngSmem), st>>>(a, 1, 3);
        ++launches;
        search_end(round);
    };
    if (f.hme_l0) l0_round(0);
    if (f.hme_l0_round_b) {
        scalar_kernel<S_L0A, S_L0A><<<sg, 128, 0, st>>>(a);
        l0_round(1);
        scalar_kernel<S_L0B, S_L0B><<<sg, 128, 0, st>>>(a);
    } else
        scalar_kernel<S_L0A, S_L0B><<<sg, 128, 0, st>>>(a), ++fused;
    mark(); // 3: HME level 0
    if (f.hme_l1) {
        search_begin(0);
        if (f.refine_small)
            refine_kernel<8><<<N, 128, sizeof(RefineSmem<8>), st>>>(a, 1, 4);
        else
            engine_kernel<<<N, SVTME_HME_THREADS, sizeof(EngSmem), st>>>(a, 2, 4);
        ++launches;
        search_end(0);
    }
    if (f.hme_l2) {
        scalar_kernel<S_L1, S_L1><<<sg, 128, 0, st>>>(a);
        mark(); // 4: HME level 1
        search_begin(0);
        if (f.refine_small)
            refine_kernel<16><<<N, 128, sizeof(RefineSmem<16>), st>>>(a, 0, 5);
        else
            engine_kernel<<<N, SVTME_HME_THREADS, sizeof(EngSmem), st>>>(a, 3, 5);
        ++launches;
        search_end(0);
        scalar_kernel<S_L2, S_L2><<<sg, 128, 0, st>>>(a);
    } else {
        mark(); // 4: HME level 1 (its post step runs fused with the search-centre step below)
        scalar_kernel<S_L1, S_L2><<<sg, 128, 0, st>>>(a), ++fused;
    }
    mark(); // 5: HME level 2 + search centre + HME pruning
    {
        // all references of a b64 in one CTA, in the reference's loop order (which also serves the dependence of
        // the other references on (list 0, ref 0), :1345-1350)
        // With parked searches the dependence needs (list 0, ref 0) complete before the others plan: two rounds.
        IntSlots round[2] = {};
        for (int li = 0; li < NL; ++li)
            for (int ri = 0; ri < f.max_refs[li]; ++ri) {
                IntSlots &t = round[(f.int_dependent && !f.small_areas && (li || ri)) ? 1 : 0];
                t.slot[t.n++] = (int8_t)(li * NR + ri);
            }
        search_begin(0);
        for (const IntSlots &t : round) {
            if (!t.n) continue;
            int_kernel<true><<<N, SVTME_INT_SMALL_THREADS, sizeof(IntSmem<true>), st>>>(a, t, f.small_areas ? INT_ALL : INT_DEFER);
            ++launches;
            if (!f.small_areas) {
                int_kernel<false><<<N, SVTME_THREADS, sizeof(IntSmem<false>), st>>>(a, t, INT_RESUME);
                ++launches;
            }
        }
        search_end(0);
    }
    mark(); // 6: integer search
    search_begin(0);
    finish_kernel<<<N, 128, sizeof(FinSmem), st>>>(a);
    search_end(0);
    mark(); // 7: candidates + statistics
    if (sev_mask) *sev_mask = smask;
    g_svtme_launches += 8 - fused + 1 + launches; // (the state clear counts itself)
}

// ---------------------------------------------------------------- dynamic-GOP detector HME
// dg_detector_hme_level0 (Codec/pd_process.c:491-589): for every b64 of the picture one early_hme_b64 (:389-487) — a
// full-SAD svt_sad_loop_kernel of the 16x16 block of the 1/16 SOURCE picture over an S x S window of the 1/16 plane of
// another picture (S = 16 / 64 / 128 by input resolution), clipped like hme_level_0 (block may move 15 px into the pad) —
// and four running metrics. One CTA per b64, the same engine as every HME round (hme_search_multi: the 128 x 128 window
// takes its tiled path); the metrics are integer sums, so atomics give the reference's totals whatever the order.
struct DgAcc { // zero-filled by the host before the launch
    unsigned long long tot_dist;
    uint32_t tot_cplx, tot_active;
    int sum_in_vectors;
    uint32_t _pad[3]; // 32 bytes: the clear works in 16-byte units
};
static_assert(sizeof(DgAcc) % 16 == 0, "svtme_launch_zero_fill clears whole 16-byte units");
__global__ void __launch_bounds__(SVTME_HME_THREADS, SVTME_HME_CTAS_PER_SM) dg_hme_kernel(DevPlane src16, DevPlane ref16, int cols, int rows,
                                                                                         int sa, DgAcc *acc) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    EngSmem *sm = reinterpret_cast<EngSmem *>(smem_raw);
    const int tid = threadIdx.x, bx = blockIdx.x % cols, by = blockIdx.x / cols;
    const int16_t org_x = (int16_t)((bx * 64) >> 2), org_y = (int16_t)((by * 64) >> 2);
    for (int i = tid; i < 16; i += SVTME_HME_THREADS) // the 16x16 source block, rows of 16 bytes (16-byte aligned in the plane)
        *reinterpret_cast<uint4 *>(&sm->src[i * 16]) = *reinterpret_cast<const uint4 *>(src16.at(org_x, org_y + i));
    if (tid == 0) {
        int16_t w = (int16_t)((sa + 7) & ~7), h = (int16_t)sa, ox = (int16_t)(-(w >> 1)), oy = (int16_t)(-(h >> 1));
        clip_hme(org_x, org_y, (int16_t)ref16.width, (int16_t)ref16.height, 15, 15, w, h, ox, oy); // (:408-448: the clipping of hme_level_0)
        HJob &j = sm->jobs[0];
        j.ref = ref16.at(org_x + ox, org_y + oy);
        j.stride = ref16.stride;
        j.sa_w = w;
        j.sa_h = h;
        j.skip_even = 0;
        j.key = HME_KEY_INIT;
        sm->cur.org_x = ox; // (scratch: the applied origin, for the metrics below)
        sm->cur.org_y = oy;
    }
    __syncthreads();
    const HBlk blk = make_hblk(sm->src, 16, 16, 16, 1); // FULL_SAD_SEARCH (:507): every row
    hme_search_multi(blk, sm->jobs, 1, &sm->plan, sm->arena);
    if (tid == 0) {
        const unsigned long long key = sm->jobs[0].key;
        const unsigned long long sad = key >> 32;
        const int16_t col = (int16_t)((int16_t)((int16_t)(key & 0xffff) + (int16_t)sm->cur.org_x) * 4);
        const int16_t row = (int16_t)((int16_t)((int16_t)((key >> 16) & 0xffff) + (int16_t)sm->cur.org_y) * 4);
        int in = 0; // (:548-577) does the vector point towards the picture centre?
        if (by < rows / 2)
            in += row > 0 ? -1 : (row < 0 ? 1 : 0);
        else if (by > rows / 2)
            in += row > 0 ? 1 : (row < 0 ? -1 : 0);
        if (bx < cols / 2)
            in += col > 0 ? -1 : (col < 0 ? 1 : 0);
        else if (bx > cols / 2)
            in += col > 0 ? 1 : (col < 0 ? -1 : 0);
        atomicAdd(&acc->tot_dist, sad);
        if (sad > 16 * 16 * 30) atomicAdd(&acc->tot_cplx, 1u);
        if (col != 0 || row != 0) atomicAdd(&acc->tot_active, 1u);
        if (in) atomicAdd(&acc->sum_in_vectors, in);
    }
}

size_t svtme_dg_acc_bytes() { return sizeof(DgAcc); }
void svtme_dg_acc_unpack(const void *host, SvtMeDgMetrics *out) {
    const DgAcc *a = static_cast<const DgAcc *>(host);
    out->tot_dist = a->tot_dist;
    out->tot_cplx = a->tot_cplx;
    out->tot_active = a->tot_active;
    out->sum_in_vectors = a->sum_in_vectors;
}
// acc: 16-byte aligned device scratch of svtme_dg_acc_bytes(); cleared here
cudaError_t svtme_launch_dg_hme(const DevPlane &src16, const DevPlane &ref16, int b64_cols, int b64_rows, int sa, void *acc, cudaStream_t st) {
    svtme_configure_stage_kernels();
    cudaError_t e = svtme_launch_zero_fill(acc, sizeof(DgAcc), st);
    if (e != cudaSuccess) return e;
    dg_hme_kernel<<<b64_cols * b64_rows, SVTME_HME_THREADS, sizeof(EngSmem), st>>>(src16, ref16, b64_cols, b64_rows, sa, static_cast<DgAcc *>(acc));
    ++g_svtme_launches;
    return cudaGetLastError();
}

// ---------------------------------------------------------------- kernel-test entry points
struct TestSmem {
    HJob job;
    HPlan plan;
    FpTables tab;
    __align__(16) uint8_t src[64 * 64];
    __align__(16) uint8_t arena[SVTME_ARENA_BYTES];
};

// svt_sad_loop_kernel on device buffers: src block (any size up to 64x64, given stride) is copied
// into shared memory first. ref_row_step_raw = src_stride_raw of the RTCD signature.
__global__ void __launch_bounds__(SVTME_THREADS) sad_loop_test_kernel(const uint8_t *src, int src_stride, const uint8_t *ref,
                                                                      int ref_stride, int bh, int bw, int raw, int skip,
                                                                      int sa_w, int sa_h, unsigned long long *out) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    TestSmem *sm = reinterpret_cast<TestSmem *>(smem_raw);
    for (int i = threadIdx.x; i < 64 * 64; i += SVTME_THREADS) {
        const int r = i >> 6, c = i & 63;
        sm->src[i] = (r < bh && c < bw) ? src[(size_t)r * src_stride + c] : 0;
    }
    const int row_step = ref_stride / raw; // the RTCD caller multiplies both strides to sub-sample rows
    if (threadIdx.x == 0) {
        sm->job.ref = ref;
        sm->job.stride = raw;
        sm->job.sa_w = sa_w;
        sm->job.sa_h = sa_h;
        sm->job.skip_even = (bw == 16 && bh <= 16 && skip) ? 1 : 0;
        sm->job.key = HME_KEY_INIT;
    }
    __syncthreads();
    const HBlk b2 = make_hblk(sm->src, 64, bw, bh, row_step);
    hme_search_multi(b2, &sm->job, 1, &sm->plan, sm->arena);
    if (threadIdx.x == 0) *out = sm->job.key;
}

void svtme_launch_sad_loop_test(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int bh, int bw,
                                int raw, int skip, int sa_w, int sa_h, unsigned long long *d_key, cudaStream_t st) {
    cudaFuncSetAttribute(sad_loop_test_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TestSmem));
    sad_loop_test_kernel<<<1, SVTME_THREADS, sizeof(TestSmem), st>>>(d_src, src_stride, d_ref, ref_stride, bh, bw, raw, skip,
                                                                      sa_w, sa_h, d_key);
    ++g_svtme_launches;
}

__global__ void __launch_bounds__(SVTME_THREADS) fullpel_test_kernel(const uint8_t *src, int src_stride, const uint8_t *ref,
                                                                     int ref_stride, int ox, int oy, int sa_w, int sa_h,
                                                                     int sub_sad, uint32_t *best_sad, uint32_t *best_mv) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    TestSmem *sm = reinterpret_cast<TestSmem *>(smem_raw);
    for (int i = threadIdx.x; i < 64 * 64; i += SVTME_THREADS) sm->src[i] = src[(size_t)(i >> 6) * src_stride + (i & 63)];
    if (threadIdx.x < SVTME_PU_COUNT) {
        sm->tab.key[threadIdx.x] = (unsigned long long)best_sad[threadIdx.x] << 32; // order 0: existing entries win ties
        sm->tab.mv[threadIdx.x] = best_mv[threadIdx.x];
    }
    __syncthreads();
    fullpel_search(sm->src, ref, ref_stride, ox, oy, sa_w, sa_h, sub_sad != 0, 1, &sm->tab, sm->arena, (int)sizeof(sm->arena));
    if (threadIdx.x < SVTME_PU_COUNT) {
        best_sad[threadIdx.x] = (uint32_t)(sm->tab.key[threadIdx.x] >> 32);
        best_mv[threadIdx.x] = sm->tab.mv[threadIdx.x];
    }
}

void svtme_launch_fullpel_test(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int ox, int oy,
                               int sa_w, int sa_h, int sub_sad, uint32_t *d_best_sad, uint32_t *d_best_mv, cudaStream_t st) {
    cudaFuncSetAttribute(fullpel_test_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TestSmem));
    fullpel_test_kernel<<<1, SVTME_THREADS, sizeof(TestSmem), st>>>(d_src, src_stride, d_ref, ref_stride, ox, oy, sa_w, sa_h,
                                                                     sub_sad, d_best_sad, d_best_mv);
    ++g_svtme_launches;
}


// ---------------------------------------------------------------- kernel-level batched forms ("*_batch(n, args[])")
// n full-pel searches of 64x64 blocks in ONE launch, one CTA per item, the same engine as the stage (fullpel_search).
// Everything is in device memory; the best tables are read-modify-write like p_sb_best_sad / p_sb_best_mv.
__global__ void __launch_bounds__(SVTME_THREADS, SVTME_CTAS_PER_SM)
    fullpel_batch_kernel(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride, const SvtMeFullpelItem *items,
                         int sub_sad, uint32_t *best_sad, uint32_t *best_mv) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    TestSmem *sm = reinterpret_cast<TestSmem *>(smem_raw);
    const SvtMeFullpelItem it = items[blockIdx.x];
    const uint8_t *s = src + it.src_offset;
    if (((reinterpret_cast<uintptr_t>(s) | (uintptr_t)src_stride) & 15) == 0) {
        for (int i = threadIdx.x; i < 256; i += SVTME_THREADS) cp_async16(&sm->src[(i >> 2) * 64 + (i & 3) * 16], s + (size_t)(i >> 2) * src_stride + (i & 3) * 16);
        cp_async_wait_all();
    } else
        for (int i = threadIdx.x; i < 64 * 64; i += SVTME_THREADS) sm->src[i] = s[(size_t)(i >> 6) * src_stride + (i & 63)];
    uint32_t *bs = best_sad + (size_t)blockIdx.x * SVTME_PU_COUNT, *bm = best_mv + (size_t)blockIdx.x * SVTME_PU_COUNT;
    if (threadIdx.x < SVTME_PU_COUNT) {
        sm->tab.key[threadIdx.x] = (unsigned long long)bs[threadIdx.x] << 32; // order 0: existing entries win ties
        sm->tab.mv[threadIdx.x] = bm[threadIdx.x];
    }
    __syncthreads();
    fullpel_search(sm->src, ref + it.ref_offset, ref_stride, it.x_search_area_origin, it.y_search_area_origin, it.search_area_width,
                   it.search_area_height, sub_sad != 0, 1, &sm->tab, sm->arena, (int)sizeof(sm->arena));
    if (threadIdx.x < SVTME_PU_COUNT) {
        bs[threadIdx.x] = (uint32_t)(sm->tab.key[threadIdx.x] >> 32);
        bm[threadIdx.x] = sm->tab.mv[threadIdx.x];
    }
}

cudaError_t svtme_launch_fullpel_batch(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride,
                                       const SvtMeFullpelItem *d_items, int n, int sub_sad, uint32_t *d_best_sad, uint32_t *d_best_mv,
                                       cudaStream_t st) {
    static std::mutex mu;
    static bool done[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (dev >= 0 && dev < 64 && !done[dev]) {
            e = cudaFuncSetAttribute(fullpel_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TestSmem));
            if (e != cudaSuccess) return e;
            done[dev] = true;
        }
    }
    fullpel_batch_kernel<<<n, SVTME_THREADS, sizeof(TestSmem), st>>>(d_src, src_stride, d_ref, ref_stride, d_items, sub_sad, d_best_sad,
                                                                    d_best_mv);
    ++g_svtme_launches;
    return cudaGetLastError();
}

// svtme_search.cuh — CTA-wide SAD search engines (sm_100a), shared by the stage kernel and the
// kernel-test entry points.
//
// Both engines stage the reference search window in shared memory as FOUR byte-phase copies
// (copy c holds the window shifted left by c bytes), so that a search position x = 4q + c reads
// its reference pixels as ALIGNED 16-byte LDS.128 from copy c: the inner loop is nothing but
// LDS.128 + VABSDIFF4.U8.ACC (4 byte-SADs per instruction), no funnel shifts. A thread owns one
// 16-pixel-wide, 8-row (row-sub-sampled) source sub-block in REGISTERS and G search positions of
// the same byte phase (x, x+4, ..., x+4(G-1)), so G+3 loaded words feed 4G VABSDIFF4.
//
//   hme_search   : svt_sad_loop_kernel (C_DEFAULT/compute_sad_c.c:58-101) — one best (SAD,x,y)
//   fullpel_search: open_loop_me_fullpel_search_sblock (Codec/motion_estimation.c:781-817) incl.
//                  svt_ext_all_sad_calculation_8x8_16x16 (:337) and
//                  svt_ext_eight_sad_calculation_32x32_64x64 (:369) — best (SAD,MV) of 85 PUs
//
// Tie-breaks: the reference keeps the FIRST strict minimum in (y, x) raster order. Here every
// candidate carries a key (sad << k | raster index) and minima are taken over keys, which picks
// the same winner independent of evaluation order.
#pragma once
#include "svtme_common.cuh"

#define SVTME_THREADS 256
#define SVTME_ARENA_BYTES (80 * 1024)

__device__ __forceinline__ uint32_t vsad4(uint32_t a, uint32_t b, uint32_t acc) {
    uint32_t r;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(acc));
    return r;
}

// Stage rows [0,rows) x bytes [0, m*16) of the window whose top-left byte is `g` (global, any
// alignment) into 4 phase copies. copy c, row R, chunk k lives at
//     arena + c*copy_bytes + rowoff(R)*16 + k*16,   rowoff(R) = R*m + skew*(R>>4)
// (skew chunks are inserted every 16 rows by the full-pel engine to spread 16-row-apart accesses
// over all banks). Reads up to 3+16 bytes past the last needed byte of a row: planes carry pads.
__device__ __forceinline__ void stage_window4(uint8_t *arena, const uint8_t *__restrict__ g, int gstride, int rows,
                                              int m, int skew, int copy_bytes) {
    // One warp per row, lanes over the words of the row: no integer division anywhere. Each lane
    // reads three ALIGNED global words (the row base is rounded down to 4 bytes) and derives the
    // four phase words with funnel shifts; neighbouring lanes hit the same L1 lines.
    const int wpr = m * 4; // words per copy row
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int R = warp; R < rows; R += SVTME_THREADS / 32) {
        const uintptr_t ai = reinterpret_cast<uintptr_t>(g + (size_t)R * gstride);
        const uint32_t *al = reinterpret_cast<const uint32_t *>(ai & ~(uintptr_t)3);
        const uint32_t mis = (uint32_t)(ai & 3);
        uint8_t *dst = arena + ((size_t)(R * m + skew * (R >> 4)) << 4);
        for (int kw = lane; kw < wpr; kw += 32) {
            const uint32_t w0 = __ldg(al + kw), w1 = __ldg(al + kw + 1), w2 = __ldg(al + kw + 2);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint32_t o = c + mis; // byte offset of phase c inside (w0,w1,w2)
                const uint32_t lo = o >= 4 ? w1 : w0, hi = o >= 4 ? w2 : w1;
                *reinterpret_cast<uint32_t *>(dst + (size_t)c * copy_bytes + 4 * kw) = __funnelshift_r(lo, hi, (o & 3) * 8);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// HME engine
// ------------------------------------------------------------------------------------------------
struct HBlk {            // the source block of this call (uniform over the CTA)
    const uint8_t *src;  // shared memory, compared row r at src + r*src_pitch
    int src_pitch;       // bytes between two COMPARED source rows
    int bw;              // block width in pixels (any value >= 1)
    int nrows;           // number of (sub-sampled) rows actually compared
    int row_step;        // reference rows between two compared rows (2 = SUB_SAD_SEARCH)
};

struct HJob {
    const uint8_t *ref;  // global: reference pixel under block(0,0) at search position (0,0)
    int stride;          // bytes between consecutive SEARCH rows == plane stride
    int sa_w, sa_h;
    int skip_even;       // skip_search_line: even search rows are not evaluated
    unsigned long long key; // result: sad << 32 | y << 16 | x ; init 0xffffff << 32
};

#define HME_KEY_INIT (0xffffffULL << 32)

// Geometry of one staged search tile (a whole job, or one band/column tile of a big job).
struct HGeo {
    uint32_t arena_off;  // byte offset of the 4 phase copies inside the arena
    uint32_t sad_off;    // byte offset of the per-position SAD buffer (multi-sub-block blocks)
    int copy_bytes, m;
    int sa_w, bh;        // positions in this tile
    int x0, y0;          // tile origin inside the job's search area
    int xt_n, n_items, G, skip_even;
    uint32_t mg_xt;
};

struct SrcRegs {         // one 16x8 source sub-block in registers
    uint32_t s[8][4];
    uint32_t msk[4];
    int cached_b, rows_here;
};

__device__ __forceinline__ uint32_t magic_div(uint32_t d) { return 0xFFFFFFFFu / d + 1u; } // exact for n < 2^16

__device__ __forceinline__ void load_src_regs(const HBlk &blk, int b, int bxi, int byi, SrcRegs &sr) {
    sr.cached_b = b;
    sr.rows_here = min(8, blk.nrows - 8 * byi);
    const int vb = min(16, blk.bw - 16 * bxi);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = min(max(vb - 4 * i, 0), 4);
        sr.msk[i] = n >= 4 ? ~0u : ((1u << (8 * n)) - 1u);
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        if (r < sr.rows_here) {
            const uint8_t *sp = blk.src + (size_t)(8 * byi + r) * blk.src_pitch + 16 * bxi;
            if (vb == 16) {
                const uint4 v = *reinterpret_cast<const uint4 *>(sp);
                sr.s[r][0] = v.x; sr.s[r][1] = v.y; sr.s[r][2] = v.z; sr.s[r][3] = v.w;
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    sr.s[r][i] = sr.msk[i] ? (*reinterpret_cast<const uint32_t *>(sp + 4 * i) & sr.msk[i]) : 0u;
            }
        }
    }
}

// One work item = G same-phase positions of one row against one 16x8 source sub-block.
template <int G>
__device__ __forceinline__ void hme_item(const HBlk &blk, const HGeo &g, const uint8_t *arena, int it, int nb, int nbx,
                                         uint32_t mg_nb, uint32_t mg_nbx, SrcRegs &sr, unsigned long long &best) {
    int t = nb == 1 ? it : (int)__umulhi((uint32_t)it, mg_nb);
    const int b = it - t * nb;
    const int c = t & 3;
    t >>= 2;
    const int y = g.xt_n == 1 ? t : (int)__umulhi((uint32_t)t, g.mg_xt);
    const int xt = t - y * g.xt_n;
    const int byi = nbx == 1 ? b : (int)__umulhi((uint32_t)b, mg_nbx);
    const int bxi = b - byi * nbx;
    if (b != sr.cached_b) load_src_regs(blk, b, bxi, byi, sr);
    const bool masked = (blk.bw & 15) != 0;
    uint32_t acc[G];
#pragma unroll
    for (int j = 0; j < G; ++j) acc[j] = 0;
    const int q0 = xt * G + 4 * bxi; // first word of copy c this item needs
    const uint8_t *base = arena + g.arena_off + (size_t)c * g.copy_bytes + 4 * q0;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        if (r < sr.rows_here) {
            const int R = y + (8 * byi + r) * blk.row_step;
            const uint8_t *rp = base + ((size_t)R * g.m << 4);
            uint32_t w[G + 4];
            if (G == 2) {
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    uint2 v = *reinterpret_cast<const uint2 *>(rp + 8 * k);
                    w[2 * k] = v.x; w[2 * k + 1] = v.y;
                }
            } else {
#pragma unroll
                for (int k = 0; k < (G + 3 + 3) / 4; ++k) {
                    uint4 v = *reinterpret_cast<const uint4 *>(rp + 16 * k);
                    if (4 * k + 0 < G + 4) w[4 * k + 0] = v.x;
                    if (4 * k + 1 < G + 4) w[4 * k + 1] = v.y;
                    if (4 * k + 2 < G + 4) w[4 * k + 2] = v.z;
                    if (4 * k + 3 < G + 4) w[4 * k + 3] = v.w;
                }
            }
            if (masked) {
#pragma unroll
                for (int j = 0; j < G; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[j] = vsad4(sr.s[r][i], w[j + i] & sr.msk[i], acc[j]);
            } else {
#pragma unroll
                for (int j = 0; j < G; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[j] = vsad4(sr.s[r][i], w[j + i], acc[j]);
            }
        }
    }
    const int yg = g.y0 + y;
    if (g.skip_even && !(yg & 1)) return;
    uint32_t *sadbuf = reinterpret_cast<uint32_t *>(const_cast<uint8_t *>(arena) + g.sad_off);
#pragma unroll
    for (int j = 0; j < G; ++j) {
        const int x = xt * 4 * G + 4 * j + c;
        if (x < g.sa_w) {
            if (nb == 1) {
                unsigned long long k = ((unsigned long long)acc[j] << 32) | ((unsigned)yg << 16) | (unsigned)(g.x0 + x);
                best = k < best ? k : best;
            } else {
                atomicAdd(&sadbuf[y * g.sa_w + x], acc[j]);
            }
        }
    }
}

__device__ __forceinline__ void hme_item_any(const HBlk &blk, const HGeo &g, const uint8_t *arena, int it, int nb, int nbx,
                                             uint32_t mg_nb, uint32_t mg_nbx, SrcRegs &sr, unsigned long long &best) {
    if (g.G == 8)
        hme_item<8>(blk, g, arena, it, nb, nbx, mg_nb, mg_nbx, sr, best);
    else if (g.G == 4)
        hme_item<4>(blk, g, arena, it, nb, nbx, mg_nb, mg_nbx, sr, best);
    else
        hme_item<2>(blk, g, arena, it, nb, nbx, mg_nb, mg_nbx, sr, best);
}

// Tile geometry for a (sub)area of wb x hb positions. Returns the arena bytes it needs.
__device__ __forceinline__ uint32_t hme_geo(const HBlk &blk, int wb, int hb, int skip_even, HGeo &g) {
    const int nbx = (blk.bw + 15) >> 4, nb = nbx * ((blk.nrows + 7) >> 3);
    const int G = wb <= 8 ? 2 : (wb <= 16 ? 4 : 8);
    const int tile_w = 4 * G;
    const int vec_words = G == 2 ? 6 : ((G + 3 + 3) / 4) * 4;
    const int xt_n = (wb + tile_w - 1) / tile_w;
    const int words = (xt_n - 1) * G + 4 * (nbx - 1) + vec_words;
    const int m = ((words + 3) / 4) | 1; // odd chunk stride: consecutive rows start in different bank groups
    const int rows = hb + (blk.nrows - 1) * blk.row_step;
    int copy_chunks = rows * m;
    copy_chunks += (10 - (copy_chunks & 7)) & 7; // copy stride == 2 (mod 8) chunks
    g.copy_bytes = copy_chunks * 16;
    g.m = m;
    g.sa_w = wb;
    g.bh = hb;
    g.xt_n = xt_n;
    g.n_items = hb * xt_n * 4 * nb;
    g.G = G;
    g.skip_even = skip_even;
    g.mg_xt = magic_div((uint32_t)xt_n);
    g.sad_off = 0;
    return (uint32_t)g.copy_bytes * 4u + (nb > 1 ? (uint32_t)(hb * wb * 4 + 15) & ~15u : 0u);
}

__device__ __forceinline__ void hme_reduce_store(unsigned long long best, HJob *job) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long v = __shfl_xor_sync(0xffffffffu, best, o);
        best = v < best ? v : best;
    }
    if ((threadIdx.x & 31) == 0 && best < HME_KEY_INIT) atomicMin(&job->key, best);
}

// One svt_sad_loop_kernel of ANY size: the search area is tiled so that each tile fits the arena.
// All threads of the CTA must call it with identical arguments; job lives in shared memory and
// job->key must be HME_KEY_INIT on entry.
__device__ void hme_search(const HBlk &blk, HJob *job, uint8_t *arena) {
    const int sa_w = job->sa_w, sa_h = job->sa_h;
    if (sa_w <= 0 || sa_h <= 0) return;
    const int nbx = (blk.bw + 15) >> 4, nb = nbx * ((blk.nrows + 7) >> 3);
    const uint32_t mg_nb = magic_div((uint32_t)nb), mg_nbx = magic_div((uint32_t)nbx);
    int wt = sa_w, bh = sa_h;
    HGeo g;
    for (;;) {
        const uint32_t need = hme_geo(blk, wt, bh, job->skip_even, g);
        if (need <= SVTME_ARENA_BYTES && g.n_items < 65536) break;
        const int tile_w = 4 * g.G;
        if (bh > 8) bh = (bh + 1) / 2;
        else if (wt > tile_w) wt = ((wt / 2 + tile_w - 1) / tile_w) * tile_w;
        else if (bh > 1) bh = (bh + 1) / 2;
        else break; // one row x one G-tile always fits
    }
    unsigned long long best = ~0ULL;
    SrcRegs sr;
    sr.cached_b = -1;
    sr.rows_here = 0;
    for (int x0 = 0; x0 < sa_w; x0 += wt) {
        const int wb = min(wt, sa_w - x0);
        for (int y0 = 0; y0 < sa_h; y0 += bh) {
            const int hb = min(bh, sa_h - y0);
            hme_geo(blk, wb, hb, job->skip_even, g);
            g.arena_off = 0;
            g.sad_off = (uint32_t)g.copy_bytes * 4u;
            g.x0 = x0;
            g.y0 = y0;
            uint32_t *sadbuf = reinterpret_cast<uint32_t *>(arena + g.sad_off);
            __syncthreads(); // previous users of the arena are done
            stage_window4(arena, job->ref + (size_t)y0 * job->stride + x0, job->stride, hb + (blk.nrows - 1) * blk.row_step,
                          g.m, 0, g.copy_bytes);
            if (nb > 1)
                for (int i = threadIdx.x; i < hb * wb; i += SVTME_THREADS) sadbuf[i] = 0;
            __syncthreads();
            for (int it = threadIdx.x; it < g.n_items; it += SVTME_THREADS)
                hme_item_any(blk, g, arena, it, nb, nbx, mg_nb, mg_nbx, sr, best);
            if (nb > 1) {
                __syncthreads();
                const uint32_t mg_wb = magic_div((uint32_t)wb);
                for (int i = threadIdx.x; i < hb * wb; i += SVTME_THREADS) {
                    const int yy = (int)__umulhi((uint32_t)i, mg_wb);
                    const int y = y0 + yy, x = x0 + (i - yy * wb);
                    if (job->skip_even && !(y & 1)) continue;
                    unsigned long long k = ((unsigned long long)sadbuf[i] << 32) | ((unsigned)y << 16) | (unsigned)x;
                    best = k < best ? k : best;
                }
            }
        }
    }
    hme_reduce_store(best, job);
    __syncthreads();
}

// SAD of the block at ONE position per job (zz-SAD, check_00_center): no staging, one warp per
// job, lanes over (row, 4-byte column) with unaligned reference words assembled from two aligned
// loads. All threads call with identical arguments.
__device__ void hme_points(const HBlk &blk, HJob *jobs, int n) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wcols = (blk.bw + 3) >> 2, items = blk.nrows * wcols;
    const uint32_t mg_wc = magic_div((uint32_t)wcols);
    for (int j = warp; j < n; j += SVTME_THREADS / 32) {
        const HJob &job = jobs[j];
        uint32_t acc = 0;
        for (int i = lane; i < items; i += 32) {
            const int r = (int)__umulhi((uint32_t)i, mg_wc), cw = i - r * wcols;
            const int nbytes = min(4, blk.bw - 4 * cw);
            const uint32_t mask = nbytes >= 4 ? ~0u : ((1u << (8 * nbytes)) - 1u);
            const uint32_t s = *reinterpret_cast<const uint32_t *>(blk.src + (size_t)r * blk.src_pitch + 4 * cw) & mask;
            const uintptr_t ai = reinterpret_cast<uintptr_t>(job.ref + (size_t)r * blk.row_step * job.stride + 4 * cw);
            const uint32_t *al = reinterpret_cast<const uint32_t *>(ai & ~(uintptr_t)3);
            const uint32_t sh = (uint32_t)(ai & 3) * 8;
            const uint32_t lo = __ldg(al), hi = sh ? __ldg(al + 1) : 0u;
            acc = vsad4(s, __funnelshift_r(lo, hi, sh) & mask, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0 && !job.skip_even) jobs[j].key = (unsigned long long)acc << 32;
    }
    __syncthreads();
}

// A round of independent jobs on the same source block: jobs are packed into the arena as many
// at a time as fit, staged together and searched as ONE flattened item list, so small jobs
// (8x3 refinement windows, quadrant searches) share the barriers and fill the CTA.
// `geo` is scratch in shared memory (one entry per job). All threads call with identical arguments.
__device__ void hme_search_multi(const HBlk &blk, HJob *jobs, int n, HGeo *geo, uint8_t *arena) {
    if (n <= 0) return;
    bool all_points = true;
    for (int j = 0; j < n; ++j) all_points = all_points && jobs[j].sa_w == 1 && jobs[j].sa_h == 1;
    if (all_points) {
        hme_points(blk, jobs, n);
        return;
    }
    const int nbx = (blk.bw + 15) >> 4, nb = nbx * ((blk.nrows + 7) >> 3);
    const uint32_t mg_nb = magic_div((uint32_t)nb), mg_nbx = magic_div((uint32_t)nbx);
    const int extra_rows = (blk.nrows - 1) * blk.row_step;
    int j0 = 0;
    while (j0 < n) {
        // ---- pack jobs [j0, j1) (uniform: every thread computes the same plan)
        uint32_t used = 0;
        int j1 = j0, total_items = 0;
        __syncthreads(); // arena + geo are free
        while (j1 < n) {
            HGeo g;
            const HJob &job = jobs[j1];
            if (job.sa_w <= 0 || job.sa_h <= 0) {
                if (threadIdx.x == 0) { g.n_items = 0; g.bh = 0; g.sa_w = 0; g.G = 2; g.xt_n = 1; geo[j1] = g; }
                ++j1;
                continue;
            }
            const uint32_t need = hme_geo(blk, job.sa_w, job.sa_h, job.skip_even, g);
            if (need > SVTME_ARENA_BYTES || g.n_items >= 65536) break; // needs tiling: single-job path
            if (used + need > SVTME_ARENA_BYTES || total_items + g.n_items >= (1 << 20)) break;
            g.arena_off = used;
            g.sad_off = used + (uint32_t)g.copy_bytes * 4u;
            g.x0 = 0;
            g.y0 = 0;
            if (threadIdx.x == 0) geo[j1] = g;
            used += need;
            total_items += g.n_items;
            ++j1;
        }
        if (j1 == j0) { // the next job does not fit as a whole
            hme_search(blk, &jobs[j0], arena);
            ++j0;
            continue;
        }
        __syncthreads(); // geo published
        // ---- stage every window of the group, clear the SAD buffers
        for (int j = j0; j < j1; ++j) {
            const HGeo &g = geo[j];
            if (g.n_items == 0) continue;
            stage_window4(arena + g.arena_off, jobs[j].ref, jobs[j].stride, g.bh + extra_rows, g.m, 0, g.copy_bytes);
            if (nb > 1) {
                uint32_t *sadbuf = reinterpret_cast<uint32_t *>(arena + g.sad_off);
                for (int i = threadIdx.x; i < g.bh * g.sa_w; i += SVTME_THREADS) sadbuf[i] = 0;
            }
        }
        __syncthreads();
        // ---- one flattened item list over the group
        {
            SrcRegs sr;
            sr.cached_b = -1;
            sr.rows_here = 0;
            unsigned long long best = ~0ULL;
            int cj = j0, cbase = 0;
            for (int it = threadIdx.x; it < total_items; it += SVTME_THREADS) {
                while (it >= cbase + geo[cj].n_items) {
                    if (nb == 1 && best != ~0ULL) {
                        if (best < HME_KEY_INIT) atomicMin(&jobs[cj].key, best);
                        best = ~0ULL;
                    }
                    cbase += geo[cj].n_items;
                    ++cj;
                }
                hme_item_any(blk, geo[cj], arena, it - cbase, nb, nbx, mg_nb, mg_nbx, sr, best);
            }
            if (nb == 1 && best < HME_KEY_INIT) atomicMin(&jobs[cj].key, best);
        }
        if (nb > 1) {
            __syncthreads();
            for (int j = j0; j < j1; ++j) {
                const HGeo &g = geo[j];
                if (g.n_items == 0) continue;
                const uint32_t *sadbuf = reinterpret_cast<const uint32_t *>(arena + g.sad_off);
                const uint32_t mg_wb = magic_div((uint32_t)g.sa_w);
                unsigned long long best = ~0ULL;
                for (int i = threadIdx.x; i < g.bh * g.sa_w; i += SVTME_THREADS) {
                    const int y = (int)__umulhi((uint32_t)i, mg_wb), x = i - y * g.sa_w;
                    if (g.skip_even && !(y & 1)) continue;
                    unsigned long long k = ((unsigned long long)sadbuf[i] << 32) | ((unsigned)y << 16) | (unsigned)x;
                    best = k < best ? k : best;
                }
                if (best < HME_KEY_INIT) atomicMin(&jobs[j].key, best);
            }
        }
        j0 = j1;
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// Full-pel engine: best (SAD, MV) of the 85 square PUs of one 64x64 block over a search area.
// ------------------------------------------------------------------------------------------------
// Thread map (256 threads): b = tid & 15 is the 16x16 block in the reference's z-order slot
// numbering (slot = (by>>1)*8 + (bx>>1)*4 + (by&1)*2 + (bx&1), motion_estimation.c:343), c = byte
// phase, yl = row lane. A half-warp therefore holds the 16 blocks of the same positions, and the
// 32x32 / 64x64 sums are a 4-step reduce-scatter butterfly over those lanes.
//
// PU tables are kept as 64-bit keys  sad << 32 | order  where `order` is the position's rank in
// the reference's evaluation order (the one-point probe ranks before the main search); an update
// happens iff the new key is smaller, which is exactly the reference's sequential strict `<`.
#define FP_TW 64
#define FP_TH 32
#define FP_IDX_BITS 11

struct FpTables {               // shared memory
    unsigned long long key[SVTME_PU_COUNT]; // sad << 32 | order
    uint32_t mv[SVTME_PU_COUNT];            // packed (y << 16) | (x & 0xffff), search-order index
    uint32_t tile_best[SVTME_PU_COUNT];     // per-tile scratch
};

template <int G>
__device__ __forceinline__ void fp_butterfly(uint32_t (&v)[G], int b, int &joff32, int &n32, uint32_t (&v32)[G],
                                             int &joff64, uint32_t &v64) {
    // reduce-scatter over lane bits 0..3; v[j] = 16x16 SAD of position j for block b
    int joff = 0;
    constexpr int N1 = G > 1 ? G / 2 : 1;
    {   // bit 0
        const bool up = b & 1;
        if (G > 1) {
#pragma unroll
            for (int k = 0; k < N1; ++k) {
                uint32_t send = up ? v[k] : v[k + N1], keep = up ? v[k + N1] : v[k];
                v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
            }
            joff += up ? N1 : 0;
        } else
            v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    }
    constexpr int N2 = N1 > 1 ? N1 / 2 : 1;
    {   // bit 1
        const bool up = b & 2;
        if (N1 > 1) {
#pragma unroll
            for (int k = 0; k < N2; ++k) {
                uint32_t send = up ? v[k] : v[k + N2], keep = up ? v[k + N2] : v[k];
                v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            joff += up ? N2 : 0;
        } else
            v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
    }
    joff32 = joff;
    n32 = N2;
#pragma unroll
    for (int k = 0; k < N2; ++k) v32[k] = v[k];
    constexpr int N3 = N2 > 1 ? N2 / 2 : 1;
    {   // bit 2
        const bool up = b & 4;
        if (N2 > 1) {
#pragma unroll
            for (int k = 0; k < N3; ++k) {
                uint32_t send = up ? v[k] : v[k + N3], keep = up ? v[k + N3] : v[k];
                v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            joff += up ? N3 : 0;
        } else
            v[0] += __shfl_xor_sync(0xffffffffu, v[0], 4);
    }
    // bit 3: N3 is 1 for every supported G (<= 8)
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 8);
    joff64 = joff;
    v64 = v[0];
}

template <int G>
__device__ __forceinline__ void fp_tile(const uint8_t *src64, const uint8_t *arena, int copy_bytes, int m, int tw, int th,
                                        bool sub_sad, FpTables *tab) {
    const int tid = threadIdx.x, b = tid & 15, c = (tid >> 4) & 3, yl = tid >> 6;
    const int bx = ((b >> 2) & 1) * 2 + (b & 1), by = ((b >> 3) & 1) * 2 + ((b >> 1) & 1);
    const int xt_n = (tw + 4 * G - 1) / (4 * G), npass = (th + 3) >> 2;
    uint32_t k8[4] = {~0u, ~0u, ~0u, ~0u}, k16 = ~0u, k32 = ~0u, k64 = ~0u;
    const int npar = sub_sad ? 1 : 2;
    for (int pass = 0; pass < npass; ++pass) {
        const int y = 4 * pass + yl;
        for (int xt = 0; xt < xt_n; ++xt) {
            uint32_t acc[G][4];
#pragma unroll
            for (int j = 0; j < G; ++j)
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[j][k] = 0;
            for (int par = 0; par < npar; ++par) {
                uint32_t s[8][4];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    uint4 v = *reinterpret_cast<const uint4 *>(src64 + (16 * by + 2 * r + par) * 64 + 16 * bx);
                    s[r][0] = v.x; s[r][1] = v.y; s[r][2] = v.z; s[r][3] = v.w;
                }
                const int q0 = xt * G + 4 * bx;
                const uint8_t *base = arena + (size_t)c * copy_bytes + 4 * q0;
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const int R = y + 16 * by + 2 * r + par;
                    const uint8_t *rp = base + ((size_t)(R * m + 4 * (R >> 4)) << 4);
                    uint32_t w[G + 4];
                    if (G == 2) {
#pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            uint2 v = *reinterpret_cast<const uint2 *>(rp + 8 * k);
                            w[2 * k] = v.x; w[2 * k + 1] = v.y;
                        }
                    } else {
#pragma unroll
                        for (int k = 0; k < (G + 3 + 3) / 4; ++k) {
                            uint4 v = *reinterpret_cast<const uint4 *>(rp + 16 * k);
                            if (4 * k + 0 < G + 4) w[4 * k + 0] = v.x;
                            if (4 * k + 1 < G + 4) w[4 * k + 1] = v.y;
                            if (4 * k + 2 < G + 4) w[4 * k + 2] = v.z;
                            if (4 * k + 3 < G + 4) w[4 * k + 3] = v.w;
                        }
                    }
                    const int kb = (r >= 4) ? 2 : 0; // 8x8 row half
#pragma unroll
                    for (int j = 0; j < G; ++j) {
                        acc[j][kb] = vsad4(s[r][0], w[j], acc[j][kb]);
                        acc[j][kb] = vsad4(s[r][1], w[j + 1], acc[j][kb]);
                        acc[j][kb + 1] = vsad4(s[r][2], w[j + 2], acc[j][kb + 1]);
                        acc[j][kb + 1] = vsad4(s[r][3], w[j + 3], acc[j][kb + 1]);
                    }
                }
            }
            const uint32_t yvalid = y < th;
            uint32_t v16[G];
#pragma unroll
            for (int j = 0; j < G; ++j) {
                const int x = xt * 4 * G + 4 * j + c;
                const bool ok = yvalid && x < tw;
                const uint32_t idx = (uint32_t)(y * FP_TW + x);
                uint32_t sum = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t sad = sub_sad ? acc[j][k] << 1 : acc[j][k];
                    sum += sad;
                    const uint32_t key = ok ? ((sad << FP_IDX_BITS) | idx) : ~0u;
                    k8[k] = min(k8[k], key);
                }
                v16[j] = sum;
                k16 = min(k16, ok ? ((sum << FP_IDX_BITS) | idx) : ~0u);
            }
            int joff32, n32, joff64;
            uint32_t v32[G], v64;
            fp_butterfly<G>(v16, b, joff32, n32, v32, joff64, v64);
#pragma unroll
            for (int k = 0; k < G; ++k) {
                if (k < n32) {
                    const int x = xt * 4 * G + 4 * (joff32 + k) + c;
                    const bool ok = yvalid && x < tw;
                    k32 = min(k32, ok ? ((v32[k] << FP_IDX_BITS) | (uint32_t)(y * FP_TW + x)) : ~0u);
                }
            }
            {
                const int x = xt * 4 * G + 4 * joff64 + c;
                const bool ok = yvalid && x < tw;
                k64 = min(k64, ok ? ((v64 << FP_IDX_BITS) | (uint32_t)(y * FP_TW + x)) : ~0u);
            }
        }
    }
    // cross-thread minima -> tab->tile_best (search-order PU index)
    const int lane = tid & 31;
#pragma unroll
    for (int k = 0; k < 4; ++k) k8[k] = min(k8[k], __shfl_xor_sync(0xffffffffu, k8[k], 16));
    k16 = min(k16, __shfl_xor_sync(0xffffffffu, k16, 16));
    k32 = min(k32, __shfl_xor_sync(0xffffffffu, k32, 16));
    k32 = min(k32, __shfl_xor_sync(0xffffffffu, k32, 1));
    k32 = min(k32, __shfl_xor_sync(0xffffffffu, k32, 2));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) k64 = min(k64, __shfl_xor_sync(0xffffffffu, k64, o));
    if (lane < 16) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (k8[k] != ~0u) atomicMin(&tab->tile_best[21 + 4 * b + k], k8[k]);
        if (k16 != ~0u) atomicMin(&tab->tile_best[5 + b], k16);
        if ((lane & 3) == 0 && k32 != ~0u) atomicMin(&tab->tile_best[1 + (b >> 2)], k32);
        if (lane == 0 && k64 != ~0u) atomicMin(&tab->tile_best[0], k64);
    }
}

// All threads call with identical arguments. ref0: global pointer to the reference pixel under
// block(0,0) at search position (0,0). (ox, oy): MV of search position (0,0). order_base: rank of
// position (0,0) in the reference's evaluation order (see FpTables). src64: shared memory, 64x64
// source block with row stride 64.
__device__ void fullpel_search(const uint8_t *src64, const uint8_t *ref0, int ref_stride, int ox, int oy, int sa_w,
                               int sa_h, bool sub_sad, uint32_t order_base, FpTables *tab, uint8_t *arena) {
    if (sa_w <= 0 || sa_h <= 0) return;
    const int G = sa_w <= 8 ? 2 : (sa_w <= 16 ? 4 : 8);
    const int last_row = sub_sad ? 62 : 63; // last block row that is read
    for (int ty0 = 0; ty0 < sa_h; ty0 += FP_TH) {
        const int th = min(FP_TH, sa_h - ty0);
        for (int tx0 = 0; tx0 < sa_w; tx0 += FP_TW) {
            const int tw = min(FP_TW, sa_w - tx0);
            const int span = ((tw + 4 * G - 1) / (4 * G)) * 4 * G;
            const int vec_words = G == 2 ? 6 : ((G + 3 + 3) / 4) * 4;
            const int words = (span / 4 - G) + 12 + vec_words;
            const int m = (words + 3) / 4;
            const int rows = ((th + 3) & ~3) + last_row;
            const int copy_chunks = rows * m + 4 * ((rows - 1) >> 4) + 4;
            const int copy_bytes = copy_chunks * 16;
            __syncthreads();
            stage_window4(arena, ref0 + (size_t)ty0 * ref_stride + tx0, ref_stride, rows, m, 4, copy_bytes);
            for (int i = threadIdx.x; i < SVTME_PU_COUNT; i += SVTME_THREADS) tab->tile_best[i] = ~0u;
            __syncthreads();
            if (G == 2)
                fp_tile<2>(src64, arena, copy_bytes, m, tw, th, sub_sad, tab);
            else if (G == 4)
                fp_tile<4>(src64, arena, copy_bytes, m, tw, th, sub_sad, tab);
            else
                fp_tile<8>(src64, arena, copy_bytes, m, tw, th, sub_sad, tab);
            __syncthreads();
            if (threadIdx.x < SVTME_PU_COUNT) {
                const uint32_t k = tab->tile_best[threadIdx.x];
                if (k != ~0u) {
                    const uint32_t sad = k >> FP_IDX_BITS, idx = k & ((1u << FP_IDX_BITS) - 1);
                    const int gy = ty0 + (int)(idx / FP_TW), gx = tx0 + (int)(idx % FP_TW);
                    const unsigned long long key =
                        ((unsigned long long)sad << 32) | (order_base + (uint32_t)gy * (uint32_t)sa_w + (uint32_t)gx);
                    if (key < tab->key[threadIdx.x]) {
                        tab->key[threadIdx.x] = key;
                        tab->mv[threadIdx.x] = ((uint32_t)(uint16_t)(oy + gy) << 16) | (uint16_t)(ox + gx);
                    }
                }
            }
        }
    }
    __syncthreads();
}

// One search position for all 85 PUs (the 8x8-variance probe of integer_search_b64,
// motion_estimation.c:1393-1425, i.e. open_loop_me_get_search_point_results_block): no staging.
// Thread t owns two (SUB_SAD: one) rows of 8x8 block e = t >> 2, with e = 4 * slot16 + k in the
// table's own order, so 8x8 / 16x16 sums are shuffles and only 32x32 / 64x64 go through smem.
__device__ void fullpel_probe(const uint8_t *src64, const uint8_t *ref0, int ref_stride, int mvx, int mvy, bool sub_sad,
                              uint32_t order, FpTables *tab) {
    const int tid = threadIdx.x, e = tid >> 2, q = tid & 3;
    const int b = e >> 2, k = e & 3;
    const int bx = ((b >> 2) & 1) * 2 + (b & 1), by = ((b >> 3) & 1) * 2 + ((b >> 1) & 1);
    const int x8 = 16 * bx + 8 * (k & 1), y8 = 16 * by + 8 * (k >> 1);
    uint32_t acc = 0;
    const int nr = sub_sad ? 1 : 2;
    for (int i = 0; i < nr; ++i) {
        const int y = y8 + 2 * q + i;
        const uint2 sv = *reinterpret_cast<const uint2 *>(src64 + y * 64 + x8);
        const uintptr_t ai = reinterpret_cast<uintptr_t>(ref0 + (size_t)y * ref_stride + x8);
        const uint32_t *al = reinterpret_cast<const uint32_t *>(ai & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(ai & 3) * 8;
        const uint32_t w0 = __ldg(al), w1 = __ldg(al + 1), w2 = sh ? __ldg(al + 2) : 0u;
        acc = vsad4(sv.x, __funnelshift_r(w0, w1, sh), acc);
        acc = vsad4(sv.y, __funnelshift_r(w1, w2, sh), acc);
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    const uint32_t s8 = sub_sad ? acc << 1 : acc;
    uint32_t s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 4);
    s16 += __shfl_xor_sync(0xffffffffu, s16, 8);
    if (q == 0) tab->tile_best[21 + e] = s8;
    if ((tid & 15) == 0) tab->tile_best[5 + b] = s16;
    __syncthreads();
    if (tid < 5) {
        uint32_t sum = 0;
        if (tid == 0)
            for (int i = 0; i < 16; ++i) sum += tab->tile_best[5 + i];
        else
            for (int i = 0; i < 4; ++i) sum += tab->tile_best[5 + 4 * (tid - 1) + i];
        tab->tile_best[tid] = sum;
    }
    __syncthreads();
    if (tid < SVTME_PU_COUNT) {
        const unsigned long long key = ((unsigned long long)tab->tile_best[tid] << 32) | order;
        if (key < tab->key[tid]) {
            tab->key[tid] = key;
            tab->mv[tid] = ((uint32_t)(uint16_t)mvy << 16) | (uint16_t)mvx;
        }
    }
    __syncthreads();
}

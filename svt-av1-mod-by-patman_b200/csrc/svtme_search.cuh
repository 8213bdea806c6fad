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

#ifndef SVTME_THREADS
#define SVTME_THREADS 128          // CTA size; 64-thread groups share one row lane of the full-pel engine
#endif
#ifndef SVTME_CTAS_PER_SM
#define SVTME_CTAS_PER_SM 4
#endif
#ifndef SVTME_ARENA_BYTES
#define SVTME_ARENA_BYTES (32 * 1024) // full-pel engine (four phase copies of a 32x16-position tile)
#endif
#ifndef SVTME_HME_ARENA_BYTES
#define SVTME_HME_ARENA_BYTES (32 * 1024) // HME engine: the tallest pre-HME window (8x400, 20.7 KB) plus its wide twin (7.9 KB)
#endif
#ifndef SVTME_INT_SMALL_THREADS
#define SVTME_INT_SMALL_THREADS 64 // small full-pel areas: one row lane per CTA (a 3-row area wastes no lane), more CTAs in flight
#endif
#ifndef SVTME_INT_SMALL_CTAS_PER_SM
#define SVTME_INT_SMALL_CTAS_PER_SM 14 // measured: 10 -> 0.97, 12 -> 0.90, 14 -> 0.84 ms (latency-bound: residency wins over spills)
#endif
#ifndef SVTME_HME_CTAS_PER_SM
#define SVTME_HME_CTAS_PER_SM 4
#endif
#define FP_NT ((int)blockDim.x)   // the full-pel engine adapts to the CTA size of its kernel (64 or 128 threads)
#define SVTME_FP_YL (FP_NT >> 6) // row lanes of the full-pel engine: 64-thread groups share one row lane
#ifndef SVTME_FP_SMALL_ARENA_BYTES
#define SVTME_FP_SMALL_ARENA_BYTES (8 * 1024) // raw tile of a small full-pel window: 16x5 positions = 67 rows x 112 bytes; the rare
                                              // wider tile (24x7 after the x1.5 of the variance probe) reads global memory directly
#endif
#ifndef SVTME_FP_DIRECT_MAX
#define SVTME_FP_DIRECT_MAX 96 // search areas up to this many positions skip the shared-memory staging
#endif

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) { // LDGSTS: global -> shared, no registers
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\n cp.async.wait_group 0;" ::: "memory"); }

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
__device__ __noinline__ void stage_window4(uint8_t *arena, const uint8_t *__restrict__ g, int gstride, int rows,
                                              int m, int skew, int copy_bytes, int nwords) {
    // Flat (row, word) item list over the whole CTA, four items per thread in flight so that the
    // global-load latency is paid once per four rows-worth of words, not once per row. Each item
    // reads three ALIGNED global words (the row address is rounded down to 4 bytes) and derives the
    // four phase words with funnel shifts; neighbouring lanes hit the same L1 lines.
    // nwords <= 4*m: words of a copy row that are actually read by the search.
    const int total = rows * nwords;
    const uint32_t mg = 0xFFFFFFFFu / (uint32_t)nwords + 1u; // exact for total < 2^16
    constexpr int U = 4;
    for (int base = threadIdx.x; base < total; base += U * FP_NT) {
        uint32_t w0[U], w1[U], w2[U], mis[U];
        uint8_t *dst[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * FP_NT;
            if (i < total) {
                const int R = (int)__umulhi((uint32_t)i, mg), kw = i - R * nwords;
                const uintptr_t ai = reinterpret_cast<uintptr_t>(g + (size_t)R * gstride);
                const uint32_t *al = reinterpret_cast<const uint32_t *>(ai & ~(uintptr_t)3) + kw;
                mis[u] = (uint32_t)(ai & 3);
                w0[u] = __ldg(al);
                w1[u] = __ldg(al + 1);
                w2[u] = mis[u] ? __ldg(al + 2) : 0u;
                dst[u] = arena + ((size_t)(R * m + skew * (R >> 4)) << 4) + 4 * kw;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (base + u * FP_NT < total) {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const uint32_t o = c + mis[u]; // byte offset of phase c inside (w0,w1,w2)
                    const uint32_t lo = o >= 4 ? w1[u] : w0[u], hi = o >= 4 ? w2[u] : w1[u];
                    *reinterpret_cast<uint32_t *>(dst[u] + (size_t)c * copy_bytes) = __funnelshift_r(lo, hi, (o & 3) * 8);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// HME engine
// ------------------------------------------------------------------------------------------------
#define SVTME_MAX_JOBS 32
#define HME_NT ((int)blockDim.x) // the HME engine adapts to the CTA size of its kernel
#ifndef SVTME_HME_THREADS
#define SVTME_HME_THREADS 128
#endif

__device__ __forceinline__ uint32_t magic_div(uint32_t d) { return 0xFFFFFFFFu / d + 1u; } // exact for n < 2^16

struct HBlk {            // the source block of a round (uniform over the CTA)
    const uint8_t *src;  // shared memory, compared row r at src + r*src_pitch
    int src_pitch;       // bytes between two COMPARED source rows
    int bw;              // block width in pixels (any value >= 1)
    int nrows;           // number of (sub-sampled) rows actually compared
    int row_step;        // reference rows between two compared rows (2 = SUB_SAD_SEARCH)
    int nbx, nb;         // 16x8 register sub-blocks across / in total
    uint32_t mg_nb, mg_nbx;
};

__device__ __forceinline__ HBlk make_hblk(const uint8_t *src, int src_pitch, int bw, int nrows, int row_step) {
    HBlk b;
    b.src = src;
    b.src_pitch = src_pitch;
    b.bw = bw;
    b.nrows = nrows;
    b.row_step = row_step;
    b.nbx = (bw + 15) >> 4;
    b.nb = b.nbx * ((nrows + 7) >> 3);
    b.mg_nb = magic_div((uint32_t)max(b.nb, 1));
    b.mg_nbx = magic_div((uint32_t)max(b.nbx, 1));
    return b;
}

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
    int xt_n, n_items, G, skip_even, variant;
    int ref_off;         // byte offset of the tile origin from the job's ref pointer
    int rows, cpr;       // staged rows (padded to whole item tiles), 16-byte chunks staged per row
    int rows_real;       // rows that exist in the window; staged rows beyond repeat the last one
    int stage_units;     // rows * cpr
    int scan_n;          // positions whose SAD is summed over sub-blocks in the SAD buffer (nb > 1)
    uint32_t mg_xt, mg_cpr, mg_wb;
    uint32_t need;       // arena bytes (copies + SAD buffer)
    int big;             // does not fit as a whole: tiled single-job path
    int stage_base, item_base, scan_base; // prefix sums inside the group
    int row_off, x_shift; // strip mode (strip_kernel): rows above / bytes left of the window in the shared strip's tiling
    int n_cols, col_base; // strip mode, streamed windows (strip_columns): columns of this job / prefix sum inside the group
    int col_chunks;       // ... each column cut into this many y-chunks (balances the CTA: see strip_kernel's planner)
};

struct HGroup {
    int j0, j1, big;
    int stage_total, item_total, scan_total;
};

struct HPlan {           // shared-memory scratch of one round
    HGeo geo[SVTME_MAX_JOBS];
    HGroup grp[SVTME_MAX_JOBS];
    int n_groups;
};

// The fields of HGeo a work item needs, held in registers while a thread stays inside one job.
struct ItemGeo {
    uint32_t arena_off, sad_off;
    int pitch, xt_n, sa_w, bh, y0, skip_even, variant;
    int row_off, x_shift; // window origin inside a SHARED strip (0 when the job staged its own window)
    uint32_t mg_xt;
};
__device__ __forceinline__ void load_item_geo(const HGeo &g, ItemGeo &ig) {
    ig.arena_off = g.arena_off; ig.sad_off = g.sad_off;
    ig.pitch = g.m << 4; ig.xt_n = g.xt_n; ig.sa_w = g.sa_w; ig.bh = g.bh; ig.y0 = g.y0;
    ig.skip_even = g.skip_even; ig.variant = g.variant; ig.mg_xt = g.mg_xt;
    ig.row_off = g.row_off; ig.x_shift = g.x_shift;
}

// Result sink of one position. Single-sub-block blocks (<= 16x8, SAD < 2^15) keep a 32-bit running
// key  sad << 17 | (y * sa_w + x)  (tile-local raster index < 2^16); larger blocks add their
// sub-block SAD into the per-position buffer.
__device__ __forceinline__ void hme_emit(bool single, uint32_t sad, int y, int x, const ItemGeo &ig, uint8_t *arena,
                                         uint32_t &best32) {
    const uint32_t idx = (uint32_t)(y * ig.sa_w + x);
    if (single)
        best32 = min(best32, (sad << 17) | idx);
    else
        atomicAdd(reinterpret_cast<uint32_t *>(arena + ig.sad_off) + idx, sad);
}

// One work item = G same-phase x-positions times TY y-positions (spaced row_step apart, so that
// they share reference rows) against one FULL 16x8 source sub-block held in registers. The window
// is staged once, byte-aligned to the search origin; phase c = x & 3 is applied in registers with
// funnel shifts (G+4 SHF per reference row, amortised over 4*G*TY VABSDIFF4). No predicates in
// the row loop: staging pads the window to whole y-tiles (clamped rows).
template <int G, int TY>
__device__ __forceinline__ void hme_item_fast(const HBlk &blk, const ItemGeo &ig, uint8_t *arena, int c, int xt, int yb,
                                              int bxi, int byi, uint32_t &best32) {
    const int rs = blk.row_step;
    uint32_t acc[TY][G];
#pragma unroll
    for (int ty = 0; ty < TY; ++ty)
#pragma unroll
        for (int j = 0; j < G; ++j) acc[ty][j] = 0;
    const uint8_t *base = arena + ig.arena_off + 4 * (xt * G + 4 * bxi);
    const uint8_t *sp = blk.src + (size_t)(8 * byi) * blk.src_pitch + 16 * bxi;
    const uint32_t sh = 8 * c;
    const int row0 = ig.row_off + yb + 8 * byi * rs;
    // sliding window of the TY source rows in use: reference row rho meets source row rho - ty at
    // y-position ty. Source rows come from shared memory as warp-wide broadcasts (every lane of an
    // item group reads the same address), which keeps the register budget at TY*4 instead of 32.
    uint32_t sw[TY][4];
#pragma unroll
    for (int rho = 0; rho < TY + 7; ++rho) {
        if (rho < 8) {
            const uint4 v = *reinterpret_cast<const uint4 *>(sp + rho * blk.src_pitch);
            sw[rho % TY][0] = v.x; sw[rho % TY][1] = v.y; sw[rho % TY][2] = v.z; sw[rho % TY][3] = v.w;
        }
        const int row = row0 + rho * rs;
        const uint8_t *rp = base + row * ig.pitch + ((row >> 3) << 4);
        uint32_t w[G + 5];
        if (G == 2) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const uint2 v = *reinterpret_cast<const uint2 *>(rp + 8 * k);
                w[2 * k] = v.x; w[2 * k + 1] = v.y;
            }
        } else {
#pragma unroll
            for (int k = 0; k < (G + 4) / 4; ++k) {
                const uint4 v = *reinterpret_cast<const uint4 *>(rp + 16 * k);
                w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
            }
        }
        w[G + 4] = *reinterpret_cast<const uint32_t *>(rp + 4 * (G + 4));
#pragma unroll
        for (int i = 0; i < G + 4; ++i) w[i] = __funnelshift_r(w[i], w[i + 1], sh);
#pragma unroll
        for (int ty = 0; ty < TY; ++ty) {
            const int r = rho - ty;
            if (r >= 0 && r < 8) {
#pragma unroll
                for (int j = 0; j < G; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[ty][j] = vsad4(sw[r % TY][i], w[j + i], acc[ty][j]);
            }
        }
    }
    const bool single = blk.nb == 1;
#pragma unroll
    for (int ty = 0; ty < TY; ++ty) {
        const int y = yb + ty * rs;
        if (y < ig.bh && !(ig.skip_even && !((ig.y0 + y) & 1))) {
#pragma unroll
            for (int j = 0; j < G; ++j) {
                const int x = xt * 4 * G + 4 * j + c - ig.x_shift;
                if ((unsigned)x < (unsigned)ig.sa_w) hme_emit(single, acc[ty][j], y, x, ig, arena, best32);
            }
        }
    }
}

// Flat wide windows (the W x 3 pre-HME search): ONE item = 32 x-positions of one byte phase (G = 8) times ALL three
// y-positions, both row parities together. The 17 reference rows the three positions touch are each loaded (3 LDS.128)
// and shifted (11 SHF) once; an even row meets two source rows (y = 0 and y = 2), an odd row one (y = 1): 768
// VABSDIFF4 per item, none wasted (the parity-split 2-row tiles computed four y-positions for three). Source rows
// come from shared memory as they are needed (two live at a time). Full 16x8 sub-blocks only.
__device__ __forceinline__ void hme_item_wide3(const HBlk &blk, const ItemGeo &ig, uint8_t *arena, int c, int xt, uint32_t &best32) {
    constexpr int G = 8;
    uint32_t acc[3][G];
#pragma unroll
    for (int y = 0; y < 3; ++y)
#pragma unroll
        for (int j = 0; j < G; ++j) acc[y][j] = 0;
    const uint8_t *base = arena + ig.arena_off + 4 * (xt * G);
    const uint32_t sh = 8 * c;
    uint32_t sa[4], sb[4]; // source rows R / 2 and R / 2 - 1
#pragma unroll
    for (int i = 0; i < 4; ++i) sa[i] = sb[i] = 0;
#pragma unroll
    for (int R = 0; R < 17; ++R) {
        const int row = ig.row_off + R;
        const uint8_t *rp = base + row * ig.pitch + ((row >> 3) << 4);
        uint32_t w[G + 4];
#pragma unroll
        for (int k = 0; k < (G + 4) / 4; ++k) {
            const uint4 v = *reinterpret_cast<const uint4 *>(rp + 16 * k);
            w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
        }
#pragma unroll
        for (int i = 0; i < G + 3; ++i) w[i] = __funnelshift_r(w[i], w[i + 1], sh);
        if ((R & 1) == 0) {
            if (R > 0) {
#pragma unroll
                for (int i = 0; i < 4; ++i) sb[i] = sa[i];
            }
            if (R < 16) {
                const uint4 v = *reinterpret_cast<const uint4 *>(blk.src + (R >> 1) * blk.src_pitch);
                sa[0] = v.x; sa[1] = v.y; sa[2] = v.z; sa[3] = v.w;
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < G; ++j) acc[0][j] = vsad4(sa[i], w[j + i], acc[0][j]); // y = 0, source row R / 2
            }
            if (R >= 2) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < G; ++j) acc[2][j] = vsad4(sb[i], w[j + i], acc[2][j]); // y = 2, source row R / 2 - 1
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < G; ++j) acc[1][j] = vsad4(sa[i], w[j + i], acc[1][j]); // y = 1, source row (R - 1) / 2
        }
    }
#pragma unroll
    for (int j = 0; j < G; ++j) {
        const int x = xt * 4 * G + 4 * j + c - ig.x_shift;
        if ((unsigned)x < (unsigned)ig.sa_w) {
#pragma unroll
            for (int y = 0; y < 3; ++y) best32 = min(best32, (acc[y][j] << 17) + (uint32_t)(y * ig.sa_w + x));
        }
    }
}

// Generic item: partial sub-blocks (block height not a multiple of 8 rows, width not a multiple of
// 16) — edge b64s only. Same tile shape and loads as the fast item, but rolled loops over the y-positions
// and the rows that exist (compact code), and byte masks when the sub-block is narrower than 16.
template <int G>
__device__ __forceinline__ void hme_slow_body(const HBlk &blk, const ItemGeo &ig, uint8_t *arena, int TY, int c, int xt, int yb,
                                              int bxi, int byi, uint32_t &best32) {
    const int rs = blk.row_step;
    const int rows_here = min(8, blk.nrows - 8 * byi), vb = min(16, blk.bw - 16 * bxi);
    uint32_t msk[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = min(max(vb - 4 * i, 0), 4);
        msk[i] = n >= 4 ? ~0u : ((1u << (8 * n)) - 1u);
    }
    const bool single = blk.nb == 1, narrow = vb < 16;
    const uint32_t sh = 8 * c;
    const uint8_t *base = arena + ig.arena_off + 4 * (xt * G + 4 * bxi);
    const uint8_t *sp = blk.src + (size_t)(8 * byi) * blk.src_pitch + 16 * bxi;
    for (int ty = 0; ty < TY; ++ty) {
        const int y = yb + ty * rs;
        if (y >= ig.bh || (ig.skip_even && !((ig.y0 + y) & 1))) continue;
        uint32_t acc[G];
#pragma unroll
        for (int j = 0; j < G; ++j) acc[j] = 0;
#pragma unroll 1
        for (int r = 0; r < rows_here; ++r) {
            const int row = ig.row_off + y + (8 * byi + r) * rs;
            const uint8_t *rp = base + row * ig.pitch + ((row >> 3) << 4);
            uint32_t w[G + 5];
            if (G == 2) {
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const uint2 v = *reinterpret_cast<const uint2 *>(rp + 8 * k);
                    w[2 * k] = v.x; w[2 * k + 1] = v.y;
                }
            } else {
#pragma unroll
                for (int k = 0; k < (G + 4) / 4; ++k) {
                    const uint4 v = *reinterpret_cast<const uint4 *>(rp + 16 * k);
                    w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
                }
            }
            w[G + 4] = *reinterpret_cast<const uint32_t *>(rp + 4 * (G + 4));
#pragma unroll
            for (int i = 0; i < G + 4; ++i) w[i] = __funnelshift_r(w[i], w[i + 1], sh);
            const uint4 sv = *reinterpret_cast<const uint4 *>(sp + r * blk.src_pitch);
            uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
            if (narrow) {
#pragma unroll
                for (int i = 0; i < 4; ++i) s[i] &= msk[i];
#pragma unroll
                for (int j = 0; j < G; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[j] = vsad4(s[i], w[j + i] & msk[i], acc[j]);
            } else {
#pragma unroll
                for (int j = 0; j < G; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[j] = vsad4(s[i], w[j + i], acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < G; ++j) {
            const int x = xt * 4 * G + 4 * j + c - ig.x_shift;
            if ((unsigned)x < (unsigned)ig.sa_w) hme_emit(single, acc[j], y, x, ig, arena, best32);
        }
    }
}

__device__ __noinline__ uint32_t hme_item_slow(const HBlk blk, const ItemGeo ig, uint8_t *arena, int G, int TY, int c, int xt,
                                               int yb, int bxi, int byi, uint32_t best32) {
    if (G == 2)
        hme_slow_body<2>(blk, ig, arena, TY, c, xt, yb, bxi, byi, best32);
    else if (G == 4)
        hme_slow_body<4>(blk, ig, arena, TY, c, xt, yb, bxi, byi, best32);
    else
        hme_slow_body<8>(blk, ig, arena, TY, c, xt, yb, bxi, byi, best32);
    return best32;
}

template <bool STRIP>
__device__ __forceinline__ void hme_item_any(const HBlk &blk, const ItemGeo &ig, uint8_t *arena, int it, uint32_t &best32) {
    const int nb = blk.nb, nbx = blk.nbx, rs = blk.row_step;
    if (STRIP && ig.variant == 5) { // flat W x 3 window of a full block: item = (x tile, byte phase), all three y
        hme_item_wide3(blk, ig, arena, it & 3, it >> 2, best32);
        return;
    }
    int t = nb == 1 ? it : (int)__umulhi((uint32_t)it, blk.mg_nb);
    const int b = it - t * nb;
    const int c = t & 3;
    t >>= 2;
    const int yt = ig.xt_n == 1 ? t : (int)__umulhi((uint32_t)t, ig.mg_xt);
    const int xt = t - yt * ig.xt_n;
    const int byi = nbx == 1 ? b : (int)__umulhi((uint32_t)b, blk.mg_nbx);
    const int bxi = b - byi * nbx;
    const int G = (ig.variant == 0 || ig.variant == 3) ? 2 : ((ig.variant == 1 || ig.variant == 4) ? 4 : 8);
    const int TY = (ig.variant == 2 || ig.variant == 4) ? 2 : (ig.variant == 3 ? 8 : 4);
    const int yb = (yt & (rs - 1)) + (yt >> (rs - 1)) * TY * rs; // rs is 1 or 2
    const bool full = blk.nrows - 8 * byi >= 8 && blk.bw - 16 * bxi >= 16;
    if (!full) {
        best32 = hme_item_slow(blk, ig, arena, G, TY, c, xt, yb, bxi, byi, best32);
        return;
    }
    // few variants on purpose: the unrolled bodies are large and the kernel must stay I-cache friendly
    if (ig.variant == 0)
        hme_item_fast<2, 4>(blk, ig, arena, c, xt, yb, bxi, byi, best32);
    else if (ig.variant == 1)
        hme_item_fast<4, 4>(blk, ig, arena, c, xt, yb, bxi, byi, best32);
    else if (!STRIP || ig.variant == 2)
        hme_item_fast<8, 2>(blk, ig, arena, c, xt, yb, bxi, byi, best32);
    else if (ig.variant == 4) // short 16-wide windows (HME level 0 of near references: 16x4, 16x12 positions)
        hme_item_fast<4, 2>(blk, ig, arena, c, xt, yb, bxi, byi, best32);
    else // tall narrow windows, chosen only where a CTA has enough items to stay balanced (strip mode)
        hme_item_fast<2, 8>(blk, ig, arena, c, xt, yb, bxi, byi, best32);
}

// Tile geometry for a (sub)area of wb x hb positions; g.need = the arena bytes it takes.
__device__ __forceinline__ void hme_geo(const HBlk &blk, int wb, int hb, int skip_even, HGeo &g,
                                        int arena_bytes = SVTME_HME_ARENA_BYTES) {
    const int G = wb <= 8 ? 2 : (wb <= 16 ? 4 : 8);
    const int TY = G == 8 ? 2 : 4;
    const int tile_w = 4 * G;
    const int xt_n = (wb + tile_w - 1) / tile_w;
    const int rs = blk.row_step;
    const int n_yt = (((hb + rs - 1) / rs) + TY - 1) / TY; // y-tiles per row-parity class
    g.variant = G == 2 ? 0 : (G == 4 ? 1 : 2);
    const int words = (xt_n - 1) * G + 4 * (blk.nbx - 1) + G + 5;
    const int cpr = (words + 3) >> 2;
    const int m = cpr | 1; // odd chunk pitch; plus one skew chunk every 8 rows (see hme_item_fast)
    g.rows_real = hb + (blk.nrows - 1) * rs;
    // rows the (unpredicated) items of the last y-tiles touch: whole tiles of 8-row sub-blocks
    const int rows = n_yt * TY * rs + (8 * ((blk.nrows + 7) >> 3) - 1) * rs;
    g.copy_bytes = (rows * m + (rows >> 3) + 1) * 16;
    g.m = m;
    g.sa_w = wb;
    g.bh = hb;
    g.x0 = g.y0 = 0;
    g.ref_off = 0;
    g.xt_n = xt_n;
    g.n_items = rs * n_yt * xt_n * 4 * blk.nb;
    g.G = G;
    g.skip_even = skip_even;
    g.rows = rows;
    g.cpr = cpr;
    g.stage_units = rows * cpr;
    g.scan_n = blk.nb > 1 ? hb * wb : 0;
    g.mg_xt = magic_div((uint32_t)xt_n);
    g.mg_cpr = magic_div((uint32_t)cpr);
    g.mg_wb = magic_div((uint32_t)wb);
    g.need = (uint32_t)g.copy_bytes + (uint32_t)g.scan_n * 4u;
    g.big = g.need > (uint32_t)arena_bytes || g.n_items >= 65536 || g.stage_units >= 65536 || hb * wb >= 65536;
    g.arena_off = 0;
    g.sad_off = (uint32_t)g.copy_bytes;
    g.stage_base = g.item_base = g.scan_base = 0;
    g.row_off = g.x_shift = 0;
}

// Stage one 16-byte chunk of a window row: five aligned global words -> byte-aligned chunk.
__device__ __forceinline__ void stage_unit_load(const uint8_t *__restrict__ row, int k, uint32_t (&w)[5], uint32_t &mis) {
    const uintptr_t ai = reinterpret_cast<uintptr_t>(row);
    const uint32_t *al = reinterpret_cast<const uint32_t *>(ai & ~(uintptr_t)3) + 4 * k;
    mis = (uint32_t)(ai & 3);
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = __ldg(al + i);
    w[4] = mis ? __ldg(al + 4) : 0u;
}
__device__ __forceinline__ void stage_unit_store(uint8_t *dst, const uint32_t (&w)[5], uint32_t mis) {
    const uint32_t sh = mis * 8;
    *reinterpret_cast<uint4 *>(dst) = make_uint4(__funnelshift_r(w[0], w[1], sh), __funnelshift_r(w[1], w[2], sh),
                                                 __funnelshift_r(w[2], w[3], sh), __funnelshift_r(w[3], w[4], sh));
}
__device__ __forceinline__ uint8_t *stage_dst(uint8_t *arena, const HGeo &g, int R, int k) {
    return arena + g.arena_off + ((size_t)(R * g.m + (R >> 3) + k) << 4);
}

// Minimum of 64-bit keys over the lanes that are flushing the same job, one atomic per such set.
__device__ __forceinline__ void flush_key(HJob *jobs, int j, unsigned long long best) {
    const unsigned act = __activemask();
    const unsigned peers = __match_any_sync(act, j);
    const uint32_t hi = (uint32_t)(best >> 32), lo = (uint32_t)best;
    const uint32_t mhi = __reduce_min_sync(peers, hi);
    const uint32_t mlo = __reduce_min_sync(peers, hi == mhi ? lo : 0xffffffffu);
    if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) {
        const unsigned long long key = ((unsigned long long)mhi << 32) | mlo;
        if (key < HME_KEY_INIT) atomicMin(&jobs[j].key, key);
    }
}

// SAD of the block at ONE position per job (zz-SAD, check_00_center): no staging, one warp per
// job, lanes over (row, 4-byte column) with unaligned reference words assembled from two aligned
// loads. All threads call with identical arguments.
__device__ __noinline__ void hme_points(const HBlk &blk, HJob *jobs, int n) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wcols = (blk.bw + 3) >> 2, items = blk.nrows * wcols;
    const uint32_t mg_wc = magic_div((uint32_t)wcols);
    for (int j = warp; j < n; j += (HME_NT >> 5)) {
        const HJob &job = jobs[j];
        uint32_t acc = 0;
        const size_t rstep = (size_t)blk.row_step * job.stride;
        if (blk.bw == 64 && blk.nrows <= 32 &&
            ((reinterpret_cast<uintptr_t>(blk.src) | reinterpret_cast<uintptr_t>(job.ref) | (uintptr_t)blk.src_pitch | rstep) & 15) == 0) {
            // aligned 64-wide block (zz-SAD): four lanes per row (one 16-byte chunk each), eight rows per
            // load instruction -> every cache line is touched once (one lane per row touched 32 lines
            // per instruction and was bound by the L1 tag rate)
            const int ch = lane & 3;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = (lane >> 2) + 8 * i;
                if (r < blk.nrows) {
                    const uint4 a = *(reinterpret_cast<const uint4 *>(blk.src + (size_t)r * blk.src_pitch) + ch);
                    const uint4 b = __ldg(reinterpret_cast<const uint4 *>(job.ref + (size_t)r * rstep) + ch);
                    acc = vsad4(a.x, b.x, acc);
                    acc = vsad4(a.y, b.y, acc);
                    acc = vsad4(a.z, b.z, acc);
                    acc = vsad4(a.w, b.w, acc);
                }
            }
        } else
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

// ---- group helpers (separate functions: separate register allocations)
__device__ __noinline__ void hme_group_stage(const HJob *jobs, const HGeo *geo, const HGroup &G_ref, uint8_t *arena) {
    const HGroup G = G_ref;
    constexpr int U = 4; // chunks per thread in flight
    const int tid = threadIdx.x;
    int cj = G.j0;
    for (int base = tid; base < G.stage_total; base += U * HME_NT) {
        uint32_t w[U][5], mis[U];
        uint8_t *dst[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * HME_NT;
            if (i < G.stage_total) {
                while (i >= geo[cj].stage_base + geo[cj].stage_units) ++cj;
                const HGeo &g = geo[cj];
                const int il = i - g.stage_base;
                const int R = (int)__umulhi((uint32_t)il, g.mg_cpr), k = il - R * g.cpr;
                stage_unit_load(jobs[cj].ref + g.ref_off + (size_t)min(R, g.rows_real - 1) * jobs[cj].stride, k, w[u], mis[u]);
                dst[u] = stage_dst(arena, g, R, k);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (base + u * HME_NT < G.stage_total) stage_unit_store(dst[u], w[u], mis[u]);
    }
    uint32_t *sadreg = reinterpret_cast<uint32_t *>(arena + geo[G.j0].sad_off);
    for (int i = tid; i < G.scan_total; i += HME_NT) sadreg[i] = 0;
}

// Flush of a 32-bit running key (see hme_emit): lanes flushing the same job combine with one
// redux.min; the leader rebuilds the 64-bit key  sad << 32 | y << 16 | x  in search-area coordinates.
__device__ __forceinline__ void flush_key32(HJob *jobs, const HGeo *geo, int j, uint32_t best32) {
    const unsigned act = __activemask();
    const unsigned peers = __match_any_sync(act, j);
    const uint32_t m = __reduce_min_sync(peers, best32);
    if ((int)(threadIdx.x & 31) == __ffs(peers) - 1 && m != ~0u) {
        const HGeo &g = geo[j];
        const uint32_t idx = m & 0x1ffffu;
        const uint32_t yl = __umulhi(idx, g.mg_wb), xl = idx - yl * (uint32_t)g.sa_w;
        const unsigned long long key =
            ((unsigned long long)(m >> 17) << 32) | ((unsigned)(g.y0 + (int)yl) << 16) | (unsigned)(g.x0 + (int)xl);
        if (key < HME_KEY_INIT) atomicMin(&jobs[j].key, key);
    }
}

__device__ __noinline__ void hme_group_items(const HBlk &blk_ref, HJob *jobs, const HGeo *geo, const HGroup &G_ref,
                                             uint8_t *arena) {
    const HBlk blk = blk_ref; // registers: the reference points into the caller's local memory
    const HGroup G = G_ref;
    const int nb = blk.nb;
    ItemGeo ig;
    uint32_t best32 = ~0u;
    int cj = G.j0, bj = -1, base = 0, end = 0;
    for (int it = threadIdx.x; it < G.item_total; it += HME_NT) {
        if (it >= end) {
            while (it >= geo[cj].item_base + geo[cj].n_items) ++cj;
            if (nb == 1 && bj >= 0) flush_key32(jobs, geo, bj, best32);
            best32 = ~0u;
            bj = cj;
            load_item_geo(geo[cj], ig);
            base = geo[cj].item_base;
            end = base + geo[cj].n_items;
        }
        hme_item_any<false>(blk, ig, arena, it - base, best32);
    }
    if (nb == 1 && bj >= 0) flush_key32(jobs, geo, bj, best32);
}

// Strip mode (strip_kernel): the jobs [j0, j1) are windows of DIFFERENT source blocks inside ONE staged
// strip of the reference plane (a column or a row of b64s searches almost the same pixels). Every job
// is a single 16x8 sub-block (nb == 1); geo[j] places its window in the strip (row_off, x_shift,
// arena_off) and sj[j] is its source block.
struct StripJob {
    const uint8_t *src; // shared memory
    int bw, nrows;
};
__device__ __noinline__ void strip_group_items(const StripJob *sj, int src_pitch, int row_step, HJob *jobs, const HGeo *geo,
                                               int j0, int item_total, uint8_t *arena) {
    ItemGeo ig;
    HBlk blk;
    blk.src = nullptr;
    blk.src_pitch = src_pitch;
    blk.row_step = row_step;
    blk.bw = blk.nrows = 0;
    blk.nbx = blk.nb = 1;
    blk.mg_nb = blk.mg_nbx = 0;
    uint32_t best32 = ~0u;
    int cj = j0, bj = -1, base = 0, end = 0;
    for (int it = threadIdx.x; it < item_total; it += HME_NT) {
        if (it >= end) {
            while (it >= geo[cj].item_base + geo[cj].n_items) ++cj;
            if (bj >= 0) flush_key32(jobs, geo, bj, best32);
            best32 = ~0u;
            bj = cj;
            load_item_geo(geo[cj], ig);
            base = geo[cj].item_base;
            end = base + geo[cj].n_items;
            blk.src = sj[cj].src;
            blk.bw = sj[cj].bw;
            blk.nrows = sj[cj].nrows;
        }
        hme_item_any<true>(blk, ig, arena, it - base, best32);
    }
    if (bj >= 0) flush_key32(jobs, geo, bj, best32);
}

// Streamed columns (tall strip windows: the 8 x H pre-HME search, prehme_core :1641). A COLUMN is the positions
// x = 8 xg + 4 j + c (j = 0, 1), y = p, p + 2, p + 4, ... of one job: same byte phase c, same row parity p. One thread
// walks down the staged strip once per column: every reference row it loads (6 words) is shifted to phase c (5 SHF)
// and meets the 8 source rows of the 8 positions in flight — position i = t - r takes source row r at step t — i.e.
// 64 VABSDIFF4 per step against 16 accumulators held in a ring (slot = step mod 8), one position pair finished per
// step. The source block stays in 32 registers for the whole column. Against the tiled items (hme_item_fast) this
// removes the per-item decode, the y-tile padding and most loads / shifts: ~85 % of the issue cycles are VABSDIFF4.
// The first / last 7 steps of a column carry positions that do not exist (their sums are dropped): 7 / (n + 7) of
// the column's work. Keys and tie-breaks are those of hme_emit / flush_key32.
template <bool FULL>
__device__ __forceinline__ void col_step(int K, const uint32_t (&s)[8][4], uint32_t (&acc)[8][2], const uint32_t (&w)[5], int nrows) {
    // K = step mod 8, a literal once the caller's step loop is unrolled: every register index below is static
    // word index outermost: consecutive VABSDIFF4 go to 16 different accumulators (no back-to-back dependent pairs)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (FULL || r < nrows) { // (nrows < 8 only for the blocks of a ragged bottom picture edge)
                const int slot = (K - r) & 7;
#pragma unroll
                for (int j = 0; j < 2; ++j) acc[slot][j] = vsad4(s[r][i], w[j + i], acc[slot][j]);
            }
        }
}

template <bool FULL>
__device__ __forceinline__ uint32_t col_walk(const uint32_t (&s)[8][4], int nrows, const uint8_t *base, int pitch, int R, uint32_t sh,
                                             uint32_t idx, int idx_step, int T) {
    uint32_t best32 = ~0u;
    uint32_t acc[8][2];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k][0] = acc[k][1] = 0;
    // software pipeline: the row of step t + 1 is loaded while step t computes (the planner stages two rows more than the
    // last step reads, so the last prefetch stays inside the strip); without it every step began with a wait for its LDS
    uint32_t nxt[6];
    {
        const uint8_t *rp = base + R * pitch + ((R >> 3) << 4);
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const uint2 v = *reinterpret_cast<const uint2 *>(rp + 8 * q);
            nxt[2 * q] = v.x; nxt[2 * q + 1] = v.y;
        }
    }
    for (int t0 = 0; t0 < T; t0 += 8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (t0 + k < T) {
                uint32_t w[5];
#pragma unroll
                for (int q = 0; q < 5; ++q) w[q] = __funnelshift_r(nxt[q], nxt[q + 1], sh);
                R += 2;
                const uint8_t *rp = base + R * pitch + ((R >> 3) << 4);
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const uint2 v = *reinterpret_cast<const uint2 *>(rp + 8 * q);
                    nxt[2 * q] = v.x; nxt[2 * q + 1] = v.y;
                }
                col_step<FULL>(k, s, acc, w, nrows);
                const int done = (k + 1) & 7; // the position started 7 steps ago is complete
                if (t0 + k >= 7) {
                    best32 = min(best32, (acc[done][0] << 17) + idx);
                    best32 = min(best32, (acc[done][1] << 17) + idx + 4);
                }
                acc[done][0] = acc[done][1] = 0;
                idx += idx_step;
            }
        }
    }
    return best32;
}

__device__ __noinline__ void strip_columns(const StripJob *sj, int src_pitch, HJob *jobs, const HGeo *geo, int j0, int col_total,
                                           uint8_t *arena) {
    int cj = j0;
    for (int col = threadIdx.x; col < col_total; col += HME_NT) {
        while (col >= geo[cj].col_base + geo[cj].n_cols) ++cj;
        const HGeo &g = geo[cj];
        const int local = col - g.col_base;
        const int sa_w = g.sa_w, n_xg = sa_w >> 3;                              // x groups of 8 positions
        const int c = local & 3, p = (local >> 2) & 1, u = local >> 3, ch = u / n_xg, xg = u - ch * n_xg;
        const int n_all = (g.bh - p + 1) >> 1;                                  // y = p, p + 2, ... < bh
        const int i0 = (ch * n_all) / g.col_chunks, n_p = ((ch + 1) * n_all) / g.col_chunks - i0; // this chunk's positions
        uint32_t best32 = ~0u;
        if (n_p > 0 && !(g.skip_even && p == 0)) {
            uint32_t s[8][4];
            const int nrows = sj[cj].nrows;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const uint4 v = *reinterpret_cast<const uint4 *>(sj[cj].src + r * src_pitch);
                s[r][0] = v.x; s[r][1] = v.y; s[r][2] = v.z; s[r][3] = v.w;
            }
            const uint8_t *base = arena + g.arena_off + 8 * xg;
            const int y0 = p + 2 * i0;
            const uint32_t idx = (uint32_t)((y0 - 14) * sa_w + 8 * xg + c); // raster index of the position finishing at step t, j = 0
            if (nrows == 8)
                best32 = col_walk<true>(s, 8, base, g.m << 4, g.row_off + y0, 8 * c, idx, 2 * sa_w, n_p + 7);
            else
                best32 = col_walk<false>(s, nrows, base, g.m << 4, g.row_off + y0, 8 * c, idx, 2 * sa_w, n_p + 7);
        }
        flush_key32(jobs, geo, cj, best32);
    }
}

__device__ __noinline__ void hme_group_scan(HJob *jobs, const HGeo *geo, const HGroup &G_ref, uint8_t *arena) {
    const HGroup G = G_ref;
    const uint32_t *sadreg = reinterpret_cast<const uint32_t *>(arena + geo[G.j0].sad_off);
    unsigned long long best = ~0ULL;
    int cj = G.j0, bj = -1;
    for (int i = threadIdx.x; i < G.scan_total; i += HME_NT) {
        while (i >= geo[cj].scan_base + geo[cj].scan_n) ++cj;
        if (cj != bj) {
            if (bj >= 0) flush_key(jobs, bj, best);
            best = ~0ULL;
            bj = cj;
        }
        const HGeo &g = geo[cj];
        const int il = i - g.scan_base;
        const int yl = (int)__umulhi((uint32_t)il, g.mg_wb);
        const int y = g.y0 + yl, x = g.x0 + il - yl * g.sa_w;
        if (g.skip_even && !(y & 1)) continue;
        const unsigned long long k = ((unsigned long long)sadreg[i] << 32) | ((unsigned)y << 16) | (unsigned)x;
        best = k < best ? k : best;
    }
    if (bj >= 0) flush_key(jobs, bj, best);
}

// One svt_sad_loop_kernel too large for the arena as a whole: its search area is tiled (row bands,
// then columns) and every tile runs through the group helpers as a one-job group.
// All threads of the CTA call it with identical arguments; jobs[J].key must be HME_KEY_INIT on entry.
__device__ __noinline__ void hme_search_tiled(const HBlk &blk, HJob *jobs, int J, HPlan *pl, uint8_t *arena,
                                              int arena_bytes = SVTME_HME_ARENA_BYTES) {
    const HJob &job = jobs[J];
    const int sa_w = job.sa_w, sa_h = job.sa_h;
    if (sa_w <= 0 || sa_h <= 0) return;
    int wt = sa_w, bh = sa_h;
    HGeo g;
    for (;;) {
        hme_geo(blk, wt, bh, job.skip_even, g, arena_bytes);
        if (!g.big) break;
        const int tile_w = 4 * g.G;
        if (bh > 8) bh = (bh + 1) / 2;
        else if (wt > tile_w) wt = ((wt / 2 + tile_w - 1) / tile_w) * tile_w;
        else if (bh > 1) bh = (bh + 1) / 2;
        else break; // one row x one G-tile always fits
    }
    for (int x0 = 0; x0 < sa_w; x0 += wt) {
        const int wb = min(wt, sa_w - x0);
        for (int y0 = 0; y0 < sa_h; y0 += bh) {
            const int hb = min(bh, sa_h - y0);
            __syncthreads(); // previous tile is done with the arena and with pl->geo[J]
            if (threadIdx.x == 0) {
                hme_geo(blk, wb, hb, job.skip_even, g, arena_bytes);
                g.x0 = x0;
                g.y0 = y0;
                g.ref_off = y0 * job.stride + x0;
                pl->geo[J] = g;
            }
            __syncthreads();
            const HGroup G = HGroup{J, J + 1, 0, pl->geo[J].stage_units, pl->geo[J].n_items, pl->geo[J].scan_n};
            hme_group_stage(jobs, pl->geo, G, arena);
            __syncthreads();
            hme_group_items(blk, jobs, pl->geo, G, arena);
            if (blk.nb > 1) {
                __syncthreads();
                hme_group_scan(jobs, pl->geo, G, arena);
            }
        }
    }
    __syncthreads();
}

// A round of independent jobs on the same source block. Jobs are packed into the arena as many
// at a time as fit ("group"); staging, search items and the SAD-buffer scan of a group are each
// ONE flattened list over the CTA, so twenty 16x6 refinement windows cost three barriers, not
// sixty, and every global load of the group is in flight at once.
// `pl` is scratch in shared memory. All threads call with identical arguments.
__device__ __noinline__ void hme_search_multi(const HBlk &blk, HJob *jobs, int n, HPlan *pl, uint8_t *arena) {
    if (n <= 0) return;
    const int tid = threadIdx.x, lane = tid & 31;
    const int nb = blk.nb;
    // ---- geometry: one thread per job
    bool is_point = true;
    if (tid < n) {
        const HJob &job = jobs[tid];
        is_point = job.sa_w == 1 && job.sa_h == 1;
        HGeo g;
        if (job.sa_w <= 0 || job.sa_h <= 0) {
            hme_geo(blk, 1, 1, 0, g);
            g.n_items = g.stage_units = g.scan_n = 0;
            g.need = 0;
            g.big = 0;
        } else {
            hme_geo(blk, job.sa_w, job.sa_h, job.skip_even, g);
        }
        pl->geo[tid] = g;
    }
    if (__syncthreads_and(is_point)) {
        if (tid == 0) pl->n_groups = 0;
        hme_points(blk, jobs, n);
        return;
    }
    // ---- grouping: warp 0, lane j <-> job j
    if (tid < 32) {
        const unsigned full = 0xffffffffu;
        const bool valid = lane < n;
        HGeo *g = &pl->geo[valid ? lane : 0];
        const uint32_t need = valid ? g->need : 0u, copies = valid ? (uint32_t)g->copy_bytes : 0u;
        const bool big = valid && g->big;
        const int su = valid ? g->stage_units : 0, ni = valid ? g->n_items : 0, sc = valid ? g->scan_n : 0;
        int start = 0, ng = 0;
        while (start < n) {
            const bool from = valid && lane >= start;
            uint32_t ps = from ? need : 0u;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(full, ps, o);
                if (lane >= o) ps += t;
            }
            const unsigned brk = __ballot_sync(full, from && (big || ps > SVTME_HME_ARENA_BYTES));
            int end = brk ? __ffs(brk) - 1 : n;
            const int first_big = end == start;
            if (first_big) end = start + 1;
            const bool in = valid && lane >= start && lane < end;
            uint32_t a0 = in ? copies : 0u;
            int a1 = in ? su : 0, a2 = in ? ni : 0, a3 = in ? sc : 0;
            uint32_t p0 = a0;
            int p1 = a1, p2 = a2, p3 = a3;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t0 = __shfl_up_sync(full, p0, o);
                const int t1 = __shfl_up_sync(full, p1, o), t2 = __shfl_up_sync(full, p2, o), t3 = __shfl_up_sync(full, p3, o);
                if (lane >= o) { p0 += t0; p1 += t1; p2 += t2; p3 += t3; }
            }
            const uint32_t tot0 = __shfl_sync(full, p0, 31);
            const int tot1 = __shfl_sync(full, p1, 31), tot2 = __shfl_sync(full, p2, 31), tot3 = __shfl_sync(full, p3, 31);
            if (in && !first_big) {
                g->arena_off = p0 - a0;
                g->stage_base = p1 - a1;
                g->item_base = p2 - a2;
                g->scan_base = p3 - a3;
                g->sad_off = tot0 + 4u * (uint32_t)(p3 - a3);
            }
            if (lane == 0) pl->grp[ng] = HGroup{start, end, first_big, tot1, tot2, tot3};
            ++ng;
            start = end;
        }
        if (lane == 0) pl->n_groups = ng;
    }
    __syncthreads();
    const int n_groups = pl->n_groups;
    for (int gi = 0; gi < n_groups; ++gi) {
        const HGroup G = pl->grp[gi];
        if (G.big) {
            hme_search_tiled(blk, jobs, G.j0, pl, arena);
            continue;
        }
        if (G.item_total == 0) continue;
        if (gi) __syncthreads(); // the previous group is done with the arena
        hme_group_stage(jobs, pl->geo, G, arena);
        __syncthreads();
        hme_group_items(blk, jobs, pl->geo, G, arena);
        if (nb > 1) { // multi-sub-block blocks: positions' SADs were summed in the SAD buffer
            __syncthreads();
            hme_group_scan(jobs, pl->geo, G, arena);
        }
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
#if SVTME_ARENA_BYTES >= 64 * 1024
#define FP_TW 64
#define FP_TH 32
#else
#define FP_TW 32
#define FP_TH 16
#endif
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

// MODE FP_COPIES: the window was staged as four byte-phase copies (stage_window4), aligned vector loads.
// MODE FP_RAW   : small windows — the rows were staged as RAW aligned 16-byte chunks (fp_stage_raw): `arena`
//                 is the tile in shared memory, `m` its row pitch in bytes and `copy_bytes` the byte offset
//                 of the window inside the first chunk; each row's words are read with scalar LDS and shifted
//                 to the thread's byte phase in registers. (Reading the rows straight from global memory,
//                 MODE FP_DIRECT, made every load touch ~8 cache lines: ncu showed l1tex at 85 % of peak.)
// MODE FP_DIRECT: no staging — `arena` is the GLOBAL address of the tile's top-left reference pixel and `m`
//                 the plane stride. Fallback only (a window that fits neither arena).
#define FP_COPIES 0
#define FP_DIRECT 1
#define FP_RAW 2
// Row placement of the raw tile. The 16 blocks of a half-warp read the same word column of rows 16*by
// apart (a multiple of 32 banks for every pitch) and of columns 16*bx bytes apart, two adjacent byte
// phases each: banks 4*bx + {0,1}. The skew moves the four block rows to bank offsets {0, 2, 16, 18}
// -> all 32 lanes on distinct banks. Rows stay 8-byte aligned (staging stores 8 bytes at a time) and
// never overlap (skews are cumulative per 32 rows, 8 < 64).
__device__ __forceinline__ int fp_raw_rowoff(int R, int pitch) { return R * pitch + ((R >> 4) & 1) * 8 + (R >> 5) * 64; }
template <int G, int MODE>
__device__ __noinline__ void fp_tile(const uint8_t *src64, const uint8_t *arena, int copy_bytes, int m, int tw, int th,
                                        bool sub_sad, FpTables *tab) {
    const int tid = threadIdx.x, b = tid & 15, c = (tid >> 4) & 3, yl = tid >> 6;
    const int bx = ((b >> 2) & 1) * 2 + (b & 1), by = ((b >> 3) & 1) * 2 + ((b >> 1) & 1);
    const int xt_n = (tw + 4 * G - 1) / (4 * G), npass = (th + SVTME_FP_YL - 1) / SVTME_FP_YL;
    uint32_t k8[4] = {~0u, ~0u, ~0u, ~0u}, k16 = ~0u, k32 = ~0u, k64 = ~0u;
    const int npar = sub_sad ? 1 : 2;
    for (int pass = 0; pass < npass; ++pass) {
        const int y = SVTME_FP_YL * pass + yl;
        for (int xt = 0; xt < xt_n; ++xt) {
            // Positions outside the tile (x >= tw, y >= th) still read staged memory; their 8x8 accumulators start
            // at 2^18, above every real 8x8 / 16x16 SAD, so their keys lose every minimum without a select per key.
            const uint32_t yvalid = y < th;
            uint32_t acc[G][4];
#pragma unroll
            for (int j = 0; j < G; ++j) {
                const uint32_t init = (yvalid && xt * 4 * G + 4 * j + c < tw) ? 0u : (1u << 18);
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[j][k] = init;
            }
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
                    uint32_t w[G + 4];
                    if (MODE == FP_DIRECT) {
                        const uintptr_t ai = reinterpret_cast<uintptr_t>(arena + (size_t)R * m + 4 * q0 + c);
                        const uint32_t *al = reinterpret_cast<const uint32_t *>(ai & ~(uintptr_t)3);
                        const uint32_t sh = (uint32_t)(ai & 3) * 8;
                        uint32_t t[G + 5];
#pragma unroll
                        for (int k = 0; k < G + 4; ++k) t[k] = __ldg(al + k);
                        t[G + 4] = sh ? __ldg(al + G + 4) : 0u;
#pragma unroll
                        for (int k = 0; k < G + 4; ++k) w[k] = __funnelshift_r(t[k], t[k + 1], sh);
                    } else if (MODE == FP_RAW) {
                        const int off = copy_bytes + 4 * q0 + c; // byte offset inside the staged row
                        const uint32_t *al = reinterpret_cast<const uint32_t *>(arena + fp_raw_rowoff(R, m) + (off & ~3));
                        const uint32_t sh = (uint32_t)(off & 3) * 8;
                        uint32_t t[G + 5];
#pragma unroll
                        for (int k = 0; k < G + 5; ++k) t[k] = al[k];
#pragma unroll
                        for (int k = 0; k < G + 4; ++k) w[k] = __funnelshift_r(t[k], t[k + 1], sh);
                    } else {
                    const uint8_t *rp = base + ((size_t)(R * m + 4 * (R >> 4)) << 4);
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
            // keys carry the SAD as accumulated (under SUB_SAD_SEARCH the doubling is applied when the tile's minima
            // are merged: x2 keeps the order): one multiply-add and one min per PU and position
            uint32_t v16[G];
#pragma unroll
            for (int j = 0; j < G; ++j) {
                const uint32_t idx = (uint32_t)(y * FP_TW + xt * 4 * G + 4 * j + c);
#pragma unroll
                for (int k = 0; k < 4; ++k) k8[k] = min(k8[k], acc[j][k] * (1u << FP_IDX_BITS) + idx);
                const uint32_t sum = (acc[j][0] + acc[j][1]) + (acc[j][2] + acc[j][3]);
                v16[j] = sum;
                k16 = min(k16, sum * (1u << FP_IDX_BITS) + idx);
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

// Raw staging for small windows: rows [0, rows) x chunks [0, nch) of the 16-byte aligned region that
// contains the tile go to shared memory as they are (LDG.128 -> STS.128, every line read once).
__device__ __forceinline__ void fp_stage_raw(uint8_t *arena, const uint8_t *__restrict__ g16, int gstride, int rows, int nch,
                                             int pitch) {
    const int total = rows * nch;
    const uint32_t mg = 0xFFFFFFFFu / (uint32_t)nch + 1u;
    for (int i = threadIdx.x; i < total; i += FP_NT) {
        const int R = (int)__umulhi((uint32_t)i, mg), ch = i - R * nch;
        // rows are 8-byte aligned in the tile (fp_raw_rowoff): two 8-byte asynchronous copies per chunk
        const uint8_t *gsrc = g16 + (size_t)R * gstride + 16 * ch;
        uint8_t *d = arena + fp_raw_rowoff(R, pitch) + 16 * ch;
        cp_async8(d, gsrc);
        cp_async8(d + 8, gsrc + 8);
    }
    cp_async_wait_all();
}

// All threads call with identical arguments. ref0: global pointer to the reference pixel under
// block(0,0) at search position (0,0). (ox, oy): MV of search position (0,0). order_base: rank of
// position (0,0) in the reference's evaluation order (see FpTables). src64: shared memory, 64x64
// source block with row stride 64. arena: shared memory of arena_bytes bytes.
__device__ __noinline__ void fullpel_search(const uint8_t *src64, const uint8_t *ref0, int ref_stride, int ox, int oy, int sa_w,
                               int sa_h, bool sub_sad, uint32_t order_base, FpTables *tab, uint8_t *arena, int arena_bytes) {
    if (sa_w <= 0 || sa_h <= 0) return;
    const int G = sa_w <= 8 ? 2 : (sa_w <= 16 ? 4 : 8);
    // small areas (the 8x3 .. 24x7 windows of the fast presets): staging four copies of a 71x66 window
    // for a few dozen positions costs more than the search; stage the rows once, raw
    const bool small = sa_w * sa_h <= SVTME_FP_DIRECT_MAX;
    const int last_row = sub_sad ? 62 : 63; // last block row that is read
    for (int ty0 = 0; ty0 < sa_h; ty0 += FP_TH) {
        const int th = min(FP_TH, sa_h - ty0);
        for (int tx0 = 0; tx0 < sa_w; tx0 += FP_TW) {
            const int tw = min(FP_TW, sa_w - tx0);
            const int span = ((tw + 4 * G - 1) / (4 * G)) * 4 * G;
            const int vec_words = G == 2 ? 6 : ((G + 3 + 3) / 4) * 4;
            const int words = (span / 4 - G) + 12 + vec_words;
            const int m = (words + 3) / 4;
            const int rows = ((th + SVTME_FP_YL - 1) / SVTME_FP_YL) * SVTME_FP_YL + last_row;
            const int copy_chunks = rows * m + 4 * ((rows - 1) >> 4) + 4;
            const int copy_bytes = copy_chunks * 16;
            const uint8_t *tile0 = ref0 + (size_t)ty0 * ref_stride + tx0;
            // raw tile: bytes [mis, mis + span + 64 + 4) of every row, from the 16-byte aligned address below tile0
            const int mis = (int)(reinterpret_cast<uintptr_t>(tile0) & 15);
            const int nch = (mis + span + 64 + 4 + 15) >> 4, pitch = nch * 16;
            const int mode = (small && fp_raw_rowoff(rows, pitch) + 64 <= arena_bytes) ? FP_RAW
                : (4 * copy_bytes <= arena_bytes ? FP_COPIES : FP_DIRECT);
            __syncthreads();
            if (mode == FP_COPIES) stage_window4(arena, tile0, ref_stride, rows, m, 4, copy_bytes, words);
            if (mode == FP_RAW) fp_stage_raw(arena, tile0 - mis, ref_stride, rows, nch, pitch);
            for (int i = threadIdx.x; i < SVTME_PU_COUNT; i += FP_NT) tab->tile_best[i] = ~0u;
            __syncthreads();
            if (mode == FP_DIRECT) {
                if (G == 2)
                    fp_tile<2, FP_DIRECT>(src64, tile0, 0, ref_stride, tw, th, sub_sad, tab);
                else if (G == 4)
                    fp_tile<4, FP_DIRECT>(src64, tile0, 0, ref_stride, tw, th, sub_sad, tab);
                else
                    fp_tile<8, FP_DIRECT>(src64, tile0, 0, ref_stride, tw, th, sub_sad, tab);
            } else if (mode == FP_RAW) {
                if (G == 2)
                    fp_tile<2, FP_RAW>(src64, arena, mis, pitch, tw, th, sub_sad, tab);
                else if (G == 4)
                    fp_tile<4, FP_RAW>(src64, arena, mis, pitch, tw, th, sub_sad, tab);
                else
                    fp_tile<8, FP_RAW>(src64, arena, mis, pitch, tw, th, sub_sad, tab);
            } else if (G == 2)
                fp_tile<2, FP_COPIES>(src64, arena, copy_bytes, m, tw, th, sub_sad, tab);
            else if (G == 4)
                fp_tile<4, FP_COPIES>(src64, arena, copy_bytes, m, tw, th, sub_sad, tab);
            else
                fp_tile<8, FP_COPIES>(src64, arena, copy_bytes, m, tw, th, sub_sad, tab);
            __syncthreads();
            for (int pu = threadIdx.x; pu < SVTME_PU_COUNT; pu += FP_NT) {
                const uint32_t k = tab->tile_best[pu];
                if (k != ~0u) {
                    const uint32_t sad = (k >> FP_IDX_BITS) << (sub_sad ? 1 : 0), idx = k & ((1u << FP_IDX_BITS) - 1);
                    const int gy = ty0 + (int)(idx / FP_TW), gx = tx0 + (int)(idx % FP_TW);
                    const unsigned long long key =
                        ((unsigned long long)sad << 32) | (order_base + (uint32_t)gy * (uint32_t)sa_w + (uint32_t)gx);
                    if (key < tab->key[pu]) {
                        tab->key[pu] = key;
                        tab->mv[pu] = ((uint32_t)(uint16_t)(oy + gy) << 16) | (uint16_t)(ox + gx);
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
__device__ __noinline__ void fullpel_probe(const uint8_t *src64, const uint8_t *ref0, int ref_stride, int mvx, int mvy, bool sub_sad,
                              uint32_t order, FpTables *tab) {
    const int tid = threadIdx.x;
    for (int vt = tid; vt < 256; vt += FP_NT) { // 256 virtual threads: 64 8x8 blocks x 4 row pairs
        const int e = vt >> 2, q = vt & 3;
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
        if ((vt & 15) == 0) tab->tile_best[5 + b] = s16;
    }
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
    for (int pu = tid; pu < SVTME_PU_COUNT; pu += FP_NT) {
        const unsigned long long key = ((unsigned long long)tab->tile_best[pu] << 32) | order;
        if (key < tab->key[pu]) {
            tab->key[pu] = key;
            tab->mv[pu] = ((uint32_t)(uint16_t)mvy << 16) | (uint16_t)mvx;
        }
    }
    __syncthreads();
}

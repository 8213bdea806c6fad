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
    const int wpr = m * 4; // words per row
    const int total = rows * wpr * 4;
    for (int i = threadIdx.x; i < total; i += SVTME_THREADS) {
        const int kw = i % wpr;
        const int t = i / wpr;
        const int R = t % rows;
        const int c = t / rows;
        const uint8_t *a = g + (size_t)R * gstride + 4 * kw + c;
        const uintptr_t ai = reinterpret_cast<uintptr_t>(a);
        const uint32_t *al = reinterpret_cast<const uint32_t *>(ai & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(ai & 3) * 8;
        const uint32_t lo = __ldg(al);
        const uint32_t hi = sh ? __ldg(al + 1) : 0u;
        const uint32_t w = __funnelshift_r(lo, hi, sh);
        *reinterpret_cast<uint32_t *>(arena + (size_t)c * copy_bytes + ((size_t)(R * m + skew * (R >> 4)) << 4) + 4 * kw) = w;
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

template <int G>
__device__ __forceinline__ void hme_items(const HBlk &blk, int sa_w, int skip_even, const uint8_t *arena, int copy_bytes,
                                          int m, int x0, int y0, int bh, uint32_t *sadbuf, unsigned long long &best) {
    const int nbx = (blk.bw + 15) >> 4, nby = (blk.nrows + 7) >> 3, nb = nbx * nby;
    const int tile_w = 4 * G;
    const int xt_n = (sa_w + tile_w - 1) / tile_w;
    const int n_items = bh * xt_n * 4 * nb;
    const bool masked = (blk.bw & 15) != 0;
    int cached_b = -1;
    uint32_t s[8][4];
    uint32_t msk[4] = {~0u, ~0u, ~0u, ~0u};
    int rows_here = 0;
    for (int it = threadIdx.x; it < n_items; it += SVTME_THREADS) {
        const int b = it % nb;
        int t = it / nb;
        const int c = t & 3;
        t >>= 2;
        const int xt = t % xt_n, y = t / xt_n;
        const int bxi = b % nbx, byi = b / nbx;
        if (b != cached_b) { // (re)load the source sub-block into registers
            cached_b = b;
            rows_here = min(8, blk.nrows - 8 * byi);
            const int vb = min(16, blk.bw - 16 * bxi);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int n = min(max(vb - 4 * i, 0), 4);
                msk[i] = n >= 4 ? ~0u : ((1u << (8 * n)) - 1u);
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                if (r < rows_here) {
                    const uint8_t *sp = blk.src + (size_t)(8 * byi + r) * blk.src_pitch + 16 * bxi;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint32_t w = 0;
                        if (msk[i]) w = *reinterpret_cast<const uint32_t *>(sp + 4 * i) & msk[i];
                        s[r][i] = w;
                    }
                }
            }
        }
        uint32_t acc[G];
#pragma unroll
        for (int j = 0; j < G; ++j) acc[j] = 0;
        const int q0 = xt * G + 4 * bxi; // first word of copy c this item needs
        const uint8_t *base = arena + (size_t)c * copy_bytes + 4 * q0;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r < rows_here) {
                const int R = y + (8 * byi + r) * blk.row_step;
                const uint8_t *rp = base + ((size_t)R * m << 4);
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
                        for (int i = 0; i < 4; ++i) acc[j] = vsad4(s[r][i], w[j + i] & msk[i], acc[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < G; ++j)
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[j] = vsad4(s[r][i], w[j + i], acc[j]);
                }
            }
        }
        const int yg = y0 + y;
        if (skip_even && !(yg & 1)) continue;
#pragma unroll
        for (int j = 0; j < G; ++j) {
            const int x = xt * tile_w + 4 * j + c;
            if (x < sa_w) {
                if (nb == 1) {
                    unsigned long long k = ((unsigned long long)acc[j] << 32) | ((unsigned)yg << 16) | (unsigned)(x0 + x);
                    best = k < best ? k : best;
                } else {
                    atomicAdd(&sadbuf[y * sa_w + x], acc[j]);
                }
            }
        }
    }
}

// One svt_sad_loop_kernel: all threads of the CTA must call it with identical arguments.
// `job` lives in shared memory; job->key must be HME_KEY_INIT on entry (set by the planner).
__device__ void hme_search(const HBlk &blk, HJob *job, uint8_t *arena) {
    const int sa_w = job->sa_w, sa_h = job->sa_h;
    if (sa_w <= 0 || sa_h <= 0) return;
    const int nbx = (blk.bw + 15) >> 4;
    const int nb = nbx * ((blk.nrows + 7) >> 3);
    const int G = sa_w <= 8 ? 2 : (sa_w <= 16 ? 4 : 8);
    const int tile_w = 4 * G;
    const int vec_words = G == 2 ? 6 : ((G + 3 + 3) / 4) * 4;
    const int extra_rows = (blk.nrows - 1) * blk.row_step;
    // Tile the search area so that 4 phase copies of (bh + extra_rows) rows x (wt + bw - 1) bytes
    // (+ the per-position SAD buffer when the block spans several sub-blocks) fit the arena.
    int wt = ((sa_w + tile_w - 1) / tile_w) * tile_w, bh = sa_h, m = 0;
    for (;;) {
        const int words = (wt / 4 - G) + 4 * (nbx - 1) + vec_words;
        m = ((words + 3) / 4) | 1; // odd chunk stride: consecutive rows start in different bank groups
        int copy_chunks = (bh + extra_rows) * m;
        copy_chunks += (10 - (copy_chunks & 7)) & 7; // copy stride == 2 (mod 8) chunks
        const size_t need = (size_t)copy_chunks * 16 * 4 + (nb > 1 ? (size_t)bh * wt * 4 : 0);
        if (need <= SVTME_ARENA_BYTES) break;
        if (bh > 8) bh = (bh + 1) / 2;
        else if (wt > tile_w) { wt = ((wt / 2 + tile_w - 1) / tile_w) * tile_w; bh = sa_h; }
        else if (bh > 1) bh = (bh + 1) / 2;
        else break; // 1 row x one G-tile always fits (<= 64 rows x 9 chunks x 4 copies)
    }
    unsigned long long best = ~0ULL;
    for (int x0 = 0; x0 < sa_w; x0 += wt) {
        const int wb = min(wt, sa_w - x0);
        for (int y0 = 0; y0 < sa_h; y0 += bh) {
            const int hb = min(bh, sa_h - y0);
            const int rows = hb + extra_rows;
            int copy_chunks = rows * m;
            copy_chunks += (10 - (copy_chunks & 7)) & 7;
            const int copy_bytes = copy_chunks * 16;
            uint32_t *sadbuf = reinterpret_cast<uint32_t *>(arena + (size_t)copy_bytes * 4);
            __syncthreads(); // previous users of the arena are done
            stage_window4(arena, job->ref + (size_t)y0 * job->stride + x0, job->stride, rows, m, 0, copy_bytes);
            if (nb > 1)
                for (int i = threadIdx.x; i < hb * wb; i += SVTME_THREADS) sadbuf[i] = 0;
            __syncthreads();
            if (G == 2)
                hme_items<2>(blk, wb, job->skip_even, arena, copy_bytes, m, x0, y0, hb, sadbuf, best);
            else if (G == 4)
                hme_items<4>(blk, wb, job->skip_even, arena, copy_bytes, m, x0, y0, hb, sadbuf, best);
            else
                hme_items<8>(blk, wb, job->skip_even, arena, copy_bytes, m, x0, y0, hb, sadbuf, best);
            if (nb > 1) {
                __syncthreads();
                for (int i = threadIdx.x; i < hb * wb; i += SVTME_THREADS) {
                    const int y = y0 + i / wb, x = x0 + i % wb;
                    if (job->skip_even && !(y & 1)) continue;
                    unsigned long long k = ((unsigned long long)sadbuf[i] << 32) | ((unsigned)y << 16) | (unsigned)x;
                    best = k < best ? k : best;
                }
            }
        }
    }
    // CTA-wide minimum
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long v = __shfl_xor_sync(0xffffffffu, best, o);
        best = v < best ? v : best;
    }
    if ((threadIdx.x & 31) == 0 && best < HME_KEY_INIT) atomicMin(&job->key, best);
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

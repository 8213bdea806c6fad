// svtme_pyramid.cu — pre-ME pyramid on the device (sm_100a).
//
// Replaces, bit-exactly:
//   svt_aom_generate_padding                    Source/Lib/Codec/pic_operators.c:338-382
//   svt_aom_downsample_2d_c                     Source/Lib/Codec/pic_analysis_process.c:131-161
//   svt_aom_downsample_filtering_input_picture  Source/Lib/Codec/pic_analysis_process.c:1922-1979
//
// HBM-bound byte work: algorithmic traffic per picture = W*H read + 5/16 W*H written (+ pads).
// One thread owns a 16x4 full-resolution patch (one 16-byte load per row, fully coalesced across
// the warp), produces the 8x2 quarter pixels and the 4x1 sixteenth pixels from registers, so the
// quarter plane is never re-read. Padding is "pixel = interior(clamp(x), clamp(y))", which is what
// the reference's horizontal-then-vertical replicate produces.
#include "svtme_common.cuh"

unsigned long long g_svtme_launches = 0;

// ------------------------------------------------------------------ padding
// One thread per pad byte. Pad region = top band, bottom band, left/right bands of interior rows.
__global__ void __launch_bounds__(256) pad_plane_kernel(uint8_t *org, int stride, int W, int H, int pl, int pr,
                                                         int pt, int pb) {
    const long long full_w = (long long)pl + W + pr;
    const long long n_top = full_w * pt, n_bot = full_w * pb, n_side = (long long)(pl + pr) * H;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int x, y;
    if (i < n_top) {
        y = -pt + (int)(i / full_w);
        x = -pl + (int)(i % full_w);
    } else if ((i -= n_top) < n_bot) {
        y = H + (int)(i / full_w);
        x = -pl + (int)(i % full_w);
    } else if ((i -= n_bot) < n_side) {
        y = (int)(i / (pl + pr));
        int k = (int)(i % (pl + pr));
        x = k < pl ? -pl + k : W + (k - pl);
    } else
        return;
    const int sx = min(max(x, 0), W - 1), sy = min(max(y, 0), H - 1);
    org[(long long)y * stride + x] = org[(long long)sy * stride + sx];
}

void svtme_launch_pad_plane(const DevPlane &p, int pl, int pr, int pt, int pb, cudaStream_t st) {
    const long long full_w = (long long)pl + p.width + pr;
    const long long n = full_w * (pt + pb) + (long long)(pl + pr) * p.height;
    if (n <= 0) return;
    uint8_t *org = p.base + (size_t)p.org_y * p.stride + p.org_x;
    pad_plane_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(org, p.stride, p.width, p.height, pl, pr, pt, pb);
    ++g_svtme_launches;
}

// ------------------------------------------------------------------ fused 1/4 + 1/16 decimation
__device__ __forceinline__ uint32_t avg4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    return (a + b + c + d + 2) >> 2;
}

// Horizontal pair sums of the 4 bytes of a word: returns (b0+b1) | (b2+b3) << 16.
__device__ __forceinline__ uint32_t hpair(uint32_t w) { return (w & 0x00ff00ffu) + ((w >> 8) & 0x00ff00ffu); }

// 2x2 box of two rows of 4 pixels each -> 2 output pixels (low 16 bits: out0, high: out1).
__device__ __forceinline__ uint32_t box2(uint32_t top, uint32_t bot) {
    uint32_t s = hpair(top) + hpair(bot) + 0x00020002u;
    return (s >> 2) & 0x00ff00ffu;
}

// full(16x4 patch) -> quarter(8x2) -> sixteenth(4x1). W, H multiples of 8.
// level1_on == 0: sixteenth comes straight from the full plane with decimation step 4 (the 2x2 at
// rows {1,2}, cols {1,2} of each 4x4), pic_analysis_process.c:1958-1969.
__global__ void __launch_bounds__(256)
    downsample_fused_kernel(const uint8_t *__restrict__ f0, int fs, int W, int H, uint8_t *__restrict__ q0, int qs,
                            uint8_t *__restrict__ s0, int ss, int level1_on) {
    const int px = blockIdx.x * blockDim.x + threadIdx.x; // 16-pixel column group
    const int py = blockIdx.y * blockDim.y + threadIdx.y; // 4-row group
    const int x0 = px * 16, y0 = py * 4;
    if (x0 >= W || y0 >= H) return;
    const bool wide = (x0 + 16 <= W); // else 8 valid pixels (W % 8 == 0)
    uint32_t r[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint8_t *row = f0 + (size_t)(y0 + k) * fs + x0;
        if (wide) {
            uint4 v = *reinterpret_cast<const uint4 *>(row);
            r[k][0] = v.x; r[k][1] = v.y; r[k][2] = v.z; r[k][3] = v.w;
        } else {
            uint2 v = *reinterpret_cast<const uint2 *>(row);
            r[k][0] = v.x; r[k][1] = v.y; r[k][2] = 0; r[k][3] = 0;
        }
    }
    // quarter rows 0,1: 8 pixels each, packed 2 per box2 result
    uint32_t q[2][2]; // q[row][word] bytes = 4 quarter pixels
#pragma unroll
    for (int qr = 0; qr < 2; ++qr) {
        uint32_t b0 = box2(r[2 * qr][0], r[2 * qr + 1][0]); // px 0,1
        uint32_t b1 = box2(r[2 * qr][1], r[2 * qr + 1][1]); // px 2,3
        uint32_t b2 = box2(r[2 * qr][2], r[2 * qr + 1][2]);
        uint32_t b3 = box2(r[2 * qr][3], r[2 * qr + 1][3]);
        q[qr][0] = (b0 & 0xff) | ((b0 >> 16) << 8) | ((b1 & 0xff) << 16) | ((b1 >> 16) << 24);
        q[qr][1] = (b2 & 0xff) | ((b2 >> 16) << 8) | ((b3 & 0xff) << 16) | ((b3 >> 16) << 24);
    }
    if (level1_on) {
        uint8_t *qrow = q0 + (size_t)(y0 >> 1) * qs + (x0 >> 1);
        if (wide) {
            *reinterpret_cast<uint2 *>(qrow) = make_uint2(q[0][0], q[0][1]);
            *reinterpret_cast<uint2 *>(qrow + qs) = make_uint2(q[1][0], q[1][1]);
        } else {
            *reinterpret_cast<uint32_t *>(qrow) = q[0][0];
            *reinterpret_cast<uint32_t *>(qrow + qs) = q[1][0];
        }
    }
    uint32_t s4;
    if (level1_on) {
        uint32_t a = box2(q[0][0], q[1][0]), b = box2(q[0][1], q[1][1]);
        s4 = (a & 0xff) | ((a >> 16) << 8) | ((b & 0xff) << 16) | ((b >> 16) << 24);
    } else {
        s4 = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) { // 4x4 group k: rows 1,2 cols 1,2 -> bytes 1,2 of word k
            uint32_t t = r[1][k], u = r[2][k];
            uint32_t v = avg4((t >> 8) & 0xff, (t >> 16) & 0xff, (u >> 8) & 0xff, (u >> 16) & 0xff);
            s4 |= v << (8 * k);
        }
    }
    uint8_t *srow = s0 + (size_t)(y0 >> 2) * ss + (x0 >> 2);
    if (wide)
        *reinterpret_cast<uint32_t *>(srow) = s4;
    else
        *reinterpret_cast<uint16_t *>(srow) = (uint16_t)s4;
}

void svtme_launch_downsample(const DevPlane &f, const DevPlane &q, const DevPlane &s, int hme_level1,
                             cudaStream_t st) {
    dim3 blk(32, 8);
    dim3 grd((f.width + 16 * 32 - 1) / (16 * 32), (f.height + 4 * 8 - 1) / (4 * 8));
    downsample_fused_kernel<<<grd, blk, 0, st>>>(f.at(0, 0), f.stride, f.width, f.height,
                                                 const_cast<uint8_t *>(q.at(0, 0)), q.stride,
                                                 const_cast<uint8_t *>(s.at(0, 0)), s.stride, hme_level1);
    ++g_svtme_launches;
}

// ------------------------------------------------------------------ batched build (n pictures, 2 launches)
// One launch decimates n pictures (blockIdx.z = picture): the 16x4 patch is read from the caller's
// luma buffer (or from the level-0 plane when the copy was done by the host->device transfer) and
// written to the level-0 interior, the quarter and the sixteenth planes. A second launch pads all
// three levels of all pictures.
__global__ void __launch_bounds__(256) downsample_batch_kernel(const SvtmePutItem *__restrict__ items) {
    const SvtmePutItem &it = items[blockIdx.z];
    const DevPlane &f = it.pyr.lvl[0], &q = it.pyr.lvl[1], &x = it.pyr.lvl[2];
    const int W = f.width, H = f.height;
    const int px = blockIdx.x * blockDim.x + threadIdx.x, py = blockIdx.y * blockDim.y + threadIdx.y;
    const int x0 = px * 16, y0 = py * 4;
    if (x0 >= W || y0 >= H) return;
    const bool wide = (x0 + 16 <= W);
    const uint8_t *src = it.src ? it.src : f.at(0, 0);
    const int sstride = it.src ? it.src_stride : f.stride;
    uint32_t r[4][4];
    const bool ten = it.src && it.src_bits == 10;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (ten) { // 16-bit samples: the 8 MSBs of the 10 go to the ME planes, (uint8_t)(sample >> 2)
            const uint4 *row16 = reinterpret_cast<const uint4 *>(src + (size_t)(y0 + k) * sstride + 2 * x0);
            const uint4 a = row16[0];
            r[k][0] = __byte_perm(a.x >> 2, a.y >> 2, 0x6420);
            r[k][1] = __byte_perm(a.z >> 2, a.w >> 2, 0x6420);
            r[k][2] = r[k][3] = 0;
            if (wide) {
                const uint4 b = row16[1];
                r[k][2] = __byte_perm(b.x >> 2, b.y >> 2, 0x6420);
                r[k][3] = __byte_perm(b.z >> 2, b.w >> 2, 0x6420);
            }
        } else {
        const uint8_t *row = src + (size_t)(y0 + k) * sstride + x0;
        if (wide) {
            const uint4 v = *reinterpret_cast<const uint4 *>(row);
            r[k][0] = v.x; r[k][1] = v.y; r[k][2] = v.z; r[k][3] = v.w;
        } else {
            const uint2 v = *reinterpret_cast<const uint2 *>(row);
            r[k][0] = v.x; r[k][1] = v.y; r[k][2] = 0; r[k][3] = 0;
        }
        }
        if (it.src) {
            uint8_t *drow = const_cast<uint8_t *>(f.at(x0, y0 + k));
            if (wide)
                *reinterpret_cast<uint4 *>(drow) = make_uint4(r[k][0], r[k][1], r[k][2], r[k][3]);
            else
                *reinterpret_cast<uint2 *>(drow) = make_uint2(r[k][0], r[k][1]);
        }
    }
    uint32_t qq[2][2];
#pragma unroll
    for (int qr = 0; qr < 2; ++qr) {
        const uint32_t b0 = box2(r[2 * qr][0], r[2 * qr + 1][0]), b1 = box2(r[2 * qr][1], r[2 * qr + 1][1]);
        const uint32_t b2 = box2(r[2 * qr][2], r[2 * qr + 1][2]), b3 = box2(r[2 * qr][3], r[2 * qr + 1][3]);
        qq[qr][0] = (b0 & 0xff) | ((b0 >> 16) << 8) | ((b1 & 0xff) << 16) | ((b1 >> 16) << 24);
        qq[qr][1] = (b2 & 0xff) | ((b2 >> 16) << 8) | ((b3 & 0xff) << 16) | ((b3 >> 16) << 24);
    }
    uint8_t *qrow = const_cast<uint8_t *>(q.at(x0 >> 1, y0 >> 1));
    if (wide) {
        *reinterpret_cast<uint2 *>(qrow) = make_uint2(qq[0][0], qq[0][1]);
        *reinterpret_cast<uint2 *>(qrow + q.stride) = make_uint2(qq[1][0], qq[1][1]);
    } else {
        *reinterpret_cast<uint32_t *>(qrow) = qq[0][0];
        *reinterpret_cast<uint32_t *>(qrow + q.stride) = qq[1][0];
    }
    const uint32_t a = box2(qq[0][0], qq[1][0]), b = box2(qq[0][1], qq[1][1]);
    const uint32_t s4 = (a & 0xff) | ((a >> 16) << 8) | ((b & 0xff) << 16) | ((b >> 16) << 24);
    uint8_t *srow = const_cast<uint8_t *>(x.at(x0 >> 2, y0 >> 2));
    if (wide)
        *reinterpret_cast<uint32_t *>(srow) = s4;
    else
        *reinterpret_cast<uint16_t *>(srow) = (uint16_t)s4;
}

// Padding of all three levels of n pictures in one launch: one thread per aligned 4-byte word that
// holds at least one pad byte (words are counted from the left edge of the padded row; a word that
// straddles the right picture edge rewrites its interior bytes with their own values). Every byte is
// the clamped interior pixel, read from interior bytes only, so there is no ordering between threads.
__device__ __forceinline__ int pad_words_per_level(int W, int H, int pl, int pr, int pt, int pb, int &fw4, int &wi0) {
    fw4 = (pl + W + pr + 3) >> 2;
    wi0 = (pl + W) >> 2;
    return fw4 * (pt + pb) + ((pl >> 2) + (fw4 - wi0)) * H;
}

__global__ void __launch_bounds__(256) pad_batch_kernel(const SvtmePutItem *__restrict__ items, SvtmePadSpec ps) {
    const SvtmePutItem &it = items[blockIdx.z];
    int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
#pragma unroll 1
    for (int lvl = 0; lvl < 3; ++lvl) {
        const DevPlane &p = it.pyr.lvl[lvl];
        const int W = p.width, H = p.height, pl = ps.l[lvl], pr = ps.r[lvl], pt = ps.t[lvl], pb = ps.b[lvl];
        int fw4, wi0;
        const int n_lvl = pad_words_per_level(W, H, pl, pr, pt, pb, fw4, wi0);
        if (i >= n_lvl) {
            i -= n_lvl;
            continue;
        }
        const int n_top = fw4 * pt, n_bot = fw4 * pb, side4 = (pl >> 2) + (fw4 - wi0);
        int wi, y;
        if (i < n_top) {
            const int r = i / fw4;
            y = -pt + r;
            wi = i - r * fw4;
        } else if (i - n_top < n_bot) {
            const int k = i - n_top, r = k / fw4;
            y = H + r;
            wi = k - r * fw4;
        } else {
            const int k = i - n_top - n_bot, r = k / side4, c = k - r * side4;
            y = r;
            wi = c < (pl >> 2) ? c : wi0 + (c - (pl >> 2));
        }
        const int x = -pl + 4 * wi;
        uint8_t *org = const_cast<uint8_t *>(p.at(0, 0));
        const uint8_t *srow = org + (long long)min(max(y, 0), H - 1) * p.stride;
        uint32_t v;
        if (x + 3 < 0)
            v = srow[0] * 0x01010101u;
        else if (x >= W)
            v = srow[W - 1] * 0x01010101u;
        else if (x >= 0 && x + 3 < W)
            v = *reinterpret_cast<const uint32_t *>(srow + x);
        else {
            v = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) v |= (uint32_t)srow[min(max(x + b, 0), W - 1)] << (8 * b);
        }
        *reinterpret_cast<uint32_t *>(org + (long long)y * p.stride + x) = v;
        return;
    }
}

void svtme_launch_build_batch(const SvtmePutItem *d_items, int n, const DevPyramid &geom, const SvtmePadSpec &ps,
                              cudaStream_t st) {
    if (n <= 0) return;
    const DevPlane &f = geom.lvl[0];
    dim3 blk(32, 8);
    dim3 grd((f.width + 16 * 32 - 1) / (16 * 32), (f.height + 4 * 8 - 1) / (4 * 8), n);
    downsample_batch_kernel<<<grd, blk, 0, st>>>(d_items);
    long long total = 0;
    for (int l = 0; l < 3; ++l) {
        const DevPlane &p = geom.lvl[l];
        const int fw4 = (ps.l[l] + p.width + ps.r[l] + 3) >> 2, wi0 = (ps.l[l] + p.width) >> 2;
        total += (long long)fw4 * (ps.t[l] + ps.b[l]) + (long long)((ps.l[l] >> 2) + (fw4 - wi0)) * p.height;
    }
    pad_batch_kernel<<<dim3((unsigned)((total + 255) / 256), 1, n), 256, 0, st>>>(d_items, ps);
    g_svtme_launches += 2;
}

// Generic svt_aom_downsample_2d (any size / step 2,4,8): one thread per output pixel.
// Kernel-test ABI only (PictureOperatorTest.cc:225-335 shapes).
__global__ void __launch_bounds__(256)
    downsample_generic_kernel(const uint8_t *__restrict__ in, int is, int w, int h, uint8_t *__restrict__ out, int os,
                              int step) {
    const int ox = blockIdx.x * blockDim.x + threadIdx.x, oy = blockIdx.y * blockDim.y + threadIdx.y;
    const int half = step >> 1;
    const int hx = half + ox * step, vy = half + oy * step;
    if (hx >= w || vy >= h) return;
    const uint8_t *lo = in + (size_t)vy * is, *hi = lo - is;
    out[(size_t)oy * os + ox] = (uint8_t)avg4(hi[hx - 1], hi[hx], lo[hx - 1], lo[hx]);
}

void svtme_launch_downsample_generic(const uint8_t *in, int is, int w, int h, uint8_t *out, int os, int step,
                                     cudaStream_t st) {
    const int half = step >> 1;
    const int ow = w > half ? (w - half + step - 1) / step : 0, oh = h > half ? (h - half + step - 1) / step : 0;
    if (ow <= 0 || oh <= 0) return;
    dim3 blk(32, 8), grd((ow + 31) / 32, (oh + 7) / 8);
    downsample_generic_kernel<<<grd, blk, 0, st>>>(in, is, w, h, out, os, step);
    ++g_svtme_launches;
}

// ---- small transfers without the copy engines
// The main stream carries only kernels: its job / descriptor arrays are READ from pinned host memory by
// a kernel, and its buffers are cleared by a kernel. A cudaMemcpyAsync / cudaMemsetAsync here would
// queue on a copy engine behind the bulk picture uploads of the upload stream (measured: the first
// pipeline of a look-ahead window then starts only after ALL queued uploads, not after its own).
__global__ void __launch_bounds__(256) fetch_words_kernel(uint4 *__restrict__ dst, const uint4 *__restrict__ src_host, size_t n16) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src_host[i];
}
__global__ void __launch_bounds__(256) zero_fill_kernel(uint4 *__restrict__ dst, size_t n16) {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = z;
}

// bytes: multiple of 16; dst 16-byte aligned device memory; src: pinned host memory (its device alias is resolved here)
cudaError_t svtme_launch_fetch_from_host(void *dst, const void *src_pinned, size_t bytes, cudaStream_t st) {
    void *alias = nullptr;
    cudaError_t e = cudaHostGetDevicePointer(&alias, const_cast<void *>(src_pinned), 0);
    if (e != cudaSuccess) return e;
    const size_t n16 = bytes / 16;
    if (!n16) return cudaSuccess;
    const unsigned blocks = (unsigned)((n16 + 255) / 256 < 148 ? (n16 + 255) / 256 : 148);
    fetch_words_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<uint4 *>(dst), reinterpret_cast<const uint4 *>(alias), n16);
    ++g_svtme_launches;
    return cudaGetLastError();
}

cudaError_t svtme_launch_zero_fill(void *dst, size_t bytes, cudaStream_t st) {
    const size_t n16 = bytes / 16;
    if (!n16) return cudaSuccess;
    const size_t want = (n16 + 255) / 256;
    const unsigned blocks = (unsigned)(want < 148 * 8 ? want : 148 * 8);
    zero_fill_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<uint4 *>(dst), n16);
    ++g_svtme_launches;
    return cudaGetLastError();
}

// svtme_testabi.cu — kernel-test ABI: the RTCD signatures of Source/Lib/Codec/aom_dsp_rtcd.h as
// synchronous host wrappers (H2D, launch, D2H), so SadTest.cc / PictureOperatorTest.cc shaped
// harnesses can be pointed at CUDA. The two hot ones (svt_sad_loop_kernel, the full-pel sblock
// search) run the SAME engines as the stage kernel; the small per-call kernels (8 positions of
// one block, table init, ...) are plain one-thread-per-output kernels: at this granularity a call
// is launch-latency bound and exists for parity checking only.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "svtme_common.cuh"

#define CK(call)                                                                        \
    do {                                                                                \
        cudaError_t e_ = (call);                                                        \
        if (e_ != cudaSuccess) {                                                        \
            fprintf(stderr, "svtme_b200: %s failed: %s\n", #call, cudaGetErrorString(e_)); \
            abort(); /* the RTCD signatures return void: there is no CPU fallback */    \
        }                                                                               \
    } while (0)

namespace {
struct DevBuf {
    uint8_t *p = nullptr;
    DevBuf(size_t n, const void *src = nullptr, size_t ncopy = 0, size_t front_slack = 0) {
        CK(cudaMalloc(&p, n + front_slack));
        CK(cudaMemset(p, 0, n + front_slack));
        if (src) CK(cudaMemcpy(p + front_slack, src, ncopy, cudaMemcpyHostToDevice));
        p += front_slack;
        slack = front_slack;
    }
    ~DevBuf() { cudaFree(p - slack); }
    size_t slack = 0;
};
} // namespace

extern "C" void svt_sad_loop_kernel_b200(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                                         uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                                         int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                                         uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height) {
    *best_sad = 0xffffff;
    if (search_area_width <= 0 || search_area_height <= 0) return;
    if (block_width > 64 || block_height > 64 || src_stride_raw == 0 || ref_stride % src_stride_raw) {
        fprintf(stderr, "svt_sad_loop_kernel_b200: unsupported geometry (block <= 64x64, ref_stride = k * raw)\n");
        abort();
    }
    const size_t src_bytes = (size_t)(block_height - 1) * src_stride + block_width;
    const size_t ref_bytes = (size_t)(search_area_height - 1) * src_stride_raw + (size_t)(block_height - 1) * ref_stride +
        search_area_width + block_width - 1;
    DevBuf dsrc(src_bytes + 64, src, src_bytes), dref(ref_bytes + 4 * src_stride_raw + 256, ref, ref_bytes, 16);
    DevBuf dkey(8);
    svtme_launch_sad_loop_test(dsrc.p, (int)src_stride, dref.p, (int)ref_stride, (int)block_height, (int)block_width,
                               (int)src_stride_raw, skip_search_line, search_area_width, search_area_height,
                               (unsigned long long *)dkey.p, 0);
    unsigned long long key = 0;
    CK(cudaMemcpy(&key, dkey.p, 8, cudaMemcpyDeviceToHost));
    if ((key >> 32) < 0xffffff) {
        *best_sad = key >> 32;
        *x_search_center = (int16_t)(key & 0xffff);
        *y_search_center = (int16_t)((key >> 16) & 0xffff);
    }
}

extern "C" void svt_open_loop_me_fullpel_search_sblock_b200(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                                            uint32_t ref_stride, int16_t x_search_area_origin,
                                                            int16_t y_search_area_origin, uint32_t search_area_width,
                                                            uint32_t search_area_height, uint8_t sub_sad,
                                                            uint32_t *p_sb_best_sad, uint32_t *p_sb_best_mv) {
    if (!search_area_width || !search_area_height) return;
    const size_t src_bytes = (size_t)63 * src_stride + 64;
    const size_t ref_bytes = (size_t)(search_area_height - 1 + 63) * ref_stride + search_area_width + 63;
    DevBuf dsrc(src_bytes + 64, src, src_bytes), dref(ref_bytes + 8 * ref_stride + 256, ref, ref_bytes, 16);
    DevBuf dsad(85 * 4, p_sb_best_sad, 85 * 4), dmv(85 * 4, p_sb_best_mv, 85 * 4);
    svtme_launch_fullpel_test(dsrc.p, (int)src_stride, dref.p, (int)ref_stride, x_search_area_origin, y_search_area_origin,
                              (int)search_area_width, (int)search_area_height, sub_sad, (uint32_t *)dsad.p, (uint32_t *)dmv.p, 0);
    CK(cudaMemcpy(p_sb_best_sad, dsad.p, 85 * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_sb_best_mv, dmv.p, 85 * 4, cudaMemcpyDeviceToHost));
}

void svtme_set_error(const char *what); // svtme_api.cu
extern "C" int svt_open_loop_me_fullpel_search_sblock_b200_batch(void *cuda_stream, uint32_t n, const uint8_t *d_src, uint32_t src_stride,
                                                                 const uint8_t *d_ref, uint32_t ref_stride,
                                                                 const SvtMeFullpelItem *d_items, uint8_t sub_sad, uint32_t *d_best_sad,
                                                                 uint32_t *d_best_mv) {
    if (n == 0) return SVTME_OK;
    if (!d_src || !d_ref || !d_items || !d_best_sad || !d_best_mv) {
        svtme_set_error("svt_open_loop_me_fullpel_search_sblock_b200_batch: null pointer");
        return SVTME_ERR_INVALID;
    }
    const cudaError_t e = svtme_launch_fullpel_batch(d_src, (int)src_stride, d_ref, (int)ref_stride, d_items, (int)n, sub_sad, d_best_sad,
                                                     d_best_mv, (cudaStream_t)cuda_stream);
    if (e != cudaSuccess) {
        svtme_set_error(cudaGetErrorString(e));
        return SVTME_ERR_CUDA;
    }
    return SVTME_OK;
}

// ---------------------------------------------------------------- small per-call kernels
__device__ uint32_t sad8xn(const uint8_t *s, int ss, const uint8_t *r, int rs, int rows) {
    uint32_t acc = 0;
    for (int y = 0; y < rows; ++y) {
        const uint2 a = make_uint2(0, 0);
        (void)a;
        for (int x = 0; x < 8; x += 4) {
            uint32_t ws = s[y * ss + x] | (s[y * ss + x + 1] << 8) | (s[y * ss + x + 2] << 16) | ((uint32_t)s[y * ss + x + 3] << 24);
            uint32_t wr = r[y * rs + x] | (r[y * rs + x + 1] << 8) | (r[y * rs + x + 2] << 16) | ((uint32_t)r[y * rs + x + 3] << 24);
            acc = __vsadu4(ws, wr) + acc;
        }
    }
    return acc;
}
__device__ uint32_t sad8x8_mode(const uint8_t *s, int ss, const uint8_t *r, int rs, int sub) {
    return sub ? (sad8xn(s, 2 * ss, r, 2 * rs, 4) << 1) : sad8xn(s, ss, r, rs, 8);
}
__device__ uint32_t mv_add_x(uint32_t mv, int k) {
    int16_t x = (int16_t)(mv & 0xffff), y = (int16_t)(mv >> 16);
    x = (int16_t)(x + k);
    return ((uint32_t)(uint16_t)y << 16) | (uint16_t)x;
}

// motion_estimation.c:210-362 — threads 0..63: one 8x8 PU each; 64..79: one 16x16 PU each.
__global__ void all_sad_8x8_16x16_kernel(const uint8_t *src, int ss, const uint8_t *ref, int rs, uint32_t mv,
                                         uint32_t *bs8, uint32_t *bs16, uint32_t *bm8, uint32_t *bm16, uint32_t *eight16,
                                         int sub) {
    const int t = threadIdx.x;
    if (t < 64) {
        const int slot = t >> 2, q = t & 3;
        const int bx = ((slot >> 2) & 1) * 2 + (slot & 1), by = ((slot >> 3) & 1) * 2 + ((slot >> 1) & 1);
        const int ox = 16 * bx + (q & 1) * 8, oy = 16 * by + (q >> 1) * 8;
        uint32_t best = bs8[t], bmv = bm8[t];
        for (int k = 0; k < 8; ++k) {
            uint32_t sad = sad8x8_mode(src + oy * ss + ox, ss, ref + oy * rs + ox + k, rs, sub);
            if (sad < best) { best = sad; bmv = mv_add_x(mv, k); }
        }
        bs8[t] = best; bm8[t] = bmv;
    } else if (t < 80) {
        const int slot = t - 64;
        const int bx = ((slot >> 2) & 1) * 2 + (slot & 1), by = ((slot >> 3) & 1) * 2 + ((slot >> 1) & 1);
        uint32_t best = bs16[slot], bmv = bm16[slot];
        for (int k = 0; k < 8; ++k) {
            uint32_t sum = 0;
            for (int q = 0; q < 4; ++q) {
                const int ox = 16 * bx + (q & 1) * 8, oy = 16 * by + (q >> 1) * 8;
                sum += sad8x8_mode(src + oy * ss + ox, ss, ref + oy * rs + ox + k, rs, sub);
            }
            eight16[slot * 8 + k] = sum;
            if (sum < best) { best = sum; bmv = mv_add_x(mv, k); }
        }
        bs16[slot] = best; bm16[slot] = bmv;
    }
}

extern "C" void svt_ext_all_sad_calculation_8x8_16x16_b200(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                                           uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                                           uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                                           uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                                           uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad) {
    (void)p_eight_sad8x8;
    const size_t sb = (size_t)63 * src_stride + 64, rb = (size_t)63 * ref_stride + 64 + 7;
    DevBuf ds(sb, src, sb), dr(rb, ref, rb);
    DevBuf b8(256, p_best_sad_8x8, 256), b16(64, p_best_sad_16x16, 64), m8(256, p_best_mv8x8, 256), m16(64, p_best_mv16x16, 64);
    DevBuf e16(16 * 8 * 4);
    all_sad_8x8_16x16_kernel<<<1, 96>>>(ds.p, (int)src_stride, dr.p, (int)ref_stride, mv, (uint32_t *)b8.p, (uint32_t *)b16.p,
                                        (uint32_t *)m8.p, (uint32_t *)m16.p, (uint32_t *)e16.p, sub_sad);
    ++g_svtme_launches;
    CK(cudaGetLastError());
    CK(cudaMemcpy(p_best_sad_8x8, b8.p, 256, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_sad_16x16, b16.p, 64, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv8x8, m8.p, 256, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv16x16, m16.p, 64, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_eight_sad16x16, e16.p, 16 * 8 * 4, cudaMemcpyDeviceToHost));
}

// motion_estimation.c:369-425 — threads 0..3: 32x32 PUs; thread 4: 64x64.
__global__ void eight_sad_32x32_64x64_kernel(const uint32_t *s16, uint32_t *bs32, uint32_t *bs64, uint32_t *bm32,
                                             uint32_t *bm64, uint32_t mv, uint32_t *s32) {
    const int t = threadIdx.x;
    if (t < 4) {
        uint32_t best = bs32[t], bmv = bm32[t];
        for (int k = 0; k < 8; ++k) {
            uint32_t s = s16[(4 * t) * 8 + k] + s16[(4 * t + 1) * 8 + k] + s16[(4 * t + 2) * 8 + k] + s16[(4 * t + 3) * 8 + k];
            s32[t * 8 + k] = s;
            if (s < best) { best = s; bmv = mv_add_x(mv, k); }
        }
        bs32[t] = best; bm32[t] = bmv;
    } else if (t == 4) {
        uint32_t best = bs64[0], bmv = bm64[0];
        for (int k = 0; k < 8; ++k) {
            uint32_t s = 0;
            for (int i = 0; i < 16; ++i) s += s16[i * 8 + k];
            if (s < best) { best = s; bmv = mv_add_x(mv, k); }
        }
        bs64[0] = best; bm64[0] = bmv;
    }
}

extern "C" void svt_ext_eight_sad_calculation_32x32_64x64_b200(uint32_t p_sad16x16[16][8], uint32_t *p_best_sad_32x32,
                                                               uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                                               uint32_t *p_best_mv64x64, uint32_t mv,
                                                               uint32_t p_sad32x32[4][8]) {
    DevBuf s16(512, p_sad16x16, 512), b32(16, p_best_sad_32x32, 16), b64(4, p_best_sad_64x64, 4), m32(16, p_best_mv32x32, 16),
        m64(4, p_best_mv64x64, 4), s32(128);
    eight_sad_32x32_64x64_kernel<<<1, 32>>>((uint32_t *)s16.p, (uint32_t *)b32.p, (uint32_t *)b64.p, (uint32_t *)m32.p,
                                            (uint32_t *)m64.p, mv, (uint32_t *)s32.p);
    ++g_svtme_launches;
    CK(cudaGetLastError());
    CK(cudaMemcpy(p_best_sad_32x32, b32.p, 16, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_sad_64x64, b64.p, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv32x32, m32.p, 16, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv64x64, m64.p, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_sad32x32, s32.p, 128, cudaMemcpyDeviceToHost));
}

// motion_estimation.c:98-165 — one 16x16 block, one position.
__global__ void sad_8x8_16x16_kernel(const uint8_t *src, int ss, const uint8_t *ref, int rs, uint32_t *bs8, uint32_t *bs16,
                                     uint32_t *bm8, uint32_t *bm16, uint32_t mv, uint32_t *s16, uint32_t *s8, int sub) {
    __shared__ uint32_t sad[4];
    const int t = threadIdx.x;
    if (t < 4) {
        const int ox = (t & 1) * 8, oy = (t >> 1) * 8;
        sad[t] = sad8x8_mode(src + oy * ss + ox, ss, ref + oy * rs + ox, rs, sub);
        s8[t] = sad[t];
        if (sad[t] < bs8[t]) { bs8[t] = sad[t]; bm8[t] = mv; }
    }
    __syncthreads();
    if (t == 0) {
        const uint32_t sum = sad[0] + sad[1] + sad[2] + sad[3];
        if (sum < bs16[0]) { bs16[0] = sum; bm16[0] = mv; }
        *s16 = sum;
    }
}

extern "C" void svt_ext_sad_calculation_8x8_16x16_b200(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                                                       uint32_t *p_best_sad_8x8, uint32_t *p_best_sad_16x16,
                                                       uint32_t *p_best_mv8x8, uint32_t *p_best_mv16x16, uint32_t mv,
                                                       uint32_t *p_sad16x16, uint32_t *p_sad8x8, uint8_t sub_sad) {
    const size_t sb = (size_t)15 * src_stride + 16, rb = (size_t)15 * ref_stride + 16;
    DevBuf ds(sb, src, sb), dr(rb, ref, rb), b8(16, p_best_sad_8x8, 16), b16(4, p_best_sad_16x16, 4), m8(16, p_best_mv8x8, 16),
        m16(4, p_best_mv16x16, 4), s16(4), s8(16);
    sad_8x8_16x16_kernel<<<1, 32>>>(ds.p, (int)src_stride, dr.p, (int)ref_stride, (uint32_t *)b8.p, (uint32_t *)b16.p,
                                    (uint32_t *)m8.p, (uint32_t *)m16.p, mv, (uint32_t *)s16.p, (uint32_t *)s8.p, sub_sad);
    ++g_svtme_launches;
    CK(cudaGetLastError());
    CK(cudaMemcpy(p_best_sad_8x8, b8.p, 16, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_sad_16x16, b16.p, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv8x8, m8.p, 16, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv16x16, m16.p, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_sad16x16, s16.p, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_sad8x8, s8.p, 16, cudaMemcpyDeviceToHost));
}

// motion_estimation.c:172-205
__global__ void sad_32x32_64x64_kernel(const uint32_t *s16, uint32_t *bs32, uint32_t *bs64, uint32_t *bm32, uint32_t *bm64,
                                       uint32_t mv, uint32_t *s32) {
    __shared__ uint32_t q[4];
    const int t = threadIdx.x;
    if (t < 4) {
        q[t] = s16[4 * t] + s16[4 * t + 1] + s16[4 * t + 2] + s16[4 * t + 3];
        s32[t] = q[t];
        if (q[t] < bs32[t]) { bs32[t] = q[t]; bm32[t] = mv; }
    }
    __syncthreads();
    if (t == 0) {
        const uint32_t tot = q[0] + q[1] + q[2] + q[3];
        if (tot < bs64[0]) { bs64[0] = tot; bm64[0] = mv; }
    }
}

extern "C" void svt_ext_sad_calculation_32x32_64x64_b200(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                                         uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                                         uint32_t *p_best_mv64x64, uint32_t mv, uint32_t *p_sad32x32) {
    DevBuf s16(64, p_sad16x16, 64), b32(16, p_best_sad_32x32, 16), b64(4, p_best_sad_64x64, 4), m32(16, p_best_mv32x32, 16),
        m64(4, p_best_mv64x64, 4), s32(16);
    sad_32x32_64x64_kernel<<<1, 32>>>((uint32_t *)s16.p, (uint32_t *)b32.p, (uint32_t *)b64.p, (uint32_t *)m32.p,
                                      (uint32_t *)m64.p, mv, (uint32_t *)s32.p);
    ++g_svtme_launches;
    CK(cudaGetLastError());
    CK(cudaMemcpy(p_best_sad_32x32, b32.p, 16, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_sad_64x64, b64.p, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv32x32, m32.p, 16, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_best_mv64x64, m64.p, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(p_sad32x32, s32.p, 16, cudaMemcpyDeviceToHost));
}

__global__ void fill32_kernel(uint32_t *p, uint32_t n, uint32_t v) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

extern "C" void svt_initialize_buffer_32bits_b200(uint32_t *pointer, uint32_t count128, uint32_t count32, uint32_t value) {
    const uint32_t n = count128 * 4 + count32;
    if (!n) return;
    DevBuf d((size_t)n * 4);
    fill32_kernel<<<(n + 255) / 256, 256>>>((uint32_t *)d.p, n, value);
    ++g_svtme_launches;
    CK(cudaGetLastError());
    CK(cudaMemcpy(pointer, d.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
}

// compute_sad_c.c:20-37 — one CTA, row-strided partial sums + block reduce.
__global__ void nxm_sad_kernel(const uint8_t *src, int ss, const uint8_t *ref, int rs, int h, int w, uint32_t *out) {
    uint32_t acc = 0;
    for (int i = threadIdx.x; i < h * w; i += blockDim.x) {
        const int y = i / w, x = i % w;
        const int a = src[(size_t)y * ss + x], b = ref[(size_t)y * rs + x];
        acc += (uint32_t)abs(a - b);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

extern "C" uint32_t svt_nxm_sad_kernel_b200(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                            uint32_t height, uint32_t width) {
    if (!height || !width) return 0;
    const size_t sb = (size_t)(height - 1) * src_stride + width, rb = (size_t)(height - 1) * ref_stride + width;
    DevBuf ds(sb, src, sb), dr(rb, ref, rb), o(4);
    nxm_sad_kernel<<<1, 256>>>(ds.p, (int)src_stride, dr.p, (int)ref_stride, (int)height, (int)width, (uint32_t *)o.p);
    ++g_svtme_launches;
    CK(cudaGetLastError());
    uint32_t r = 0;
    CK(cudaMemcpy(&r, o.p, 4, cudaMemcpyDeviceToHost));
    return r;
}

extern "C" void svt_aom_downsample_2d_b200(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                                           uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                                           uint32_t decim_step) {
    const uint32_t half = decim_step >> 1;
    if (input_area_width <= half || input_area_height <= half) return;
    const uint32_t ow = (input_area_width - half + decim_step - 1) / decim_step,
                   oh = (input_area_height - half + decim_step - 1) / decim_step;
    const size_t ib = (size_t)(input_area_height - 1) * input_stride + input_area_width;
    DevBuf di(ib, input_samples, ib), dout((size_t)oh * ow);
    svtme_launch_downsample_generic(di.p, (int)input_stride, (int)input_area_width, (int)input_area_height, dout.p, (int)ow,
                                    (int)decim_step, 0);
    CK(cudaGetLastError());
    CK(cudaMemcpy2D(decim_samples, decim_stride, dout.p, ow, ow, oh, cudaMemcpyDeviceToHost));
}

extern "C" void svt_aom_generate_padding_b200(uint8_t *src_pic, uint32_t src_stride, uint32_t original_src_width,
                                              uint32_t original_src_height, uint32_t padding_width, uint32_t padding_height) {
    if (!src_pic) return;
    const size_t rows = original_src_height + 2 * (size_t)padding_height, bytes = rows * src_stride;
    DevBuf d(bytes, src_pic, bytes);
    DevPlane pl;
    pl.base = d.p;
    pl.stride = (int)src_stride;
    pl.org_x = (int)padding_width;
    pl.org_y = (int)padding_height;
    pl.width = (int)original_src_width;
    pl.height = (int)original_src_height;
    // the reference copies whole strides vertically; only [0, width + 2*pad) is defined content
    const int pr = (int)min((size_t)padding_width, (size_t)src_stride - padding_width - original_src_width);
    svtme_launch_pad_plane(pl, (int)padding_width, pr, (int)padding_height, (int)padding_height, 0);
    CK(cudaGetLastError());
    CK(cudaMemcpy(src_pic, d.p, bytes, cudaMemcpyDeviceToHost));
}

// ---------------------------------------------------------------- integer-SAD roofline probe
// Register-only VABSDIFF4.U8.ACC chains on every SM: the measured issue rate of the instruction the
// search engines are built on (the roofline denominator MEASURED_PEAKS.json does not carry).
__global__ void __launch_bounds__(256) vsad_rate_kernel(uint32_t *out, uint32_t seed, int iters) {
    uint32_t a[8], acc[8];
    const uint32_t b = seed * 0x9E3779B9u + threadIdx.x;
#pragma unroll
    for (int c = 0; c < 8; ++c) { a[c] = b * (c + 1) + 12345u; acc[c] = c; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < 8; ++c) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(acc[c]) : "r"(a[c]), "r"(b));
    }
    uint32_t s = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) s += acc[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" double svtme_measure_vabsdiff4_ginst_per_s(void) {
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = sms * 8, iters = 4096;
    DevBuf out((size_t)grid * 256 * 4);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 8; ++r) {
        CK(cudaEventRecord(e0));
        vsad_rate_kernel<<<grid, 256>>>((uint32_t *)out.p, (uint32_t)r, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r >= 3 && ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return (double)grid * 256 * iters * 8 / (best * 1e-3) / 1e9;
}

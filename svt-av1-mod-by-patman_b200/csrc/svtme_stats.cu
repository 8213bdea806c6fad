// svtme_stats.cu — picture statistics of a resident picture: the device twin of
// svt_aom_gathering_picture_statistics (Source/Lib/Codec/pic_analysis_process.c:1560-1578), which the reference's
// picture-analysis kernel runs right after the pyramid build (:2137) on the padded luma and the 1/16 plane — the two
// planes that are in HBM (and mostly still in L2) when svtme_put_picture* returns.
//
//   stats_variance_kernel   compute_picture_spatial_statistics / compute_block_mean_compute_variance (:307-1378,
//                           :1530-1552): one CTA per b64, one thread per 8x8 block; the sums and sums of squares come
//                           from two DP4A per 4 samples; 16x16 / 32x32 / 64x64 are the >> 2 averages of the level below.
//   stats_histogram_kernel  sub_sample_luma_generate_pixel_intensity_histogram_bins / calculate_histogram (:166-184,
//                           :1462-1528): one CTA per (region, band of sampled rows), shared-memory bins, one global
//                           atomic per non-empty bin and CTA.
//   stats_finalize_kernel   the scalar tails: bins start at 1 and are scaled by 16 * decim^2, average intensity per
//                           region, avg_luma, pic_avg_variance.
// All arithmetic is the reference's, in the reference's integer types (uint64_t means, uint16_t variance): bit-exact,
// including the incomplete b64 at the right / bottom edge, which reads the replicated padding like the reference.
#include "svtme_common.cuh"

namespace {
#define VARIANCE_PRECISION 16 // definitions.h

__device__ __forceinline__ uint16_t var_of(unsigned long long msq, unsigned long long mean) {
    return (uint16_t)((msq - mean * mean) >> VARIANCE_PRECISION);
}

// One call covers n pictures of one geometry: blockIdx.z (or .y) is the picture, its plane bases come from `items`
// (device memory, 16 bytes per picture), its outputs go to scratch + pic * pitch.
struct StatsItem {
    const uint8_t *full, *sixteenth;
};
struct StatsLayout {
    size_t pitch, o_hist, o_rsum, o_vsum, o_tail;
};

__global__ void __launch_bounds__(64) stats_variance_kernel(DevPlane full, const StatsItem *items, int cols, int prec_sub, int all_sizes,
                                                            uint8_t *scratch, StatsLayout L) {
    __shared__ unsigned long long m8[64], q8[64], m16[16], q16[16], m32[4], q32[4];
    full.base = const_cast<uint8_t *>(items[blockIdx.y].full);
    uint16_t *var = reinterpret_cast<uint16_t *>(scratch + L.pitch * blockIdx.y);
    unsigned long long *var_sum = reinterpret_cast<unsigned long long *>(scratch + L.pitch * blockIdx.y + L.o_vsum);
    const int b = blockIdx.x, t = threadIdx.x, by = t >> 3, bx = t & 7;
    const uint8_t *p = full.at((b % cols) * 64 + 8 * bx, (b / cols) * 64 + 8 * by); // 8-byte aligned (plane rows are 128-byte aligned)
    uint32_t s = 0, q = 0;
    const int step = prec_sub ? 2 : 1; // svt_compute_interm_var_four8x8 reads rows 0, 2, 4, 6 (:234-272)
    for (int r = 0; r < 8; r += step) {
        const uint2 v = *reinterpret_cast<const uint2 *>(p + (size_t)r * full.stride);
        s = __dp4a(v.x, 0x01010101u, s);
        s = __dp4a(v.y, 0x01010101u, s);
        q = __dp4a(v.x, v.x, q);
        q = __dp4a(v.y, v.y, q);
    }
    // (:197-232) mean = (sum << 8) / 64, mean of squares = (sum << 16) / 64; (:234-272) sum << 3 and sum << 11 of the half block
    const unsigned long long mean = prec_sub ? (unsigned long long)s << 3 : (unsigned long long)s << 2;
    const unsigned long long msq = prec_sub ? (unsigned long long)q << 11 : (unsigned long long)q << 10;
    m8[t] = mean;
    q8[t] = msq;
    __syncthreads();
    if (t < 16) {
        const int k = (2 * (t >> 2)) * 8 + 2 * (t & 3);
        m16[t] = (m8[k] + m8[k + 1] + m8[k + 8] + m8[k + 9]) >> 2;
        q16[t] = (q8[k] + q8[k + 1] + q8[k + 8] + q8[k + 9]) >> 2;
    }
    __syncthreads();
    if (t < 4) {
        const int k = (2 * (t >> 1)) * 4 + 2 * (t & 1);
        m32[t] = (m16[k] + m16[k + 1] + m16[k + 4] + m16[k + 5]) >> 2;
        q32[t] = (q16[k] + q16[k + 1] + q16[k + 4] + q16[k + 5]) >> 2;
    }
    __syncthreads();
    const int n = all_sizes ? 85 : 1;
    uint16_t *o = var + (size_t)b * n;
    if (all_sizes) { // raster-scan indices (me_context.h:54-140): 64x64 = 0, 32x32 = 1..4, 16x16 = 5..20, 8x8 = 21..84
        o[21 + t] = var_of(msq, mean);
        if (t < 16) o[5 + t] = var_of(q16[t], m16[t]);
        if (t < 4) o[1 + t] = var_of(q32[t], m32[t]);
    }
    if (t == 0) {
        const unsigned long long m64 = (m32[0] + m32[1] + m32[2] + m32[3]) >> 2, q64 = (q32[0] + q32[1] + q32[2] + q32[3]) >> 2;
        const uint16_t v = var_of(q64, m64);
        o[0] = v;
        atomicAdd(var_sum, (unsigned long long)v);
    }
}

#define HIST_ROWS_PER_CTA 16 // sampled rows
__global__ void __launch_bounds__(256) stats_histogram_kernel(DevPlane pl, const StatsItem *items, int nw, int nh, int decim, uint8_t *scratch,
                                                              StatsLayout L) {
    __shared__ uint32_t bins[256];
    __shared__ unsigned long long sum_sh;
    pl.base = const_cast<uint8_t *>(items[blockIdx.z].sixteenth);
    uint32_t *hist = reinterpret_cast<uint32_t *>(scratch + L.pitch * blockIdx.z + L.o_hist);                          // [4][4][256]
    unsigned long long *region_sum = reinterpret_cast<unsigned long long *>(scratch + L.pitch * blockIdx.z + L.o_rsum); // [4][4]
    const int region = blockIdx.x, ix = region / nh, iy = region % nh, band = blockIdx.y, t = threadIdx.x;
    const int rw = pl.width / nw, rh = pl.height / nh;
    const int w = rw + (ix == nw - 1 ? pl.width - nw * rw : 0), h = rh + (iy == nh - 1 ? pl.height - nh * rh : 0);
    const int n_cols = (w + decim - 1) / decim, n_rows = (h + decim - 1) / decim;
    const int r0 = band * HIST_ROWS_PER_CTA, r1 = min(n_rows, r0 + HIST_ROWS_PER_CTA);
    if (r0 >= n_rows) return;
    bins[t] = 0;
    if (t == 0) sum_sh = 0;
    __syncthreads();
    unsigned long long sum = 0;
    const int total = (r1 - r0) * n_cols;
    for (int i = t; i < total; i += 256) {
        const int r = r0 + i / n_cols, c = i % n_cols;
        const uint8_t v = *pl.at(ix * rw + c * decim, iy * rh + r * decim);
        atomicAdd(&bins[v], 1u);
        sum += v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((t & 31) == 0 && sum) atomicAdd(&sum_sh, sum);
    __syncthreads();
    if (bins[t]) atomicAdd(&hist[(ix * 4 + iy) * 256 + t], bins[t]);
    if (t == 0 && sum_sh) atomicAdd(&region_sum[ix * 4 + iy], sum_sh);
}

struct StatsTail { // written by stats_finalize_kernel, fetched by the host wrapper
    unsigned long long avg_intensity[16];
    unsigned long long avg_luma;
    uint32_t pic_avg_variance;
    uint32_t _pad;
};

__global__ void __launch_bounds__(256) stats_finalize_kernel(int plane_w, int plane_h, int nw, int nh, int decim, int calc_hist,
                                                             int calc_var, int n_b64, uint8_t *scratch, StatsLayout L) {
    uint8_t *mine = scratch + L.pitch * blockIdx.x;
    uint32_t *hist = reinterpret_cast<uint32_t *>(mine + L.o_hist);
    const unsigned long long *region_sum = reinterpret_cast<const unsigned long long *>(mine + L.o_rsum);
    const unsigned long long *var_sum = reinterpret_cast<const unsigned long long *>(mine + L.o_vsum);
    StatsTail *tail = reinterpret_cast<StatsTail *>(mine + L.o_tail);
    const int t = threadIdx.x;
    const uint32_t scale = 4u * 4u * (uint32_t)decim * (uint32_t)decim;
    if (calc_hist)
        for (int region = 0; region < nw * nh; ++region) {
            const int ix = region / nh, iy = region % nh;
            uint32_t *hp = hist + (ix * 4 + iy) * 256;
            hp[t] = (hp[t] + 1u) * scale; // "Initialize bins to 1" (:1481), then the scaling loop (:1516-1522)
        }
    if (t == 0) {
        unsigned long long total = 0;
        if (calc_hist) {
            const int rw = plane_w / nw, rh = plane_h / nh;
            for (int ix = 0; ix < nw; ++ix)
                for (int iy = 0; iy < nh; ++iy) {
                    const uint32_t w = (uint32_t)(rw + (ix == nw - 1 ? plane_w - nw * rw : 0));
                    const uint32_t h = (uint32_t)(rh + (iy == nh - 1 ? plane_h - nh * rh : 0));
                    const unsigned long long sum = region_sum[ix * 4 + iy] * (unsigned long long)(decim * decim);
                    tail->avg_intensity[ix * 4 + iy] = (uint8_t)((sum + ((w * h) >> 1)) / (w * h));
                    total += sum;
                }
            tail->avg_luma = total / (unsigned long long)((uint32_t)plane_w * (uint32_t)plane_h);
        } else
            tail->avg_luma = 256; // INVALID_LUMA (:1563)
        tail->pic_avg_variance = calc_var ? (uint16_t)(*var_sum / (unsigned long long)n_b64) : 0;
    }
}
} // namespace

static StatsLayout stats_layout(int n_b64) {
    StatsLayout L;
    L.o_hist = ((size_t)n_b64 * 85 * sizeof(uint16_t) + 255) & ~(size_t)255; // variance rows first
    L.o_rsum = L.o_hist + 16 * 256 * sizeof(uint32_t);
    L.o_vsum = L.o_rsum + 16 * sizeof(unsigned long long);
    L.o_tail = L.o_vsum + 256;
    L.pitch = (L.o_tail + sizeof(StatsTail) + 255) & ~(size_t)255;
    return L;
}
size_t svtme_stats_scratch_bytes(int n_b64) { return stats_layout(n_b64).pitch; }
size_t svtme_stats_item_bytes() { return sizeof(StatsItem); }
void svtme_stats_set_item(void *items_host, int i, const DevPyramid &pyr) {
    static_cast<StatsItem *>(items_host)[i] = StatsItem{pyr.lvl[0].base, pyr.lvl[2].base};
}

namespace {
__global__ void __launch_bounds__(256) stats_clear_kernel(uint8_t *scratch, StatsLayout L) { // bins, sums and tail of every picture
    uint4 *p = reinterpret_cast<uint4 *>(scratch + L.pitch * blockIdx.x + L.o_hist);
    const int n16 = (int)((L.pitch - L.o_hist) / 16);
    for (int i = threadIdx.x; i < n16; i += 256) p[i] = make_uint4(0, 0, 0, 0);
}
} // namespace

// Enqueues the statistics of n resident pyramids of one geometry (`geom`: any of them) on `st`: four launches whatever n
// is. items: n StatsItem in device memory (filled on the host by svtme_stats_set_item and copied over); scratch: n * svtme_stats_scratch_bytes(), 256-byte
// aligned; picture i's variance rows start at scratch + i * pitch, its histogram at + *off_hist, its tail at + *off_tail.
int svtme_launch_picture_statistics(const DevPyramid &geom, const void *items, int n, const SvtMeStatsParams &prm, uint8_t *scratch,
                                    size_t *off_hist, size_t *off_tail, cudaStream_t st) {
    const DevPlane &full = geom.lvl[0], &six = geom.lvl[2];
    const int cols = (full.width + 63) / 64, rows = (full.height + 63) / 64, n_b64 = cols * rows;
    const StatsLayout L = stats_layout(n_b64);
    const StatsItem *it = static_cast<const StatsItem *>(items);
    int launches = 1;
    stats_clear_kernel<<<n, 256, 0, st>>>(scratch, L);
    const int decim = prm.scene_change_detection ? 1 : 4;
    if (prm.calculate_variance) {
        stats_variance_kernel<<<dim3(n_b64, n), 64, 0, st>>>(full, it, cols, prm.block_mean_prec_sub ? 1 : 0, prm.all_block_sizes ? 1 : 0, scratch, L);
        ++launches;
    }
    if (prm.calc_hist) {
        const int nw = (int)prm.regions_per_width, nh = (int)prm.regions_per_height;
        const int max_h = six.height / nh + six.height % nh, bands = ((max_h + decim - 1) / decim + HIST_ROWS_PER_CTA - 1) / HIST_ROWS_PER_CTA;
        stats_histogram_kernel<<<dim3(nw * nh, bands, n), 256, 0, st>>>(six, it, nw, nh, decim, scratch, L);
        ++launches;
    }
    stats_finalize_kernel<<<n, 256, 0, st>>>(six.width, six.height, (int)prm.regions_per_width, (int)prm.regions_per_height, decim,
                                             prm.calc_hist ? 1 : 0, prm.calculate_variance ? 1 : 0, n_b64, scratch, L);
    ++launches;
    *off_hist = L.o_hist;
    *off_tail = L.o_tail;
    return launches;
}

size_t svtme_stats_tail_bytes() { return sizeof(StatsTail); }
void svtme_stats_tail_unpack(const void *tail_host, unsigned long long *avg_intensity16, unsigned long long *avg_luma, uint16_t *pic_avg_variance) {
    const StatsTail *t = static_cast<const StatsTail *>(tail_host);
    if (avg_intensity16)
        for (int i = 0; i < 16; ++i) avg_intensity16[i] = t->avg_intensity[i];
    if (avg_luma) *avg_luma = t->avg_luma;
    if (pic_avg_variance) *pic_avg_variance = (uint16_t)t->pic_avg_variance;
}

// Issue-rate micro-benchmark for the integer-SAD roofline denominator.
// Register-only dependent chains, 8 independent chains per thread, all SMs.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipebench tools/pipebench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
#define CHAINS 8

template <int OP>
__global__ void __launch_bounds__(256) bench(uint32_t *out, uint32_t seed) {
    uint32_t a[CHAINS], acc[CHAINS];
    uint32_t b = seed * 0x9E3779B9u + threadIdx.x;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { a[c] = b * (c + 1) + 12345u; acc[c] = c; }
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (OP == 0) {  // VABSDIFF4.U8.ACC
                asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(acc[c]) : "r"(a[c]), "r"(b));
            } else if (OP == 1) {  // vabsdiff4 (per-byte) + dp4a
                uint32_t d;
                asm volatile("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a[c]), "r"(acc[c]), "r"(0));
                asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(acc[c]) : "r"(d), "r"(0x01010101));
            } else if (OP == 2) {  // SHF (funnel shift)
                asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(acc[c]) : "r"(a[c]), "r"(8));
            } else if (OP == 3) {  // PRMT
                asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(acc[c]) : "r"(a[c]), "r"(0x4321));
            } else if (OP == 4) {  // IADD3
                asm volatile("add.u32 %0, %0, %1;" : "+r"(acc[c]) : "r"(a[c]));
            } else if (OP == 5) {  // umin
                asm volatile("min.u32 %0, %0, %1;" : "+r"(acc[c]) : "r"(a[c]));
            } else if (OP == 6) {  // dp4a alone
                asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(acc[c]) : "r"(a[c]), "r"(b));
            } else if (OP == 7) {  // IMAD
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(acc[c]) : "r"(a[c]), "r"(b));
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += acc[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// shared-memory load rate: LDS.32 / LDS.64 / LDS.128, conflict-free
template <int W>
__global__ void __launch_bounds__(256) lds_bench(uint32_t *out) {
    __shared__ __align__(16) uint32_t sm[256 * 4 + 64];
    for (int i = threadIdx.x; i < 256 * 4 + 64; i += 256) sm[i] = i * 2654435761u;
    __syncthreads();
    uint32_t s = 0;
    int off = 0;
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            if (W == 1) {
                s += sm[threadIdx.x + ((off + c) & 15)];
            } else if (W == 2) {
                uint2 v = *reinterpret_cast<uint2 *>(&sm[threadIdx.x * 2 + 2 * ((off + c) & 7)]);
                s += v.x ^ v.y;
            } else {
                uint4 v = *reinterpret_cast<uint4 *>(&sm[threadIdx.x * 4 + 4 * ((off + c) & 7)]);
                s += v.x ^ v.y ^ v.z ^ v.w;
            }
        }
        off += s & 1;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); f(); f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    int grid = sms * 8;  // 8 CTAs x 256 thr = 2048 thr/SM
    uint32_t *out; cudaMalloc(&out, (size_t)grid * 256 * 4);
    const char *names[] = {"vabsdiff4.acc", "vabsdiff4+dp4a", "shf", "prmt", "iadd", "umin", "dp4a", "imad"};
    double inst_per_launch = (double)grid * 256 * ITERS * CHAINS;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"max_clock_mhz\": %d, \"results\": {", p.name, sms, clk_khz / 1000);
#define RUN(OP) { double ms = time_ms([&] { bench<OP><<<grid, 256>>>(out, 1); }); \
    double gops = inst_per_launch / (ms * 1e-3) / 1e9; \
    printf("\"%s\": {\"ms\": %.4f, \"ginst_per_s\": %.1f, \"inst_per_clk_per_sm_at_max\": %.2f}, ", names[OP], ms, gops, gops * 1e9 / (sms * (clk_khz * 1e3))); }
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7)
#define RUNL(W, nm) { double ms = time_ms([&] { lds_bench<W><<<grid, 256>>>(out); }); \
    double g = (double)grid * 256 * ITERS * 8 / (ms * 1e-3) / 1e9; \
    printf("\"%s\": {\"ms\": %.4f, \"gload_per_s\": %.1f, \"bytes_per_clk_per_sm_at_max\": %.1f}, ", nm, ms, g, g * 1e9 * 4 * W / (sms * (clk_khz * 1e3))); }
    RUNL(1, "lds32") RUNL(2, "lds64") RUNL(4, "lds128")
    printf("\"_\": 0}}\n");
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { fprintf(stderr, "cuda error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}

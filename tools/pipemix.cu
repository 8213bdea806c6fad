// pipemix.cu — which instructions share the issue pipe of VABSDIFF4.U8.ACC on B200 (sm_100a)?
// Every kernel runs N_V VABSDIFF4 and N_X "other" instructions per loop iteration on independent register
// chains; if the other instruction executes on another pipe the time stays that of the VABSDIFF4 stream alone.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipemix tools/pipemix.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048
enum { X_NONE = 0, X_SHF, X_IMAD, X_IMADHI, X_IMADWIDE, X_IADD, X_UMIN, X_LOP3, X_PRMT, X_LDS, X_LDS128, X_ISETP_SEL, X_SHL, X_VIADD_IMM, X_N };
static const char *kNames[] = {"none", "shf", "imad", "imad.hi", "imad.wide", "iadd3", "umin", "lop3", "prmt", "lds32", "lds128", "isetp+sel", "shl", "imad.iadd"};

template <int X, int NV, int NX>
__global__ void __launch_bounds__(256) mix(uint32_t *out, uint32_t seed) {
    __shared__ __align__(16) uint32_t sm[256 * 4 + 64];
    for (int i = threadIdx.x; i < 256 * 4 + 64; i += 256) sm[i] = i * 2654435761u;
    __syncthreads();
    uint32_t a[8], acc[8], xs[8];
    uint32_t b = seed * 0x9E3779B9u + threadIdx.x;
#pragma unroll
    for (int c = 0; c < 8; ++c) { a[c] = b * (c + 1) + 12345u; acc[c] = c; xs[c] = b ^ (c * 77u); }
    const uint32_t m = (seed & 3) ? 256u : 65536u;
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            if (c < NV) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(acc[c]) : "r"(a[c]), "r"(b));
            if (c < NX) {
                if (X == X_SHF) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(xs[c]) : "r"(a[c]), "r"(8));
                else if (X == X_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(xs[c]) : "r"(m), "r"(b));
                else if (X == X_IMADHI) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(xs[c]) : "r"(m), "r"(b));
                else if (X == X_IMADWIDE) { unsigned long long w; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(xs[c]), "r"(m)); xs[c] = (uint32_t)(w >> 32) + (uint32_t)w; }
                else if (X == X_IADD) asm volatile("add.u32 %0, %0, %1;" : "+r"(xs[c]) : "r"(a[c]));
                else if (X == X_UMIN) asm volatile("min.u32 %0, %0, %1;" : "+r"(xs[c]) : "r"(a[c]));
                else if (X == X_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(xs[c]) : "r"(a[c]), "r"(b));
                else if (X == X_PRMT) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(xs[c]) : "r"(a[c]), "r"(0x4321));
                else if (X == X_LDS) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"((uint32_t)__cvta_generic_to_shared(&sm[(threadIdx.x + c + (xs[0] & 1)) & 1023]))); xs[c] ^= v; }
                else if (X == X_LDS128) { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"((uint32_t)__cvta_generic_to_shared(&sm[4 * ((threadIdx.x + c + (xs[0] & 1)) & 255)]))); xs[c] ^= v.x ^ v.w; }
                else if (X == X_ISETP_SEL) { asm volatile("{ .reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %1, %0, p; }" : "+r"(xs[c]) : "r"(a[c])); }
                else if (X == X_SHL) asm volatile("shl.b32 %0, %0, 3;" : "+r"(xs[c]));
                else if (X == X_VIADD_IMM) asm volatile("mad.lo.u32 %0, %0, 1, %1;" : "+r"(xs[c]) : "r"(a[c]));
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) s += acc[c] + xs[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

template <int X>
static void run_x(uint32_t *out, int grid, double base) {
    const double only = time_ms([&] { mix<X, 0, 8><<<grid, 256>>>(out, 1); });
    const double m88 = time_ms([&] { mix<X, 8, 8><<<grid, 256>>>(out, 1); });
    const double m84 = time_ms([&] { mix<X, 8, 4><<<grid, 256>>>(out, 1); });
    const double m82 = time_ms([&] { mix<X, 8, 2><<<grid, 256>>>(out, 1); });
    printf("  \"%s\": {\"alone_8\": %.4f, \"vabs8_x8\": %.4f, \"vabs8_x4\": %.4f, \"vabs8_x2\": %.4f, \"vabs8_alone\": %.4f},\n", kNames[X], only, m88,
           m84, m82, base);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int grid = p.multiProcessorCount * 8;
    uint32_t *out; cudaMalloc(&out, (size_t)grid * 256 * 4);
    const double base = time_ms([&] { mix<X_NONE, 8, 0><<<grid, 256>>>(out, 1); });
    printf("{\"gpu\": \"%s\", \"note\": \"ms per launch; 8 VABSDIFF4 + k other instructions per iteration and thread\",\n", p.name);
    run_x<X_SHF>(out, grid, base); run_x<X_IMAD>(out, grid, base); run_x<X_IMADHI>(out, grid, base); run_x<X_IMADWIDE>(out, grid, base);
    run_x<X_IADD>(out, grid, base); run_x<X_UMIN>(out, grid, base); run_x<X_LOP3>(out, grid, base); run_x<X_PRMT>(out, grid, base);
    run_x<X_LDS>(out, grid, base); run_x<X_LDS128>(out, grid, base); run_x<X_ISETP_SEL>(out, grid, base); run_x<X_SHL>(out, grid, base);
    run_x<X_VIADD_IMM>(out, grid, base);
    printf("  \"_\": 0}\n");
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { fprintf(stderr, "cuda error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}

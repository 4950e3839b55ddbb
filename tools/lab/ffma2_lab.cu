// ffma2_lab.cu — does fma.rn.f32x2 (FFMA2) double the fp32 FMAs per issue slot on sm_100a?
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{ unsigned long long d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
template <int MODE> __global__ void k(float *out, float g, int iters)
{
    float a[8]; unsigned long long p[8];
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 0.001f + i; p[i] = ((unsigned long long)__float_as_uint(a[i]) << 32) | __float_as_uint(a[i] + 0.5f); }
    const unsigned long long gg = ((unsigned long long)__float_as_uint(g) << 32) | __float_as_uint(g);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) a[i] = __fmaf_rn(a[i], g, 0.25f);              // imm-ish
            if (MODE == 1) a[i] = __fmaf_rn(a[i], g, a[(i + 1) & 7]);     // 3-reg
            if (MODE == 2) p[i] = fma2(p[i], gg, p[(i + 1) & 7]);
        }
    }
    float s = 0; for (int i = 0; i < 8; i++) s += a[i] + __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32));
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main()
{
    float *d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096;
    for (int mode = 0; mode < 3; mode++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(d, 1.0001f, iters);
            if (mode == 1) k<1><<<148 * 8, 256>>>(d, 1.0001f, iters);
            if (mode == 2) k<2><<<148 * 8, 256>>>(d, 1.0001f, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double inst = 148.0 * 8 * 256 * iters * 8;
        printf("mode %d: %.3f ms  %.1f G warp-lane-instr/s  (%.1f lane-instr/clk/SM @1.965GHz)  fma/s %.1f G\n", mode, ms, inst / ms / 1e6,
               inst / ms / 1e6 / 148 / 1.965, inst * (mode == 2 ? 2 : 1) / ms / 1e6);
    }
    return 0;
}

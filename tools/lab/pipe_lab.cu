// pipe_lab.cu — which sm_100a instructions share the ALU pipe with VIMNMX / PRMT / LOP3, and which run beside it?
// k_fast is ALU-pipe-bound (ncu: pipe_alu 78 %, pipe_fma 11 %): an op that does the same work on the FMA pipe is free issue room.
// Each mode runs 8 independent dependency chains per thread, 8 warps x 8 CTAs per SM; reports lane-instructions / clk / SM.
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>

__device__ __forceinline__ uint32_t vmin2(uint32_t a, uint32_t b) { return __vminu2(a, b); }
__device__ __forceinline__ uint32_t vmin3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t hmin2(uint32_t a, uint32_t b)
{ uint32_t d; asm("min.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t hmin2_bf(uint32_t a, uint32_t b)
{ uint32_t d; asm("min.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t hadd2(uint32_t a, uint32_t b)
{ uint32_t d; asm("add.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t hfma2(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ float fmnmx(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) { return a * b + c; }
__device__ __forceinline__ uint32_t dp4a(uint32_t a, uint32_t b, uint32_t c) { return __dp4a(a, b, c); }
__device__ __forceinline__ uint32_t vabsdiff4(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t vset4_gt(uint32_t a, uint32_t b) { return __vcmpgtu4(a, b); }
__device__ __forceinline__ uint32_t vmax4(uint32_t a, uint32_t b) { return __vmaxu4(a, b); }
__device__ __forceinline__ uint32_t vabsd4(uint32_t a, uint32_t b) { return __vabsdiffu4(a, b); }

enum { M_VMIN2, M_VMIN3, M_HMIN2, M_HMIN2BF, M_HADD2, M_HFMA2, M_PRMT, M_LOP3, M_IMAD, M_IADD3, M_FMNMX, M_DP4A, M_VCMP4, M_VMAX4, M_VABSD4,
       M_MIX_VMIN_HMIN, M_MIX_VMIN_IMAD, M_MIX_VMIN_HFMA, M_MIX_PRMT_IMAD, M_MIX_VMIN3_HMIN, M_MIX_LOP_IMAD, M_SHF, M_MIX_VMIN_FMNMX, M_I2F, M_MIX_VMIN_DP4A,
       M_COUNT };
static const char *names[M_COUNT] = {"vimnmx.u16x2", "vimnmx3.u16x2", "hmnmx2 (min.f16x2)", "min.bf16x2", "hadd2", "hfma2", "prmt", "lop3", "imad", "iadd3", "fmnmx",
                                     "idp.4a", "vcmpgtu4 (emulated?)", "vmaxu4 (emulated?)", "vabsdiffu4", "MIX vimnmx+hmnmx2", "MIX vimnmx+imad", "MIX vimnmx+hfma2",
                                     "MIX prmt+imad", "MIX vimnmx3+hmnmx2", "MIX lop3+imad", "shf (funnel)", "MIX vimnmx+fmnmx", "i2f.u8 -> f32 (cvt)", "MIX vimnmx+idp4a"};
static const int ops_per_iter[M_COUNT] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 1, 2, 1, 2};

template <int MODE> __global__ void __launch_bounds__(256) k(uint32_t *out, uint32_t seed, int iters)
{
    uint32_t a[8], b[8];
    float f[8];
    for (int i = 0; i < 8; i++) { a[i] = seed * (threadIdx.x + 1) + i * 0x01010101u; b[i] = a[i] ^ 0x3c003c00u; f[i] = (float)(a[i] & 255); }
    const uint32_t s = seed | 0x3210u, one = (seed & 1) | 1u;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int j = (i + 1) & 7;
            if (MODE == M_VMIN2) a[i] = vmin2(a[i], b[j]);
            if (MODE == M_VMIN3) a[i] = vmin3(a[i], b[j], b[i]);
            if (MODE == M_HMIN2) a[i] = hmin2(a[i], b[j]);
            if (MODE == M_HMIN2BF) a[i] = hmin2_bf(a[i], b[j]);
            if (MODE == M_HADD2) a[i] = hadd2(a[i], b[j]);
            if (MODE == M_HFMA2) a[i] = hfma2(a[i], b[j], b[i]);
            if (MODE == M_PRMT) a[i] = prmt(a[i], b[j], s);
            if (MODE == M_LOP3) asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(a[i]) : "r"(a[i]), "r"(b[j]), "r"(b[i]));
            if (MODE == M_IMAD) a[i] = imad(a[i], s, b[j]);
            if (MODE == M_IADD3) asm("add.u32 %0, %1, %2;" : "=r"(a[i]) : "r"(a[i]), "r"(b[j]));
            if (MODE == M_FMNMX) f[i] = fmnmx(f[i], f[j] * 0.0f + __uint_as_float(b[j]));
            if (MODE == M_DP4A) a[i] = dp4a(a[i], b[j], b[i]);
            if (MODE == M_VCMP4) a[i] = vset4_gt(a[i], b[j]);
            if (MODE == M_VMAX4) a[i] = vmax4(a[i], b[j]);
            if (MODE == M_VABSD4) a[i] = vabsd4(a[i], b[j]);
            if (MODE == M_SHF) a[i] = __funnelshift_r(a[i], b[j], s);
            if (MODE == M_I2F) f[i] = (float)(uint8_t)__float_as_uint(f[i]) + f[j];
            if (MODE == M_MIX_VMIN_HMIN) { a[i] = vmin2(a[i], b[j]); b[i] = hmin2(b[i], a[j]); }
            if (MODE == M_MIX_VMIN_IMAD) { a[i] = vmin2(a[i], b[j]); b[i] = imad(b[i], s, a[j]); }
            if (MODE == M_MIX_VMIN_HFMA) { a[i] = vmin2(a[i], b[j]); b[i] = hfma2(b[i], s, a[j]); }
            if (MODE == M_MIX_PRMT_IMAD) { a[i] = prmt(a[i], b[j], s); b[i] = imad(b[i], s, a[j]); }
            if (MODE == M_MIX_VMIN3_HMIN) { a[i] = vmin3(a[i], b[j], b[i]); b[i] = hmin2(b[i], a[j]); }
            if (MODE == M_MIX_LOP_IMAD) { asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(a[i]) : "r"(a[i]), "r"(b[j]), "r"(one)); b[i] = imad(b[i], s, a[j]); }
            if (MODE == M_MIX_VMIN_FMNMX) { a[i] = vmin2(a[i], b[j]); f[i] = fmnmx(f[i], f[j]); }
            if (MODE == M_MIX_VMIN_DP4A) { a[i] = vmin2(a[i], b[j]); b[i] = dp4a(b[i], s, a[j]); }
        }
    }
    uint32_t r = 0;
    for (int i = 0; i < 8; i++) r += a[i] ^ b[i] ^ __float_as_uint(f[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE> static void run(uint32_t *d, cudaEvent_t e0, cudaEvent_t e1, double ghz)
{
    const int iters = 2048, blocks = 148 * 8;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(e0);
        k<MODE><<<blocks, 256>>>(d, 12345u, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
    }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double inst = (double)blocks * 256 * iters * 8 * ops_per_iter[MODE];
    printf("%-26s %8.3f ms  %6.1f lane-ops/clk/SM (at %.3f GHz)\n", names[MODE], ms, inst / (ms * 1e-3) / 148 / (ghz * 1e9), ghz);
    if constexpr (MODE + 1 < M_COUNT) run<MODE + 1>(d, e0, e1, ghz);
}

int main()
{
    uint32_t *d;
    cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    run<0>(d, e0, e1, khz * 1e-6);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}

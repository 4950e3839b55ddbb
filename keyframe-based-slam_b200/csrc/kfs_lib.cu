// kfs_lib.cu — library-level entry points: version, thread-local error string, device probe and the
// integer-pipe microbenchmarks that give the matcher its roofline denominator (SURVEY.md §8d:
// "POPC peak is unmeasured — write the microbenchmark first").
#include "kfs_common.cuh"
#include <string.h>

static thread_local char g_err[512] = "";

void kfs_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" {

int kfs_version(void) { return 100; } /* 0.1.0 */

const char *kfs_last_error(void) { return g_err; }

int kfs_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

} // extern "C"

// ---- pipe-rate probes -----------------------------------------------------------------------------
// Each thread runs ILP independent chains; the chain op is either POPC (what=0) or LOP3 (what=1).
// ops counted = threads * iters * ILP; other instructions (one LOP/IADD per POPC to keep the chain
// data-dependent and un-hoistable) are NOT counted, so the figure is the rate of the probed op under
// the issue mix a real XOR+POPC+ADD loop has.
template <int WHAT>
__global__ void __launch_bounds__(256) k_pipe_probe(uint32_t *out, int iters, uint32_t seed)
{
    constexpr int ILP = 8;
    uint32_t x[ILP], acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) {
        x[i] = seed * (threadIdx.x + 1u) + 0x9e3779b9u * (i + 1u) + blockIdx.x;
        acc[i] = 0;
    }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (WHAT == 0) {
                acc[i] += __popc(x[i]);          // POPC + IADD
                x[i] ^= acc[i];                  // LOP  (keeps the next POPC dependent on this one)
            } else {
                // two dependent 3-input logic ops per step; counted as 2 LOP3
                uint32_t t;
                asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(t) : "r"(x[i]), "r"(acc[i]), "r"(seed));
                asm volatile("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(acc[i]) : "r"(t), "r"(x[i]), "r"(acc[i]));
                x[i] = t;
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) r += acc[i] ^ x[i];
    if (r == 0x12345678u) out[0] = r; // never true in practice; defeats dead-code elimination
}

extern "C" kfs_status kfs_pipe_peak(int device, int what, double *gops)
{
    if (!gops || what < 0 || what > 1) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_pipe_peak: bad arguments");
    if (kfs_device_count() <= device || device < 0) KFS_FAIL(KFS_ERR_NO_DEVICE, "kfs_pipe_peak: no CUDA device %d", device);
    KFS_DEVICE_GUARD(device);
    cudaDeviceProp prop;
    KFS_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    uint32_t *d = nullptr;
    KFS_CUDA_TRY(cudaMalloc(&d, 64));
    cudaEvent_t e0, e1;
    KFS_CUDA_TRY(cudaEventCreate(&e0));
    KFS_CUDA_TRY(cudaEventCreate(&e1));
    const int iters = 4096, blocks = prop.multiProcessorCount * 8, threads = 256;
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        KFS_CUDA_TRY(cudaEventRecord(e0));
        if (what == 0) k_pipe_probe<0><<<blocks, threads>>>(d, iters, 12345u + rep);
        else k_pipe_probe<1><<<blocks, threads>>>(d, iters, 12345u + rep);
        KFS_CUDA_TRY(cudaEventRecord(e1));
        KFS_CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0;
        KFS_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        double ops = (double)blocks * threads * iters * 8.0 * (what == 1 ? 2.0 : 1.0);
        double g = ops / (ms * 1e-3) / 1e9;
        if (rep > 0 && g > best) best = g;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *gops = best;
    return KFS_OK;
}

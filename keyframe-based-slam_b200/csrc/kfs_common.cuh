// kfs_common.cuh — error plumbing shared by the translation units of libkfslam_frontend.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/kfslam_frontend.h"

void kfs_set_error(const char *fmt, ...);

#define KFS_CUDA_TRY(expr)                                                                         \
    do {                                                                                           \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            kfs_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__));  \
            return KFS_ERR_CUDA;                                                                   \
        }                                                                                          \
    } while (0)

#define KFS_FAIL(code, ...)            \
    do {                               \
        kfs_set_error(__VA_ARGS__);    \
        return (code);                 \
    } while (0)

static inline int kfs_div_up(int a, int b) { return (a + b - 1) / b; }
static inline size_t kfs_align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// Selects `device` for the calling thread and remembers the previous one.
struct KfsDeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit KfsDeviceGuard(int device)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; }
        if (prev != device) ok = (cudaSetDevice(device) == cudaSuccess);
    }
    ~KfsDeviceGuard()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// ---- library-internal interfaces between the translation units (not part of the C ABI) -----------------
// One (query, train) matching problem as the kernels see it (device pointers).
struct BfPairDev {
    const uint8_t *query; const uint8_t *train;
    const int *d_nq; const int *d_nt;
    kfs_dmatch *out; int *counts;
    int nq, nt;
};
kfs_status kfs_orb_enqueue_host_batch(kfs_orb *o, const uint8_t *gray, size_t step, size_t frame_stride, int w, int h, int batch,
                                      const uint8_t *mask, size_t mask_step, size_t mask_frame_stride,
                                      kfs_keypoint *kp_out, uint8_t *desc_out, int capacity);
kfs_status kfs_orb_finish_host_batch(kfs_orb *o, int batch, int *n_out);
void kfs_orb_internal_streams(kfs_orb *o, cudaStream_t *s_in, cudaStream_t *s_out);
void kfs_orb_internal_start_fence(kfs_orb *o, bool on);
kfs_status kfs_bf_launch_table(kfs_bf *b, const BfPairDev *d_table, int npairs, int max_nq, int k, cudaStream_t st);
const BfPairDev *kfs_bf_device_table(const kfs_bf *b); // device copy of the pair table of the last kfs_bf_knn_match_pairs_device
kfs_status kfs_bf_launch_ratio_table(kfs_bf *b, const BfPairDev *d_table, int npairs, float ratio, kfs_dmatch *good_base, long long good_stride,
                                     int *n_good, cudaStream_t st);

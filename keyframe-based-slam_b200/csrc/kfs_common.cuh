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

// Every entry point that returns a status selects its handle's device through this: a failed cudaSetDevice is an error, not a
// silent run on whatever device happens to be current.
#define KFS_DEVICE_GUARD(dev)                                                             \
    KfsDeviceGuard guard(dev);                                                            \
    if (!guard.ok) KFS_FAIL(KFS_ERR_CUDA, "cudaSetDevice(%d) failed", (int)(dev))

// ---- library-internal interfaces between the translation units (not part of the C ABI) -----------------
// One (query, train) matching problem as the kernels see it (device pointers).
struct BfPairDev {
    const uint8_t *query; const uint8_t *train;
    const int *d_nq; const int *d_nt;
    kfs_dmatch *out; int *counts;
    int nq, nt;
};
// One batch of host frames through the ORB handle's three-stream pipeline (H2D | kernel chain | D2H).
struct KfsHostBatch {
    const uint8_t *image; size_t step, frame_stride; int w, h, channels, batch; // channels 1 (gray) or 3 (R,G,B bytes: converted on the device)
    const uint8_t *mask; size_t mask_step, mask_frame_stride;                    // nullptr = cv::noArray()
    kfs_keypoint *kp_out; uint8_t *desc_out; int capacity;                       // [batch][capacity] host rows
    int out_rows;              // rows per frame copied right behind the kernels (<= capacity; 0 = capacity): nfeatures covers every
                               // frame whose selection thresholds had no ties; the rest of a longer frame is fetched by finish()
    const uint8_t *skip_rows;  // [batch] or nullptr: non-zero = this frame's keypoints / descriptors are not wanted on the host
};
kfs_status kfs_orb_enqueue_host_batch(kfs_orb *o, const KfsHostBatch &hb);
const uint8_t *kfs_orb_alias_desc(const kfs_orb *o, const uint8_t *host, int rows, int device);
kfs_status kfs_launch_rgb2gray(const uint8_t *src, long long sstep, long long sfs, uint8_t *dst, int dpitch, long long dfs, int w, int h, int nframes,
                               cudaStream_t st);
kfs_status kfs_orb_finish_host_batch(kfs_orb *o, int batch, int *n_out);
void kfs_orb_internal_streams(kfs_orb *o, cudaStream_t *s_in, cudaStream_t *s_out);
void kfs_orb_internal_start_fence(kfs_orb *o, bool on);
kfs_status kfs_bf_launch_table(kfs_bf *b, const BfPairDev *d_table, int npairs, int max_nq, int k, cudaStream_t st);
const BfPairDev *kfs_bf_device_table(const kfs_bf *b); // device copy of the pair table of the last kfs_bf_knn_match_pairs_device
kfs_status kfs_bf_launch_ratio_table(kfs_bf *b, const BfPairDev *d_table, int npairs, float ratio, kfs_dmatch *good_base, long long good_stride,
                                     int *n_good, cudaStream_t st);

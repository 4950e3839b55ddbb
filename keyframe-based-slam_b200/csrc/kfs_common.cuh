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

// kfs_math.cuh — per-pixel / per-keypoint arithmetic of the ORB front-end, written once as
// __host__ __device__ inline functions.  The CUDA kernels (kfs_orb.cu) call them on the device; the CPU
// test shim (tests/cpu_shim.cpp, g++ only) calls the very same source on the host so the bit-level
// contracts of SURVEY.md §8(a) (a1.3 FAST score, a1.5 Harris, a1.7 fastAtan2, a1.9 blur order, a1.10
// rBRIEF rounding) are checked against the oracle without a GPU.  No FMA contraction is allowed anywhere
// in here: every float op goes through an explicit _rn intrinsic (device) / volatile-free plain op
// compiled with -ffp-contract=off (host).
#pragma once
#include <stdint.h>
#if defined(__CUDACC__)
#define KFS_HD __host__ __device__ __forceinline__
#else
#define KFS_HD static inline
#include <math.h>
#endif

// ---- float ops with pinned rounding / no contraction ------------------------------------------------
#if defined(__CUDA_ARCH__)
KFS_HD float kfs_mul(float a, float b) { return __fmul_rn(a, b); }
KFS_HD float kfs_add(float a, float b) { return __fadd_rn(a, b); }
KFS_HD float kfs_sub(float a, float b) { return __fsub_rn(a, b); }
KFS_HD float kfs_div(float a, float b) { return __fdiv_rn(a, b); }
KFS_HD float kfs_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
// round-half-even through the 1.5*2^23 magic add (exact for |a| < 2^22): keeps the XU pipe (F2I) out of the loops
KFS_HD int kfs_rint(float a) { return __float_as_int(__fadd_rn(a, 12582912.f)) - 0x4B400000; }
#else
KFS_HD float kfs_mul(float a, float b) { return a * b; }
KFS_HD float kfs_add(float a, float b) { return a + b; }
KFS_HD float kfs_sub(float a, float b) { return a - b; }
KFS_HD float kfs_div(float a, float b) { return a / b; }
KFS_HD float kfs_fma(float a, float b, float c) { return fmaf(a, b, c); }
KFS_HD int kfs_rint(float a) { return (int)lrintf(a); }
#endif

// ---- packed u16x2 min / max (VIMNMX[3].U16x2 on sm_100a) --------------------------------------------
#if defined(__CUDA_ARCH__)
KFS_HD uint32_t kfs_min2(uint32_t a, uint32_t b) { return __vminu2(a, b); }
KFS_HD uint32_t kfs_max2(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
KFS_HD uint32_t kfs_min3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
KFS_HD uint32_t kfs_max3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }
#else
KFS_HD uint32_t kfs_min2(uint32_t a, uint32_t b)
{
    uint32_t al = a & 0xffffu, bl = b & 0xffffu, ah = a >> 16, bh = b >> 16;
    return (al < bl ? al : bl) | ((ah < bh ? ah : bh) << 16);
}
KFS_HD uint32_t kfs_max2(uint32_t a, uint32_t b)
{
    uint32_t al = a & 0xffffu, bl = b & 0xffffu, ah = a >> 16, bh = b >> 16;
    return (al > bl ? al : bl) | ((ah > bh ? ah : bh) << 16);
}
KFS_HD uint32_t kfs_min3(uint32_t a, uint32_t b, uint32_t c) { return kfs_min2(kfs_min2(a, b), c); }
KFS_HD uint32_t kfs_max3(uint32_t a, uint32_t b, uint32_t c) { return kfs_max2(kfs_max2(a, b), c); }
#endif

// ---- a1.3 FAST-9/16 ---------------------------------------------------------------------------------
// Ring order (dx,dy), k = 0..15:
//  (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
//
// Both polarities are evaluated at once in one u16x2 word per ring pixel:
//   E_k = (v - p_k + 256) | (p_k - v + 256) << 16  = C_v + p_k * 65535,  C_v = (v + 256) + (256 - v) * 65536
// (each half stays in [1, 511], so the single 32-bit multiply-add never borrows across halves).
// m = max over the 16 arcs of 9 contiguous k of min over the arc of E, taken per half, then max of the two
// halves, minus the bias.  Corner iff m > threshold; OpenCV's score is m - 1.
KFS_HD uint32_t kfs_fast_cv(int v) { return (uint32_t)(v + 256) + ((uint32_t)(256 - v) << 16); }
KFS_HD uint32_t kfs_fast_e(uint32_t cv, int p) { return cv + (uint32_t)p * 65535u; }

// Arc k = ring positions k .. k+8 = min3(t[k], t[k+3], t[k+6]) with t[k] = min3(e[k], e[k+1], e[k+2]).  Arcs k and k+3 share
// t[k+3] and t[k+6], and max(min(X, c), min(X, d)) = min(X, max(c, d)): their maximum is min3(t[k+3], t[k+6], max(t[k], t[k+9]))
// — two operations for two arcs.  k -> k+3 walks all 16 arcs in one cycle, so they pair up perfectly: 16 + 8 x 2 + 4 = 36
// three-input operations instead of 16 + 16 + 8 = 40.
KFS_HD int kfs_fast9_m(const uint32_t e[16])
{
    uint32_t t[16], best = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) t[k] = kfs_min3(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        const int k0 = (6 * j) & 15, k1 = (6 * (j + 1)) & 15; // pairs (k, k+3) for k = 0, 6, 12, 2, 8, 14, 4, 10
        const uint32_t p0 = kfs_min3(t[(k0 + 3) & 15], t[(k0 + 6) & 15], kfs_max2(t[k0], t[(k0 + 9) & 15]));
        const uint32_t p1 = kfs_min3(t[(k1 + 3) & 15], t[(k1 + 6) & 15], kfs_max2(t[k1], t[(k1 + 9) & 15]));
        best = kfs_max3(best, p0, p1);
    }
    uint32_t lo = best & 0xffffu, hi = best >> 16;
    return (int)(lo > hi ? lo : hi) - 256;
}

// Necessary condition on two horizontally adjacent pixels at once (halves = pixel a, pixel b):
// a 9-arc contains at least one pixel of every opposite pair, so a bright (dark) corner needs
// max(p0,p8) > v+t and max(p4,p12) > v+t   (min(p0,p8) < v-t and min(p4,p12) < v-t).
// Inputs are u16x2 words holding the two pixels' values (0..255 per half).  Returns a word whose bit 15
// (bit 31) is set iff pixel a (b) MAY be a corner.
KFS_HD uint32_t kfs_fast_quick2m(uint32_t V, uint32_t U, uint32_t D, uint32_t R, uint32_t L, int thr, uint32_t mask)
{
    // per half: q + (0x8000 - t - 1) - v  >= 0x8000  <=>  q > v + t ;  v + (0x8000 - t - 1) - r >= 0x8000  <=>  r < v - t
    // (each half stays in [0x8000 - 511, 0x8000 + 254]: no borrow/carry crosses the halves)
    const uint32_t K = 0x80008000u - (uint32_t)(thr + 1) * 0x10001u;
    uint32_t q = kfs_min2(kfs_max2(U, D), kfs_max2(R, L));   // bright side
    uint32_t r = kfs_max2(kfs_min2(U, D), kfs_min2(R, L));   // dark side
    uint32_t X = q + K - V;
    uint32_t Y = V + K - r;
    return (X | Y) & mask;   // mask: subset of 0x80008000 (pixel validity), folded into the same LOP3
}
KFS_HD uint32_t kfs_fast_quick2(uint32_t V, uint32_t U, uint32_t D, uint32_t R, uint32_t L, int thr)
{
    return kfs_fast_quick2m(V, U, D, R, L, thr, 0x80008000u);
}

// ---- a1.5 Harris response from the integer sums -----------------------------------------------------
KFS_HD float kfs_harris_response(int a, int b, int c)
{
    const float scale = 1.f / (4 * 7 * 255.f);
    const float s2 = scale * scale;
    const float s3 = s2 * scale;
    const float s4 = s3 * scale;            // constant-folded in IEEE fp32, same as ((s*s)*s)*s at run time
    float fa = (float)a, fb = (float)b, fc = (float)c;
    float t0 = kfs_mul(fa, fb);
    float t1 = kfs_mul(fc, fc);
    float sum = kfs_add(fa, fb);
    float t2 = kfs_mul(kfs_mul(0.04f, sum), sum);
    return kfs_mul(kfs_sub(kfs_sub(t0, t1), t2), s4);
}

// ---- a1.7 OpenCV fastAtan2 (scalar atan_f32 path), degrees in [0,360) -------------------------------
KFS_HD float kfs_fast_atan2(float y, float x)
{
    const float k = (float)(180 / 3.1415926535897932384626433832795);
    const float p1 = 0.9997878412794807f * k, p3 = -0.3258083974640975f * k;
    const float p5 = 0.1555786518463281f * k, p7 = -0.04432655554792128f * k;
    const float eps = 2.2204460492503131e-16f; // (float)DBL_EPSILON
    float ax = x < 0 ? -x : x, ay = y < 0 ? -y : y, a, c, c2;
    if (ax >= ay) {
        c = kfs_div(ay, kfs_add(ax, eps));
        c2 = kfs_mul(c, c);
        a = kfs_mul(kfs_add(kfs_mul(kfs_add(kfs_mul(kfs_add(kfs_mul(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = kfs_div(ax, kfs_add(ay, eps));
        c2 = kfs_mul(c, c);
        a = kfs_sub(90.f, kfs_mul(kfs_add(kfs_mul(kfs_add(kfs_mul(kfs_add(kfs_mul(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = kfs_sub(180.f, a);
    if (y < 0) a = kfs_sub(360.f, a);
    return a;
}

// ---- a1.9 7x7 sigma=2 blur taps (getGaussianKernel(7, 2, CV_32F)) -----------------------------------
#define KFS_G0 0x1.1f5f62p-4f
#define KFS_G1 0x1.0c70fcp-3f
#define KFS_G2 0x1.869472p-3f
#define KFS_G3 0x1.ba95c0p-3f

// row pass: r = g0*p0; r = fma(g_k, p_k, r), k = 1..6, left to right
KFS_HD float kfs_blur_row(float p0, float p1, float p2, float p3, float p4, float p5, float p6)
{
    float r = kfs_mul(KFS_G0, p0);
    r = kfs_fma(KFS_G1, p1, r);
    r = kfs_fma(KFS_G2, p2, r);
    r = kfs_fma(KFS_G3, p3, r);
    r = kfs_fma(KFS_G2, p4, r);
    r = kfs_fma(KFS_G1, p5, r);
    r = kfs_fma(KFS_G0, p6, r);
    return r;
}
// column pass: s = g3*r0; s = fma(g_{3+k}, r_{+k} + r_{-k}, s), k = 1..3; round half even; saturate
KFS_HD int kfs_blur_col(float rm3, float rm2, float rm1, float r0, float rp1, float rp2, float rp3)
{
    float s = kfs_mul(KFS_G3, r0);
    s = kfs_fma(KFS_G2, kfs_add(rp1, rm1), s);
    s = kfs_fma(KFS_G1, kfs_add(rp2, rm2), s);
    s = kfs_fma(KFS_G0, kfs_add(rp3, rm3), s);
    int v = kfs_rint(s);
    return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// same, result left as the bits of (s + 1.5*2^23): byte 0 is the rounded pixel (s is a convex combination of u8 values,
// so 0 <= rint(s) <= 255 and no saturation can trigger)
KFS_HD uint32_t kfs_blur_col_bits(float rm3, float rm2, float rm1, float r0, float rp1, float rp2, float rp3)
{
    float s = kfs_mul(KFS_G3, r0);
    s = kfs_fma(KFS_G2, kfs_add(rp1, rm1), s);
    s = kfs_fma(KFS_G1, kfs_add(rp2, rm2), s);
    s = kfs_fma(KFS_G0, kfs_add(rp3, rm3), s);
    s = kfs_add(s, 12582912.f);
#if defined(__CUDA_ARCH__)
    return __float_as_uint(s);
#else
    union { float f; uint32_t u; } c; c.f = s; return c.u;
#endif
}

// ---- a1.10 rotated BRIEF sample offset ----------------------------------------------------------------
// (x,y) pattern point, a = cos, b = sin: X = x*a - y*b, Y = x*b + y*a (mul, mul, sub/add; no FMA), rint.
KFS_HD void kfs_brief_rot(int px, int py, float a, float b, int *ix, int *iy)
{
    float fx = (float)px, fy = (float)py;
    *ix = kfs_rint(kfs_sub(kfs_mul(fx, a), kfs_mul(fy, b)));
    *iy = kfs_rint(kfs_add(kfs_mul(fx, b), kfs_mul(fy, a)));
}

KFS_HD void kfs_brief_rotf(float fx, float fy, float a, float b, int *ix, int *iy)
{
    *ix = kfs_rint(kfs_sub(kfs_mul(fx, a), kfs_mul(fy, b)));
    *iy = kfs_rint(kfs_add(kfs_mul(fx, b), kfs_mul(fy, a)));
}

// order-preserving map fp32 -> u32 (for radix select on the Harris response, a1.6)
KFS_HD uint32_t kfs_float_key(float f)
{
#if defined(__CUDA_ARCH__)
    uint32_t u = __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; uint32_t u = c.u;
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
#if defined(__CUDACC__)
__device__ __forceinline__ float kfs_key_float(uint32_t k) // inverse of kfs_float_key (bit-exact, -0.f and NaN payloads included)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
#endif

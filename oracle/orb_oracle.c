/*
 * oracle/orb_oracle.c — TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * Plain-C CPU restatement of the arithmetic behind the reference's feature-extraction call
 *     this->ptrExtractor->detectAndCompute(gray, mask, kpCurrImg, dpCurrImg);
 *         /root/reference/src/lib/VisualOdometry.cpp:30   (ORB built at :10-16)
 * The reference holds no arithmetic of its own for this path: the call goes into OpenCV
 * (find_package(OpenCV) at /root/reference/CMakeLists.txt:19, unpinned, not vendored).  The algorithm
 * restated here is OpenCV's published ORB (features2d orb.cpp / fast.cpp / keypoint.cpp, imgproc
 * resize INTER_LINEAR_EXACT and sepFilter2D) as pinned stage by stage in SURVEY.md §8(a) rows
 * a1.1–a1.10, and it is PINNED against opencv-python-headless 4.13.0.92 (the same C++ code the
 * reference links): tests/test_oracle_golden.py checks it against golden vectors that
 * tools/make_golden.py generated from cv2 itself, and tests/test_oracle_vs_cv2.py re-checks live
 * wherever cv2 is importable.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call
 * this file.  Nothing under keyframe-based-slam_b200/ links or imports it.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -march=x86-64-v3; fmaf() must be a real FMA
 * because OpenCV's AVX2 sepFilter2D path is, see a1.9).
 */
#include <math.h>
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define KFO_MAX_LEVELS 16

typedef struct {
    int nfeatures;
    float scale_factor;
    int nlevels;
    int edge_threshold;
    int first_level;
    int wta_k;
    int score_type; /* 0 = HARRIS_SCORE */
    int patch_size;
    int fast_threshold;
} kfo_orb_params;

typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} kfo_keypoint; /* 28 bytes == cv::KeyPoint */

static const int kfo_pattern[1024] = {
#include "../keyframe-based-slam_b200/csrc/orb_pattern.inc"
};

/* cvRound on a double: round-half-even (SSE2 cvtsd2si under the default rounding mode). */
static inline int kfo_cvround(double v) { return (int)lrint(v); }

/* ---- a1.1 geometry: per-level scale, size, quota (SURVEY §8 header + a1.1) ------------------- */
float kfo_level_scale(float scale_factor, int level, int first_level)
{
    return (float)pow((double)scale_factor, (double)(level - first_level));
}

void kfo_level_size(int w, int h, float scale_factor, int level, int *lw, int *lh)
{
    float inv = 1.f / kfo_level_scale(scale_factor, level, 0);
    float fw = (float)w * inv, fh = (float)h * inv;
    *lw = kfo_cvround((double)fw);
    *lh = kfo_cvround((double)fh);
}

void kfo_level_quotas(int nfeatures, float scale_factor, int nlevels, int *quota)
{
    float factor = (float)(1.0 / scale_factor);
    float ndes = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        quota[l] = kfo_cvround((double)ndes);
        sum += quota[l];
        ndes *= factor;
    }
    int last = nfeatures - sum;
    quota[nlevels - 1] = last > 0 ? last : 0;
}

/* ---- a1.1 resize INTER_LINEAR_EXACT (8.8 fixed point, chained level to level) ---------------- */
static void kfo_axis_table(int src, int dst, int *ofs, int *m1)
{
    double scale = 1.0 / ((double)dst / (double)src);
    for (int d = 0; d < dst; d++) {
        double f = scale * ((double)d + 0.5) - 0.5;
        int i = (int)floor(f);
        if (i < 0 || src <= 1) { ofs[d] = 0; m1[d] = 0; }
        else if (i >= src - 1) { ofs[d] = src - 1; m1[d] = 0; }
        else { ofs[d] = i; m1[d] = (int)lrint((f - (double)i) * 256.0); }
    }
}

void kfo_resize_linear_exact(const uint8_t *src, int sw, int sh, size_t sstep,
                             uint8_t *dst, int dw, int dh, size_t dstep)
{
    int *xo = (int *)malloc(sizeof(int) * 2 * (size_t)dw);
    int *xm = xo + dw;
    int *yo = (int *)malloc(sizeof(int) * 2 * (size_t)dh);
    int *ym = yo + dh;
    kfo_axis_table(sw, dw, xo, xm);
    kfo_axis_table(sh, dh, yo, ym);
    for (int y = 0; y < dh; y++) {
        const uint8_t *r0 = src + (size_t)yo[y] * sstep;
        const uint8_t *r1 = src + (size_t)(yo[y] + 1 < sh ? yo[y] + 1 : sh - 1) * sstep;
        int n1 = ym[y], n0 = 256 - n1;
        uint8_t *o = dst + (size_t)y * dstep;
        for (int x = 0; x < dw; x++) {
            int a = xo[x], b = a + 1 < sw ? a + 1 : sw - 1;
            int k1 = xm[x], k0 = 256 - k1;
            uint32_t h0 = (uint32_t)(k0 * r0[a] + k1 * r0[b]);
            uint32_t h1 = (uint32_t)(k0 * r1[a] + k1 * r1[b]);
            o[x] = (uint8_t)((n0 * h0 + n1 * h1 + 32768u) >> 16);
        }
    }
    free(xo);
    free(yo);
}

/* a1.2 mask pyramid step: resize then threshold(254, TOZERO) i.e. keep only 255. */
void kfo_mask_binarise(const uint8_t *src, int w, int h, size_t sstep, uint8_t *dst, size_t dstep)
{
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) dst[y * dstep + x] = src[y * sstep + x] ? 255 : 0;
}
void kfo_mask_tozero254(uint8_t *m, int w, int h, size_t step)
{
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) if (m[y * step + x] <= 254) m[y * step + x] = 0;
}

/* ---- a1.3 FAST-9/16 score + 3x3 NMS ----------------------------------------------------------- */
static const int kfo_circle[16][2] = {
    {0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
    {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

/* m = max over the 16 arcs of 9 contiguous ring pixels of max(min d, min -d); corner iff m > thr. */
int kfo_fast9_m(const uint8_t *p, const int *off)
{
    int d[25];
    int v = p[0];
    for (int k = 0; k < 16; k++) d[k] = v - p[off[k]];
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int best = -256;
    for (int s = 0; s < 16; s++) {
        int mn = d[s], mx = d[s];
        for (int k = 1; k < 9; k++) {
            if (d[s + k] < mn) mn = d[s + k];
            if (d[s + k] > mx) mx = d[s + k];
        }
        if (mn > best) best = mn;
        if (-mx > best) best = -mx;
    }
    return best;
}

/* Dense score map: score[y][x] = m-1 if m > thr else 0 (x in [3,w-4], y in [3,h-4]). */
void kfo_fast9_scoremap(const uint8_t *img, int w, int h, size_t step, int thr, uint8_t *score, size_t sstep)
{
    int off[16];
    for (int k = 0; k < 16; k++) off[k] = kfo_circle[k][1] * (int)step + kfo_circle[k][0];
    for (int y = 0; y < h; y++) memset(score + y * sstep, 0, (size_t)w);
    for (int y = 3; y < h - 3; y++) {
        const uint8_t *row = img + (size_t)y * step;
        for (int x = 3; x < w - 3; x++) {
            const uint8_t *p = row + x;
            int v = p[0], hi = v + thr, lo = v - thr;
            /* cheap necessary condition: every opposite pair must hold one arc member */
            int a = p[off[0]], b = p[off[8]], c = p[off[4]], e = p[off[12]];
            int bright = (a > hi || b > hi) && (c > hi || e > hi);
            int dark = (a < lo || b < lo) && (c < lo || e < lo);
            if (!bright && !dark) continue;
            int m = kfo_fast9_m(p, off);
            if (m > thr) score[y * sstep + x] = (uint8_t)(m - 1);
        }
    }
}

/* FAST with nms=true: emits (x,y,score) row-major. Returns count (may exceed cap; only cap written). */
int kfo_fast9_detect(const uint8_t *img, int w, int h, size_t step, int thr,
                     int *xs, int *ys, int *scores, int cap)
{
    uint8_t *sc = (uint8_t *)malloc((size_t)w * (size_t)h);
    kfo_fast9_scoremap(img, w, h, step, thr, sc, (size_t)w);
    int n = 0;
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            int s = sc[y * w + x];
            if (!s) continue;
            const uint8_t *q = sc + y * w + x;
            if (s > q[-1] && s > q[1] && s > q[-w - 1] && s > q[-w] && s > q[-w + 1] &&
                s > q[w - 1] && s > q[w] && s > q[w + 1]) {
                if (n < cap) { xs[n] = x; ys[n] = y; scores[n] = s; }
                n++;
            }
        }
    free(sc);
    return n;
}

/* ---- a1.5 Harris response (block 7, k 0.04) on the un-blurred level -------------------------- */
float kfo_harris(const uint8_t *img, size_t step, int x, int y)
{
    const int bs = 7, r = 3;
    int a = 0, b = 0, c = 0;
    for (int i = 0; i < bs; i++)
        for (int j = 0; j < bs; j++) {
            const uint8_t *p = img + (size_t)(y - r + i) * step + (x - r + j);
            int s = (int)step;
            int Ix = (p[1] - p[-1]) * 2 + (p[-s + 1] - p[-s - 1]) + (p[s + 1] - p[s - 1]);
            int Iy = (p[s] - p[-s]) * 2 + (p[s - 1] - p[-s - 1]) + (p[s + 1] - p[-s + 1]);
            a += Ix * Ix;
            b += Iy * Iy;
            c += Ix * Iy;
        }
    float scale = 1.f / ((1 << 2) * bs * 255.f);
    float s4 = scale * scale * scale * scale;
    float fa = (float)a, fb = (float)b, fc = (float)c;
    float t0 = fa * fb;
    float t1 = fc * fc;
    float sum = fa + fb;
    float t2 = 0.04f * sum;
    t2 = t2 * sum;
    float r0 = t0 - t1;
    r0 = r0 - t2;
    return r0 * s4;
}

/* ---- a1.7 intensity-centroid angle ----------------------------------------------------------- */
static const int kfo_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

float kfo_fast_atan2(float y, float x)
{
    const float k = (float)(180 / 3.1415926535897932384626433832795);
    const float p1 = 0.9997878412794807f * k, p3 = -0.3258083974640975f * k;
    const float p5 = 0.1555786518463281f * k, p7 = -0.04432655554792128f * k;
    float ax = fabsf(x), ay = fabsf(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

void kfo_ic_moments(const uint8_t *img, size_t step, int x, int y, int *m10, int *m01)
{
    const uint8_t *c = img + (size_t)y * step + x;
    int s = (int)step, a10 = 0, a01 = 0;
    for (int u = -15; u <= 15; u++) a10 += u * c[u];
    for (int v = 1; v <= 15; v++) {
        int vs = 0, d = kfo_umax[v];
        for (int u = -d; u <= d; u++) {
            int vp = c[u + v * s], vm = c[u - v * s];
            vs += vp - vm;
            a10 += u * (vp + vm);
        }
        a01 += v * vs;
    }
    *m10 = a10;
    *m01 = a01;
}

float kfo_ic_angle(const uint8_t *img, size_t step, int x, int y)
{
    int m10, m01;
    kfo_ic_moments(img, step, x, y, &m10, &m01);
    return kfo_fast_atan2((float)m01, (float)m10);
}

/* ---- a1.9 7x7 sigma=2 blur, sepFilter2D float path with OpenCV's FMA order, REFLECT_101 ------ */
static const float kfo_g[7] = {0x1.1f5f62p-4f, 0x1.0c70fcp-3f, 0x1.869472p-3f, 0x1.ba95c0p-3f,
                               0x1.869472p-3f, 0x1.0c70fcp-3f, 0x1.1f5f62p-4f};

static inline int kfo_reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * n - 2 - i;
    }
    return i;
}

/* use_fma = 1: OpenCV AVX2/AVX-512 dispatch (default); 0: separate mul+add (non-FMA host). */
void kfo_blur7(const uint8_t *src, int w, int h, size_t sstep, uint8_t *dst, size_t dstep, int use_fma)
{
    float *rows = (float *)malloc(sizeof(float) * (size_t)w * (size_t)h);
    for (int y = 0; y < h; y++) {
        const uint8_t *s = src + (size_t)y * sstep;
        float *r = rows + (size_t)y * w;
        for (int x = 0; x < w; x++) {
            float acc = kfo_g[0] * (float)s[kfo_reflect101(x - 3, w)];
            for (int k = 1; k < 7; k++) {
                float p = (float)s[kfo_reflect101(x - 3 + k, w)];
                if (use_fma) acc = fmaf(kfo_g[k], p, acc);
                else { float t = kfo_g[k] * p; acc = acc + t; }
            }
            r[x] = acc;
        }
    }
    for (int y = 0; y < h; y++) {
        uint8_t *o = dst + (size_t)y * dstep;
        const float *r0 = rows + (size_t)y * w;
        for (int x = 0; x < w; x++) {
            float acc = kfo_g[3] * r0[x];
            for (int k = 1; k <= 3; k++) {
                float up = rows[(size_t)kfo_reflect101(y + k, h) * w + x];
                float dn = rows[(size_t)kfo_reflect101(y - k, h) * w + x];
                float sm = up + dn;
                if (use_fma) acc = fmaf(kfo_g[3 + k], sm, acc);
                else { float t = kfo_g[3 + k] * sm; acc = acc + t; }
            }
            long v = lrintf(acc);
            o[x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
    free(rows);
}

/* ---- a1.10 rotated BRIEF-256 on the blurred level -------------------------------------------- */
void kfo_brief(const uint8_t *blurred, size_t step, int cx, int cy, float angle_deg, uint8_t *desc)
{
    float ang = angle_deg * (float)(3.1415926535897932384626433832795 / 180.f);
    float a = (float)cos((double)ang), b = (float)sin((double)ang);
    const uint8_t *c = blurred + (size_t)cy * step + cx;
    for (int j = 0; j < 32; j++) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const int *pt = kfo_pattern + (j * 8 + k) * 4;
            float x0 = (float)pt[0] * a, t0 = (float)pt[1] * b;
            float y0 = (float)pt[0] * b, u0 = (float)pt[1] * a;
            float X0 = x0 - t0, Y0 = y0 + u0;
            float x1 = (float)pt[2] * a, t1 = (float)pt[3] * b;
            float y1 = (float)pt[2] * b, u1 = (float)pt[3] * a;
            float X1 = x1 - t1, Y1 = y1 + u1;
            int ix0 = (int)lrintf(X0), iy0 = (int)lrintf(Y0);
            int ix1 = (int)lrintf(X1), iy1 = (int)lrintf(Y1);
            int v0 = c[iy0 * (int)step + ix0], v1 = c[iy1 * (int)step + ix1];
            val |= (v0 < v1) << k;
        }
        desc[j] = (uint8_t)val;
    }
}

/* ---- selection (a1.4, a1.6): keep everything >= the n-th largest (ties kept) ----------------- */
static int kfo_cmp_desc_f(const void *a, const void *b)
{
    float x = *(const float *)a, y = *(const float *)b;
    return (x < y) - (x > y);
}

static float kfo_nth_largest(const float *v, int n, int nth /* 1-based */)
{
    float *t = (float *)malloc(sizeof(float) * (size_t)n);
    memcpy(t, v, sizeof(float) * (size_t)n);
    qsort(t, (size_t)n, sizeof(float), kfo_cmp_desc_f);
    float r = t[nth - 1];
    free(t);
    return r;
}

typedef struct { int x, y, lvl; float fast, harris, angle; } kfo_cand;

static int kfo_cmp_cand(const void *pa, const void *pb)
{
    const kfo_cand *a = (const kfo_cand *)pa, *b = (const kfo_cand *)pb;
    if (a->lvl != b->lvl) return a->lvl - b->lvl;
    if (a->y != b->y) return a->y - b->y;
    return a->x - b->x;
}

static int kfo_retain_best(kfo_cand *c, int n, int keep, int by_harris)
{
    if (keep <= 0) return 0;
    if (n <= keep) return n;
    float *v = (float *)malloc(sizeof(float) * (size_t)n);
    for (int i = 0; i < n; i++) v[i] = by_harris ? c[i].harris : c[i].fast;
    float thr = kfo_nth_largest(v, n, keep);
    int m = 0;
    for (int i = 0; i < n; i++)
        if (v[i] >= thr) c[m++] = c[i];
    free(v);
    return m;
}

/* ---- full pipeline ---------------------------------------------------------------------------- */
/* Optional stage taps for tests: if pyr_out != NULL, pyr_out[l] must hold w_l*h_l bytes (tight). */
int kfo_orb_detect_and_compute_ex(const kfo_orb_params *P, const uint8_t *gray, size_t step, int w, int h,
                                  const uint8_t *mask, size_t mask_step,
                                  kfo_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out,
                                  uint8_t **pyr_out, int blur_fma)
{
    *n_out = 0;
    if (!P || P->nlevels < 1 || P->nlevels > KFO_MAX_LEVELS || P->wta_k != 2 || P->score_type != 0 ||
        P->patch_size != 31 || P->first_level != 0 || P->edge_threshold != 31)
        return -1;
    if (w <= 0 || h <= 0 || !gray) return 0;
    int L = P->nlevels, lw[KFO_MAX_LEVELS], lh[KFO_MAX_LEVELS], quota[KFO_MAX_LEVELS];
    float sc[KFO_MAX_LEVELS];
    uint8_t *img[KFO_MAX_LEVELS], *msk[KFO_MAX_LEVELS];
    kfo_level_quotas(P->nfeatures, P->scale_factor, L, quota);
    for (int l = 0; l < L; l++) {
        sc[l] = kfo_level_scale(P->scale_factor, l, 0);
        kfo_level_size(w, h, P->scale_factor, l, &lw[l], &lh[l]);
        if (lw[l] < 1 || lh[l] < 1) { L = l; break; }
        img[l] = (uint8_t *)malloc((size_t)lw[l] * lh[l]);
        msk[l] = mask ? (uint8_t *)malloc((size_t)lw[l] * lh[l]) : NULL;
        if (l == 0) {
            for (int y = 0; y < h; y++) memcpy(img[0] + (size_t)y * w, gray + (size_t)y * step, (size_t)w);
            if (mask) kfo_mask_binarise(mask, w, h, mask_step, msk[0], (size_t)w);
        } else {
            kfo_resize_linear_exact(img[l - 1], lw[l - 1], lh[l - 1], (size_t)lw[l - 1], img[l], lw[l], lh[l], (size_t)lw[l]);
            if (mask) {
                kfo_resize_linear_exact(msk[l - 1], lw[l - 1], lh[l - 1], (size_t)lw[l - 1], msk[l], lw[l], lh[l], (size_t)lw[l]);
                kfo_mask_tozero254(msk[l], lw[l], lh[l], (size_t)lw[l]);
            }
        }
        if (pyr_out && pyr_out[l]) memcpy(pyr_out[l], img[l], (size_t)lw[l] * lh[l]);
    }

    int total_cap = 0, ntot = 0;
    kfo_cand *all = NULL;
    const int border = P->edge_threshold;
    for (int l = 0; l < L; l++) {
        int W = lw[l], H = lh[l];
        if (W <= 2 * border || H <= 2 * border || W < 7 || H < 7) continue;
        int cap = (W * H) / 4 + 16;
        int *xs = (int *)malloc(sizeof(int) * 3 * (size_t)cap), *ys = xs + cap, *ss = ys + cap;
        int n = kfo_fast9_detect(img[l], W, H, (size_t)W, P->fast_threshold, xs, ys, ss, cap);
        if (n > cap) n = cap;
        kfo_cand *c = (kfo_cand *)malloc(sizeof(kfo_cand) * (size_t)(n + 1));
        int m = 0;
        for (int i = 0; i < n; i++) {
            int x = xs[i], y = ys[i];
            if (msk[l] && msk[l][(size_t)y * W + x] == 0) continue;                     /* runByPixelsMask */
            if (!(x >= border && x < W - border && y >= border && y < H - border)) continue; /* runByImageBorder */
            c[m].x = x; c[m].y = y; c[m].lvl = l; c[m].fast = (float)ss[i]; c[m].harris = 0; c[m].angle = 0;
            m++;
        }
        free(xs);
        m = kfo_retain_best(c, m, 2 * quota[l], 0);
        for (int i = 0; i < m; i++) c[i].harris = kfo_harris(img[l], (size_t)W, c[i].x, c[i].y);
        m = kfo_retain_best(c, m, quota[l], 1);
        for (int i = 0; i < m; i++) c[i].angle = kfo_ic_angle(img[l], (size_t)W, c[i].x, c[i].y);
        if (ntot + m > total_cap) {
            total_cap = (ntot + m) * 2 + 64;
            all = (kfo_cand *)realloc(all, sizeof(kfo_cand) * (size_t)total_cap);
        }
        memcpy(all + ntot, c, sizeof(kfo_cand) * (size_t)m);
        ntot += m;
        free(c);
    }
    if (ntot) qsort(all, (size_t)ntot, sizeof(kfo_cand), kfo_cmp_cand);

    /* descriptors: blur each level that holds keypoints, then rBRIEF */
    int idx = 0, rc = 0;
    for (int l = 0; l < L && idx < ntot; l++) {
        int first = idx;
        while (idx < ntot && all[idx].lvl == l) idx++;
        if (idx == first) continue;
        uint8_t *bl = (uint8_t *)malloc((size_t)lw[l] * lh[l]);
        kfo_blur7(img[l], lw[l], lh[l], (size_t)lw[l], bl, (size_t)lw[l], blur_fma);
        for (int i = first; i < idx; i++) {
            if (i >= capacity) { rc = 1; continue; }
            kfo_keypoint *k = kp_out + i;
            k->x = (float)all[i].x * sc[l];
            k->y = (float)all[i].y * sc[l];
            k->size = (float)P->patch_size * sc[l];
            k->angle = all[i].angle;
            k->response = all[i].harris;
            k->octave = l;
            k->class_id = -1;
            float inv = 1.f / sc[l];
            int cx = kfo_cvround((double)(k->x * inv)), cy = kfo_cvround((double)(k->y * inv));
            kfo_brief(bl, (size_t)lw[l], cx, cy, k->angle, desc_out + (size_t)i * 32);
        }
        free(bl);
    }
    *n_out = ntot;
    free(all);
    for (int l = 0; l < L; l++) { free(img[l]); free(msk[l]); }
    return rc; /* 1 = capacity too small (n_out holds the needed count) */
}

int kfo_orb_detect_and_compute(const kfo_orb_params *P, const uint8_t *gray, size_t step, int w, int h,
                               const uint8_t *mask, size_t mask_step,
                               kfo_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out)
{
    return kfo_orb_detect_and_compute_ex(P, gray, step, w, h, mask, mask_step, kp_out, desc_out, capacity, n_out, NULL, 1);
}

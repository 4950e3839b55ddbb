// tests/cpu_shim.cpp — compiles keyframe-based-slam_b200/csrc/kfs_math.cuh (the SAME source the CUDA kernels
// call on the device) with g++ so its bit-level contracts can be checked against the oracle on a CPU-only
// box.  Mirrors the per-pixel dataflow of k_fast / k_harris / k_orient / k_describe; it is test scaffolding,
// not a product path (nothing in the package loads it).
#include <stdint.h>
#include <string.h>
#include <math.h>
#include "../keyframe-based-slam_b200/csrc/kfs_math.cuh"

static const int8_t pattern[1024] = {
#include "../keyframe-based-slam_b200/csrc/orb_pattern.inc"
};
static const int RING[16][2] = {{0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
                                {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

extern "C" {

// score map through the kernel's two-step logic: quick2 on pixel pairs (x, x+2), then the exact packed score
// returns the number of pixels the quick test rejected
long shim_scoremap(const uint8_t *img, int w, int h, long step, int thr, uint8_t *out, long ostep, int use_quick)
{
    long rejected = 0;
    for (int y = 0; y < h; y++) memset(out + y * ostep, 0, (size_t)w);
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            const uint8_t *c = img + y * step + x;
            if (use_quick) {
                // second half of the pair is a dummy copy of the same pixel
                uint32_t V = c[0] | (c[0] << 16), U = c[-3 * step] | (c[-3 * step] << 16), D = c[3 * step] | (c[3 * step] << 16);
                uint32_t R = c[3] | (c[3] << 16), L = c[-3] | (c[-3] << 16);
                uint32_t m = kfs_fast_quick2(V, U, D, R, L, thr);
                if (!(m & 0x8000u)) { rejected++; continue; }
                if (((m >> 16) & 0x8000u) == 0) return -1; // halves must agree
            }
            uint32_t cv = kfs_fast_cv(c[0]), e[16];
            for (int k = 0; k < 16; k++) e[k] = kfs_fast_e(cv, c[RING[k][1] * step + RING[k][0]]);
            int m = kfs_fast9_m(e);
            if (m > thr) out[y * ostep + x] = (uint8_t)(m - 1);
        }
    return rejected;
}

float shim_harris(const uint8_t *img, long step, int x, int y)
{
    int a = 0, b = 0, c = 0;
    for (int i = -3; i <= 3; i++)
        for (int j = -3; j <= 3; j++) {
            const uint8_t *p = img + (y + i) * step + x + j;
            int Ix = (p[1] - p[-1]) * 2 + (p[-step + 1] - p[-step - 1]) + (p[step + 1] - p[step - 1]);
            int Iy = (p[step] - p[-step]) * 2 + (p[step - 1] - p[-step - 1]) + (p[step + 1] - p[-step + 1]);
            a += Ix * Ix; b += Iy * Iy; c += Ix * Iy;
        }
    return kfs_harris_response(a, b, c);
}

float shim_atan2(float y, float x) { return kfs_fast_atan2(y, x); }

// column-wise moments exactly as k_orient's lanes do
float shim_ic_angle(const uint8_t *img, long step, int x, int y)
{
    int m10 = 0, m01 = 0;
    for (int lane = 0; lane < 31; lane++) {
        int u = lane - 15, au = u < 0 ? -u : u;
        int d = (int)((0x3689ABCDDEEEFFFFull >> (4 * au)) & 15ull);
        int colsum = 0;
        for (int v = -d; v <= d; v++) { int I = img[(y + v) * step + x + u]; colsum += I; m01 += v * I; }
        m10 += u * colsum;
    }
    return kfs_fast_atan2((float)m01, (float)m10);
}

// 37x37 blurred patch around (cx, cy) the way k_describe builds it
void shim_blur_patch(const uint8_t *img, long step, int cx, int cy, uint8_t *out /*37*37*/)
{
    static float row[43][37];
    for (int r = 0; r < 43; r++)
        for (int c = 0; c < 37; c++) {
            const uint8_t *p = img + (cy - 21 + r) * step + (cx - 21 + c);
            row[r][c] = kfs_blur_row((float)p[0], (float)p[1], (float)p[2], (float)p[3], (float)p[4], (float)p[5], (float)p[6]);
        }
    for (int o = 0; o < 37; o++)
        for (int c = 0; c < 37; c++)
            out[o * 37 + c] = (uint8_t)kfs_blur_col(row[o][c], row[o + 1][c], row[o + 2][c], row[o + 3][c], row[o + 4][c], row[o + 5][c], row[o + 6][c]);
}

void shim_brief(const uint8_t *patch /*37*37 blurred, centre (18,18)*/, float angle_deg, uint8_t *desc)
{
    float rad = kfs_mul(angle_deg, (float)(3.1415926535897932384626433832795 / 180.f));
    float a = (float)cos((double)rad), b = (float)sin((double)rad);
    uint32_t words[8];
    for (int wd = 0; wd < 8; wd++) {
        uint32_t v = 0;
        for (int i = 0; i < 32; i++) {
            int pair = 32 * wd + i, ix0, iy0, ix1, iy1;
            kfs_brief_rot(pattern[4 * pair], pattern[4 * pair + 1], a, b, &ix0, &iy0);
            kfs_brief_rot(pattern[4 * pair + 2], pattern[4 * pair + 3], a, b, &ix1, &iy1);
            int v0 = patch[(18 + iy0) * 37 + 18 + ix0], v1 = patch[(18 + iy1) * 37 + 18 + ix1];
            v |= (uint32_t)(v0 < v1) << i;
        }
        words[wd] = v;
    }
    memcpy(desc, words, 32);
}

uint32_t shim_float_key(float f) { return kfs_float_key(f); }
}

// kfs_orb.cu — ORB detect-and-describe on sm_100a: the replacement for
//     ptrExtractor->detectAndCompute(gray, mask, kp, desc)   /root/reference/src/lib/VisualOdometry.cpp:30
// Stage contracts are SURVEY.md §8(a) rows a1.1–a1.10 (OpenCV's ORB, which the reference links but does
// not vendor).  Every kernel carries a leading batch dimension (blockIdx.y / .z = frame) so that one
// launch covers a whole batch of frames; nothing on this path ever goes back to the host for a count.
//
// Kernel chain per batch:
//   k_mask_binarise*  k_resize x(L-1) [x2 with a mask]  k_blur7 (side stream) || k_fast<62>  k_select_fast  k_harris
//   k_select_sort  k_describe (orientation + rBRIEF on the blurred levels)
// Calls of a few frames (the tracker's one frame per call): k_pyramid instead of the k_resize chain, k_fast<31>, programmatic
// dependent launch along the chain, the whole chain replayed as a CUDA graph.
#include "kfs_common.cuh"
#include "kfs_math.cuh"
#include <cuda.h> // CUtensorMap + the cuTensorMapEncodeTiled prototype only: the entry point is looked up at run time, libcuda is not linked
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <vector>
#include <algorithm>
#include <new>
#include <mutex>

#ifndef KFS_DK
#define KFS_DK 2
#endif
#ifndef KFS_KPW
#define KFS_KPW 8
#endif
#define KFS_MAX_LEVELS 16
#define KFS_BORDER 31

struct LevelDev {
    const uint8_t *img;      // frame f of this level at img + f*img_fs
    const uint8_t *mask;     // nullptr when no mask
    uint8_t *blur;           // 7x7-blurred level (a1.9), frame f at blur + f*img_fs_blur
    long long img_fs_blur;
    int pitch_blur;
    long long img_fs, mask_fs;
    int pitch, mask_pitch;
    int w, h, quota;
    float scale;
    int list_off, list_cap;  // this level's slice inside a frame's list segment
    int tiles_x, tiles_y, tile_base;
};

struct OrbDev {
    int nlevels, batch, fast_thr, cap_out, list_fs, total_tiles;
    int frame0;          // first frame of this launch (chunked host pipeline); frame = blockIdx + frame0
    uint32_t *raw;       // [batch][list_fs]  FAST corners after NMS/border/mask: score<<24 | y<<12 | x
    uint32_t *cand_xy;   // [batch][list_fs]  survivors of retainBest(2*quota) on the FAST score
    float *cand_resp;    // [batch][list_fs]  their Harris responses
    uint32_t *kept_xy;   // [batch][list_fs]  survivors of retainBest(quota) on Harris, sorted by (y,x)
    float *kept_resp;
    int *raw_count, *cand_count, *kept_count; // [batch][KFS_MAX_LEVELS]
    int *n_out;          // [batch]
    kfs_keypoint *kp;    // [batch][cap_out]
    uint8_t *desc;       // [batch][cap_out][32]
    float2 *rot;         // [batch][cap_out]  (cos, sin) of the keypoint angle
    uint32_t *slot_meta; // [batch][cap_out]  x | y<<12 | level<<24
    LevelDev lv[KFS_MAX_LEVELS];
};


// ======================================================================================================
// a1.1 / a1.2  pyramid: INTER_LINEAR_EXACT, 8.8 fixed point, chained level to level
// ======================================================================================================
// xtab/ytab entry: source offset (low 16 bits) | m1 (high 16 bits, 0..256).  xtab is padded to a multiple
// of 4 entries with (sw-1, 0) so the pad bytes of a row replicate the last column.
// mask_mode: after the resize keep only ==255 (threshold(254, TOZERO) of a 0/255 mask).
__global__ void __launch_bounds__(256)
k_resize_generic(const uint8_t *__restrict__ src, long long sfs, int spitch, int sw, int sh,
         uint8_t *__restrict__ dst, long long dfs, int dpitch, int dw, int dh,
         const uint32_t *__restrict__ xtab, const uint32_t *__restrict__ ytab, int mask_mode)
{
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int f = blockIdx.z;
    if (x4 >= dw || y >= dh) return;
    const uint32_t yt = __ldg(ytab + y);
    const int y0 = yt & 0xffff, n1 = yt >> 16, n0 = 256 - n1;
    const int y1 = min(y0 + 1, sh - 1);
    const uint8_t *r0 = src + (long long)f * sfs + (long long)y0 * spitch;
    const uint8_t *r1 = src + (long long)f * sfs + (long long)y1 * spitch;
    const uint4 xt = __ldg(reinterpret_cast<const uint4 *>(xtab + x4));
    const uint32_t t[4] = {xt.x, xt.y, xt.z, xt.w};
    uint32_t outw = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int a = t[j] & 0xffff, m1 = t[j] >> 16, m0 = 256 - m1;
        const int b = min(a + 1, sw - 1);
        const uint32_t h0 = m0 * __ldg(r0 + a) + m1 * __ldg(r0 + b);
        const uint32_t h1 = m0 * __ldg(r1 + a) + m1 * __ldg(r1 + b);
        uint32_t v = (n0 * h0 + n1 * h1 + 32768u) >> 16;
        if (mask_mode) v = (v == 255u) ? 255u : 0u;
        outw |= v << (8 * j);
    }
    *reinterpret_cast<uint32_t *>(dst + (long long)f * dfs + (long long)y * dpitch + x4) = outw;
}

// Fast path (scale factor <= 2, i.e. every level ratio ORB is used with): one thread = 4 adjacent output
// pixels x RZ_ROWS output rows.  The 4 pixels' taps span <= 8 source bytes starting at a0, so per source row
// three aligned 32-bit loads + two funnel shifts give a 64-bit window from which PRMT (run-time selector from
// the table) picks each pixel's (a, b) pair; the horizontal lerp m0*a + m1*b is one IDP.2A (u16x2 . u8x2), the
// vertical one two IMADs; byte 2 of n0*h0 + n1*h1 + 32768 is the output pixel, so three PRMTs pack the word.
// Table per 4-pixel group (2 x uint4, built on the host): {a0, sel[4 x 8 bit: rel_a | rel_b << 4], -, -},
// {M0..M3} with M = m0 | m1 << 16.
constexpr int RZ_ROWS = 4;
#ifndef KFS_RZ_YT
#define KFS_RZ_YT 2
#endif
constexpr int RZ_YT = KFS_RZ_YT; // k_resize_tma: a CTA's tile is 8 * RZ_ROWS * RZ_YT output rows high
__device__ __forceinline__ uint32_t kfs_prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
__global__ void __launch_bounds__(256)
k_resize(const uint8_t *__restrict__ src, long long sfs, int spitch, int sh,
         uint8_t *__restrict__ dst, long long dfs, int dpitch, int dw, int dh,
         const uint4 *__restrict__ xt, const uint32_t *__restrict__ ytab, int mask_mode)
{
    const int gx = blockIdx.x * 32 + threadIdx.x;
    const int ybase = (blockIdx.y * 8 + threadIdx.y) * RZ_ROWS;
    if (gx * 4 >= dw || ybase >= dh) return;
    const uint4 ta = __ldg(xt + 2 * gx), tm = __ldg(xt + 2 * gx + 1);
    const uint32_t spw = (uint32_t)spitch >> 2, wmax = spw - 1u;
    const uint32_t w0 = ta.x >> 2;
    const int shb = (int)(ta.x & 3u) * 8;
    const uint32_t w1 = min(w0 + 1u, wmax), w2 = min(w0 + 2u, wmax); // a clamped word is never one whose bytes are needed
    // selectors: nibbles (rel_a, rel_b) of pixel j in bits [8j, 8j+8); prmt reads 4 nibbles, the upper two pick
    // don't-care bytes (every nibble is <= 7, so no sign-replicate mode is ever selected)
    const uint32_t s0 = ta.y, s1 = ta.y >> 8, s2 = ta.y >> 16, s3 = ta.y >> 24;
    const uint32_t *sf = reinterpret_cast<const uint32_t *>(src + (long long)blockIdx.z * sfs); // CTA-uniform base
    uint32_t *df = reinterpret_cast<uint32_t *>(dst + (long long)blockIdx.z * dfs);
    const uint32_t dpw = (uint32_t)dpitch >> 2;
#pragma unroll
    for (int r = 0; r < RZ_ROWS; r++) {
        const int y = ybase + r;
        if (y >= dh) break;
        const uint32_t yt = __ldg(ytab + y);
        const uint32_t y0 = yt & 0xffffu, y1 = min(y0 + 1u, (uint32_t)sh - 1u);
        const uint32_t n1 = yt >> 16, n0 = 256u - n1;
        const uint32_t o0 = y0 * spw, o1 = y1 * spw; // 32-bit word offsets from the uniform base
        const uint32_t a0 = __ldg(sf + (o0 + w0)), a1 = __ldg(sf + (o0 + w1)), a2 = __ldg(sf + (o0 + w2));
        const uint32_t b0 = __ldg(sf + (o1 + w0)), b1 = __ldg(sf + (o1 + w1)), b2 = __ldg(sf + (o1 + w2));
        const uint32_t lo0 = __funnelshift_r(a0, a1, shb), hi0 = __funnelshift_r(a1, a2, shb);
        const uint32_t lo1 = __funnelshift_r(b0, b1, shb), hi1 = __funnelshift_r(b1, b2, shb);
        const uint32_t S0 = n0 * __dp2a_lo(tm.x, kfs_prmt(lo0, hi0, s0), 0u) + (n1 * __dp2a_lo(tm.x, kfs_prmt(lo1, hi1, s0), 0u) + 32768u);
        const uint32_t S1 = n0 * __dp2a_lo(tm.y, kfs_prmt(lo0, hi0, s1), 0u) + (n1 * __dp2a_lo(tm.y, kfs_prmt(lo1, hi1, s1), 0u) + 32768u);
        const uint32_t S2 = n0 * __dp2a_lo(tm.z, kfs_prmt(lo0, hi0, s2), 0u) + (n1 * __dp2a_lo(tm.z, kfs_prmt(lo1, hi1, s2), 0u) + 32768u);
        const uint32_t S3 = n0 * __dp2a_lo(tm.w, kfs_prmt(lo0, hi0, s3), 0u) + (n1 * __dp2a_lo(tm.w, kfs_prmt(lo1, hi1, s3), 0u) + 32768u);
        uint32_t outw = kfs_prmt(kfs_prmt(S0, S1, 0x0062u), kfs_prmt(S2, S3, 0x0062u), 0x5410u);
        if (mask_mode) { // keep only bytes == 255
            const uint32_t t = ((outw & 0x7f7f7f7fu) + 0x01010101u) & outw & 0x80808080u;
            outw = (t >> 7) * 255u;
        }
        df[(uint32_t)y * dpw + (uint32_t)gx] = outw;
    }
}

// a1.2 level 0 of the mask pyramid: mask != 0 -> 255
__global__ void __launch_bounds__(256)
k_mask_binarise(const uint8_t *__restrict__ src, long long sfs, long long sstep,
                uint8_t *__restrict__ dst, long long dfs, int dpitch, int w, int h)
{
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int f = blockIdx.z;
    if (x4 >= w || y >= h) return;
    const uint8_t *s = src + (long long)f * sfs + (long long)y * sstep;
    uint32_t outw = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int x = min(x4 + j, w - 1);
        outw |= (__ldg(s + x) ? 255u : 0u) << (8 * j);
    }
    *reinterpret_cast<uint32_t *>(dst + (long long)f * dfs + (long long)y * dpitch + x4) = outw;
}

// ======================================================================================================
// a1.3 / a1.4  FAST-9/16 score + 3x3 NMS + border / mask filter -> per-(frame, level) corner list
// ======================================================================================================
// One CTA = one 124x62 tile of the keepable region [31, W-31) x [31, H-31) of one level of one frame
// (tile -> (level, tix, tiy) through a table, so one launch covers every level of every frame).
//   score  tile 128 x 64 bytes: column sx <-> image x = tx0 - 1 + sx, row sy <-> y = ty0 - 1 + sy.  Columns 0 .. FT_W + 1 = 125 are
//               needed (the owned 124 + one NMS neighbour on each side): 126 of the 128 columns a warp computes are useful
//               (FT_W = 112 with its 16-byte aligned tile origins left 114 of 128: 388 instead of 345 tiles per 720p frame)
//   source tile 160 x 70 bytes, one TMA box whose x origin xa is the 16-byte aligned address at or below x0 = tx0 - 5; the tile's
//               own column 0 sits sh = x0 - xa bytes into the box (sh is a multiple of 4 because FAST_X0 = 5 mod 4 and FT_W = 0
//               mod 4, so every lane's centre word stays 4-byte aligned).  Score column sx has its centre at box column sx + 4 + sh.
// phase 1: each warp owns score rows w, w+8, ...; a lane tests 4 horizontally adjacent pixels with the
//          necessary condition kfs_fast_quick2 (u16x2 SIMD, two pixels per op, conflict-free LDS.32); the
//          pass flags of a row leave the warp as 4 ballots = one 128-bit mask row in shared memory
// scan   : one mask word per thread -> popc -> block exclusive scan -> the set bits are expanded into a
//          dense queue of pixel indices (no atomics, deterministic order)
// phase 2: exact score of the queued pixels (both polarities in one u16x2 word, VIMNMX3 trees)
// phase 3: 3x3 NMS on the score tile, border/mask predicate; survivors are gathered in shared memory and
//          appended to the (frame, level) list with ONE global atomic per tile
constexpr int FT_W = 124, FT_H = 62;
constexpr int FS_ROWB = 160, FS_P = 160, FS_H = FT_H + 8; // box row = 4 + 128 + 4 columns + up to 12 bytes of alignment slack
constexpr int SC_W = 128, SC_H = FT_H + 2;
constexpr int FAST_X0 = 29; // interior x origin of tile column 0 (owned columns below 31 fall to the border filter)
static_assert(FT_W % 4 == 0 && (FAST_X0 - 5) % 4 == 0 && FT_W + 2 <= SC_W, "4-byte aligned centre words; owned columns + NMS neighbours fit the score tile");
static_assert(SC_W + 8 + 12 <= FS_P && FS_P % 16 == 0, "the source box holds the tile at every alignment shift");
constexpr int FAST_THREADS = 256;
constexpr int FAST_MAXOUT = (FT_W / 2) * (FT_H / 2); // strict 3x3 maxima: at most one per 2x2 block
constexpr int FT_HH = 31;      // tile height of the low-latency variant (k_fast<FT_HH>)
constexpr int FAST_HALF_MAX_FRAMES = 2;
static_assert(SC_H * 4 == FAST_THREADS, "one mask word per thread");
static_assert(FAST_MAXOUT * 4 <= FS_P * FS_H, "survivor list aliases the source tile");

// ---- TMA (cp.async.bulk.tensor) staging of the source tile ------------------------------------------------------
// One 3-D tensor map per level over [frame][row][byte] (u8, row pitch and frame stride in bytes, both multiples of 16);
// a tile is one box of FS_P x FS_H x 1 bytes landing densely in shared memory (rows FS_P = 160 B apart); bytes outside the
// level arrive as zeros (they only feed scores the border filter drops).  One thread arms an mbarrier with the box size
// and issues the copy; every thread waits on the barrier's phase 0.
struct FastMaps {
    CUtensorMap m[KFS_MAX_LEVELS];
    int enabled; // 0: the driver entry point is unavailable -> the tile is loaded with LDG.128 / STS.128
};
__device__ __forceinline__ void kfs_mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void kfs_mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void kfs_tma_load_3d(void *smem, const CUtensorMap *map, int x, int y, int z, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(map), "r"(x), "r"(y), "r"(z), "r"((uint32_t)__cvta_generic_to_shared(bar))
                 : "memory");
}
__device__ __forceinline__ void kfs_mbar_wait(uint64_t *bar, uint32_t phase)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "KFS_MBAR_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra KFS_MBAR_DONE;\n"
                 "bra KFS_MBAR_WAIT;\n"
                 "KFS_MBAR_DONE:\n"
                 "}" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(phase) : "memory");
}

// Programmatic dependent launch (griddepcontrol): a kernel launched with the programmatic-serialization attribute may be set up
// while its predecessor in the stream drains, and parks at `wait` until that grid has completed and its writes are visible
// (a no-op for a plain launch).  Rule kept by every kernel here: nothing that another kernel produced is read, and nothing global
// is written, before kfs_pdl_wait().  No kernel triggers its dependents early (griddepcontrol.launch_dependents): with the trigger
// at the top of k_pyramid / k_fast the parked CTAs of the next kernels cost more than the overlap gains (one frame: 65.5 us
// against 57.7 us without PDL and 54.5 us with the implicit trigger at grid completion).
__device__ __forceinline__ void kfs_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Shared-memory accesses through a 32-bit shared address + immediate offset: the address arithmetic of the hot loops stays one
// register (a generic pointer into a __shared__ array makes the compiler rebuild the window base inside the loop).  volatile keeps
// the accesses on their side of the barriers; ptxas still schedules them freely in between.
template <int OFF> __device__ __forceinline__ uint32_t kfs_lds_u8(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(OFF));
    return v;
}
__device__ __forceinline__ uint32_t kfs_lds_u16(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void kfs_sts_u8(uint32_t a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void kfs_sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void kfs_sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
// the empty asm makes the address an opaque register value: the compiler keeps it instead of rebuilding the shared-window base
// (S2R SR_CgaCtaId + LEA) next to every use
__device__ __forceinline__ uint32_t kfs_smem_u32(const void *p)
{
    uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("" : "+r"(a));
    return a;
}

__device__ __forceinline__ int kfs_atoms_add(int *p, int v)
{
    int r;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(r) : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    return r;
}

// k_resize_tma: k_resize's arithmetic on a source tile that ONE TMA box brings into shared memory (k_resize's top stall is
// long_scoreboard on its 4-byte LDGs).  CTA = 128 x 32 output pixels as before; the box starts at the 16-byte aligned byte at or
// below the first tap of the tile's first pixel group and at the top source row of the tile's first output row; its size (bw x bh,
// host-computed from the tables, kfs_orb_plan_resize_boxes) covers every tap of every tile of every level.  Bytes outside the level
// arrive as zeros and are never among the taps that count (the clamped second row / clamped words of k_resize).
__global__ void __launch_bounds__(256)
k_resize_tma(const __grid_constant__ CUtensorMap map, int frame0, int sh, uint8_t *__restrict__ dst, long long dfs, int dpitch, int dw, int dh,
             const uint4 *__restrict__ xt, const uint32_t *__restrict__ ytab, int mask_mode, int bw, int bh)
{
    extern __shared__ __align__(128) uint8_t s_rt[];
    __shared__ __align__(8) uint64_t s_bar;
    const int tid = threadIdx.y * 32 + threadIdx.x;
    const int gx0 = blockIdx.x * 32, yb0 = blockIdx.y * 8 * RZ_ROWS * RZ_YT;
    const int xa0 = (int)(__ldg(xt + 2 * gx0).x) & ~15;        // CTA-uniform: first tap byte of the tile, aligned down
    const int ya0 = (int)(__ldg(ytab + yb0) & 0xffffu);          // top source row of the tile's first output row
    if (tid == 0) kfs_mbar_init(&s_bar, 1);
    __syncthreads();
    kfs_pdl_wait(); // the source level is complete from here on
    if (tid == 0) {
        kfs_mbar_expect_tx(&s_bar, (uint32_t)(bw * bh));
        kfs_tma_load_3d(s_rt, &map, xa0, ya0, (int)blockIdx.z + frame0, &s_bar);
    }
    const int gx = gx0 + threadIdx.x;
    constexpr int NR = RZ_ROWS * RZ_YT;                          // output rows per thread: warp w owns rows yb0 + NR * w ...
    const int ybase = yb0 + threadIdx.y * NR;
    const bool live = gx * 4 < dw && ybase < dh;
    // table entries travel while the box is in flight
    uint4 ta = make_uint4(0, 0, 0, 0), tm = ta;
    uint32_t yt[NR] = {};
    if (live) {
        ta = __ldg(xt + 2 * gx); tm = __ldg(xt + 2 * gx + 1);
#pragma unroll
        for (int r = 0; r < NR; r++) yt[r] = __ldg(ytab + min(ybase + r, dh - 1));
    }
    kfs_mbar_wait(&s_bar, 0);
    if (!live) return;
    const uint32_t off = ta.x - (uint32_t)xa0;                   // first tap byte inside the box
    const int shb = (int)(off & 3u) * 8;
    const uint32_t a_t = kfs_smem_u32(s_rt) + (off & ~3u);
    const uint32_t s0 = ta.y, s1 = ta.y >> 8, s2 = ta.y >> 16, s3 = ta.y >> 24;
    uint32_t *df = reinterpret_cast<uint32_t *>(dst + (long long)blockIdx.z * dfs);
    const uint32_t dpw = (uint32_t)dpitch >> 2;
#pragma unroll
    for (int r = 0; r < NR; r++) {
        const int y = ybase + r;
        if (y >= dh) break;
        const uint32_t y0 = yt[r] & 0xffffu, y1 = min(y0 + 1u, (uint32_t)sh - 1u);
        const uint32_t n1 = yt[r] >> 16, n0 = 256u - n1;
        const uint32_t pa = a_t + (y0 - (uint32_t)ya0) * (uint32_t)bw, pb = a_t + (y1 - (uint32_t)ya0) * (uint32_t)bw;
        uint32_t a0, a1, a2, b0, b1, b2;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a0) : "r"(pa));
        asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(a1) : "r"(pa));
        asm volatile("ld.shared.u32 %0, [%1+8];" : "=r"(a2) : "r"(pa));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(b0) : "r"(pb));
        asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(b1) : "r"(pb));
        asm volatile("ld.shared.u32 %0, [%1+8];" : "=r"(b2) : "r"(pb));
        const uint32_t lo0 = __funnelshift_r(a0, a1, shb), hi0 = __funnelshift_r(a1, a2, shb);
        const uint32_t lo1 = __funnelshift_r(b0, b1, shb), hi1 = __funnelshift_r(b1, b2, shb);
        const uint32_t S0 = n0 * __dp2a_lo(tm.x, kfs_prmt(lo0, hi0, s0), 0u) + (n1 * __dp2a_lo(tm.x, kfs_prmt(lo1, hi1, s0), 0u) + 32768u);
        const uint32_t S1 = n0 * __dp2a_lo(tm.y, kfs_prmt(lo0, hi0, s1), 0u) + (n1 * __dp2a_lo(tm.y, kfs_prmt(lo1, hi1, s1), 0u) + 32768u);
        const uint32_t S2 = n0 * __dp2a_lo(tm.z, kfs_prmt(lo0, hi0, s2), 0u) + (n1 * __dp2a_lo(tm.z, kfs_prmt(lo1, hi1, s2), 0u) + 32768u);
        const uint32_t S3 = n0 * __dp2a_lo(tm.w, kfs_prmt(lo0, hi0, s3), 0u) + (n1 * __dp2a_lo(tm.w, kfs_prmt(lo1, hi1, s3), 0u) + 32768u);
        uint32_t outw = kfs_prmt(kfs_prmt(S0, S1, 0x0062u), kfs_prmt(S2, S3, 0x0062u), 0x5410u);
        if (mask_mode) { // keep only bytes == 255
            const uint32_t t = ((outw & 0x7f7f7f7fu) + 0x01010101u) & outw & 0x80808080u;
            outw = (t >> 7) * 255u;
        }
        df[(uint32_t)y * dpw + (uint32_t)gx] = outw;
    }
}

// ======================================================================================================
// a1.1 / a1.2  the whole pyramid in ONE launch (used when every level's region fits shared memory)
// ======================================================================================================
// A CTA owns one 32 x 16 tile of the COARSEST level and, on every finer level, the proportional tile.  It loads the level-0
// region all of that depends on with one TMA box, then walks down the levels in shared memory: level l is computed from the
// CTA's own copy of level l-1 (same 8.8 fixed-point arithmetic and tables as k_resize_generic), the part it owns is written to
// global memory, the rest (a 1-2 px halo the next level needs) stays private.  The chained resize reads each level from
// shared memory instead of HBM/L2 (2.85 instead of 4.71 MB per 720p frame) and a frame costs one launch instead of seven.
// Regions (host-computed from the same tables, kfs_orb_plan_pyramid): per (tile, level) C = computed box, O = owned box.
struct PyrDev {
    int nlevels, box_bytes;
    int soff[KFS_MAX_LEVELS], spitch[KFS_MAX_LEVELS];      // shared-memory offset / row pitch of level l's region
    int sw[KFS_MAX_LEVELS], sh[KFS_MAX_LEVELS];            // level sizes
    uint8_t *dst[KFS_MAX_LEVELS];                          // level buffers (l >= 1)
    long long dfs[KFS_MAX_LEVELS];
    int dpitch[KFS_MAX_LEVELS];
    const uint32_t *ytab[KFS_MAX_LEVELS];
    const uint4 *xtab4[KFS_MAX_LEVELS];                    // k_resize's 4-pixel group tables
    int tab_off, xt_off;                                   // shared-memory offset of the table slices; of the x tables inside them
    int ytoff[KFS_MAX_LEVELS], xtoff[KFS_MAX_LEVELS];      // level l's slice: first y entry / first 4-pixel group
    const short *rects;                                    // [tiles][nlevels][8]: C.x0, C.y0, C.w, C.h, O.x0, O.y0, O.w, O.h
};

struct PyrMap { CUtensorMap m; };
// One launch instead of six dependent ones is what a single frame wants (20 instead of 36 us for the 720p pyramid); a full batch
// keeps the per-level kernels, whose PRMT tap gather is 2.2x cheaper per pixel than this kernel's byte loads from shared memory.
constexpr int PYR_FUSED_MAX_FRAMES = 4;

constexpr int PYR_THREADS = 512;
__global__ void __launch_bounds__(PYR_THREADS)
k_pyramid(const __grid_constant__ PyrDev P, const __grid_constant__ PyrMap map0, int frame0, int mask_mode)
{
    extern __shared__ __align__(128) uint8_t s_pyr[];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ short s_R[KFS_MAX_LEVELS * 8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, f = blockIdx.y + frame0;
    const int L = P.nlevels;
    if (tid == 0) kfs_mbar_init(&s_bar, 1);
    if (tid < L * 8) s_R[tid] = __ldg(P.rects + (size_t)blockIdx.x * L * 8 + tid);
    __syncthreads();
    if (tid == 0) {
        kfs_mbar_expect_tx(&s_bar, (uint32_t)P.box_bytes);
        kfs_tma_load_3d(s_pyr + P.soff[0], &map0.m, s_R[0], s_R[1], f, &s_bar); // s_R[0] is a multiple of 16 (plan)
    }
    // While the level-0 box is in flight: this tile's slices of every level's resize tables go to shared memory, one warp per
    // (level, table), so that the level loop below waits on no global load (it used to pay three dependent ones per level)
    uint32_t *s_yt = reinterpret_cast<uint32_t *>(s_pyr + P.tab_off);
    uint4 *s_xt = reinterpret_cast<uint4 *>(s_pyr + P.tab_off + P.xt_off);
    for (int job = warp; job < 2 * (L - 1); job += PYR_THREADS / 32) {
        const int l = 1 + (job >> 1);
        const short *Rc = s_R + 8 * l;
        if (job & 1) {
            const uint4 *src = P.xtab4[l] + 2 * (Rc[0] >> 2);
            uint4 *dst = s_xt + 2 * P.xtoff[l];
            for (int j = lane; j < 2 * (Rc[2] >> 2); j += 32) dst[j] = __ldg(src + j);
        } else {
            const uint32_t *src = P.ytab[l] + Rc[1];
            uint32_t *dst = s_yt + P.ytoff[l];
            for (int j = lane; j < Rc[3]; j += 32) dst[j] = __ldg(src + j);
        }
    }
    kfs_mbar_wait(&s_bar, 0);
    __syncthreads();
    for (int l = 1; l < L; l++) {
        const short *Rc = s_R + 8 * l, *Rp = Rc - 8;
        const int cx0 = Rc[0], cy0 = Rc[1], cw4 = Rc[2] >> 2, ch = Rc[3];
        const int ox0 = Rc[4], oy0 = Rc[5], ox1 = ox0 + Rc[6], oy1 = oy0 + Rc[7];
        const int px0 = Rp[0], py0 = Rp[1];
        const uint8_t *src = s_pyr + P.soff[l - 1];
        uint8_t *dsm = s_pyr + P.soff[l];
        const int sp = P.spitch[l - 1], dp = P.spitch[l];
        const int shm = P.sh[l - 1] - 1;
        const uint32_t *yt = s_yt + P.ytoff[l];
        const uint4 *xt4 = s_xt + 2 * P.xtoff[l];
        uint8_t *dg = P.dst[l] + (long long)f * P.dfs[l];
        const int dgp = P.dpitch[l];
        // lane = 4-pixel group (its table entry, selectors and word offset stay in registers), warp = every 16th output row:
        // the arithmetic of k_resize (three words per source row, funnel shift, PRMT tap select, IDP.2A, two IMADs), the
        // words coming from this CTA's copy of level l-1
        for (int g4 = lane; g4 < cw4; g4 += 32) {
            const int x = cx0 + 4 * g4;
            const uint4 ta = xt4[2 * g4], tm = xt4[2 * g4 + 1];
            const uint32_t off = ta.x - (uint32_t)px0, w0 = off >> 2;
            const int shb = (int)(off & 3u) * 8;
            const uint32_t s0 = ta.y, s1 = ta.y >> 8, s2 = ta.y >> 16, s3 = ta.y >> 24;
            const bool xown = x >= ox0 && x < ox1;
            const uint32_t *sw = reinterpret_cast<const uint32_t *>(src) + w0;
            uint32_t *dw = reinterpret_cast<uint32_t *>(dsm) + g4;
            uint32_t *gw = reinterpret_cast<uint32_t *>(dg + x);
            const int spw = sp >> 2, dpw = dp >> 2, gpw = dgp >> 2;
            for (int gy = warp; gy < ch; gy += PYR_THREADS / 32) {
                const int y = cy0 + gy;
                const uint32_t ye = yt[gy];
                const int y0 = ye & 0xffffu, y1 = min(y0 + 1, shm);
                const uint32_t n1 = ye >> 16, n0 = 256u - n1;
                const uint32_t *r0 = sw + (y0 - py0) * spw, *r1 = sw + (y1 - py0) * spw;
                const uint32_t a0 = r0[0], a1 = r0[1], a2 = r0[2]; // words past the region's last tap are never selected
                const uint32_t b0 = r1[0], b1 = r1[1], b2 = r1[2];
                const uint32_t lo0 = __funnelshift_r(a0, a1, shb), hi0 = __funnelshift_r(a1, a2, shb);
                const uint32_t lo1 = __funnelshift_r(b0, b1, shb), hi1 = __funnelshift_r(b1, b2, shb);
                const uint32_t S0 = n0 * __dp2a_lo(tm.x, kfs_prmt(lo0, hi0, s0), 0u) + (n1 * __dp2a_lo(tm.x, kfs_prmt(lo1, hi1, s0), 0u) + 32768u);
                const uint32_t S1 = n0 * __dp2a_lo(tm.y, kfs_prmt(lo0, hi0, s1), 0u) + (n1 * __dp2a_lo(tm.y, kfs_prmt(lo1, hi1, s1), 0u) + 32768u);
                const uint32_t S2 = n0 * __dp2a_lo(tm.z, kfs_prmt(lo0, hi0, s2), 0u) + (n1 * __dp2a_lo(tm.z, kfs_prmt(lo1, hi1, s2), 0u) + 32768u);
                const uint32_t S3 = n0 * __dp2a_lo(tm.w, kfs_prmt(lo0, hi0, s3), 0u) + (n1 * __dp2a_lo(tm.w, kfs_prmt(lo1, hi1, s3), 0u) + 32768u);
                uint32_t outw = kfs_prmt(kfs_prmt(S0, S1, 0x0062u), kfs_prmt(S2, S3, 0x0062u), 0x5410u);
                if (mask_mode) { // keep only bytes == 255
                    const uint32_t t = ((outw & 0x7f7f7f7fu) + 0x01010101u) & outw & 0x80808080u;
                    outw = (t >> 7) * 255u;
                }
                dw[gy * dpw] = outw;
                if (xown && y >= oy0 && y < oy1) gw[(long long)y * gpw] = outw;
            }
        }
        __syncthreads();
    }
}

// TH = tile height: 62 for batches; 31 for calls of one or two frames, whose 388 tiles per 720p frame leave the SMs with 2-3 CTAs
// of long dependent phases each (twice the CTAs of half the work finish sooner: one wave either way).
template <int TH>
__global__ void __launch_bounds__(FAST_THREADS)
k_fast(const __grid_constant__ OrbDev g, const uint32_t *__restrict__ tile_tab, const __grid_constant__ FastMaps maps)
{
    constexpr int FT_H = TH, FS_H = TH + 8, SC_H = TH + 2; // shadow the batch geometry above
    static_assert(SC_H * 4 <= FAST_THREADS, "one mask word per thread");
    __shared__ __align__(128) uint8_t s_src[FS_P * FS_H + 16];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ __align__(16) uint8_t s_sc[SC_W * SC_H];
    __shared__ uint16_t s_q[SC_W * SC_H];
    __shared__ __align__(16) uint32_t s_mask[FAST_THREADS]; // rows past SC_H stay zero
    __shared__ int s_wsum[FAST_THREADS / 32];
    __shared__ int s_nout, s_gbase;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int f = blockIdx.y + g.frame0;
    const uint32_t tt = __ldg(tile_tab + blockIdx.x); // host-written table: not another kernel's output
    const int l = tt >> 24, tiy = (tt >> 12) & 0xfff, tix = tt & 0xfff;
    const LevelDev &L = g.lv[l];
    const int W = L.w, H = L.h, pitch = L.pitch;
    const int tx0 = FAST_X0 + FT_W * tix, ty0 = KFS_BORDER + FT_H * tiy;
    const int ox = tx0 - 1, sy0 = ty0 - 4;          // image x of score column 0; image y of source row 0
    const int xa = (tx0 - 5) & ~15;                 // 16-byte aligned x origin of the source box
    const uint8_t *s_tile = s_src + (tx0 - 5 - xa) + 4; // + r * FS_P + sx = centre of score column sx on source row r
    const uint8_t *img = L.img + (long long)f * L.img_fs;
    const int thr = g.fast_thr;

    // ---- load the source tile (one TMA box, or uint4 per thread), clear the score tile and the mask rows meanwhile
    const bool tma = maps.enabled != 0;
    if (tma && tid == 0) kfs_mbar_init(&s_bar, 1);
    __syncthreads();
    kfs_pdl_wait(); // the pyramid is complete from here on
    if (tma) {
        if (tid == 0) {
            kfs_mbar_expect_tx(&s_bar, FS_P * FS_H);
            kfs_tma_load_3d(s_src, &maps.m[l], xa, sy0, f, &s_bar);
        }
    } else {
        for (int i = tid; i < FS_H * (FS_ROWB / 16); i += FAST_THREADS) {
            const int r = i / (FS_ROWB / 16), c = i % (FS_ROWB / 16);
            const int gy = min(max(sy0 + r, 0), H - 1), gx = xa + 16 * c;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (gx + 16 <= pitch) v = __ldg(reinterpret_cast<const uint4 *>(img + (long long)gy * pitch + gx));
            *reinterpret_cast<uint4 *>(s_src + r * FS_P + 16 * c) = v;
        }
    }
    for (int i = tid; i < SC_W * SC_H / 16; i += FAST_THREADS) reinterpret_cast<uint4 *>(s_sc)[i] = make_uint4(0, 0, 0, 0);
    s_mask[tid] = 0;
    if (tid == 0) s_nout = 0;
    if (tma) kfs_mbar_wait(&s_bar, 0);
    __syncthreads();

    // ---- phase 1: quick reject, one score row per warp-iteration, 4 pixels per lane
    // x range whose score this tile needs: [tx0-1, tx0+FT_W] clipped to [30, W-31]; y likewise
    const int xlo = max(tx0 - 1, KFS_BORDER - 1), xhi = min(tx0 + FT_W, W - KFS_BORDER);
    const int ylo = max(ty0 - 1, KFS_BORDER - 1), yhi = min(ty0 + FT_H, H - KFS_BORDER);
    uint32_t vme = 0, vmo = 0; // validity of this lane's 4 pixels, in kfs_fast_quick2's result bit positions
    {
        const int x = ox + 4 * lane;
        if (x + 0 >= xlo && x + 0 <= xhi) vme |= 0x00008000u;
        if (x + 2 >= xlo && x + 2 <= xhi) vme |= 0x80000000u;
        if (x + 1 >= xlo && x + 1 <= xhi) vmo |= 0x00008000u;
        if (x + 3 >= xlo && x + 3 <= xhi) vmo |= 0x80000000u;
    }
    const int sy_lo = max(ylo - (ty0 - 1), 0), sy_hi = min(yhi - (ty0 - 1), SC_H - 1);
#pragma unroll 2
    for (int sy = sy_lo + warp; sy <= sy_hi; sy += FAST_THREADS / 32) {
        const uint8_t *rowc = s_tile + (sy + 3) * FS_P + 4 * lane;
        const uint32_t B = *reinterpret_cast<const uint32_t *>(rowc);
        const uint32_t A = *reinterpret_cast<const uint32_t *>(rowc - 4);
        const uint32_t C = *reinterpret_cast<const uint32_t *>(rowc + 4);
        const uint32_t U = *reinterpret_cast<const uint32_t *>(rowc - 3 * FS_P);
        const uint32_t D = *reinterpret_cast<const uint32_t *>(rowc + 3 * FS_P);
        const uint32_t Rw = __funnelshift_r(B, C, 24); // pixels x+3 .. x+6
        const uint32_t Lw = __funnelshift_r(A, B, 8);  // pixels x-3 .. x
        // even bytes -> (px0, px2), odd bytes -> (px1, px3) as u16x2
        const uint32_t me = kfs_fast_quick2m(__byte_perm(B, 0, 0x4240), __byte_perm(U, 0, 0x4240), __byte_perm(D, 0, 0x4240),
                                             __byte_perm(Rw, 0, 0x4240), __byte_perm(Lw, 0, 0x4240), thr, vme);
        const uint32_t mo = kfs_fast_quick2m(__byte_perm(B, 0, 0x4341), __byte_perm(U, 0, 0x4341), __byte_perm(D, 0, 0x4341),
                                             __byte_perm(Rw, 0, 0x4341), __byte_perm(Lw, 0, 0x4341), thr, vmo);
        const uint32_t b0 = __ballot_sync(0xffffffffu, me & 0x00008000u), b1 = __ballot_sync(0xffffffffu, mo & 0x00008000u);
        const uint32_t b2 = __ballot_sync(0xffffffffu, (int)me < 0), b3 = __ballot_sync(0xffffffffu, (int)mo < 0);
        if (lane == 0) *reinterpret_cast<uint4 *>(s_mask + 4 * sy) = make_uint4(b0, b1, b2, b3);
    }
    __syncthreads();

    // ---- scan: mask word tid = (row tid>>2, pixel phase tid&3); bit b of it <-> score column 4*b + phase
    int nq;
    {
        uint32_t w = s_mask[tid];
        const int cnt = __popc(w);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        int pos = incl - cnt;
        nq = 0;
#pragma unroll
        for (int j = 0; j < FAST_THREADS / 32; j++) {
            const int t = s_wsum[j];
            if (j < warp) pos += t;
            nq += t;
        }
        const int base_pi = (tid >> 2) * SC_W + (tid & 3);
        while (w) {
            const int b = __ffs(w) - 1;
            w &= w - 1;
            s_q[pos++] = (uint16_t)(base_pi + 4 * b);
        }
    }
    __syncthreads();

    // ---- phase 2: exact score of queued pixels; the corners among them (about half) go to a second queue so that phase 3 walks
    //      corners only.  Every WARP owns a private segment of that queue (upper half of s_q, CQ_SEG entries per warp: a warp never
    //      sees more than ceil(nq / 256) * 32 <= CQ_SEG entries when nq <= CQ_MAX) and keeps its fill count in a register: one ballot,
    //      two POPC and a store per iteration instead of a shared-memory atomic + shuffle round trip (the order inside a tile is
    //      free, the lists are re-ordered by k_select_sort).  Tiles with more than half of their pixels queued (dense-corner images)
    //      keep the full queue for phase 3.
    constexpr int CQ_OFF = SC_W * SC_H / 2;
    constexpr int CQ_SEG = (SC_W * SC_H - CQ_OFF) / (FAST_THREADS / 32) / 32 * 32; // 512 (TH = 62), 256 (TH = 31)
    constexpr int CQ_MAX = CQ_SEG / 32 * FAST_THREADS;                             // queue lengths the segments can take
    static_assert(CQ_OFF + (FAST_THREADS / 32) * CQ_SEG <= SC_W * SC_H && CQ_MAX <= CQ_OFF, "per-warp corner segments fit the upper half of the queue");
    const bool compact = nq <= CQ_MAX;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint16_t *s_cq = s_q + CQ_OFF + warp * CQ_SEG;
    const uint32_t a_q = kfs_smem_u32(s_q), a_sc = kfs_smem_u32(s_sc), a_cq = kfs_smem_u32(s_cq);
    const uint32_t a_ctr = kfs_smem_u32(s_tile) + 3 * FS_P; // centre of score pixel (sx, sy) = a_ctr + sy * FS_P + sx
    int wc = 0; // corners this warp has queued (warp-uniform)
    for (int e0 = warp * 32; e0 < nq; e0 += FAST_THREADS) { // warp-uniform bounds: the ballot below sees whole warps
        const int e = e0 + lane;
        bool corner = false;
        uint32_t pi = 0;
        if (e < nq) {
            pi = kfs_lds_u16(a_q + 2 * e);
            const uint32_t c = a_ctr + pi + (pi >> 7) * (FS_P - SC_W);
            const uint32_t cv = kfs_fast_cv((int)kfs_lds_u8<0>(c));
            uint32_t ek[16];
            ek[0] = kfs_fast_e(cv, kfs_lds_u8<3 * FS_P + 0>(c));   ek[1] = kfs_fast_e(cv, kfs_lds_u8<3 * FS_P + 1>(c));
            ek[2] = kfs_fast_e(cv, kfs_lds_u8<2 * FS_P + 2>(c));   ek[3] = kfs_fast_e(cv, kfs_lds_u8<1 * FS_P + 3>(c));
            ek[4] = kfs_fast_e(cv, kfs_lds_u8<3>(c));              ek[5] = kfs_fast_e(cv, kfs_lds_u8<-1 * FS_P + 3>(c));
            ek[6] = kfs_fast_e(cv, kfs_lds_u8<-2 * FS_P + 2>(c));  ek[7] = kfs_fast_e(cv, kfs_lds_u8<-3 * FS_P + 1>(c));
            ek[8] = kfs_fast_e(cv, kfs_lds_u8<-3 * FS_P + 0>(c));  ek[9] = kfs_fast_e(cv, kfs_lds_u8<-3 * FS_P - 1>(c));
            ek[10] = kfs_fast_e(cv, kfs_lds_u8<-2 * FS_P - 2>(c)); ek[11] = kfs_fast_e(cv, kfs_lds_u8<-1 * FS_P - 3>(c));
            ek[12] = kfs_fast_e(cv, kfs_lds_u8<-3>(c));            ek[13] = kfs_fast_e(cv, kfs_lds_u8<1 * FS_P - 3>(c));
            ek[14] = kfs_fast_e(cv, kfs_lds_u8<2 * FS_P - 2>(c));  ek[15] = kfs_fast_e(cv, kfs_lds_u8<3 * FS_P - 1>(c));
            const int m = kfs_fast9_m(ek);
            corner = m > thr;
            if (corner) kfs_sts_u8(a_sc + pi, (uint32_t)(m - 1));
        }
        if (compact) {
            const uint32_t bc = __ballot_sync(0xffffffffu, corner);
            if (corner) kfs_sts_u16(a_cq + 2 * (wc + __popc(bc & lt_mask)), pi);
            wc += __popc(bc);
        }
    }
    __syncthreads();

    // ---- phase 3: NMS + border/mask filter.  compact: a warp walks its own corner segment and gathers its survivors in a private
    //      segment as well (warps 0-3 in the dead lower half of s_q, warps 4-7 in the dead source tile), so the phase needs no
    //      atomic either; the eight counts meet in s_wsum.  Otherwise: the full queue, block-wide, survivors through a shared counter.
    const uint8_t *mask = L.mask ? L.mask + (long long)f * L.mask_fs : nullptr;
    // keepable score coordinates: owned (1 <= sx <= FT_W, 1 <= sy <= FT_H) and inside the border ([31, W-31) x [31, H-31)), folded
    // into one unsigned range test per axis; the 8 neighbours are then read back to back and reduced with three max3 + one max
    const int nx_lo = max(1, KFS_BORDER - ox), nx_n = min(FT_W, W - KFS_BORDER - 1 - ox) - nx_lo;
    const int ny_lo = max(1, KFS_BORDER - (ty0 - 1)), ny_n = min(FT_H, H - KFS_BORDER - 1 - (ty0 - 1)) - ny_lo;
    const uint32_t a_nms = kfs_smem_u32(s_sc);
    auto nms_keep = [&](uint32_t pi, uint32_t &packed) -> bool {
        const int sy = pi >> 7, sx = pi & (SC_W - 1);
        const int x = ox + sx, y = ty0 - 1 + sy;
        packed = 0;
        if (!((uint32_t)(sx - nx_lo) <= (uint32_t)nx_n && (uint32_t)(sy - ny_lo) <= (uint32_t)ny_n && (nx_n | ny_n) >= 0)) return false;
        const uint32_t p = a_nms + pi; // 1 <= sx <= FT_W, 1 <= sy <= FT_H: all eight neighbours are inside the score tile
        const uint32_t s = kfs_lds_u8<0>(p);
        const uint32_t n0 = kfs_lds_u8<-1>(p), n1 = kfs_lds_u8<1>(p), n2 = kfs_lds_u8<-SC_W - 1>(p), n3 = kfs_lds_u8<-SC_W>(p);
        const uint32_t n4 = kfs_lds_u8<-SC_W + 1>(p), n5 = kfs_lds_u8<SC_W - 1>(p), n6 = kfs_lds_u8<SC_W>(p), n7 = kfs_lds_u8<SC_W + 1>(p);
        const uint32_t nmax = max(__vimax3_u32(n0, n1, n2), __vimax3_u32(n3, n4, __vimax3_u32(n5, n6, n7)));
        bool keep = s > nmax; // implies s > 0
        if (keep && mask) keep = __ldg(mask + (long long)y * L.mask_pitch + x) != 0;
        packed = (s << 24) | ((uint32_t)y << 12) | (uint32_t)x;
        return keep;
    };
    uint32_t *out = g.raw + (long long)f * g.list_fs + L.list_off;
    if (compact) {
        static_assert((FAST_THREADS / 64) * CQ_SEG * 4 <= CQ_OFF * 2 && (FAST_THREADS / 64) * CQ_SEG * 4 <= FS_P * FS_H, "survivor segments fit");
        uint32_t *s_seg = (warp < FAST_THREADS / 64 ? reinterpret_cast<uint32_t *>(s_q) : reinterpret_cast<uint32_t *>(s_src)) + (warp & (FAST_THREADS / 64 - 1)) * CQ_SEG;
        int wk = 0; // survivors of this warp (warp-uniform)
        for (int e0 = 0; e0 < wc; e0 += 32) {
            const int e = e0 + lane;
            uint32_t packed = 0;
            const bool keep = e < wc && nms_keep(s_cq[e], packed);
            const uint32_t bk = __ballot_sync(0xffffffffu, keep);
            if (keep) s_seg[wk + __popc(bk & lt_mask)] = packed;
            wk += __popc(bk);
        }
        if (lane == 0) s_wsum[warp] = wk;
        __syncthreads();
        int before = 0, nout = 0;
#pragma unroll
        for (int j = 0; j < FAST_THREADS / 32; j++) {
            const int t = s_wsum[j];
            if (j < warp) before += t;
            nout += t;
        }
        if (nout == 0) return;
        if (tid == 0) s_gbase = atomicAdd(g.raw_count + f * KFS_MAX_LEVELS + l, nout);
        __syncthreads();
        const int gbase = s_gbase + before;
        for (int i = lane; i < wk; i += 32)
            if (gbase + i < L.list_cap) out[gbase + i] = s_seg[i];
        return;
    }
    uint32_t *s_out = reinterpret_cast<uint32_t *>(s_src);
    for (int e0 = 0; e0 < nq; e0 += FAST_THREADS) {
        const int e = e0 + tid;
        uint32_t packed = 0;
        const bool keep = e < nq && nms_keep(s_q[e], packed);
        const uint32_t bk = __ballot_sync(0xffffffffu, keep);
        if (bk) {
            int base = 0;
            if (lane == 0) base = kfs_atoms_add(&s_nout, __popc(bk));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (keep) s_out[base + __popc(bk & lt_mask)] = packed;
        }
    }
    __syncthreads();
    const int nout = s_nout;
    if (nout == 0) return;
    if (tid == 0) s_gbase = atomicAdd(g.raw_count + f * KFS_MAX_LEVELS + l, nout);
    __syncthreads();
    const int gbase = s_gbase;
    for (int i = tid; i < nout; i += FAST_THREADS)
        if (gbase + i < L.list_cap) out[gbase + i] = s_out[i];
}

// ======================================================================================================
// a1.4  retainBest(2*quota) on the FAST score: keep every corner whose score >= the (2*quota)-th largest
// ======================================================================================================
// grid (level, frame); scores are integers in [thr, 254] so a 256-bin histogram select is exact.
constexpr int SEL_THREADS = 1024;
__global__ void __launch_bounds__(SEL_THREADS)
k_select_fast(const __grid_constant__ OrbDev g)
{
    __shared__ int s_hist[256];
    __shared__ int s_thr, s_ccount;
    const int l = blockIdx.x, f = blockIdx.y + g.frame0, tid = threadIdx.x, lane = tid & 31;
    const LevelDev &L = g.lv[l];
    if (L.list_cap == 0) return;
    if (tid == 0) s_ccount = 0;
    kfs_pdl_wait();
    const int n = min(g.raw_count[f * KFS_MAX_LEVELS + l], L.list_cap);
    const int target = 2 * L.quota;
    const uint32_t *raw = g.raw + (long long)f * g.list_fs + L.list_off;
    int thr = 0;
    if (n > target) {
        if (tid < 256) s_hist[tid] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += SEL_THREADS) atomicAdd(&s_hist[raw[i] >> 24], 1);
        __syncthreads();
        if (tid < 32) {
            // warp scan from the top bin down, 8 bins per lane
            int loc[8], sum = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) { loc[k] = s_hist[255 - (lane * 8 + k)]; sum += loc[k]; }
            int incl = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += v;
            }
            int cum = incl - sum, found = -1;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                cum += loc[k];
                if (found < 0 && cum >= target) found = 255 - (lane * 8 + k);
            }
            const uint32_t b = __ballot_sync(0xffffffffu, found >= 0);
            const int src = __ffs(b) - 1;
            const int t = __shfl_sync(0xffffffffu, found, src < 0 ? 0 : src);
            if (lane == 0) s_thr = (b ? t : 0);
        }
        __syncthreads();
        thr = s_thr;
    }
    if (target <= 0) thr = 256; // nothing is kept for a zero quota
    uint32_t *cand = g.cand_xy + (long long)f * g.list_fs + L.list_off;
    // this CTA is the only writer of the level's candidate list: slots come from a shared-memory counter (a warp waits ~30
    // cycles for its base instead of a global-atomic round trip per 32 corners); the count goes out once at the end
    __syncthreads();
    for (int i0 = 0; i0 < n; i0 += SEL_THREADS) {
        const int i = i0 + tid;
        uint32_t v = 0;
        bool keep = false;
        if (i < n) { v = raw[i]; keep = (int)(v >> 24) >= thr; }
        const uint32_t b = __ballot_sync(0xffffffffu, keep);
        if (b) {
            int base = 0;
            if (lane == 0) base = atomicAdd(&s_ccount, __popc(b));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (keep) cand[base + __popc(b & ((1u << lane) - 1u))] = v;
        }
    }
    __syncthreads();
    if (tid == 0) g.cand_count[f * KFS_MAX_LEVELS + l] = s_ccount;
}

// ======================================================================================================
// a1.5  Harris response (block 7, k 0.04) on the un-blurred level, one thread per candidate
// ======================================================================================================
__global__ void __launch_bounds__(256)
k_harris(const __grid_constant__ OrbDev g)
{
    const int l = blockIdx.y, f = blockIdx.z + g.frame0;
    const LevelDev &L = g.lv[l];
    if (L.list_cap == 0) return;
    kfs_pdl_wait();
    const int n = min(g.cand_count[f * KFS_MAX_LEVELS + l], L.list_cap);
    const uint32_t *cand = g.cand_xy + (long long)f * g.list_fs + L.list_off;
    float *resp = g.cand_resp + (long long)f * g.list_fs + L.list_off;
    const uint8_t *img = L.img + (long long)f * L.img_fs;
    const int pitch = L.pitch;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t v = cand[i];
        const int x = v & 0xfff, y = (v >> 12) & 0xfff;
        const uint8_t *p = img + (long long)(y - 4) * pitch + (x - 4);
        // the 9 x 9 bytes arrive as 9 x 3 aligned words, all in flight at once (a keypoint is >= 31 px from every border, so the
        // up to 3 bytes before / after the 9 of a row are inside the image); a row is then shifted into place
        uint32_t w[9][3];
        int sh[9];
#pragma unroll
        for (int r = 0; r < 9; r++) {
            const uintptr_t a = (uintptr_t)(p + (long long)r * pitch);
            const uint32_t *q = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
            sh[r] = (int)(a & 3) * 8;
            w[r][0] = __ldg(q); w[r][1] = __ldg(q + 1); w[r][2] = __ldg(q + 2);
        }
        int r0[9], r1[9], r2[9];
        auto unpack = [&](int r, int *o) {
            const uint32_t lo = __funnelshift_r(w[r][0], w[r][1], sh[r]), hi = __funnelshift_r(w[r][1], w[r][2], sh[r]);
#pragma unroll
            for (int j = 0; j < 4; j++) { o[j] = (lo >> (8 * j)) & 0xff; o[4 + j] = (hi >> (8 * j)) & 0xff; }
            o[8] = (w[r][2] >> sh[r]) & 0xff;
        };
        unpack(0, r0); unpack(1, r1);
        int a = 0, b = 0, c = 0;
#pragma unroll
        for (int i7 = 0; i7 < 7; i7++) {
            unpack(i7 + 2, r2);
#pragma unroll
            for (int j = 0; j < 7; j++) {
                const int Ix = (r1[j + 2] - r1[j]) * 2 + (r0[j + 2] - r0[j]) + (r2[j + 2] - r2[j]);
                const int Iy = (r2[j + 1] - r0[j + 1]) * 2 + (r2[j] - r0[j]) + (r2[j + 2] - r0[j + 2]);
                a += Ix * Ix; b += Iy * Iy; c += Ix * Iy;
            }
#pragma unroll
            for (int j = 0; j < 9; j++) { r0[j] = r1[j]; r1[j] = r2[j]; }
        }
        resp[i] = kfs_harris_response(a, b, c);
    }
}

// ======================================================================================================
// a1.6  retainBest(quota) on the Harris response (ties kept) + deterministic (y,x) order
// ======================================================================================================
// grid (level, frame).  4-pass radix select on the order-preserving u32 key of the fp32 response gives the
// exact quota-th largest; everything >= it is kept; the rank of a kept entry among kept entries by
// (y<<12|x) is its output position.
constexpr int SORT_THREADS = 512;
// dynamic shared memory: key[N] | pos[N] | cpos[N] | cidx[N]  (u32 each), N = smem_n (host: 2*max quota + slack)
__global__ void __launch_bounds__(SORT_THREADS)
k_select_sort(const __grid_constant__ OrbDev g, int smem_n, int rows_cap)
{
    extern __shared__ __align__(16) uint32_t s_dyn[];
    uint32_t *s_key = s_dyn, *s_pos = s_dyn + smem_n, *s_cpos = s_dyn + 2 * smem_n, *s_cidx = s_dyn + 3 * smem_n;
    int *s_rowa = reinterpret_cast<int *>(s_dyn + 4 * smem_n), *s_rowb = s_rowa + rows_cap; // per image row: start / fill cursor
    __shared__ int s_wsum[SORT_THREADS / 32];
    __shared__ int s_hist[256];
    __shared__ uint32_t s_prefix;
    __shared__ int s_need, s_nkept, s_small, s_ngather;
    __shared__ uint32_t s_gather[32];
    const int l = blockIdx.x, f = blockIdx.y + g.frame0, tid = threadIdx.x, lane = tid & 31;
    const LevelDev &L = g.lv[l];
    if (L.list_cap == 0) return;
    kfs_pdl_wait();
    const int n = min(g.cand_count[f * KFS_MAX_LEVELS + l], L.list_cap);
    const uint32_t *cand = g.cand_xy + (long long)f * g.list_fs + L.list_off;
    const float *resp = g.cand_resp + (long long)f * g.list_fs + L.list_off;
    const bool in_smem = n <= smem_n;
    if (in_smem)
        for (int i = tid; i < n; i += SORT_THREADS) { s_key[i] = kfs_float_key(resp[i]); s_pos[i] = cand[i] & 0xffffffu; }
    if (tid == 0) { s_prefix = 0; s_need = L.quota; s_nkept = 0; s_small = 0; s_ngather = 0; }
    __syncthreads();
    uint32_t thr_key = 0;
    if (L.quota <= 0) thr_key = 0xffffffffu;
    else if (n > L.quota) {
        for (int pass = 0; pass < 4; pass++) {
            const int shift = 24 - 8 * pass;
            const uint32_t hi_mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
            if (tid < 256) s_hist[tid] = 0;
            __syncthreads();
            const uint32_t prefix = s_prefix;
            // the responses of one level share their top byte(s): lanes with the same bin add once (MATCH.ANY) instead of
            // queueing on one shared-memory word
            for (int i0 = 0; i0 < n; i0 += SORT_THREADS) {
                const int i = i0 + tid;
                uint32_t k = 0;
                if (i < n) k = in_smem ? s_key[i] : kfs_float_key(resp[i]);
                const int bin = (i < n && (k & hi_mask) == prefix) ? (int)((k >> shift) & 255u) : -1;
                const uint32_t peers = __match_any_sync(0xffffffffu, bin);
                if (bin >= 0 && lane == __ffs(peers) - 1) atomicAdd(&s_hist[bin], __popc(peers));
            }
            __syncthreads();
            if (tid < 32) {
                // warp scan from the top bin down, 8 bins per lane: first bin where the running count reaches `need`
                const int need = s_need;
                int loc[8], sum = 0;
#pragma unroll
                for (int k = 0; k < 8; k++) { loc[k] = s_hist[255 - (lane * 8 + k)]; sum += loc[k]; }
                int incl = sum;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += v;
                }
                int cum = incl - sum, found = -1, before = 0;
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    if (found < 0 && cum + loc[k] >= need) { found = 255 - (lane * 8 + k); before = cum; }
                    cum += loc[k];
                }
                const uint32_t b = __ballot_sync(0xffffffffu, found >= 0);
                const int src = __ffs(b) - 1;
                if (b && lane == src) {
                    s_prefix = prefix | ((uint32_t)found << shift);
                    s_need = need - before;
                    s_small = (pass < 3 && s_hist[found] <= 32) ? s_hist[found] : 0;
                }
            }
            __syncthreads();
            const int small = s_small;
            if (small > 0) {
                // the bin that holds the quota-th response has at most 32 keys: one warp finishes the select on the full keys
                // instead of one more histogram pass per remaining byte
                const uint32_t pfx = s_prefix, msk = 0xffffffffu << shift;
                for (int i = tid; i < n; i += SORT_THREADS) {
                    const uint32_t k = in_smem ? s_key[i] : kfs_float_key(resp[i]);
                    if ((k & msk) == pfx) s_gather[atomicAdd(&s_ngather, 1)] = k;
                }
                __syncthreads();
                if (tid < 32) {
                    const uint32_t mine = lane < small ? s_gather[lane] : 0u;
                    const int need = s_need;
                    int greater = 0, ge = 0;
                    for (int j = 0; j < small; j++) {
                        const uint32_t kj = __shfl_sync(0xffffffffu, mine, j);
                        greater += kj > mine; ge += kj >= mine;
                    }
                    if (lane < small && greater < need && need <= ge) s_prefix = mine; // equal keys write the same value
                }
                __syncthreads();
                break;
            }
        }
        thr_key = s_prefix;
    }
    uint32_t *kxy = g.kept_xy + (long long)f * g.list_fs + L.list_off;
    float *kresp = g.kept_resp + (long long)f * g.list_fs + L.list_off;
    if (in_smem) {
        // compact the kept entries (position key + source index), then rank each by (y<<12|x) among the kept ones
        for (int i0 = 0; i0 < n; i0 += SORT_THREADS) {
            const int i = i0 + tid;
            const bool keep = i < n && s_key[i] >= thr_key;
            const uint32_t b = __ballot_sync(0xffffffffu, keep);
            if (b) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&s_nkept, __popc(b));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (keep) {
                    const int slot = base + __popc(b & ((1u << lane) - 1u));
                    s_cpos[slot] = s_pos[i];
                    s_cidx[slot] = (uint32_t)i;
                }
            }
        }
        __syncthreads();
        const int nk = s_nkept;
        // (y, x) order without the nk^2 compares: count the kept entries per image row, prefix-sum the rows, drop every entry
        // into its row's segment (any order), then rank it inside the segment (a handful of entries) by x
        const int nrows = L.h + 1; // entry L.h of the scan = nk: closes the last row's segment
        if (nrows <= rows_cap) {
            for (int i = tid; i < nrows; i += SORT_THREADS) s_rowa[i] = 0;
            __syncthreads();
            for (int e = tid; e < nk; e += SORT_THREADS) atomicAdd(&s_rowa[s_cpos[e] >> 12], 1);
            __syncthreads();
            {   // exclusive scan of s_rowa[0 .. nrows): thread = contiguous chunk, warp scan of the chunk sums, scan of the warp sums
                const int chunk = (nrows + SORT_THREADS - 1) / SORT_THREADS, i0 = tid * chunk, i1 = min(i0 + chunk, nrows);
                int sum = 0;
                for (int i = i0; i < i1; i++) sum += s_rowa[i];
                int incl = sum;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += v;
                }
                if (lane == 31) s_wsum[tid >> 5] = incl;
                __syncthreads();
                if (tid < 32) {
                    const int w = tid < SORT_THREADS / 32 ? s_wsum[tid] : 0;
                    int wi = w;
#pragma unroll
                    for (int d = 1; d < SORT_THREADS / 32; d <<= 1) {
                        const int v = __shfl_up_sync(0xffffffffu, wi, d);
                        if (lane >= d) wi += v;
                    }
                    if (tid < SORT_THREADS / 32) s_wsum[tid] = wi - w;
                }
                __syncthreads();
                int run = s_wsum[tid >> 5] + incl - sum;
                for (int i = i0; i < i1; i++) { const int c = s_rowa[i]; s_rowa[i] = run; s_rowb[i] = run; run += c; }
            }
            __syncthreads();
            for (int e = tid; e < nk; e += SORT_THREADS) s_pos[atomicAdd(&s_rowb[s_cpos[e] >> 12], 1)] = (uint32_t)e; // s_pos is free by now
            __syncthreads();
            for (int t = tid; t < nk; t += SORT_THREADS) {
                const int e = (int)s_pos[t];
                const uint32_t pi = s_cpos[e];
                const int y = (int)(pi >> 12), seg0 = s_rowa[y], seg1 = s_rowa[y + 1];
                int rank = seg0;
                for (int u = seg0; u < seg1; u++) rank += s_cpos[s_pos[u]] < pi;
                kxy[rank] = pi;
                kresp[rank] = kfs_key_float(s_key[s_cidx[e]]);
            }
        } else { // more rows than the row arrays hold (cannot happen for heights the position encoding allows): nk^2 compares
            for (int e = tid; e < nk; e += SORT_THREADS) {
                const uint32_t pi = s_cpos[e];
                int rank = 0;
                for (int j = 0; j < nk; j++) rank += s_cpos[j] < pi;
                kxy[rank] = pi;
                kresp[rank] = kfs_key_float(s_key[s_cidx[e]]);
            }
        }
        if (tid == 0) g.kept_count[f * KFS_MAX_LEVELS + l] = nk;
        return;
    }
    // oversized candidate list (ties far beyond the slack): same result straight from global memory
    int mine = 0;
    for (int i = tid; i < n; i += SORT_THREADS) {
        const uint32_t ki = kfs_float_key(resp[i]);
        if (ki < thr_key) continue;
        const uint32_t pi = cand[i] & 0xffffffu;
        int rank = 0;
        for (int j = 0; j < n; j++) rank += (kfs_float_key(resp[j]) >= thr_key && (cand[j] & 0xffffffu) < pi) ? 1 : 0;
        kxy[rank] = pi;
        kresp[rank] = resp[i];
        mine++;
    }
    if (mine) atomicAdd(&s_nkept, mine);
    __syncthreads();
    if (tid == 0) g.kept_count[f * KFS_MAX_LEVELS + l] = s_nkept;
}

// ======================================================================================================
// a1.9  7x7 sigma=2 blur of every level (OpenCV's sepFilter2D float path, exact FMA order: kfs_blur_row / _col)
// ======================================================================================================
// One CTA = one 128 x 58 tile of one level of one frame (tile table like k_fast's), 256 threads.
//   source tile 160 x 64 bytes (origin (x0-16, y0-3): rows are 16-byte aligned -> LDG.128)
//   row pass   : thread = (row PAIR, 16-pixel segment): 2 x 22 source bytes -> I2F.U8 with byte select; the two rows ride
//                in the two halves of packed fp32x2 registers, so the 16 x 7 taps are 112 FFMA2 (fma.rn.f32x2: two exact
//                IEEE FMAs per issue slot — the kernel is issue-bound, the FMA pipe itself is not) -> 2 x 4 STS.128
//                into the 64 x 128 float tile
//   column pass: thread = (4 adjacent columns, 7..8-row segment): LDS.128 down the column block = two packed column
//                pairs, sliding 7-row window in registers, the symmetric taps as ADD2 + FMA2, round-half-even by the
//                1.5*2^23 magic add (no F2I), 4 pixels packed -> one STG.32; a warp writes 128 contiguous bytes per row
// Every lane of a packed operation evaluates exactly kfs_blur_row / kfs_blur_col_bits (kfs_math.cuh, checked on the CPU
// against the oracle): same operands, same order, same rounding.
// Borders: source rows/columns outside the image are clamped / zero-filled instead of REFLECT_101: a descriptor
// sample lies >= 31-19 = 12 px inside the level and its 7x7 support >= 9 px inside, so those pixels of the blurred
// level are never read (rotated BRIEF reaches <= 18.4 px from a keypoint that is >= 31 px inside).
constexpr int BT_W = 128, BT_H = 58, BT_THREADS = 256; // 58 + 6 halo rows = 64 = 32 row pairs x 8 segments = 256 row-pass items
constexpr int BT_Y0 = 12;                               // first blurred row / rows [12, H-12) are the only ones ever sampled
constexpr int BS_P = 160, BS_H = BT_H + 6;
static_assert(BS_H == 64 && BT_THREADS == 256, "one row-pass item per thread");

typedef unsigned long long kfs_f2; // two fp32 lanes in one 64-bit register pair
__device__ __forceinline__ kfs_f2 kfs_pk2(float lo, float hi) { kfs_f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void kfs_unpk2(kfs_f2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ kfs_f2 kfs_mul2(kfs_f2 a, kfs_f2 b) { kfs_f2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ kfs_f2 kfs_add2(kfs_f2 a, kfs_f2 b) { kfs_f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ kfs_f2 kfs_fma2(kfs_f2 a, kfs_f2 b, kfs_f2 c) { kfs_f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
// g*p + r as OpenCV's FMA dispatch evaluates it (one rounding), or as its plain-C++ build does (cv::setUseOptimized(false) or
// a host without FMA: product rounded, then the sum) - SURVEY a1.9 "dispatch-dependent"
template <bool FMA> __device__ __forceinline__ kfs_f2 kfs_mad2(kfs_f2 g, kfs_f2 p, kfs_f2 r)
{
    if (FMA) return kfs_fma2(g, p, r);
    // scalar on purpose: ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2 (even with --fmad=false), which would turn
    // this variant back into the fused one; scalar __fmul_rn / __fadd_rn are never contracted
    float gl, gh, pl, ph, rl, rh;
    kfs_unpk2(g, gl, gh); kfs_unpk2(p, pl, ph); kfs_unpk2(r, rl, rh);
    return kfs_pk2(__fadd_rn(__fmul_rn(gl, pl), rl), __fadd_rn(__fmul_rn(gh, ph), rh));
}
// kfs_blur_row on both lanes
template <bool FMA> __device__ __forceinline__ kfs_f2 kfs_blur_row2(kfs_f2 g0, kfs_f2 g1, kfs_f2 g2, kfs_f2 g3, kfs_f2 p0, kfs_f2 p1, kfs_f2 p2, kfs_f2 p3,
                                                kfs_f2 p4, kfs_f2 p5, kfs_f2 p6)
{
    kfs_f2 r = kfs_mul2(g0, p0);
    r = kfs_mad2<FMA>(g1, p1, r);
    r = kfs_mad2<FMA>(g2, p2, r);
    r = kfs_mad2<FMA>(g3, p3, r);
    r = kfs_mad2<FMA>(g2, p4, r);
    r = kfs_mad2<FMA>(g1, p5, r);
    r = kfs_mad2<FMA>(g0, p6, r);
    return r;
}
// kfs_blur_col_bits on both lanes: the bits of (s + 1.5*2^23), byte 0 of each half is the rounded pixel
template <bool FMA> __device__ __forceinline__ kfs_f2 kfs_blur_col2(kfs_f2 g0, kfs_f2 g1, kfs_f2 g2, kfs_f2 g3, kfs_f2 magic, kfs_f2 rm3, kfs_f2 rm2, kfs_f2 rm1,
                                                kfs_f2 r0, kfs_f2 rp1, kfs_f2 rp2, kfs_f2 rp3)
{
    kfs_f2 s = kfs_mul2(g3, r0);
    s = kfs_mad2<FMA>(g2, kfs_add2(rp1, rm1), s);
    s = kfs_mad2<FMA>(g1, kfs_add2(rp2, rm2), s);
    s = kfs_mad2<FMA>(g0, kfs_add2(rp3, rm3), s);
    return kfs_add2(s, magic);
}

template <bool FMA>
__global__ void __launch_bounds__(BT_THREADS)
k_blur7(const __grid_constant__ OrbDev g, const uint32_t *__restrict__ tile_tab, const __grid_constant__ FastMaps maps)
{
    __shared__ __align__(128) uint8_t s_src[BS_P * BS_H];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ __align__(16) float s_row[BS_H * BT_W];
    const int tid = threadIdx.x;
    const int f = blockIdx.y + g.frame0;
    const uint32_t tt = __ldg(tile_tab + blockIdx.x);
    const int l = tt >> 24, tiy = (tt >> 12) & 0xfff, tix = tt & 0xfff;
    const LevelDev &L = g.lv[l];
    const int H = L.h, pitch = L.pitch;
    const int x0 = BT_W * tix, y0 = BT_Y0 + BT_H * tiy;
    const uint8_t *img = L.img + (long long)f * L.img_fs;
    const kfs_f2 G0 = kfs_pk2(KFS_G0, KFS_G0), G1 = kfs_pk2(KFS_G1, KFS_G1), G2 = kfs_pk2(KFS_G2, KFS_G2), G3 = kfs_pk2(KFS_G3, KFS_G3);
    if (maps.enabled) { // one TMA box (bytes outside the level arrive as zeros: they only feed outputs that are never sampled)
        if (tid == 0) kfs_mbar_init(&s_bar, 1);
        __syncthreads();
        if (tid == 0) {
            kfs_mbar_expect_tx(&s_bar, BS_P * BS_H);
            kfs_tma_load_3d(s_src, &maps.m[l], x0 - 16, y0 - 3, f, &s_bar);
        }
        kfs_mbar_wait(&s_bar, 0);
    } else {
        for (int i = tid; i < BS_H * (BS_P / 16); i += BT_THREADS) {
            const int r = i / (BS_P / 16), c = i - r * (BS_P / 16);
            const int gy = min(max(y0 - 3 + r, 0), H - 1), gx = x0 - 16 + 16 * c;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (gx >= 0 && gx + 16 <= pitch) v = __ldg(reinterpret_cast<const uint4 *>(img + (long long)gy * pitch + gx));
            *reinterpret_cast<uint4 *>(s_src + r * BS_P + 16 * c) = v;
        }
        __syncthreads();
    }
    // ---- row pass: rows 2m (low lanes) and 2m+1 (high lanes), pixels [16 sg, 16 sg + 16)
    {
        const int m = tid >> 3, sg = tid & 7;
        const uint8_t *ba = s_src + (2 * m) * BS_P + 16 + 16 * sg, *bb = ba + BS_P;
        const uint32_t al = *reinterpret_cast<const uint32_t *>(ba - 4), ar = *reinterpret_cast<const uint32_t *>(ba + 16);
        const uint32_t bl = *reinterpret_cast<const uint32_t *>(bb - 4), br = *reinterpret_cast<const uint32_t *>(bb + 16);
        const uint4 am = *reinterpret_cast<const uint4 *>(ba), bm = *reinterpret_cast<const uint4 *>(bb);
        const uint32_t aw[6] = {al, am.x, am.y, am.z, am.w, ar}, bw[6] = {bl, bm.x, bm.y, bm.z, bm.w, br};
        kfs_f2 p[22]; // p[i] = pixel x - 3 + i of both rows = byte i + 1 of the 24-byte window
#pragma unroll
        for (int i = 0; i < 22; i++) {
            const int wi = (i + 1) >> 2, sh = 8 * ((i + 1) & 3);
            p[i] = kfs_pk2((float)((aw[wi] >> sh) & 0xffu), (float)((bw[wi] >> sh) & 0xffu));
        }
        // float4 slot c4 of a row lives at physical slot c4 ^ (c4 >> 3): the 8 lanes of a quarter warp (sg = 0..7, same k)
        // then hit 8 distinct 16-byte bank groups instead of 2; the column pass reads a permutation of a full row
        float4 *oa = reinterpret_cast<float4 *>(s_row + (2 * m) * BT_W), *ob = oa + BT_W / 4;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            float4 va, vb;
            kfs_unpk2(kfs_blur_row2<FMA>(G0, G1, G2, G3, p[4 * k + 0], p[4 * k + 1], p[4 * k + 2], p[4 * k + 3], p[4 * k + 4], p[4 * k + 5], p[4 * k + 6]), va.x, vb.x);
            kfs_unpk2(kfs_blur_row2<FMA>(G0, G1, G2, G3, p[4 * k + 1], p[4 * k + 2], p[4 * k + 3], p[4 * k + 4], p[4 * k + 5], p[4 * k + 6], p[4 * k + 7]), va.y, vb.y);
            kfs_unpk2(kfs_blur_row2<FMA>(G0, G1, G2, G3, p[4 * k + 2], p[4 * k + 3], p[4 * k + 4], p[4 * k + 5], p[4 * k + 6], p[4 * k + 7], p[4 * k + 8]), va.z, vb.z);
            kfs_unpk2(kfs_blur_row2<FMA>(G0, G1, G2, G3, p[4 * k + 3], p[4 * k + 4], p[4 * k + 5], p[4 * k + 6], p[4 * k + 7], p[4 * k + 8], p[4 * k + 9]), va.w, vb.w);
            const int slot = (4 * sg + k) ^ (sg >> 1);
            oa[slot] = va;
            ob[slot] = vb;
        }
    }
    __syncthreads();
    // ---- column pass: warp w owns output rows [r0, r0 + n): 8, 8, 7, 7, 7, 7, 7, 7; lane = 4 adjacent columns
    {
        const int cg = tid & 31, w = tid >> 5;
        const int r0 = w < 2 ? 8 * w : 16 + 7 * (w - 2), n = w < 2 ? 8 : 7;
        const ulonglong2 *col = reinterpret_cast<const ulonglong2 *>(s_row + r0 * BT_W) + (cg ^ ((cg >> 3) & 3));
        const kfs_f2 magic = kfs_pk2(12582912.f, 12582912.f);
        ulonglong2 q0 = col[0], q1 = col[BT_W / 4], q2 = col[2 * (BT_W / 4)], q3 = col[3 * (BT_W / 4)], q4 = col[4 * (BT_W / 4)],
                   q5 = col[5 * (BT_W / 4)];
        uint8_t *dst = L.blur + (long long)f * L.img_fs_blur + (long long)(y0 + r0) * L.pitch_blur + x0 + 4 * cg;
        const int nrows = min(n, H - BT_Y0 - (y0 + r0)); // rows at or below H - 12 are never sampled
        const long long dp = L.pitch_blur;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (k >= nrows) break;
            const ulonglong2 q6 = col[(k + 6) * (BT_W / 4)];
            const kfs_f2 ab = kfs_blur_col2<FMA>(G0, G1, G2, G3, magic, q0.x, q1.x, q2.x, q3.x, q4.x, q5.x, q6.x);
            const kfs_f2 cd = kfs_blur_col2<FMA>(G0, G1, G2, G3, magic, q0.y, q1.y, q2.y, q3.y, q4.y, q5.y, q6.y);
            // byte 0 of each magic-rounded float is the pixel
            const uint32_t a = (uint32_t)ab, b = (uint32_t)(ab >> 32), c = (uint32_t)cd, d = (uint32_t)(cd >> 32);
            *reinterpret_cast<uint32_t *>(dst) = __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
            dst += dp;
            q0 = q1; q1 = q2; q2 = q3; q3 = q4; q4 = q5; q5 = q6;
        }
    }
}

// ======================================================================================================
// a1.7 / a1.8 / a1.10  orientation + keypoint record + rotated BRIEF: one WARP per keypoint
// ======================================================================================================
// grid (ceil(cap_out/(DK*KPW)), frame), DK warps per CTA; a warp walks KPW keypoints (slots), software-pipelined:
//   0. the windows of keypoint i+1 (31 x 48 B un-blurred rows for the IC disc, 37 x 64 B of the blurred level for
//      BRIEF) travel HBM -> shared memory with cp.async (LDGSTS, 16-byte aligned chunks, zero-fill past the pitch)
//      while keypoint i is computed: no register staging, no exposed DRAM latency; lane = (row, 16-byte chunk) so
//      one warp instruction touches 8 rows
//   1. slot -> (level, rank) through the per-level kept counts (warp prefix), read (x, y, response) two rounds ahead
//   2. intensity-centroid moments: lane = (disc row v, chunk): 2 x 4 IDP.4A per 16 bytes against a weight table indexed
//      by (alignment shift, |v|, word): bytes = u (signed) for m10 and 0/1 for the row sum (m01 = v * row sum); the disc
//      mask |u| <= umax[|v|] is baked into the table.  Two REDUX.SUM finish the warp sums.
//      kfs_fast_atan2 (OpenCV's polynomial, no FMA)
//   3. fp64 sincos of the DK keypoints of the round by DK lanes; cv::KeyPoint record
//   4. rotated BRIEF: lane = pair 32h + lane (its 8 pairs stay in registers for all KPW keypoints), 8 ballots give the
//      8 descriptor words; sample coordinates are rounded half-even by a magic add whose constant already carries the
//      window origin and the shared-memory address, so a sample costs 4 FMUL + 2 FADD + 2 FADD + IMAD + IADD + LDS.U8
constexpr int DK = KFS_DK;    // warps (keypoints in flight) per CTA
constexpr int KPW = KFS_KPW;  // keypoints per warp, one after the other (the lane's sample pairs stay in registers); a call of fewer
                              // than KPW frames walks fewer (kernel argument kpw): one frame alone has 2000 warps instead of 250
constexpr int BW_P = 64;      // pitch of the blurred window
__device__ __align__(16) const float d_patf[1024] = {
#include "orb_pattern.inc"
};

// IC weight table [shift 0..15][|v| 0..15][word 0..11] for the 48 aligned bytes that cover one disc row:
// wu = byte-wise u (signed) where |u| <= umax[|v|], else 0; w1 = byte-wise 1 under the same mask.
struct IcTab { uint32_t wu[16 * 16 * 12], w1[16 * 16 * 12]; };
constexpr IcTab kfs_make_ictab()
{
    IcTab t{};
    const int umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
    for (int s = 0; s < 16; s++)
        for (int a = 0; a < 16; a++)
            for (int k = 0; k < 12; k++) {
                uint32_t wu = 0, w1 = 0;
                for (int j = 0; j < 4; j++) {
                    const int p = 4 * k + j - s, u = p - 15;
                    if (p < 0 || p > 30) continue;
                    if ((u < 0 ? -u : u) > umax[a]) continue;
                    wu |= (uint32_t)(u & 0xff) << (8 * j);
                    w1 |= 1u << (8 * j);
                }
                t.wu[(s * 16 + a) * 12 + k] = wu;
                t.w1[(s * 16 + a) * 12 + k] = w1;
            }
    return t;
}
__device__ __align__(16) const IcTab d_ictab = kfs_make_ictab();

__device__ __forceinline__ int kfs_dp4a_us(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ void kfs_cp_async16_zfill(void *smem, const void *gmem, bool pred)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    const int n = pred ? 16 : 0; // src-size 0: nothing is read, the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(n) : "memory");
}

constexpr int IC_BYTES = 31 * 48, WB_BYTES = 37 * BW_P;

__global__ void __launch_bounds__(32 * DK)
k_describe(const __grid_constant__ OrbDev g, int kpw)
{
    // per warp, double-buffered: the 31 x 48 B un-blurred rows of the IC disc and the 37 x 64 B blurred window
    __shared__ __align__(16) uint8_t s_win[DK][2][IC_BYTES + WB_BYTES];
    __shared__ float s_angle[DK];
    __shared__ float2 s_rot[DK];
    kfs_pdl_wait();
    const int f = blockIdx.y + g.frame0, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // per-level kept counts -> prefix
    int cnt = 0;
    if (lane < g.nlevels) cnt = min(g.kept_count[f * KFS_MAX_LEVELS + lane], g.lv[lane].list_cap);
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < KFS_MAX_LEVELS; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    const int total = __shfl_sync(0xffffffffu, incl, KFS_MAX_LEVELS - 1);
    if (blockIdx.x == 0 && threadIdx.x == 0) g.n_out[f] = total;
    const int nvalid = min(total, g.cap_out);
    if (blockIdx.x * (DK * kpw) >= nvalid) return; // whole CTA beyond the last keypoint
    // this lane's 8 sample pairs (pair 32h + lane), kept in registers for all KPW keypoints of the warp
    float4 pat[8];
#pragma unroll
    for (int h = 0; h < 8; h++) pat[h] = __ldg(reinterpret_cast<const float4 *>(d_patf) + 32 * h + lane);
    const int q = lane & 3, rq = lane >> 2;

    // keypoint of round `it` of this warp: slot -> (level, rank) -> (x, y, response); xy = ~0 when there is none
    auto load_meta = [&](int it, uint32_t &xy, float &resp, int &level) {
        xy = 0xffffffffu; resp = 0.f; level = 0;
        const int slot = (blockIdx.x * kpw + it) * DK + warp;
        if (it < kpw && slot < nvalid) { // warp-uniform
            level = __popc(__ballot_sync(0xffffffffu, incl <= slot) & ((1u << KFS_MAX_LEVELS) - 1u));
            const int excl = __shfl_sync(0xffffffffu, incl - cnt, level);
            const long long li = (long long)f * g.list_fs + g.lv[level].list_off + (slot - excl);
            xy = g.kept_xy[li];
            resp = g.kept_resp[li];
        }
    };
    // asynchronous copies (LDGSTS) of both windows of one keypoint: lane = (row rq + 8 i, 16-byte chunk q), so a warp
    // instruction touches 8 rows; the x origins are rounded down to 16 bytes, the remainders are kept as shifts
    auto issue = [&](uint32_t xy, int level, int buf) {
        if (xy != 0xffffffffu) {
            const LevelDev &L = g.lv[level];
            const int x = xy & 0xfff, y = (xy >> 12) & 0xfff;
            uint8_t *ic = s_win[warp][buf], *wb = ic + IC_BYTES;
            const int ax = (x - 15) & ~15;
            const bool in_i = ax + 16 * q + 16 <= L.pitch;
            const uint8_t *srci = L.img + (long long)f * L.img_fs + (long long)(y - 15 + rq) * L.pitch + ax + (in_i ? 16 * q : 0);
            if (q < 3) {
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (rq + 8 * i < 31) kfs_cp_async16_zfill(ic + (rq + 8 * i) * 48 + 16 * q, srci + (long long)(8 * i) * L.pitch, in_i);
            }
            const int axb = (x - 18) & ~15;
            const bool in_b = axb + 16 * q + 16 <= L.pitch_blur;
            const uint8_t *srb = L.blur + (long long)f * L.img_fs_blur + (long long)(y - 18 + rq) * L.pitch_blur + axb + (in_b ? 16 * q : 0);
#pragma unroll
            for (int i = 0; i < 5; i++)
                if (rq + 8 * i < 37) kfs_cp_async16_zfill(wb + (rq + 8 * i) * BW_P + 16 * q, srb + (long long)(8 * i) * L.pitch_blur, in_b);
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };

    uint32_t xy_c, xy_n, xy_nn;
    float resp_c, resp_n, resp_nn;
    int level_c, level_n, level_nn;
    load_meta(0, xy_c, resp_c, level_c);
    issue(xy_c, level_c, 0);
    load_meta(1, xy_n, resp_n, level_n);

#pragma unroll 1
    for (int it = 0; it < kpw; it++) {
        const int slot0 = (blockIdx.x * kpw + it) * DK; // first slot of this round (block-uniform)
        if (slot0 >= nvalid) break;
        const int buf = it & 1;
        issue(xy_n, level_n, buf ^ 1);           // next keypoint's windows fly while this one is computed
        load_meta(it + 2, xy_nn, resp_nn, level_nn);
        asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        __syncwarp();
        const bool valid = xy_c != 0xffffffffu;
        const int slot = slot0 + warp;
        const int x = xy_c & 0xfff, y = (xy_c >> 12) & 0xfff;
        const uint8_t *ic = s_win[warp][buf], *wb = ic + IC_BYTES;
        float angle = 0.f;
        if (valid) {
            // ---- IC moments (a1.7): 2 x 4 IDP.4A per 16-byte chunk against the (shift, |v|, word) weight table
            const int shi = (x - 15) & 15;
            int m10 = 0, m01 = 0;
            const int qc = min(q, 2);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int r = rq + 8 * i, v = r - 15, a = min(v < 0 ? -v : v, 15);
                uint4 px = make_uint4(0, 0, 0, 0);
                if (q < 3 && r < 31) px = *reinterpret_cast<const uint4 *>(ic + r * 48 + 16 * q);
                const int ti = (shi * 16 + a) * 12 + 4 * qc;
                const uint4 tu = __ldg(reinterpret_cast<const uint4 *>(d_ictab.wu + ti));
                const uint4 t1 = __ldg(reinterpret_cast<const uint4 *>(d_ictab.w1 + ti));
                m10 = kfs_dp4a_us(px.x, tu.x, m10); m10 = kfs_dp4a_us(px.y, tu.y, m10);
                m10 = kfs_dp4a_us(px.z, tu.z, m10); m10 = kfs_dp4a_us(px.w, tu.w, m10);
                int rs = kfs_dp4a_us(px.x, t1.x, 0); rs = kfs_dp4a_us(px.y, t1.y, rs);
                rs = kfs_dp4a_us(px.z, t1.z, rs); rs = kfs_dp4a_us(px.w, t1.w, rs);
                m01 += v * rs; // lanes with q == 3 or row > 30 hold zero pixels
            }
            m10 = __reduce_add_sync(0xffffffffu, m10);
            m01 = __reduce_add_sync(0xffffffffu, m01);
            angle = kfs_fast_atan2((float)m01, (float)m10);
            if (lane == 0) s_angle[warp] = angle;
        }
        __syncthreads();
        if (threadIdx.x < DK && slot0 + (int)threadIdx.x < nvalid) {
            const float rad = kfs_mul(s_angle[threadIdx.x], (float)(3.1415926535897932384626433832795 / 180.f));
            double sn, cs;
            sincos((double)rad, &sn, &cs);
            s_rot[threadIdx.x] = make_float2((float)cs, (float)sn);
        }
        __syncthreads();
        if (valid) {
            const float2 ab = s_rot[warp];
            const float sc = g.lv[level_c].scale;
            const long long o = (long long)f * g.cap_out + slot;
            if (lane == 0) {
                kfs_keypoint kp;
                kp.x = kfs_mul((float)x, sc);
                kp.y = kfs_mul((float)y, sc);
                kp.size = kfs_mul(31.f, sc);
                kp.angle = angle;
                kp.response = resp_c;
                kp.octave = level_c;
                kp.class_id = -1;
                g.kp[o] = kp;
            }
            // OpenCV recomputes the integer centre from the scaled float point: cvRound(pt * (1.f/scale)).  (x*s)*(1/s)
            // is within 1e-3 of x for x < 4096, so the rounded centre is (x, y) itself.
            // Sample address = wb + (iy + 18) * 64 + (ix + 18 + shb).  rint(v) + c = bits(v + (1.5*2^23 + c)) - 0x4B400000
            // for small even integer c, so the y magic carries +18, the x magic carries +18 + shb + the shared address of
            // wb, and  by * 64 + bx = 65 * 0x4B400000 + address  (mod 2^32): one IMAD and one constant subtraction.
            // (c must be even, or the tie rule of the magic add would flip to round-half-odd: the odd bit goes into kfold)
            const int shb = (x - 18) & 15;
            const int cx = 18 + shb + (int)__cvta_generic_to_shared(wb);
            const float mx = 12582912.f + (float)(cx & ~1), my = 12582912.f + 18.f;
            const uint32_t kfold = 0x4B400000u * 65u - (uint32_t)(cx & 1);
            uint32_t word = 0;
#pragma unroll
            for (int h = 0; h < 8; h++) {
                const float4 pt = pat[h];
                const uint32_t bx0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(pt.x, ab.x), __fmul_rn(pt.y, ab.y)), mx));
                const uint32_t by0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(pt.x, ab.y), __fmul_rn(pt.y, ab.x)), my));
                const uint32_t bx1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(pt.z, ab.x), __fmul_rn(pt.w, ab.y)), mx));
                const uint32_t by1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(pt.z, ab.y), __fmul_rn(pt.w, ab.x)), my));
                uint32_t v0, v1;
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v0) : "r"(by0 * 64u + bx0 - kfold));
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v1) : "r"(by1 * 64u + bx1 - kfold));
                const uint32_t b = __ballot_sync(0xffffffffu, v0 < v1);
                if (lane == h) word = b;
            }
            if (lane < 8) reinterpret_cast<uint32_t *>(g.desc + o * 32)[lane] = word;
        }
        // buffer reuse: the windows of round it+2 are issued (by this warp, in program order) after the ballots above have
        // consumed every read of this round's buffer; s_angle / s_rot reuse is ordered by the next round's two barriers
        xy_c = xy_n; resp_c = resp_n; level_c = level_n;
        xy_n = xy_nn; resp_n = resp_nn; level_n = level_nn;
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory"); // nothing may still be landing in shared memory at exit
}

// ------------------------------------------------------------------------------------------------------
// k_describe2: the same arithmetic in TWO PASSES per warp, no block-level barrier.
//   pass A  the warp's kpw keypoints one after the other: IC moments from the un-blurred window (double-buffered cp.async); the
//           sums of keypoint i stay in lane i
//   between lanes 0 .. kpw-1 run fastAtan2 and the fp64 sincos of THEIR keypoint side by side (one latency for kpw keypoints, where
//           k_describe pays one per keypoint and parks both warps of the CTA at two barriers around it) and write the cv::KeyPoint
//   pass B  rotated BRIEF of the kpw keypoints from the blurred windows (double-buffered as well; the first two fly during the end of
//           pass A), cos / sin / centre broadcast from lane i
// Slots: warp w of CTA b owns slots (b * DK + w) * kpw .. + kpw - 1, so a warp walks neighbouring keypoints of the (level, y, x) order.
// TMA: each window is ONE box (48 x 31 bytes of the level, 64 x 37 of its blurred copy; x origin rounded down to 16 bytes as before, so the
// arithmetic does not change) issued by lane 0 on a per-(warp, buffer) mbarrier, instead of 7 + 9 16-byte cp.async per lane with their
// address arithmetic (a fifth of the kernel's instructions).  Bytes past the row pitch arrive as zeros either way.
constexpr int IC_PAD = (IC_BYTES + 127) / 128 * 128, WB_PAD = (WB_BYTES + 127) / 128 * 128; // TMA destinations are 128-byte aligned
template <bool TMA>
__global__ void __launch_bounds__(32 * DK)
k_describe2(const __grid_constant__ OrbDev g, int kpw, const __grid_constant__ FastMaps maps_i, const __grid_constant__ FastMaps maps_b)
{
    __shared__ __align__(128) uint8_t s_ic[DK][2][IC_PAD];
    __shared__ __align__(128) uint8_t s_wb[DK][2][WB_PAD];
    __shared__ __align__(8) uint64_t s_dbar[DK][4]; // [un-blurred buffer 0, 1, blurred buffer 0, 1]
    kfs_pdl_wait();
    const int f = blockIdx.y + g.frame0, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int cnt = 0;
    if (lane < g.nlevels) cnt = min(g.kept_count[f * KFS_MAX_LEVELS + lane], g.lv[lane].list_cap);
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < KFS_MAX_LEVELS; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    const int total = __shfl_sync(0xffffffffu, incl, KFS_MAX_LEVELS - 1);
    if (blockIdx.x == 0 && threadIdx.x == 0) g.n_out[f] = total;
    const int nvalid = min(total, g.cap_out);
    const int slot0 = (blockIdx.x * DK + warp) * kpw; // this warp's first slot
    if (slot0 >= nvalid) return;                      // warp-uniform; no block-level barrier below
    const int nk = min(kpw, nvalid - slot0);          // keypoints of this warp (>= 1)
    const int q = lane & 3, rq = lane >> 2;
    if (TMA) {
        if (lane == 0) {
#pragma unroll
            for (int b = 0; b < 4; b++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(kfs_smem_u32(&s_dbar[warp][b])) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }

    // lane i < nk: its keypoint's level, list entry
    int level_l = 0;
    uint32_t xy_l = 0xffffffffu;
    float resp_l = 0.f;
    {
        const int slot = slot0 + lane;
        int lv = 0, excl = 0;
        for (int l = 0; l < g.nlevels; l++) {
            const int v = __shfl_sync(0xffffffffu, incl, l);
            if (v <= slot) { lv = l + 1; excl = v; }
        }
        if (lane < nk) {
            level_l = lv;
            const long long li = (long long)f * g.list_fs + g.lv[lv].list_off + (slot - excl);
            xy_l = g.kept_xy[li];
            resp_l = g.kept_resp[li];
        }
    }
    // job s of the warp: s < nk -> un-blurred window of keypoint s, else blurred window of keypoint s - nk; one commit group per job
    auto issue = [&](int s) {
        if (s < 2 * nk) {
            const int i = s < nk ? s : s - nk;
            const uint32_t xy = __shfl_sync(0xffffffffu, xy_l, i);
            const int level = __shfl_sync(0xffffffffu, level_l, i);
            const LevelDev &L = g.lv[level];
            const int x = xy & 0xfff, y = (xy >> 12) & 0xfff;
            if (TMA) {
                if (lane == 0) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // the buffer's last readers (generic proxy) are done
                    if (s < nk) {
                        kfs_mbar_expect_tx(&s_dbar[warp][i & 1], 48 * 31);
                        kfs_tma_load_3d(s_ic[warp][i & 1], &maps_i.m[level], (x - 15) & ~15, y - 15, f, &s_dbar[warp][i & 1]);
                    } else {
                        kfs_mbar_expect_tx(&s_dbar[warp][2 + (i & 1)], BW_P * 37);
                        kfs_tma_load_3d(s_wb[warp][i & 1], &maps_b.m[level], (x - 18) & ~15, y - 18, f, &s_dbar[warp][2 + (i & 1)]);
                    }
                }
            } else if (s < nk) {
                uint8_t *ic = s_ic[warp][i & 1];
                const int ax = (x - 15) & ~15;
                const bool in_i = ax + 16 * q + 16 <= L.pitch;
                const uint8_t *srci = L.img + (long long)f * L.img_fs + (long long)(y - 15 + rq) * L.pitch + ax + (in_i ? 16 * q : 0);
                if (q < 3) {
#pragma unroll
                    for (int r = 0; r < 4; r++)
                        if (rq + 8 * r < 31) kfs_cp_async16_zfill(ic + (rq + 8 * r) * 48 + 16 * q, srci + (long long)(8 * r) * L.pitch, in_i);
                }
            } else {
                uint8_t *wb = s_wb[warp][i & 1];
                const int axb = (x - 18) & ~15;
                const bool in_b = axb + 16 * q + 16 <= L.pitch_blur;
                const uint8_t *srb = L.blur + (long long)f * L.img_fs_blur + (long long)(y - 18 + rq) * L.pitch_blur + axb + (in_b ? 16 * q : 0);
#pragma unroll
                for (int r = 0; r < 5; r++)
                    if (rq + 8 * r < 37) kfs_cp_async16_zfill(wb + (rq + 8 * r) * BW_P + 16 * q, srb + (long long)(8 * r) * L.pitch_blur, in_b);
            }
        }
        if (!TMA) asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    issue(0);
    issue(1);

    // ---- pass A: IC moments (a1.7), 2 x 4 IDP.4A per 16-byte chunk against the (shift, |v|, word) weight table
    int m10_l = 0, m01_l = 0;
#pragma unroll 1
    for (int i = 0; i < nk; i++) {
        if (TMA) kfs_mbar_wait(&s_dbar[warp][i & 1], (uint32_t)((i >> 1) & 1)); // use (i >> 1) of this buffer
        else asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        __syncwarp();
        const int x = __shfl_sync(0xffffffffu, xy_l, i) & 0xfff;
        const uint8_t *ic = s_ic[warp][i & 1];
        const int shi = (x - 15) & 15;
        int m10 = 0, m01 = 0;
        const int qc = min(q, 2);
#pragma unroll
        for (int r4 = 0; r4 < 4; r4++) {
            const int r = rq + 8 * r4, v = r - 15, a = min(v < 0 ? -v : v, 15);
            uint4 px = make_uint4(0, 0, 0, 0);
            if (q < 3 && r < 31) px = *reinterpret_cast<const uint4 *>(ic + r * 48 + 16 * q);
            const int ti = (shi * 16 + a) * 12 + 4 * qc;
            const uint4 tu = __ldg(reinterpret_cast<const uint4 *>(d_ictab.wu + ti));
            const uint4 t1 = __ldg(reinterpret_cast<const uint4 *>(d_ictab.w1 + ti));
            m10 = kfs_dp4a_us(px.x, tu.x, m10); m10 = kfs_dp4a_us(px.y, tu.y, m10);
            m10 = kfs_dp4a_us(px.z, tu.z, m10); m10 = kfs_dp4a_us(px.w, tu.w, m10);
            int rs = kfs_dp4a_us(px.x, t1.x, 0); rs = kfs_dp4a_us(px.y, t1.y, rs);
            rs = kfs_dp4a_us(px.z, t1.z, rs); rs = kfs_dp4a_us(px.w, t1.w, rs);
            m01 += v * rs; // lanes with q == 3 or row > 30 hold zero pixels
        }
        m10 = __reduce_add_sync(0xffffffffu, m10);
        m01 = __reduce_add_sync(0xffffffffu, m01);
        if (lane == i) { m10_l = m10; m01_l = m01; }
        __syncwarp();      // every lane has read this buffer before job i + 2 lands in it
        issue(i + 2);
    }

    // ---- between the passes: lane i owns keypoint i
    float cs_l = 1.f, sn_l = 0.f;
    if (lane < nk) {
        const float angle = kfs_fast_atan2((float)m01_l, (float)m10_l);
        const float rad = kfs_mul(angle, (float)(3.1415926535897932384626433832795 / 180.f));
        double sn, cs;
        sincos((double)rad, &sn, &cs);
        cs_l = (float)cs; sn_l = (float)sn;
        const float sc = g.lv[level_l].scale;
        const int x = xy_l & 0xfff, y = (xy_l >> 12) & 0xfff;
        kfs_keypoint kp;
        kp.x = kfs_mul((float)x, sc);
        kp.y = kfs_mul((float)y, sc);
        kp.size = kfs_mul(31.f, sc);
        kp.angle = angle;
        kp.response = resp_l;
        kp.octave = level_l;
        kp.class_id = -1;
        g.kp[(long long)f * g.cap_out + slot0 + lane] = kp;
    }
    __syncwarp();

    // ---- pass B: rotated BRIEF (a1.10).  This lane's 8 sample pairs (pair 32h + lane) serve all nk keypoints.
    // Packed fp32x2: a sample (x, y) rides in one register pair; (x a, y b) and (x b, y a) are two mul.rn.f32x2, X = x a - y b and
    // Y = x b + y a two scalar adds ACROSS the halves (nothing a contraction could fuse), and the two magic adds one add.rn.f32x2:
    // 5 operations per sample instead of 8, every one the same IEEE operation on the same operands as kfs_brief_rot.
    ulonglong2 pat[8];
#pragma unroll
    for (int h = 0; h < 8; h++) pat[h] = __ldg(reinterpret_cast<const ulonglong2 *>(d_patf) + 32 * h + lane);
#pragma unroll 1
    for (int i = 0; i < nk; i++) {
        if (TMA) kfs_mbar_wait(&s_dbar[warp][2 + (i & 1)], (uint32_t)((i >> 1) & 1)); // use (i >> 1) of this buffer
        else asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        __syncwarp();
        const int x = __shfl_sync(0xffffffffu, xy_l, i) & 0xfff;
        const float ca = __shfl_sync(0xffffffffu, cs_l, i), sb = __shfl_sync(0xffffffffu, sn_l, i);
        const kfs_f2 AB = kfs_pk2(ca, sb), BA = kfs_pk2(sb, ca);
        const uint8_t *wb = s_wb[warp][i & 1];
        // see k_describe: the magic-add constants carry the window origin and the shared-memory address of wb
        const int shb = (x - 18) & 15;
        const int cx = 18 + shb + (int)__cvta_generic_to_shared(wb);
        const kfs_f2 M = kfs_pk2(12582912.f + (float)(cx & ~1), 12582912.f + 18.f);
        const uint32_t kfold = 0x4B400000u * 65u - (uint32_t)(cx & 1);
        uint32_t word = 0;
        auto sample = [&](kfs_f2 p) -> uint32_t { // shared-memory address of the rotated, rounded sample
            float xa, yb, xb, ya;
            kfs_unpk2(kfs_mul2(p, AB), xa, yb);
            kfs_unpk2(kfs_mul2(p, BA), xb, ya);
            const kfs_f2 r = kfs_add2(kfs_pk2(__fsub_rn(xa, yb), __fadd_rn(xb, ya)), M);
            return (uint32_t)(r >> 32) * 64u + (uint32_t)r - kfold;
        };
#pragma unroll
        for (int h = 0; h < 8; h++) {
            uint32_t v0, v1;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v0) : "r"(sample(pat[h].x)));
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v1) : "r"(sample(pat[h].y)));
            const uint32_t b = __ballot_sync(0xffffffffu, v0 < v1);
            if (lane == h) word = b;
        }
        if (lane < 8) reinterpret_cast<uint32_t *>(g.desc + ((long long)f * g.cap_out + slot0 + i) * 32)[lane] = word;
        __syncwarp();      // (the ballots already ordered every read of this buffer before the next job lands in it)
        issue(nk + i + 2);
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory"); // nothing may still be landing in shared memory at exit
}

// ======================================================================================================
// host side
// ======================================================================================================
// One captured kernel chain: everything kfs_orb_run_eager's launches depend on, and the instantiated graph.
struct OrbGraph {
    OrbDev g;
    int nb;
    const uint8_t *mask;
    size_t mask_step, mask_fs;
    bool first;
    cudaStream_t stream;
    int seen;                 // calls with this key so far (the graph is captured on the second one)
    int nlaunch;              // kernel launches inside
    cudaGraphExec_t exec;
};

struct PyrMapKeyStore { const uint8_t *base; long long fs; int pitch, h, bw, bh; };
// Tensor maps of one kernel's tile shape + what they were encoded from
struct MapCache {
    FastMaps maps;
    const uint8_t *img[KFS_MAX_LEVELS] = {};
    long long fs[KFS_MAX_LEVELS] = {};
    int pitch[KFS_MAX_LEVELS] = {}, h[KFS_MAX_LEVELS] = {};
};

struct kfs_orb {
    kfs_orb_params p;
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    // geometry currently allocated for
    int W = 0, H = 0, B = 0;
    bool has_mask_alloc = false;
    int cap_out = 0;
    int nlevels_active = 0;
    int lw[KFS_MAX_LEVELS], lh[KFS_MAX_LEVELS], lpitch[KFS_MAX_LEVELS], quota[KFS_MAX_LEVELS];
    float lscale[KFS_MAX_LEVELS];
    long long lfs[KFS_MAX_LEVELS];
    uint8_t *pyr[KFS_MAX_LEVELS];   // internal level buffers (level 0 included)
    uint8_t *bpyr[KFS_MAX_LEVELS];  // blurred levels (a1.9)
    uint32_t *blur_tab = nullptr;   // blur tile -> level<<24 | tiy<<12 | tix
    int blur_tiles = 0;
    int sm_count = 0;
    uint8_t *mpyr[KFS_MAX_LEVELS];  // mask pyramid
    uint32_t *xtab[KFS_MAX_LEVELS], *ytab[KFS_MAX_LEVELS];
    uint4 *xtab4[KFS_MAX_LEVELS];   // per 4-pixel group tables of the fast resize path
    bool resize_fast[KFS_MAX_LEVELS];
    void *arena = nullptr;          // one cudaMalloc for everything above and below
    size_t arena_bytes = 0;
    int *counters = nullptr;        // raw_count | cand_count | kept_count | n_out
    uint32_t *tile_tab = nullptr;   // FAST tile -> level<<24 | tiy<<12 | tix
    uint32_t *tile_tab_h = nullptr; // the same for tiles of height FT_HH (calls of one or two frames)
    int total_tiles_h = 0;
    bool fast_half = !(getenv("KFS_ORB_FAST_HALF") && atoi(getenv("KFS_ORB_FAST_HALF")) == 0);
    size_t counters_bytes = 0;
    OrbDev g;
    int last_batch = 0;
    int launches = 0;
    bool profiling = false;
    // CUDA graphs of the kernel chain, one per distinct launch configuration (see kfs_orb_run)
    std::vector<OrbGraph> graphs;
    bool use_graphs = true;
    bool blur_fma = true;           // a1.9: OpenCV's default (FMA) dispatch of sepFilter2D; false = its plain mul + add build
    // TMA descriptors of the levels for k_fast (re-encoded only when a level's address / pitch / size changes)
    MapCache fmaps, fmaps_h, bmaps; // k_fast's 160 x 70 (160 x 39 for half-height tiles) boxes, k_blur7's 160 x 64 boxes
    MapCache dmaps_i, dmaps_b;      // k_describe2's window boxes: 48 x 31 over the levels, 64 x 37 over their blurred copies
    MapCache rmaps, rmaps_m;        // k_resize_tma's source boxes over the image / mask pyramid (map l = level l as a SOURCE)
    int rz_bw = 0, rz_bh = 0;       // box of k_resize_tma (0: not planned / does not fit -> k_resize)
    bool resize_tma = !(getenv("KFS_ORB_RESIZE_TMA") && atoi(getenv("KFS_ORB_RESIZE_TMA")) == 0);
    // fused pyramid (k_pyramid): host copies of the resize tables, the region plan, level-0 tensor maps of image / mask
    std::vector<uint32_t> h_xtab[KFS_MAX_LEVELS], h_ytab[KFS_MAX_LEVELS];
    bool pyr_fused = false;
    bool use_pdl = !(getenv("KFS_ORB_PDL") && atoi(getenv("KFS_ORB_PDL")) == 0);
    bool describe2 = !(getenv("KFS_ORB_DESCRIBE2") && atoi(getenv("KFS_ORB_DESCRIBE2")) == 0); // two-pass k_describe2 (0: k_describe)
    int pyr_mode = 1;               // KFS_ORB_FUSED_PYRAMID: 0 = never, 1 = calls of at most PYR_FUSED_MAX_FRAMES frames, 2 = always
    PyrDev pyr_dev;
    short *d_rects = nullptr;
    int pyr_tiles = 0, pyr_box_w = 0, pyr_box_h = 0;
    size_t pyr_smem = 0;
    PyrMap pyr_map_img, pyr_map_mask;
    PyrMapKeyStore pyr_key_img = {}, pyr_key_mask = {};
    int tma_state = 0;              // 0 = not probed yet, 1 = cuTensorMapEncodeTiled available, -1 = not (LDG path)
    cudaEvent_t ev[KFS_STAGE_COUNT + 1];
    bool ev_made = false;
    float stage_ms[KFS_STAGE_COUNT];
    int *h_counts = nullptr; // pinned staging for n_out
    // host-buffer pipeline (kfs_orb_detect_and_compute_batch): H2D / kernels / D2H of frame chunks on three streams
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaStream_t s_blur = nullptr;  // side stream of the blur (fork after the pyramid, join before k_describe)
    cudaEvent_t ev_pyr = nullptr, ev_blur = nullptr;
    std::vector<cudaEvent_t> ev_in, ev_cmp;
    cudaEvent_t ev_start = nullptr;
    int pipeline_chunk = 0;       // frames per chunk; 0 = auto
    bool start_fence = true;      // copy streams wait for work already queued on the compute stream (off: replay executor)
    bool alias_valid = false;     // the last host batch is complete and nothing has run since: its descriptors are still in g.desc
    std::vector<uint8_t> alias_sig; // per frame of the last host batch: its first and last live descriptor row as delivered (2 x 32 B)
    KfsHostBatch pending = {};    // the host batch in flight (enqueue -> finish)
    int pending_rows = 0;         // rows per frame its eager copies carried
    uint8_t *rgb_stage = nullptr; // colour frames of a host batch ([batch][h][align16(3w)]), converted into level 0 on the device
    size_t rgb_stage_bytes = 0;
    uint8_t *mask_stage = nullptr; // raw host mask staged on the device
    size_t mask_stage_bytes = 0;
};

static int kfs_cvround(double v) { return (int)lrint(v); }

static void kfs_axis_table(int src, int dst, int padded, std::vector<uint32_t> &tab)
{
    tab.assign((size_t)padded, (uint32_t)(src - 1));
    const double scale = 1.0 / ((double)dst / (double)src);
    for (int d = 0; d < dst; d++) {
        const double fv = scale * ((double)d + 0.5) - 0.5;
        const int i = (int)floor(fv);
        uint32_t ofs, m1;
        if (i < 0 || src <= 1) { ofs = 0; m1 = 0; }
        else if (i >= src - 1) { ofs = (uint32_t)(src - 1); m1 = 0; }
        else { ofs = (uint32_t)i; m1 = (uint32_t)lrint((fv - (double)i) * 256.0); }
        tab[(size_t)d] = ofs | (m1 << 16);
    }
}

static void kfs_orb_drop_graphs(kfs_orb *o)
{
    for (OrbGraph &G : o->graphs) if (G.exec) cudaGraphExecDestroy(G.exec);
    o->graphs.clear();
}

static void kfs_orb_free_arena(kfs_orb *o)
{
    kfs_orb_drop_graphs(o);
    if (o->d_rects) { cudaFree(o->d_rects); o->d_rects = nullptr; }
    o->pyr_fused = false;
    if (o->arena) { cudaFree(o->arena); o->arena = nullptr; }
    o->alias_valid = false;
    o->W = o->H = o->B = 0;
}

static kfs_status kfs_orb_plan_pyramid(kfs_orb *o);
static void kfs_orb_plan_resize_boxes(kfs_orb *o);
static void kfs_orb_tma_probe(kfs_orb *o);

static kfs_status kfs_orb_ensure(kfs_orb *o, int w, int h, int batch, bool need_mask)
{
    if (o->arena && o->W == w && o->H == h && o->B >= batch && (o->has_mask_alloc || !need_mask)) return KFS_OK;
    if (o->arena) { KFS_CUDA_TRY(cudaStreamSynchronize(o->stream)); kfs_orb_free_arena(o); }
    const kfs_orb_params &P = o->p;
    // geometry (SURVEY §8 header, a1.1): scale_l = (float)pow((double)sf, l); size_l = cvRound(W * (1.f/scale_l))
    int L = P.nlevels;
    {
        const float factor = (float)(1.0 / P.scale_factor);
        float ndes = P.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)P.nlevels));
        int sum = 0;
        for (int l = 0; l < P.nlevels - 1; l++) {
            o->quota[l] = kfs_cvround((double)ndes);
            sum += o->quota[l];
            ndes *= factor;
        }
        o->quota[P.nlevels - 1] = P.nfeatures - sum > 0 ? P.nfeatures - sum : 0;
    }
    for (int l = 0; l < P.nlevels; l++) {
        o->lscale[l] = (float)pow((double)P.scale_factor, (double)l);
        const float inv = 1.f / o->lscale[l];
        o->lw[l] = kfs_cvround((double)((float)w * inv));
        o->lh[l] = kfs_cvround((double)((float)h * inv));
        if (o->lw[l] < 1 || o->lh[l] < 1) { L = l; break; }
        o->lpitch[l] = (int)kfs_align_up((size_t)o->lw[l], 128);
        o->lfs[l] = (long long)o->lpitch[l] * o->lh[l];
    }
    o->nlevels_active = L;
    // arena layout
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t r = off; off = kfs_align_up(off + bytes, 256); return r; };
    size_t o_pyr[KFS_MAX_LEVELS], o_bpyr[KFS_MAX_LEVELS], o_mpyr[KFS_MAX_LEVELS], o_xt[KFS_MAX_LEVELS], o_xt4[KFS_MAX_LEVELS], o_yt[KFS_MAX_LEVELS];
    int list_off[KFS_MAX_LEVELS], list_cap[KFS_MAX_LEVELS], list_fs = 0;
    for (int l = 0; l < L; l++) {
        o_pyr[l] = take((size_t)o->lfs[l] * batch + 256);
        o_bpyr[l] = take((size_t)o->lfs[l] * batch + 256);
        o_mpyr[l] = need_mask ? take((size_t)o->lfs[l] * batch + 256) : 0;
        o_xt[l] = take(kfs_align_up((size_t)o->lw[l], 4) * 4);
        o_xt4[l] = take(kfs_align_up((size_t)o->lw[l], 4) * 8);
        o_yt[l] = take((size_t)o->lh[l] * 4);
        const int W = o->lw[l], H = o->lh[l];
        list_cap[l] = (W > 2 * KFS_BORDER && H > 2 * KFS_BORDER) ? ((W - 62) / 2 + 2) * ((H - 62) / 2 + 2) : 0;
        list_off[l] = list_fs;
        list_fs += (list_cap[l] + 63) / 64 * 64;
    }
    const size_t lists = (size_t)list_fs * batch;
    const size_t o_raw = take(lists * 4), o_cxy = take(lists * 4), o_cr = take(lists * 4), o_kxy = take(lists * 4), o_kr = take(lists * 4);
    const size_t cnt_bytes = ((size_t)batch * KFS_MAX_LEVELS * 3 + (size_t)batch) * sizeof(int);
    const size_t o_cnt = take(cnt_bytes);
    int n_tiles_total = 0;
    for (int l = 0; l < L; l++)
        if (list_cap[l] > 0) n_tiles_total += kfs_div_up(o->lw[l] - KFS_BORDER - FAST_X0, FT_W) * kfs_div_up(o->lh[l] - 2 * KFS_BORDER, FT_H);
    const size_t o_tt = take((size_t)(n_tiles_total + 1) * 4);
    int n_tiles_half = 0;
    for (int l = 0; l < L; l++)
        if (list_cap[l] > 0) n_tiles_half += kfs_div_up(o->lw[l] - KFS_BORDER - FAST_X0, FT_W) * kfs_div_up(o->lh[l] - 2 * KFS_BORDER, FT_HH);
    const size_t o_tth = take((size_t)(n_tiles_half + 1) * 4);
    int n_blur_tiles = 0;
    for (int l = 0; l < L; l++)
        if (list_cap[l] > 0) n_blur_tiles += kfs_div_up(o->lw[l], BT_W) * kfs_div_up(o->lh[l] - 2 * BT_Y0, BT_H);
    const size_t o_bt = take((size_t)(n_blur_tiles + 1) * 4);
    const size_t outs = (size_t)batch * o->cap_out;
    const size_t o_kp = take(outs * sizeof(kfs_keypoint)), o_desc = take(outs * 32), o_rot = take(outs * sizeof(float2)), o_meta = take(outs * 4);
    KFS_CUDA_TRY(cudaMalloc(&o->arena, off));
    o->arena_bytes = off;
    uint8_t *base = (uint8_t *)o->arena;
    KFS_CUDA_TRY(cudaMemsetAsync(base, 0, off, o->stream));
    OrbDev &g = o->g;
    memset(&g, 0, sizeof(g));
    g.nlevels = L; g.batch = batch; g.fast_thr = P.fast_threshold; g.cap_out = o->cap_out; g.list_fs = list_fs;
    g.raw = (uint32_t *)(base + o_raw); g.cand_xy = (uint32_t *)(base + o_cxy); g.cand_resp = (float *)(base + o_cr);
    g.kept_xy = (uint32_t *)(base + o_kxy); g.kept_resp = (float *)(base + o_kr);
    o->counters = (int *)(base + o_cnt); o->counters_bytes = cnt_bytes;
    g.raw_count = o->counters; g.cand_count = g.raw_count + (size_t)batch * KFS_MAX_LEVELS;
    g.kept_count = g.cand_count + (size_t)batch * KFS_MAX_LEVELS; g.n_out = g.kept_count + (size_t)batch * KFS_MAX_LEVELS;
    g.kp = (kfs_keypoint *)(base + o_kp); g.desc = base + o_desc; g.rot = (float2 *)(base + o_rot); g.slot_meta = (uint32_t *)(base + o_meta);
    int tile_base = 0;
    std::vector<uint32_t> tab, ttab, ttab_h, btab;
    o->tile_tab = (uint32_t *)(base + o_tt);
    o->tile_tab_h = (uint32_t *)(base + o_tth);
    o->blur_tab = (uint32_t *)(base + o_bt);
    for (int l = 0; l < L; l++) {
        o->pyr[l] = base + o_pyr[l];
        o->bpyr[l] = base + o_bpyr[l];
        o->mpyr[l] = need_mask ? base + o_mpyr[l] : nullptr;
        o->xtab[l] = (uint32_t *)(base + o_xt[l]);
        o->xtab4[l] = (uint4 *)(base + o_xt4[l]);
        o->resize_fast[l] = false;
        o->ytab[l] = (uint32_t *)(base + o_yt[l]);
        LevelDev &D = g.lv[l];
        D.img = o->pyr[l]; D.img_fs = o->lfs[l]; D.pitch = o->lpitch[l];
        D.mask = nullptr; D.mask_fs = o->lfs[l]; D.mask_pitch = o->lpitch[l];
        D.blur = o->bpyr[l]; D.img_fs_blur = o->lfs[l]; D.pitch_blur = o->lpitch[l];
        D.w = o->lw[l]; D.h = o->lh[l]; D.quota = o->quota[l]; D.scale = o->lscale[l];
        D.list_off = list_off[l]; D.list_cap = list_cap[l];
        if (list_cap[l] > 0) {
            D.tiles_x = kfs_div_up(D.w - KFS_BORDER - FAST_X0, FT_W);
            D.tiles_y = kfs_div_up(D.h - 2 * KFS_BORDER, FT_H);
        } else { D.tiles_x = 0; D.tiles_y = 0; }
        if (list_cap[l] > 0) // levels too small to hold a keypoint are never sampled
            for (int ty = 0; ty < kfs_div_up(D.h - 2 * BT_Y0, BT_H); ty++)
                for (int tx = 0; tx < kfs_div_up(D.w, BT_W); tx++) btab.push_back(((uint32_t)l << 24) | ((uint32_t)ty << 12) | (uint32_t)tx);
        D.tile_base = tile_base;
        tile_base += D.tiles_x * D.tiles_y;
        for (int ty = 0; ty < D.tiles_y; ty++)
            for (int tx = 0; tx < D.tiles_x; tx++) ttab.push_back(((uint32_t)l << 24) | ((uint32_t)ty << 12) | (uint32_t)tx);
        if (list_cap[l] > 0)
            for (int ty = 0; ty < kfs_div_up(D.h - 2 * KFS_BORDER, FT_HH); ty++)
                for (int tx = 0; tx < D.tiles_x; tx++) ttab_h.push_back(((uint32_t)l << 24) | ((uint32_t)ty << 12) | (uint32_t)tx);
        if (l > 0) {
            kfs_axis_table(o->lw[l - 1], o->lw[l], (int)kfs_align_up((size_t)o->lw[l], 4), tab);
            o->h_xtab[l] = tab;
            KFS_CUDA_TRY(cudaMemcpyAsync(o->xtab[l], tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, o->stream));
            KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
            {   // group tables of the fast path: {a0, sel, 0, 0} {M0..M3}; usable when every group's taps fit 8 bytes
                const int sw = o->lw[l - 1], ngroups = (int)tab.size() / 4;
                std::vector<uint32_t> t4((size_t)ngroups * 8, 0u);
                bool ok = true;
                for (int gi = 0; gi < ngroups; gi++) {
                    const int a0 = (int)(tab[(size_t)4 * gi] & 0xffffu);
                    uint32_t sel = 0;
                    for (int j = 0; j < 4; j++) {
                        const uint32_t e = tab[(size_t)4 * gi + j];
                        const int a = (int)(e & 0xffffu), m1 = (int)(e >> 16);
                        const int b = a + 1 < sw ? a + 1 : sw - 1;
                        const int ra = a - a0, rb = b - a0;
                        if (ra < 0 || rb > 7) ok = false;
                        sel |= (uint32_t)((ra & 7) | ((rb & 7) << 4)) << (8 * j);
                        t4[(size_t)8 * gi + 4 + j] = (uint32_t)(256 - m1) | ((uint32_t)m1 << 16);
                    }
                    t4[(size_t)8 * gi + 0] = (uint32_t)a0;
                    t4[(size_t)8 * gi + 1] = sel;
                }
                o->resize_fast[l] = ok;
                KFS_CUDA_TRY(cudaMemcpyAsync(o->xtab4[l], t4.data(), t4.size() * 4, cudaMemcpyHostToDevice, o->stream));
                KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
            }
            kfs_axis_table(o->lh[l - 1], o->lh[l], o->lh[l], tab);
            o->h_ytab[l] = tab;
            KFS_CUDA_TRY(cudaMemcpyAsync(o->ytab[l], tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, o->stream));
            KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
        }
    }
    g.total_tiles = tile_base;
    o->total_tiles_h = (int)ttab_h.size();
    if (!ttab_h.empty()) KFS_CUDA_TRY(cudaMemcpyAsync(o->tile_tab_h, ttab_h.data(), ttab_h.size() * 4, cudaMemcpyHostToDevice, o->stream));
    if (!ttab.empty()) {
        KFS_CUDA_TRY(cudaMemcpyAsync(o->tile_tab, ttab.data(), ttab.size() * 4, cudaMemcpyHostToDevice, o->stream));
        KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    }
    o->blur_tiles = (int)btab.size();
    if (!btab.empty()) {
        KFS_CUDA_TRY(cudaMemcpyAsync(o->blur_tab, btab.data(), btab.size() * 4, cudaMemcpyHostToDevice, o->stream));
        KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    }
    if (o->h_counts) { cudaFreeHost(o->h_counts); o->h_counts = nullptr; }
    KFS_CUDA_TRY(cudaMallocHost(&o->h_counts, sizeof(int) * (size_t)batch));
    o->W = w; o->H = h; o->B = batch; o->has_mask_alloc = need_mask;
    kfs_orb_plan_resize_boxes(o);
    o->rmaps = MapCache(); o->rmaps_m = MapCache();
    return kfs_orb_plan_pyramid(o);
}

// Launch with the programmatic-serialization attribute: the kernel may start while its predecessor in the stream still runs and
// synchronises inside (kfs_pdl_wait).  Captured into the CUDA graph as a programmatic edge.  KFS_ORB_PDL=0: plain launches.
template <typename... KArgs, typename... Args>
static void kfs_launch_pdl(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...); // the error, if any, is picked up by KFS_LAUNCH_CHECK
}

#define KFS_LAUNCH_CHECK(o)                                                                     \
    do {                                                                                        \
        cudaError_t e__ = cudaGetLastError();                                                   \
        if (e__ != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "kernel launch failed at %s:%d: %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
        (o)->launches++;                                                                        \
    } while (0)

static kfs_status kfs_stage_mark(kfs_orb *o, int idx)
{
    if (!o->profiling) return KFS_OK;
    if (!o->ev_made) {
        for (int i = 0; i <= KFS_STAGE_COUNT; i++) KFS_CUDA_TRY(cudaEventCreate(&o->ev[i]));
        o->ev_made = true;
    }
    KFS_CUDA_TRY(cudaEventRecord(o->ev[idx], o->stream));
    return KFS_OK;
}

// cuTensorMapEncodeTiled is a driver-API function: looked up through the runtime so that libcuda is not a link dependency
// (this library also has to build, and load for the ABI tests, on machines without a driver).
typedef CUresult (*kfs_encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                        const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static kfs_encode_tiled_fn kfs_encode_tiled()
{
    // function-local static: initialised exactly once, also when several host threads create their handles at the same time
    static const kfs_encode_tiled_fn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            return (kfs_encode_tiled_fn)p;
        cudaGetLastError();
        return (kfs_encode_tiled_fn) nullptr;
    }();
    return fn;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-function, per-DEVICE limit shared by every handle of the process: both
// kernels that may ask for more than 48 KB get the device's opt-in maximum once per device and it is never lowered again
// (a per-handle cache of "what I set last" let a second handle with a smaller request pull the limit below the first one's).
static cudaError_t kfs_orb_raise_smem_limits(int device)
{
    static std::mutex mu;
    static bool done[64] = {};
    std::lock_guard<std::mutex> lock(mu);
    if (device >= 0 && device < 64 && done[device]) return cudaSuccess;
    int optin = 0;
    cudaError_t e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    // the opt-in maximum covers static + dynamic shared memory of a block: what a kernel may ask for dynamically is the rest
    cudaFuncAttributes fa;
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, k_pyramid);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_pyramid, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, k_select_sort);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_select_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);
    if (e == cudaSuccess && device >= 0 && device < 64) done[device] = true;
    return e;
}

// Box of k_resize_tma: the smallest (width, height) that holds the taps of every 128 x 32 output tile of every level, read off the
// host copies of the resize tables.  A tile's box starts at the 16-byte aligned byte at or below the first tap of its first 4-pixel
// group and at the top source row of its first output row; k_resize's arithmetic reads three aligned words from the word that holds a
// group's first tap, and source rows y0 and min(y0 + 1, sh - 1).
static void kfs_orb_plan_resize_boxes(kfs_orb *o)
{
    o->rz_bw = o->rz_bh = 0;
    const int L = o->nlevels_active;
    int bw = 0, bh = 0;
    for (int l = 1; l < L; l++) {
        if (!o->resize_fast[l]) return; // scale factor > 2: the generic kernel
        const std::vector<uint32_t> &xt = o->h_xtab[l], &yt = o->h_ytab[l];
        const int ngroups = (int)xt.size() / 4, dh = o->lh[l], sh = o->lh[l - 1];
        for (int g0 = 0; g0 < ngroups; g0 += 32) {
            const int g1 = std::min(g0 + 31, ngroups - 1);
            const int xa0 = (int)(xt[(size_t)4 * g0] & 0xffffu) & ~15;
            int last = 0;
            for (int gi = g0; gi <= g1; gi++) last = std::max(last, ((int)(xt[(size_t)4 * gi] & 0xffffu) & ~3) + 12);
            bw = std::max(bw, last - xa0);
        }
        for (int y0 = 0; y0 < dh; y0 += 8 * RZ_ROWS * RZ_YT) {
            const int y1 = std::min(y0 + 8 * RZ_ROWS * RZ_YT - 1, dh - 1);
            const int ya0 = (int)(yt[(size_t)y0] & 0xffffu);
            int last = 0;
            for (int y = y0; y <= y1; y++) last = std::max(last, std::min((int)(yt[(size_t)y] & 0xffffu) + 1, sh - 1));
            bh = std::max(bh, last - ya0 + 1);
        }
    }
    bw = (int)kfs_align_up((size_t)bw, 16);
    if (bw <= 0 || bh <= 0 || bw > 256 || bh > 256 || bw * bh > 40 * 1024) return; // TMA box limits / shared memory: stay on k_resize
    o->rz_bw = bw; o->rz_bh = bh;
}

// Tensor maps of the pyramid levels AS SOURCES of the next level (image or mask pyramid), box = (rz_bw, rz_bh); a level whose buffer
// cannot be described (an unaligned caller buffer as level 0) switches the TMA resize off for the handle.
static bool kfs_orb_resize_maps(kfs_orb *o, MapCache &C, bool mask)
{
    kfs_orb_tma_probe(o);
    if (o->tma_state <= 0 || o->rz_bw <= 0) return false;
    const OrbDev &g = o->g;
    for (int l = 0; l + 1 < g.nlevels; l++) {
        const uint8_t *img = mask ? o->mpyr[l] : g.lv[l].img;
        const long long fs = mask ? o->lfs[l] : g.lv[l].img_fs;
        const int pitch = mask ? o->lpitch[l] : g.lv[l].pitch, h = o->lh[l];
        if (C.img[l] == img && C.fs[l] == fs && C.pitch[l] == pitch && C.h[l] == h) continue;
        const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)h, 65535};
        const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)fs};
        const cuuint32_t box[3] = {(cuuint32_t)o->rz_bw, (cuuint32_t)o->rz_bh, 1}, estr[3] = {1, 1, 1};
        const bool ok = img && ((uintptr_t)img % 16 == 0) && (pitch % 16 == 0) && (fs % 16 == 0) &&
                        kfs_encode_tiled()(&C.maps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void *)img, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
        if (!ok) { o->resize_tma = false; return false; }
        C.img[l] = img; C.fs[l] = fs; C.pitch[l] = pitch; C.h[l] = h;
    }
    return true;
}

// Regions of the fused pyramid kernel (k_pyramid): tile (i, j) of the coarsest level owns, on every level l, the box between
// the proportional boundaries bx(l, i) .. bx(l, i+1) (multiples of 4 in x) and by(l, j) .. by(l, j+1); what it has to COMPUTE on
// level l is the bounding box of that and of the pre-image of what it computes on level l+1 (read off the resize tables).
// Leaves o->pyr_fused false when a level-0 region does not fit one TMA box / the shared-memory budget (large scale factors).
static kfs_status kfs_orb_plan_pyramid(kfs_orb *o)
{
    o->pyr_fused = false;
    const int L = o->nlevels_active;
    const char *env = getenv("KFS_ORB_FUSED_PYRAMID");
    o->pyr_mode = env ? atoi(env) : 1;
    if (L < 2 || o->pyr_mode <= 0) return KFS_OK;
    const int TW = 48, TH = 16; // 720p: 8 x 13 tiles of the coarsest level; 40 x 20, 32 x 16 and 48 x 24 are within 10 % of it
    const int Wc = o->lw[L - 1], Hc = o->lh[L - 1];
    if (Wc < 1 || Hc < 1) return KFS_OK;
    const int nx = kfs_div_up(Wc, TW), ny = kfs_div_up(Hc, TH);
    auto bx = [&](int l, int i) { return i >= nx ? (int)kfs_align_up((size_t)o->lw[l], 4) : (int)(((long long)i * TW * o->lw[l] / Wc) & ~3LL); };
    auto by = [&](int l, int j) { return j >= ny ? o->lh[l] : (int)((long long)j * TH * o->lh[l] / Hc); };
    struct Box { int x0, y0, x1, y1; };
    std::vector<short> rects((size_t)nx * ny * L * 8, 0);
    int maxw[KFS_MAX_LEVELS] = {}, maxh[KFS_MAX_LEVELS] = {};
    for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++) {
            Box C[KFS_MAX_LEVELS], O[KFS_MAX_LEVELS];
            C[L - 1] = O[L - 1] = Box{bx(L - 1, i), by(L - 1, j), bx(L - 1, i + 1), by(L - 1, j + 1)};
            for (int l = L - 1; l >= 1; l--) {
                const std::vector<uint32_t> &xt = o->h_xtab[l], &yt = o->h_ytab[l];
                Box Pm;
                Pm.x0 = (int)(xt[(size_t)C[l].x0] & 0xffffu);
                Pm.x1 = std::min((int)(xt[(size_t)C[l].x1 - 1] & 0xffffu) + 1, o->lw[l - 1] - 1) + 1;
                Pm.y0 = (int)(yt[(size_t)C[l].y0] & 0xffffu);
                Pm.y1 = std::min((int)(yt[(size_t)C[l].y1 - 1] & 0xffffu) + 1, o->lh[l - 1] - 1) + 1;
                if (l - 1 >= 1) {
                    O[l - 1] = Box{bx(l - 1, i), by(l - 1, j), bx(l - 1, i + 1), by(l - 1, j + 1)};
                    Box U = O[l - 1];
                    if (U.x1 <= U.x0 || U.y1 <= U.y0) U = Pm;
                    U.x0 = std::min(U.x0, Pm.x0) & ~3; U.y0 = std::min(U.y0, Pm.y0);
                    U.x1 = std::min((std::max(U.x1, Pm.x1) + 3) & ~3, (int)kfs_align_up((size_t)o->lw[l - 1], 4));
                    U.y1 = std::max(U.y1, Pm.y1);
                    C[l - 1] = U;
                } else {
                    Pm.x0 &= ~15; // the TMA box has to start on a 16-byte boundary in global memory
                    C[0] = Pm; O[0] = Box{0, 0, 0, 0};
                }
            }
            for (int l = 0; l < L; l++) {
                short *r = &rects[(((size_t)j * nx + i) * L + l) * 8];
                r[0] = (short)C[l].x0; r[1] = (short)C[l].y0; r[2] = (short)(C[l].x1 - C[l].x0); r[3] = (short)(C[l].y1 - C[l].y0);
                r[4] = (short)O[l].x0; r[5] = (short)O[l].y0; r[6] = (short)std::max(O[l].x1 - O[l].x0, 0); r[7] = (short)std::max(O[l].y1 - O[l].y0, 0);
                if (r[2] <= 0 || r[3] <= 0) return KFS_OK; // degenerate geometry: keep the per-level kernels
                maxw[l] = std::max(maxw[l], (int)r[2]); maxh[l] = std::max(maxh[l], (int)r[3]);
            }
        }
    PyrDev &P = o->pyr_dev;
    memset(&P, 0, sizeof(P));
    P.nlevels = L;
    size_t off = 0;
    for (int l = 0; l < L; l++) {
        P.spitch[l] = (int)kfs_align_up((size_t)maxw[l], 16);
        P.soff[l] = (int)off;
        off = kfs_align_up(off + (size_t)P.spitch[l] * maxh[l], 128);
        P.sw[l] = o->lw[l]; P.sh[l] = o->lh[l];
        P.xtab4[l] = o->xtab4[l]; P.ytab[l] = o->ytab[l];
        if (l >= 1 && !o->resize_fast[l]) return KFS_OK; // scale factor > 2: the per-level generic kernel
    }
    if (P.spitch[0] > 256 || maxh[0] > 256 || off > 140 * 1024) return KFS_OK; // one TMA box is at most 256 x 256; leave room for 1 CTA / SM at least
    o->pyr_box_w = P.spitch[0]; o->pyr_box_h = maxh[0];
    P.box_bytes = P.spitch[0] * maxh[0];
    off += 16; // a row's third word may lie past the last region
    {   // table slices: y entries of every level, then (16-byte aligned) the 2 x uint4 per 4-pixel group
        P.tab_off = (int)off;
        int n_y = 0, n_g = 0;
        for (int l = 1; l < L; l++) { P.ytoff[l] = n_y; n_y += maxh[l]; P.xtoff[l] = n_g; n_g += maxw[l] / 4; }
        P.xt_off = (int)kfs_align_up((size_t)n_y * 4, 16);
        off += (size_t)P.xt_off + (size_t)n_g * 32;
    }
    o->pyr_smem = off;
    o->pyr_tiles = nx * ny;
    if (o->d_rects) { cudaFree(o->d_rects); o->d_rects = nullptr; }
    KFS_CUDA_TRY(cudaMalloc((void **)&o->d_rects, rects.size() * sizeof(short)));
    KFS_CUDA_TRY(cudaMemcpyAsync(o->d_rects, rects.data(), rects.size() * sizeof(short), cudaMemcpyHostToDevice, o->stream));
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    P.rects = o->d_rects;
    memset(&o->pyr_key_img, 0, sizeof(o->pyr_key_img));
    memset(&o->pyr_key_mask, 0, sizeof(o->pyr_key_mask));
    o->pyr_fused = true;
    return KFS_OK;
}

// KFS_ORB_TMA=0 or a driver without cuTensorMapEncodeTiled: every kernel stays on its LDG path (and the pyramid on k_resize)
static void kfs_orb_tma_probe(kfs_orb *o)
{
    if (o->tma_state != 0) return;
    const char *e = getenv("KFS_ORB_TMA");
    o->tma_state = (e && atoi(e) == 0) || !kfs_encode_tiled() ? -1 : 1;
}

// Level-0 tensor map of the fused pyramid kernel (image or mask source), re-encoded only when the source changes.
static bool kfs_orb_pyramid_map(kfs_orb *o, PyrMap *pmap, PyrMapKeyStore *key, const uint8_t *base, int pitch, long long fs, int h)
{
    kfs_orb_tma_probe(o);
    if (o->tma_state <= 0) return false;
    if (key->base == base && key->fs == fs && key->pitch == pitch && key->h == h && key->bw == o->pyr_box_w && key->bh == o->pyr_box_h) return true;
    if (((uintptr_t)base % 16) || (pitch % 16) || (fs % 16)) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)h, 65535};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)fs};
    const cuuint32_t box[3] = {(cuuint32_t)o->pyr_box_w, (cuuint32_t)o->pyr_box_h, 1}, estr[3] = {1, 1, 1};
    if (kfs_encode_tiled()(&pmap->m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void *)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    key->base = base; key->fs = fs; key->pitch = pitch; key->h = h; key->bw = o->pyr_box_w; key->bh = o->pyr_box_h;
    return true;
}

// (Re)builds the per-level tensor maps of one kernel's tile shape (box_w x box_h bytes).  KFS_ORB_TMA=0 forces the LDG path.
static void kfs_orb_update_maps(kfs_orb *o, MapCache &C, int box_w, int box_h, bool blurred = false)
{
    kfs_orb_tma_probe(o);
    C.maps.enabled = o->tma_state > 0;
    if (o->tma_state < 0) return;
    const OrbDev &g = o->g;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelDev &L = g.lv[l];
        if (L.list_cap <= 0) continue; // no keypoint can come from this level: no kernel has a tile or a window on it
        const uint8_t *img = blurred ? L.blur : L.img;
        const long long fs = blurred ? L.img_fs_blur : L.img_fs;
        const int pitch = blurred ? L.pitch_blur : L.pitch;
        if (C.img[l] == img && C.fs[l] == fs && C.pitch[l] == pitch && C.h[l] == L.h) continue;
        const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)L.h, 65535};
        const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)fs};
        const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1}, estr[3] = {1, 1, 1};
        const bool ok = ((uintptr_t)img % 16 == 0) && (pitch % 16 == 0) && (fs % 16 == 0) &&
                        kfs_encode_tiled()(&C.maps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void *)img, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
        if (!ok) { o->tma_state = -1; C.maps.enabled = 0; return; } // e.g. an unaligned caller buffer: stay on the LDG path for good
        C.img[l] = img; C.fs[l] = fs; C.pitch[l] = pitch; C.h[l] = L.h;
    }
}

// Runs the kernel chain on frames [f0, f0 + nb) whose level 0 is already described by g.lv[0].img / .mask.
// d_mask_in points at frame f0 of the raw mask (or nullptr).  `first` = first chunk of a call: clears the counters.
static kfs_status kfs_orb_run_eager(kfs_orb *o, int f0, int nb, const uint8_t *d_mask_in, size_t mask_step, size_t mask_fs, bool first)
{
    OrbDev &g = o->g;
    const int L = g.nlevels;
    cudaStream_t st = o->stream;
    g.frame0 = f0;
    const dim3 rb(32, 8);
    const bool has_mask = d_mask_in != nullptr;
    if (first) {
        KFS_CUDA_TRY(cudaMemsetAsync(o->counters, 0, o->counters_bytes, st));
        for (int l = 0; l < L; l++) g.lv[l].mask = has_mask ? o->mpyr[l] : nullptr;
    }
    if (has_mask) {
        k_mask_binarise<<<dim3(kfs_div_up(o->lw[0], 128), kfs_div_up(o->lh[0], 8), nb), rb, 0, st>>>(
            d_mask_in, (long long)mask_fs, (long long)mask_step, o->mpyr[0] + (size_t)f0 * o->lfs[0], o->lfs[0], o->lpitch[0], o->lw[0], o->lh[0]);
        KFS_LAUNCH_CHECK(o);
    }
    // the whole pyramid in one launch when the plan allows it (kfs_orb_plan_pyramid) and the level-0 source can be described
    // by a tensor map; otherwise level by level
    bool fused = o->pyr_fused && L >= 2 && (o->pyr_mode >= 2 || nb <= PYR_FUSED_MAX_FRAMES);
    if (fused) fused = kfs_orb_pyramid_map(o, &o->pyr_map_img, &o->pyr_key_img, g.lv[0].img, g.lv[0].pitch, g.lv[0].img_fs, o->lh[0]);
    if (fused && has_mask) fused = kfs_orb_pyramid_map(o, &o->pyr_map_mask, &o->pyr_key_mask, o->mpyr[0], o->lpitch[0], o->lfs[0], o->lh[0]);
    if (fused) {
        PyrDev P = o->pyr_dev;
        for (int l = 1; l < L; l++) { P.dst[l] = o->pyr[l]; P.dfs[l] = o->lfs[l]; P.dpitch[l] = o->lpitch[l]; }
        k_pyramid<<<dim3(o->pyr_tiles, nb), PYR_THREADS, o->pyr_smem, st>>>(P, o->pyr_map_img, f0, 0);
        KFS_LAUNCH_CHECK(o);
        if (has_mask) {
            for (int l = 1; l < L; l++) P.dst[l] = o->mpyr[l];
            k_pyramid<<<dim3(o->pyr_tiles, nb), PYR_THREADS, o->pyr_smem, st>>>(P, o->pyr_map_mask, f0, 1);
            KFS_LAUNCH_CHECK(o);
        }
    }
    const bool rz_tma = !fused && o->resize_tma && kfs_orb_resize_maps(o, o->rmaps, false);
    const bool rz_tma_m = rz_tma && has_mask && kfs_orb_resize_maps(o, o->rmaps_m, true);
    for (int l = 1; l < L && !fused; l++) {
        const long long sfs = g.lv[l - 1].img_fs;
        const uint8_t *simg = g.lv[l - 1].img + (long long)f0 * sfs;
        const int spitch = g.lv[l - 1].pitch;
        uint8_t *dimg = o->pyr[l] + (size_t)f0 * o->lfs[l];
        const uint8_t *smask = has_mask ? o->mpyr[l - 1] + (size_t)f0 * o->lfs[l - 1] : nullptr;
        uint8_t *dmask = has_mask ? o->mpyr[l] + (size_t)f0 * o->lfs[l] : nullptr;
        const bool fast = o->resize_fast[l] && ((uintptr_t)simg % 4 == 0) && (spitch % 4 == 0) && (sfs % 4 == 0);
        if (fast) {
            const dim3 grid(kfs_div_up(o->lw[l], 128), kfs_div_up(o->lh[l], 8 * RZ_ROWS), nb);
            const dim3 grid_t(kfs_div_up(o->lw[l], 128), kfs_div_up(o->lh[l], 8 * RZ_ROWS * RZ_YT), nb);
            const size_t rsm = (size_t)o->rz_bw * o->rz_bh;
            if (rz_tma) // source tile by TMA (KFS_ORB_RESIZE_TMA=0: k_resize's LDGs)
                k_resize_tma<<<grid_t, rb, rsm, st>>>(o->rmaps.maps.m[l - 1], f0, o->lh[l - 1], dimg, o->lfs[l], o->lpitch[l], o->lw[l], o->lh[l],
                                                   o->xtab4[l], o->ytab[l], 0, o->rz_bw, o->rz_bh);
            else
                k_resize<<<grid, rb, 0, st>>>(simg, sfs, spitch, o->lh[l - 1], dimg, o->lfs[l], o->lpitch[l], o->lw[l], o->lh[l],
                                              o->xtab4[l], o->ytab[l], 0);
            KFS_LAUNCH_CHECK(o);
            if (has_mask) {
                if (rz_tma_m)
                    k_resize_tma<<<grid_t, rb, rsm, st>>>(o->rmaps_m.maps.m[l - 1], f0, o->lh[l - 1], dmask, o->lfs[l], o->lpitch[l], o->lw[l], o->lh[l],
                                                       o->xtab4[l], o->ytab[l], 1, o->rz_bw, o->rz_bh);
                else
                    k_resize<<<grid, rb, 0, st>>>(smask, o->lfs[l - 1], o->lpitch[l - 1], o->lh[l - 1], dmask, o->lfs[l], o->lpitch[l],
                                                  o->lw[l], o->lh[l], o->xtab4[l], o->ytab[l], 1);
                KFS_LAUNCH_CHECK(o);
            }
        } else { // scale factor > 2 (taps of a 4-pixel group span more than 8 bytes) or an unaligned caller buffer
            const dim3 grid(kfs_div_up(o->lw[l], 128), kfs_div_up(o->lh[l], 8), nb);
            k_resize_generic<<<grid, rb, 0, st>>>(simg, sfs, spitch, o->lw[l - 1], o->lh[l - 1],
                                                  dimg, o->lfs[l], o->lpitch[l], o->lw[l], o->lh[l], o->xtab[l], o->ytab[l], 0);
            KFS_LAUNCH_CHECK(o);
            if (has_mask) {
                k_resize_generic<<<grid, rb, 0, st>>>(smask, o->lfs[l - 1], o->lpitch[l - 1], o->lw[l - 1], o->lh[l - 1],
                                                      dmask, o->lfs[l], o->lpitch[l], o->lw[l], o->lh[l], o->xtab[l], o->ytab[l], 1);
                KFS_LAUNCH_CHECK(o);
            }
        }
    }
    // the blur needs only the pyramid: outside profiling it runs on a side stream, concurrently with the
    // FAST -> select -> Harris -> select chain (whose small latency-bound kernels leave most SMs idle)
    const bool fork_blur = !o->profiling && o->blur_tiles > 0;
    const bool pdl = o->use_pdl && !o->profiling;
    if (fork_blur) {
        if (!o->s_blur) {
            KFS_CUDA_TRY(cudaStreamCreateWithFlags(&o->s_blur, cudaStreamNonBlocking));
            KFS_CUDA_TRY(cudaEventCreateWithFlags(&o->ev_pyr, cudaEventDisableTiming));
            KFS_CUDA_TRY(cudaEventCreateWithFlags(&o->ev_blur, cudaEventDisableTiming));
        }
        KFS_CUDA_TRY(cudaEventRecord(o->ev_pyr, st));
        KFS_CUDA_TRY(cudaStreamWaitEvent(o->s_blur, o->ev_pyr, 0));
        kfs_orb_update_maps(o, o->bmaps, BS_P, BS_H);
        if (o->blur_fma) k_blur7<true><<<dim3(o->blur_tiles, nb), BT_THREADS, 0, o->s_blur>>>(g, o->blur_tab, o->bmaps.maps);
        else k_blur7<false><<<dim3(o->blur_tiles, nb), BT_THREADS, 0, o->s_blur>>>(g, o->blur_tab, o->bmaps.maps);
        KFS_LAUNCH_CHECK(o);
        KFS_CUDA_TRY(cudaEventRecord(o->ev_blur, o->s_blur));
    }
    if (kfs_stage_mark(o, KFS_STAGE_FAST)) return KFS_ERR_CUDA;
    if (g.total_tiles > 0) {
        if (o->fast_half && nb <= FAST_HALF_MAX_FRAMES && o->total_tiles_h > 0) {
            kfs_orb_update_maps(o, o->fmaps_h, FS_P, FT_HH + 8);
            kfs_launch_pdl(pdl, k_fast<FT_HH>, dim3(o->total_tiles_h, nb), dim3(FAST_THREADS), 0, st, g, (const uint32_t *)o->tile_tab_h, o->fmaps_h.maps);
        } else {
            kfs_orb_update_maps(o, o->fmaps, FS_P, FS_H);
            kfs_launch_pdl(pdl, k_fast<FT_H>, dim3(g.total_tiles, nb), dim3(FAST_THREADS), 0, st, g, (const uint32_t *)o->tile_tab, o->fmaps.maps);
        }
        KFS_LAUNCH_CHECK(o);
    }
    if (kfs_stage_mark(o, KFS_STAGE_SELECT_HARRIS)) return KFS_ERR_CUDA;
    int maxq = 1;
    for (int l = 0; l < L; l++) maxq = o->quota[l] > maxq ? o->quota[l] : maxq;
    if (g.total_tiles > 0) {
        kfs_launch_pdl(pdl, k_select_fast, dim3(L, nb), dim3(SEL_THREADS), 0, st, g);
        KFS_LAUNCH_CHECK(o);
        kfs_launch_pdl(pdl, k_harris, dim3(kfs_div_up(2 * maxq + 256, 256), L, nb), dim3(256), 0, st, g);
        KFS_LAUNCH_CHECK(o);
    }
    if (kfs_stage_mark(o, KFS_STAGE_SELECT_SORT)) return KFS_ERR_CUDA;
    if (g.total_tiles > 0) {
        int smem_n = (2 * maxq + maxq / 2 + 64 + 63) / 64 * 64;
        if (smem_n > 11264) smem_n = 11264; // 16 B per entry -> 176 KB (+ <= 32 KB of row arrays), under the 227 KB opt-in limit
        const int rows_cap = o->lh[0] + 2; // two row arrays (segment start / fill cursor) for the (y, x) ordering
        const size_t dyn = (size_t)smem_n * 16 + (size_t)rows_cap * 8;
        kfs_launch_pdl(pdl, k_select_sort, dim3(L, nb), dim3(SORT_THREADS), dyn, st, g, smem_n, rows_cap);
        KFS_LAUNCH_CHECK(o);
    }
    if (kfs_stage_mark(o, KFS_STAGE_BLUR)) return KFS_ERR_CUDA;
    if (fork_blur) {
        KFS_CUDA_TRY(cudaStreamWaitEvent(st, o->ev_blur, 0));
    } else if (o->blur_tiles > 0) {
        kfs_orb_update_maps(o, o->bmaps, BS_P, BS_H);
        if (o->blur_fma) k_blur7<true><<<dim3(o->blur_tiles, nb), BT_THREADS, 0, st>>>(g, o->blur_tab, o->bmaps.maps);
        else k_blur7<false><<<dim3(o->blur_tiles, nb), BT_THREADS, 0, st>>>(g, o->blur_tab, o->bmaps.maps);
        KFS_LAUNCH_CHECK(o);
    }
    if (kfs_stage_mark(o, KFS_STAGE_DESCRIBE)) return KFS_ERR_CUDA;
    const int kpw = std::max(1, std::min(KPW, nb));
    if (o->describe2) {
        // windows by TMA when both tensor-map sets exist (KFS_ORB_TMA=0 or an undescribable buffer: the cp.async instantiation)
        kfs_orb_update_maps(o, o->dmaps_i, 48, 31);
        kfs_orb_update_maps(o, o->dmaps_b, BW_P, 37, true);
        if (o->dmaps_i.maps.enabled && o->dmaps_b.maps.enabled)
            kfs_launch_pdl(pdl, k_describe2<true>, dim3(kfs_div_up(g.cap_out, DK * kpw), nb), dim3(32 * DK), 0, st, g, kpw, o->dmaps_i.maps, o->dmaps_b.maps);
        else
            kfs_launch_pdl(pdl, k_describe2<false>, dim3(kfs_div_up(g.cap_out, DK * kpw), nb), dim3(32 * DK), 0, st, g, kpw, o->dmaps_i.maps, o->dmaps_b.maps);
    }
    else kfs_launch_pdl(pdl, k_describe, dim3(kfs_div_up(g.cap_out, DK * kpw), nb), dim3(32 * DK), 0, st, g, kpw);
    KFS_LAUNCH_CHECK(o);
    if (kfs_stage_mark(o, KFS_STAGE_COUNT)) return KFS_ERR_CUDA;
    o->last_batch = f0 + nb;
    return KFS_OK;
}

// The chain is 14+ launches, two streams and a memset: for small batches (the tracker's one frame per call) the launch
// overhead is a large part of the latency.  A launch configuration seen for the second time is captured into a CUDA graph
// (the side-stream blur becomes a parallel branch) and replayed from then on.  The key is every input of the launches: the
// whole OrbDev block (all pointers and sizes, frame0, per-level masks), the frame count, the mask source and the stream.
// Profiling (per-stage events) and a stream that is itself being captured by the caller take the eager path.
static kfs_status kfs_orb_run(kfs_orb *o, int f0, int nb, const uint8_t *d_mask_in, size_t mask_step, size_t mask_fs, bool first)
{
    cudaStream_t st = o->stream;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (!o->use_graphs || o->profiling || cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) {
        cudaGetLastError();
        return kfs_orb_run_eager(o, f0, nb, d_mask_in, mask_step, mask_fs, first);
    }
    // the fields kfs_orb_run_eager sets before launching, so that the key below is complete
    o->g.frame0 = f0;
    if (first) for (int l = 0; l < o->g.nlevels; l++) o->g.lv[l].mask = d_mask_in ? o->mpyr[l] : nullptr;
    OrbGraph *G = nullptr;
    for (OrbGraph &c : o->graphs)
        if (c.nb == nb && c.mask == d_mask_in && c.mask_step == mask_step && c.mask_fs == mask_fs && c.first == first && c.stream == st &&
            memcmp(&c.g, &o->g, sizeof(OrbDev)) == 0) { G = &c; break; }
    if (!G) {
        if (o->graphs.size() >= 32) kfs_orb_drop_graphs(o);
        OrbGraph n;
        n.g = o->g; n.nb = nb; n.mask = d_mask_in; n.mask_step = mask_step; n.mask_fs = mask_fs; n.first = first; n.stream = st;
        n.seen = 0; n.nlaunch = 0; n.exec = nullptr;
        o->graphs.push_back(n);
        G = &o->graphs.back();
    }
    G->seen++;
    if (G->exec) {
        KFS_CUDA_TRY(cudaGraphLaunch(G->exec, st));
        o->launches += G->nlaunch;
        o->last_batch = f0 + nb;
        return KFS_OK;
    }
    if (G->seen < 2) return kfs_orb_run_eager(o, f0, nb, d_mask_in, mask_step, mask_fs, first); // also creates the side stream, sets attributes
    if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError();
        o->use_graphs = false;
        return kfs_orb_run_eager(o, f0, nb, d_mask_in, mask_step, mask_fs, first);
    }
    const int before = o->launches;
    const kfs_status rs = kfs_orb_run_eager(o, f0, nb, d_mask_in, mask_step, mask_fs, first);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(st, &graph);
    cudaGraphExec_t exec = nullptr;
    if (rs == KFS_OK && ce == cudaSuccess && graph && cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess) {
        cudaGraphDestroy(graph);
        G->exec = exec;
        G->nlaunch = o->launches - before;
        KFS_CUDA_TRY(cudaGraphLaunch(exec, st));
        return KFS_OK;
    }
    // capture failed: nothing was executed; forget graphs on this handle and run the chain directly
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    o->use_graphs = false;
    o->launches = before;
    return kfs_orb_run_eager(o, f0, nb, d_mask_in, mask_step, mask_fs, first);
}

static kfs_status kfs_orb_check_frame_args(const kfs_orb *o, const void *gray, size_t step, int w, int h, int batch)
{
    if (!o) KFS_FAIL(KFS_ERR_INVALID_ARG, "orb handle is NULL");
    if (w < 0 || h < 0 || batch < 1) KFS_FAIL(KFS_ERR_INVALID_ARG, "negative image size or batch < 1");
    if (w == 0 || h == 0) return KFS_OK;
    if (!gray) KFS_FAIL(KFS_ERR_INVALID_ARG, "image pointer is NULL");
    if (step < (size_t)w) KFS_FAIL(KFS_ERR_INVALID_ARG, "step %zu < width %d", step, w);
    if (w > 4096 || h > 4096) KFS_FAIL(KFS_ERR_UNSUPPORTED, "image larger than 4096 px on a side (%dx%d)", w, h);
    if (batch > 65535) KFS_FAIL(KFS_ERR_UNSUPPORTED, "batch > 65535");
    return KFS_OK;
}

extern "C" {

void kfs_orb_default_params(kfs_orb_params *p)
{
    if (!p) return;
    p->nfeatures = 500; p->scale_factor = 1.2f; p->nlevels = 8; p->edge_threshold = 31; p->first_level = 0;
    p->wta_k = 2; p->score_type = KFS_ORB_HARRIS_SCORE; p->patch_size = 31; p->fast_threshold = 20;
}

kfs_status kfs_orb_create(const kfs_orb_params *params, int device, kfs_orb **out)
{
    if (!params || !out) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_create: NULL argument");
    *out = nullptr;
    const kfs_orb_params &P = *params;
    if (P.nfeatures < 1 || !(P.scale_factor > 1.f) || P.nlevels < 1 || P.fast_threshold < 1 || P.fast_threshold > 254)
        KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_create: nfeatures>=1, scale_factor>1, nlevels>=1, 1<=fast_threshold<=254 required");
    if (P.nlevels > KFS_MAX_LEVELS) KFS_FAIL(KFS_ERR_UNSUPPORTED, "nlevels > %d", KFS_MAX_LEVELS);
    if (P.edge_threshold != 31 || P.first_level != 0 || P.wta_k != 2 || P.score_type != KFS_ORB_HARRIS_SCORE || P.patch_size != 31)
        KFS_FAIL(KFS_ERR_UNSUPPORTED, "only edgeThreshold=31, firstLevel=0, WTA_K=2, HARRIS_SCORE, patchSize=31 (the reference's values) are built");
    const int ndev = kfs_device_count();
    if (ndev == 0) KFS_FAIL(KFS_ERR_NO_DEVICE, "no CUDA device visible; this library has no CPU path");
    if (device < 0 || device >= ndev) KFS_FAIL(KFS_ERR_INVALID_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    kfs_orb *o = new (std::nothrow) kfs_orb();
    if (!o) KFS_FAIL(KFS_ERR_INVALID_ARG, "out of host memory");
    o->p = P;
    o->device = device;
    o->cap_out = (int)kfs_align_up((size_t)(P.nfeatures + P.nfeatures / 4 + 64), 8);
    memset(&o->g, 0, sizeof(o->g));
    if (const char *eg = getenv("KFS_ORB_GRAPHS")) o->use_graphs = atoi(eg) != 0; // experiments: 0 = always launch the chain directly
    memset(o->stage_ms, 0, sizeof(o->stage_ms));
    KfsDeviceGuard guard(device);
    cudaError_t e = guard.ok ? cudaStreamCreateWithFlags(&o->own_stream, cudaStreamNonBlocking) : cudaErrorInvalidDevice;
    if (e == cudaSuccess) e = kfs_orb_raise_smem_limits(device);
    if (e != cudaSuccess) {
        if (o->own_stream) cudaStreamDestroy(o->own_stream);
        delete o;
        KFS_FAIL(KFS_ERR_CUDA, "kfs_orb_create on device %d: %s", device, cudaGetErrorString(e));
    }
    o->stream = o->own_stream;
    *out = o;
    return KFS_OK;
}

void kfs_orb_destroy(kfs_orb *o)
{
    if (!o) return;
    KfsDeviceGuard guard(o->device);
    cudaStreamSynchronize(o->stream);
    kfs_orb_free_arena(o);
    if (o->h_counts) cudaFreeHost(o->h_counts);
    if (o->ev_made) for (int i = 0; i <= KFS_STAGE_COUNT; i++) cudaEventDestroy(o->ev[i]);
    for (cudaEvent_t e : o->ev_in) cudaEventDestroy(e);
    for (cudaEvent_t e : o->ev_cmp) cudaEventDestroy(e);
    if (o->ev_start) cudaEventDestroy(o->ev_start);
    if (o->mask_stage) cudaFree(o->mask_stage);
    if (o->rgb_stage) cudaFree(o->rgb_stage);
    if (o->ev_pyr) cudaEventDestroy(o->ev_pyr);
    if (o->ev_blur) cudaEventDestroy(o->ev_blur);
    if (o->s_blur) cudaStreamDestroy(o->s_blur);
    if (o->s_in) cudaStreamDestroy(o->s_in);
    if (o->s_out) cudaStreamDestroy(o->s_out);
    if (o->own_stream) cudaStreamDestroy(o->own_stream);
    delete o;
}

int kfs_orb_capacity(const kfs_orb *o) { return o ? o->cap_out : 0; }

kfs_status kfs_orb_set_capacity(kfs_orb *o, int capacity)
{
    if (!o || capacity < 1) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_set_capacity: bad arguments");
    KFS_DEVICE_GUARD(o->device);
    const int c = (int)kfs_align_up((size_t)capacity, 8);
    if (c != o->cap_out) {
        if (o->arena) { KFS_CUDA_TRY(cudaStreamSynchronize(o->stream)); kfs_orb_free_arena(o); }
        o->cap_out = c;
    }
    return KFS_OK;
}

kfs_status kfs_orb_set_stream(kfs_orb *o, void *s)
{
    if (!o) KFS_FAIL(KFS_ERR_INVALID_ARG, "orb handle is NULL");
    KFS_DEVICE_GUARD(o->device);
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    o->stream = s ? (cudaStream_t)s : o->own_stream;
    return KFS_OK;
}

void *kfs_orb_get_stream(kfs_orb *o) { return o ? (void *)o->stream : nullptr; }

kfs_status kfs_orb_sync(kfs_orb *o)
{
    if (!o) KFS_FAIL(KFS_ERR_INVALID_ARG, "orb handle is NULL");
    KFS_DEVICE_GUARD(o->device);
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    return KFS_OK;
}

kfs_status kfs_orb_extract_device(kfs_orb *o, const uint8_t *d_gray, size_t step, size_t frame_stride, int w, int h, int batch,
                                  const uint8_t *d_mask, size_t mask_step, size_t mask_frame_stride)
{
    kfs_status s = kfs_orb_check_frame_args(o, d_gray, step, w, h, batch);
    if (s) return s;
    KFS_DEVICE_GUARD(o->device);
    o->launches = 0;
    o->alias_valid = false;
    if (w == 0 || h == 0) { o->last_batch = 0; return KFS_OK; }
    if (d_mask && mask_step < (size_t)w) KFS_FAIL(KFS_ERR_INVALID_ARG, "mask_step < width");
    s = kfs_orb_ensure(o, w, h, batch, d_mask != nullptr);
    if (s) return s;
    if (kfs_stage_mark(o, KFS_STAGE_INGEST)) return KFS_ERR_CUDA;
    LevelDev &L0 = o->g.lv[0];
    const bool alias = ((uintptr_t)d_gray % 16 == 0) && (step % 16 == 0) && (frame_stride % 16 == 0);
    if (alias) {
        L0.img = d_gray; L0.pitch = (int)step; L0.img_fs = (long long)frame_stride;
    } else {
        L0.img = o->pyr[0]; L0.pitch = o->lpitch[0]; L0.img_fs = o->lfs[0];
        if (frame_stride == step * (size_t)h) {
            KFS_CUDA_TRY(cudaMemcpy2DAsync(o->pyr[0], (size_t)o->lpitch[0], d_gray, step, (size_t)w, (size_t)h * batch, cudaMemcpyDeviceToDevice, o->stream));
        } else {
            for (int f = 0; f < batch; f++)
                KFS_CUDA_TRY(cudaMemcpy2DAsync(o->pyr[0] + (size_t)f * o->lfs[0], (size_t)o->lpitch[0], d_gray + (size_t)f * frame_stride, step,
                                               (size_t)w, (size_t)h, cudaMemcpyDeviceToDevice, o->stream));
        }
    }
    if (kfs_stage_mark(o, KFS_STAGE_PYRAMID)) return KFS_ERR_CUDA;
    return kfs_orb_run(o, 0, batch, d_mask, mask_step, mask_frame_stride, true);
}

kfs_status kfs_orb_device_results(kfs_orb *o, const kfs_keypoint **d_kp, const uint8_t **d_desc, const int **d_counts, int *capacity, int *batch)
{
    if (!o) KFS_FAIL(KFS_ERR_INVALID_ARG, "orb handle is NULL");
    if (!o->arena) KFS_FAIL(KFS_ERR_INVALID_ARG, "no extraction has run on this handle yet");
    if (d_kp) *d_kp = o->g.kp;
    if (d_desc) *d_desc = o->g.desc;
    if (d_counts) *d_counts = o->g.n_out;
    if (capacity) *capacity = o->cap_out;
    if (batch) *batch = o->last_batch;
    return KFS_OK;
}

kfs_status kfs_orb_fetch_results(kfs_orb *o, kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out)
{
    if (!o || !n_out || capacity < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_fetch_results: bad arguments");
    const int B = o->last_batch;
    if (B == 0) return KFS_OK;
    if ((!kp_out || !desc_out) && capacity > 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "output buffers are NULL");
    KFS_DEVICE_GUARD(o->device);
    const int cp = capacity < o->cap_out ? capacity : o->cap_out;
    KFS_CUDA_TRY(cudaMemcpyAsync(o->h_counts, o->g.n_out, sizeof(int) * (size_t)B, cudaMemcpyDeviceToHost, o->stream));
    if (cp > 0) {
        KFS_CUDA_TRY(cudaMemcpy2DAsync(kp_out, (size_t)capacity * sizeof(kfs_keypoint), o->g.kp, (size_t)o->cap_out * sizeof(kfs_keypoint),
                                       (size_t)cp * sizeof(kfs_keypoint), (size_t)B, cudaMemcpyDeviceToHost, o->stream));
        KFS_CUDA_TRY(cudaMemcpy2DAsync(desc_out, (size_t)capacity * 32, o->g.desc, (size_t)o->cap_out * 32,
                                       (size_t)cp * 32, (size_t)B, cudaMemcpyDeviceToHost, o->stream));
    }
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    kfs_status rc = KFS_OK;
    for (int f = 0; f < B; f++) {
        n_out[f] = o->h_counts[f];
        if (n_out[f] > cp) rc = KFS_ERR_CAPACITY;
    }
    if (rc == KFS_ERR_CAPACITY) kfs_set_error("output capacity %d too small (handle capacity %d); n_out holds the required counts", capacity, o->cap_out);
    if (o->profiling && o->ev_made) {
        for (int i = 0; i < KFS_STAGE_COUNT; i++) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, o->ev[i], o->ev[i + 1]) == cudaSuccess) o->stage_ms[i] = ms;
            else cudaGetLastError();
        }
    }
    return rc;
}

static kfs_status kfs_orb_pipeline_init(kfs_orb *o, int nchunks)
{
    if (!o->s_in) KFS_CUDA_TRY(cudaStreamCreateWithFlags(&o->s_in, cudaStreamNonBlocking));
    if (!o->s_out) KFS_CUDA_TRY(cudaStreamCreateWithFlags(&o->s_out, cudaStreamNonBlocking));
    if (!o->ev_start) KFS_CUDA_TRY(cudaEventCreateWithFlags(&o->ev_start, cudaEventDisableTiming));
    while ((int)o->ev_in.size() < nchunks) {
        cudaEvent_t a, b;
        KFS_CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        o->ev_in.push_back(a);
        KFS_CUDA_TRY(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        o->ev_cmp.push_back(b);
    }
    return KFS_OK;
}

static kfs_status kfs_h2d_frames(uint8_t *d, size_t dpitch, size_t dfs, const uint8_t *h, size_t step, size_t frame_stride,
                                 int w, int h_rows, int f0, int nb, cudaStream_t st)
{
    if (frame_stride == step * (size_t)h_rows && dfs == dpitch * (size_t)h_rows && step == (size_t)w && dpitch == (size_t)w) {
        // both sides fully contiguous: one linear copy
        KFS_CUDA_TRY(cudaMemcpyAsync(d + (size_t)f0 * dfs, h + (size_t)f0 * frame_stride, (size_t)w * h_rows * nb, cudaMemcpyHostToDevice, st));
    } else if (frame_stride == step * (size_t)h_rows && dfs == dpitch * (size_t)h_rows) {
        KFS_CUDA_TRY(cudaMemcpy2DAsync(d + (size_t)f0 * dfs, dpitch, h + (size_t)f0 * frame_stride, step, (size_t)w, (size_t)h_rows * nb,
                                       cudaMemcpyHostToDevice, st));
    } else {
        for (int f = f0; f < f0 + nb; f++)
            KFS_CUDA_TRY(cudaMemcpy2DAsync(d + (size_t)f * dfs, dpitch, h + (size_t)f * frame_stride, step, (size_t)w, (size_t)h_rows,
                                           cudaMemcpyHostToDevice, st));
    }
    return KFS_OK;
}

// Host-buffer entry point.  The batch is cut into chunks of frames that flow through three streams:
//   s_in : H2D of chunk i+1        | stream : kernel chain of chunk i        | s_out : D2H of chunk i-1's results
// so PCIe in, the SMs and PCIe out work concurrently (pinned host buffers give true overlap; pageable ones
// still work, the copies just serialise inside the driver).  The call returns when everything is on the host.
} // extern "C"  (the two functions below are library-internal: kfs_replay.cu drives them)

// Enqueues the whole host-buffer pipeline (chunked H2D | [colour -> gray] | kernel chain | D2H of keypoints, descriptors and
// the per-frame counts) WITHOUT synchronising.  Results are complete once the handle's copy-out stream has drained;
// kfs_orb_finish_host_batch then hands out the counts (and fetches the rows past out_rows of a tie-rich frame).
kfs_status kfs_orb_enqueue_host_batch(kfs_orb *o, const KfsHostBatch &hb)
{
    const int w = hb.w, h = hb.h, batch = hb.batch, capacity = hb.capacity, ch = hb.channels;
    if (ch != 1 && ch != 3) KFS_FAIL(KFS_ERR_UNSUPPORTED, "%d channels (the reference handles 1 or 3, VisualOdometry.cpp:27)", ch);
    kfs_status s = kfs_orb_check_frame_args(o, hb.image, hb.step / (size_t)ch, w, h, batch);
    if (s) return s;
    if (hb.step < (size_t)w * ch) KFS_FAIL(KFS_ERR_INVALID_ARG, "step %zu < width %d x %d channels", hb.step, w, ch);
    if (capacity < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "capacity < 0");
    if (w == 0 || h == 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "empty image");
    if (hb.mask && hb.mask_step < (size_t)w) KFS_FAIL(KFS_ERR_INVALID_ARG, "mask_step < width");
    if ((!hb.kp_out || !hb.desc_out) && capacity > 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "output buffers are NULL");
    KFS_DEVICE_GUARD(o->device);
    o->launches = 0;
    o->alias_valid = false;
    s = kfs_orb_ensure(o, w, h, batch, hb.mask != nullptr);
    if (s) return s;
    // Chunk schedule.  An explicit chunk size (kfs_orb_set_pipeline_chunk) cuts the batch evenly.  The automatic schedule is three
    // quarters of the batch in quarter-size chunks and a TAPERED tail (1/8, 1/16, 1/16): what is left to do after the last byte of the
    // upload has arrived — the kernel chain and the download of the last chunk — is then a sixteenth of the batch instead of a quarter.
    std::vector<int> cstart, ccount;
    auto add_chunk = [&](int f0, int n) { cstart.push_back(f0); ccount.push_back(n); return f0 + n; };
    if (o->profiling || batch < 8 || o->pipeline_chunk >= batch) {
        add_chunk(0, batch); // (profiling: per-stage event timing describes one unchunked pass)
    } else if (o->pipeline_chunk > 0) {
        for (int f0 = 0; f0 < batch;) f0 = add_chunk(f0, std::min(o->pipeline_chunk, batch - f0));
    } else {
        const int q = kfs_div_up(batch, 4);
        int f0 = 0;
        for (int i = 0; i < 3 && f0 < batch; i++) f0 = add_chunk(f0, std::min(q, batch - f0));
        const int taper[3] = {kfs_div_up(batch, 8), kfs_div_up(batch, 16), batch};
        for (int i = 0; i < 3 && f0 < batch; i++) f0 = add_chunk(f0, std::min(batch >= 32 ? taper[i] : batch, batch - f0));
    }
    const int nchunks = (int)cstart.size();
    s = kfs_orb_pipeline_init(o, nchunks);
    if (s) return s;
    if (hb.mask) {
        const size_t need = (size_t)o->lfs[0] * batch;
        if (need > o->mask_stage_bytes) {
            if (o->mask_stage) { cudaFree(o->mask_stage); o->mask_stage = nullptr; o->mask_stage_bytes = 0; }
            KFS_CUDA_TRY(cudaMalloc((void **)&o->mask_stage, need));
            o->mask_stage_bytes = need;
        }
    }
    const size_t ipitch = kfs_align_up((size_t)w * 3, 16), ifs = ipitch * (size_t)h; // colour staging: [batch][h][ipitch]
    if (ch == 3 && ifs * batch > o->rgb_stage_bytes) {
        if (o->rgb_stage) { KFS_CUDA_TRY(cudaStreamSynchronize(o->stream)); cudaFree(o->rgb_stage); o->rgb_stage = nullptr; o->rgb_stage_bytes = 0; }
        KFS_CUDA_TRY(cudaMalloc((void **)&o->rgb_stage, ifs * batch));
        o->rgb_stage_bytes = ifs * batch;
    }
    if (kfs_stage_mark(o, KFS_STAGE_INGEST)) return KFS_ERR_CUDA;
    LevelDev &L0 = o->g.lv[0];
    L0.img = o->pyr[0]; L0.pitch = o->lpitch[0]; L0.img_fs = o->lfs[0];
    // the copy streams must not start before work already queued on the compute stream (a caller-provided
    // stream may still be producing / consuming the buffers of an earlier call)
    if (o->start_fence) {
        KFS_CUDA_TRY(cudaEventRecord(o->ev_start, o->stream));
        KFS_CUDA_TRY(cudaStreamWaitEvent(o->s_in, o->ev_start, 0));
        KFS_CUDA_TRY(cudaStreamWaitEvent(o->s_out, o->ev_start, 0));
    }
    const int cp = capacity < o->cap_out ? capacity : o->cap_out;
    int rows = hb.out_rows > 0 ? hb.out_rows : cp;
    if (rows > cp) rows = cp;
    for (int c = 0; c < nchunks; c++) {
        const int f0 = cstart[(size_t)c], nb = ccount[(size_t)c];
        if (ch == 3) s = kfs_h2d_frames(o->rgb_stage, ipitch, ifs, hb.image, hb.step, hb.frame_stride, 3 * w, h, f0, nb, o->s_in);
        else s = kfs_h2d_frames(o->pyr[0], (size_t)o->lpitch[0], (size_t)o->lfs[0], hb.image, hb.step, hb.frame_stride, w, h, f0, nb, o->s_in);
        if (s) return s;
        if (hb.mask) {
            s = kfs_h2d_frames(o->mask_stage, (size_t)o->lpitch[0], (size_t)o->lfs[0], hb.mask, hb.mask_step, hb.mask_frame_stride, w, h, f0, nb, o->s_in);
            if (s) return s;
        }
        KFS_CUDA_TRY(cudaEventRecord(o->ev_in[c], o->s_in));
        KFS_CUDA_TRY(cudaStreamWaitEvent(o->stream, o->ev_in[c], 0));
        if (ch == 3) { // VisualOdometry.cpp:25-28 for the whole chunk in one launch, straight into level 0
            s = kfs_launch_rgb2gray(o->rgb_stage + (size_t)f0 * ifs, (long long)ipitch, (long long)ifs, o->pyr[0] + (size_t)f0 * o->lfs[0], o->lpitch[0],
                                    o->lfs[0], w, h, nb, o->stream);
            if (s) return s;
            o->launches++;
        }
        if (c == 0 && kfs_stage_mark(o, KFS_STAGE_PYRAMID)) return KFS_ERR_CUDA;
        s = kfs_orb_run(o, f0, nb, hb.mask ? o->mask_stage + (size_t)f0 * o->lfs[0] : nullptr, (size_t)o->lpitch[0], (size_t)o->lfs[0], c == 0);
        if (s) return s;
        KFS_CUDA_TRY(cudaEventRecord(o->ev_cmp[c], o->stream));
        KFS_CUDA_TRY(cudaStreamWaitEvent(o->s_out, o->ev_cmp[c], 0));
        // rows [0, rows) of every wanted frame: runs of consecutive wanted frames travel as one (2-D) copy each
        for (int a = f0; rows > 0 && a < f0 + nb;) {
            if (hb.skip_rows && hb.skip_rows[a]) { a++; continue; }
            int b = a + 1;
            while (b < f0 + nb && !(hb.skip_rows && hb.skip_rows[b])) b++;
            const size_t n = (size_t)(b - a);
            if (rows == capacity && capacity == o->cap_out) { // same row pitch on both sides, whole rows: linear copies
                KFS_CUDA_TRY(cudaMemcpyAsync(hb.kp_out + (size_t)a * capacity, o->g.kp + (size_t)a * o->cap_out,
                                             n * capacity * sizeof(kfs_keypoint), cudaMemcpyDeviceToHost, o->s_out));
                KFS_CUDA_TRY(cudaMemcpyAsync(hb.desc_out + (size_t)a * capacity * 32, o->g.desc + (size_t)a * o->cap_out * 32,
                                             n * capacity * 32, cudaMemcpyDeviceToHost, o->s_out));
            } else {
                KFS_CUDA_TRY(cudaMemcpy2DAsync(hb.kp_out + (size_t)a * capacity, (size_t)capacity * sizeof(kfs_keypoint),
                                               o->g.kp + (size_t)a * o->cap_out, (size_t)o->cap_out * sizeof(kfs_keypoint),
                                               (size_t)rows * sizeof(kfs_keypoint), n, cudaMemcpyDeviceToHost, o->s_out));
                KFS_CUDA_TRY(cudaMemcpy2DAsync(hb.desc_out + (size_t)a * capacity * 32, (size_t)capacity * 32,
                                               o->g.desc + (size_t)a * o->cap_out * 32, (size_t)o->cap_out * 32,
                                               (size_t)rows * 32, n, cudaMemcpyDeviceToHost, o->s_out));
            }
            a = b;
        }
    }
    KFS_CUDA_TRY(cudaMemcpyAsync(o->h_counts, o->g.n_out, sizeof(int) * (size_t)batch, cudaMemcpyDeviceToHost, o->s_out));
    o->pending = hb;
    o->pending_rows = rows;
    return KFS_OK;
}

// After the copy-out stream has drained: counts to the caller, and for a frame that kept more than `out_rows` keypoints (ties on
// a selection threshold) the remaining rows in a second, blocking copy — rare, and never wrong.
kfs_status kfs_orb_finish_host_batch(kfs_orb *o, int batch, int *n_out)
{
    const KfsHostBatch &hb = o->pending;
    const int capacity = hb.capacity;
    const int cp = capacity < o->cap_out ? capacity : o->cap_out;
    kfs_status rc = KFS_OK;
    bool late = false;
    for (int f = 0; f < batch; f++) {
        n_out[f] = o->h_counts[f];
        if (n_out[f] > cp) rc = KFS_ERR_CAPACITY;
        const int live = n_out[f] < cp ? n_out[f] : cp;
        if (live > o->pending_rows && !(hb.skip_rows && hb.skip_rows[f])) {
            const size_t r0 = (size_t)o->pending_rows, n = (size_t)(live - o->pending_rows);
            KFS_CUDA_TRY(cudaMemcpyAsync(hb.kp_out + (size_t)f * capacity + r0, o->g.kp + (size_t)f * o->cap_out + r0, n * sizeof(kfs_keypoint),
                                         cudaMemcpyDeviceToHost, o->s_out));
            KFS_CUDA_TRY(cudaMemcpyAsync(hb.desc_out + ((size_t)f * capacity + r0) * 32, o->g.desc + ((size_t)f * o->cap_out + r0) * 32, n * 32,
                                         cudaMemcpyDeviceToHost, o->s_out));
            late = true;
        }
    }
    if (late) KFS_CUDA_TRY(cudaStreamSynchronize(o->s_out));
    // what an attached matcher will compare the caller's rows against before it trusts the device copy (kfs_orb_alias_desc)
    o->alias_sig.assign((size_t)batch * 64, 0);
    for (int f = 0; f < batch && hb.desc_out; f++) {
        const int live = n_out[f] < cp ? n_out[f] : cp;
        if (live <= 0 || (hb.skip_rows && hb.skip_rows[f])) continue;
        const uint8_t *rows = hb.desc_out + (size_t)f * capacity * 32;
        memcpy(&o->alias_sig[(size_t)f * 64], rows, 32);
        memcpy(&o->alias_sig[(size_t)f * 64 + 32], rows + (size_t)(live - 1) * 32, 32);
    }
    o->alias_valid = true;
    o->pending.batch = batch;
    if (rc == KFS_ERR_CAPACITY) kfs_set_error("output capacity %d too small (handle capacity %d); n_out holds the required counts", capacity, o->cap_out);
    if (o->profiling && o->ev_made) {
        for (int i = 0; i < KFS_STAGE_COUNT; i++) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, o->ev[i], o->ev[i + 1]) == cudaSuccess) o->stage_ms[i] = ms;
            else cudaGetLastError();
        }
    }
    return rc;
}

// Device copy of descriptor rows the last host-buffer extraction delivered to `host`, or nullptr: lets an attached matcher
// (kfs_bf_attach_orb) skip the upload of rows that never left the device.  Recognised: row 0 of one of that call's frames with
// EXACTLY that frame's keypoint count of rows (a caller that filtered or compacted the rows in place passes fewer and is uploaded),
// whose first and last row still hold the bytes that were delivered (a buffer that was freed and reallocated at the same address, or
// that another extractor has written since, does not).
const uint8_t *kfs_orb_alias_desc(const kfs_orb *o, const uint8_t *host, int rows, int device)
{
    if (!o || !o->alias_valid || !o->arena || o->device != device || !host || rows <= 0) return nullptr;
    const KfsHostBatch &hb = o->pending;
    if (!hb.desc_out || hb.capacity <= 0 || host < hb.desc_out) return nullptr;
    const size_t off = (size_t)(host - hb.desc_out), fs = (size_t)hb.capacity * 32;
    if (off % fs) return nullptr;
    const size_t f = off / fs;
    if (f >= (size_t)hb.batch || (hb.skip_rows && hb.skip_rows[f])) return nullptr;
    const int cp = hb.capacity < o->cap_out ? hb.capacity : o->cap_out;
    const int live = o->h_counts[f] < cp ? o->h_counts[f] : cp;
    if (rows != live || o->alias_sig.size() < (f + 1) * 64) return nullptr;
    const uint8_t *sig = &o->alias_sig[f * 64];
    if (memcmp(host, sig, 32) != 0 || memcmp(host + (size_t)(live - 1) * 32, sig + 32, 32) != 0) return nullptr;
    return o->g.desc + f * (size_t)o->cap_out * 32;
}

void kfs_orb_internal_start_fence(kfs_orb *o, bool on) { o->start_fence = on; }

void kfs_orb_internal_streams(kfs_orb *o, cudaStream_t *s_in, cudaStream_t *s_out)
{
    if (s_in) *s_in = o->s_in;
    if (s_out) *s_out = o->s_out;
}

extern "C" {

// Host-buffer entry point: enqueue, wait, hand out the counts.
kfs_status kfs_orb_detect_and_compute_batch(kfs_orb *o, const uint8_t *gray, size_t step, size_t frame_stride, int w, int h, int batch,
                                            const uint8_t *mask, size_t mask_step, size_t mask_frame_stride,
                                            kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out)
{
    kfs_status s = kfs_orb_check_frame_args(o, gray, step, w, h, batch);
    if (s) return s;
    if (!n_out || capacity < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "n_out is NULL or capacity < 0");
    for (int f = 0; f < batch; f++) n_out[f] = 0;
    if (w == 0 || h == 0) return KFS_OK; // OpenCV returns early on an empty image
    KfsHostBatch hb = {};
    hb.image = gray; hb.step = step; hb.frame_stride = frame_stride; hb.w = w; hb.h = h; hb.channels = 1; hb.batch = batch;
    hb.mask = mask; hb.mask_step = mask_step; hb.mask_frame_stride = mask_frame_stride;
    hb.kp_out = kp_out; hb.desc_out = desc_out; hb.capacity = capacity;
    hb.out_rows = o->p.nfeatures; // rows past nfeatures exist only after ties: fetched by finish() when a count says so
    s = kfs_orb_enqueue_host_batch(o, hb);
    if (s) return s;
    KFS_DEVICE_GUARD(o->device);
    KFS_CUDA_TRY(cudaStreamSynchronize(o->s_out));
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    return kfs_orb_finish_host_batch(o, batch, n_out);
}

kfs_status kfs_orb_set_pipeline_chunk(kfs_orb *o, int frames)
{
    if (!o || frames < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_set_pipeline_chunk: bad arguments");
    o->pipeline_chunk = frames;
    return KFS_OK;
}

kfs_status kfs_orb_detect_and_compute(kfs_orb *o, const uint8_t *gray, size_t step, int w, int h,
                                      const uint8_t *mask, size_t mask_step,
                                      kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out)
{
    return kfs_orb_detect_and_compute_batch(o, gray, step, step * (size_t)(h > 0 ? h : 0), w, h, 1, mask, mask_step,
                                            mask_step * (size_t)(h > 0 ? h : 0), kp_out, desc_out, capacity, n_out);
}

kfs_status kfs_orb_level_info(kfs_orb *o, int level, int *w, int *h, int *quota, float *scale)
{
    if (!o || !o->arena) KFS_FAIL(KFS_ERR_INVALID_ARG, "no extraction has run on this handle yet");
    if (level < 0 || level >= o->nlevels_active) KFS_FAIL(KFS_ERR_INVALID_ARG, "level out of range");
    if (w) *w = o->lw[level];
    if (h) *h = o->lh[level];
    if (quota) *quota = o->quota[level];
    if (scale) *scale = o->lscale[level];
    return KFS_OK;
}

kfs_status kfs_orb_download_level(kfs_orb *o, int frame, int level, uint8_t *out, size_t out_step)
{
    if (!o || !o->arena || !out) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_download_level: bad arguments");
    if (level < 0 || level >= o->nlevels_active || frame < 0 || frame >= o->last_batch) KFS_FAIL(KFS_ERR_INVALID_ARG, "frame/level out of range");
    KFS_DEVICE_GUARD(o->device);
    const LevelDev &L = o->g.lv[level];
    KFS_CUDA_TRY(cudaMemcpy2DAsync(out, out_step, L.img + (size_t)frame * L.img_fs, (size_t)L.pitch, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost, o->stream));
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    return KFS_OK;
}

kfs_status kfs_orb_set_blur_fma(kfs_orb *o, int fma)
{
    if (!o) KFS_FAIL(KFS_ERR_INVALID_ARG, "orb handle is NULL");
    if (o->blur_fma != (fma != 0)) {
        KFS_DEVICE_GUARD(o->device);
        KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
        kfs_orb_drop_graphs(o); // the captured chains hold the other kernel
        o->blur_fma = fma != 0;
    }
    return KFS_OK;
}

kfs_status kfs_orb_download_blurred_level(kfs_orb *o, int frame, int level, uint8_t *out, size_t out_step)
{
    if (!o || !o->arena || !out) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_download_blurred_level: bad arguments");
    if (level < 0 || level >= o->nlevels_active || frame < 0 || frame >= o->last_batch) KFS_FAIL(KFS_ERR_INVALID_ARG, "frame/level out of range");
    KFS_DEVICE_GUARD(o->device);
    const LevelDev &L = o->g.lv[level];
    if (!L.blur) KFS_FAIL(KFS_ERR_INVALID_ARG, "level %d has no blurred copy (too small to hold a keypoint)", level);
    KFS_CUDA_TRY(cudaMemcpy2DAsync(out, out_step, L.blur + (size_t)frame * L.img_fs_blur, (size_t)L.pitch_blur, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost, o->stream));
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    return KFS_OK;
}

kfs_status kfs_orb_download_fast_corners(kfs_orb *o, int frame, int level, uint32_t *out, int capacity, int *n_out)
{
    if (!o || !o->arena || !n_out) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_download_fast_corners: bad arguments");
    if (level < 0 || level >= o->nlevels_active || frame < 0 || frame >= o->last_batch) KFS_FAIL(KFS_ERR_INVALID_ARG, "frame/level out of range");
    KFS_DEVICE_GUARD(o->device);
    const LevelDev &L = o->g.lv[level];
    int n = 0;
    KFS_CUDA_TRY(cudaMemcpyAsync(&n, o->g.raw_count + frame * KFS_MAX_LEVELS + level, sizeof(int), cudaMemcpyDeviceToHost, o->stream));
    KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    if (n > L.list_cap) n = L.list_cap;
    *n_out = n;
    const int c = n < capacity ? n : capacity;
    if (c > 0 && out) {
        KFS_CUDA_TRY(cudaMemcpyAsync(out, o->g.raw + (size_t)frame * o->g.list_fs + L.list_off, sizeof(uint32_t) * (size_t)c, cudaMemcpyDeviceToHost, o->stream));
        KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
    }
    return n > capacity ? KFS_ERR_CAPACITY : KFS_OK;
}

kfs_status kfs_orb_set_profiling(kfs_orb *o, int on)
{
    if (!o) KFS_FAIL(KFS_ERR_INVALID_ARG, "orb handle is NULL");
    o->profiling = on != 0;
    return KFS_OK;
}

kfs_status kfs_orb_stage_times(kfs_orb *o, float *ms)
{
    if (!o || !ms) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_orb_stage_times: bad arguments");
    KFS_DEVICE_GUARD(o->device);
    if (o->profiling && o->ev_made) {
        KFS_CUDA_TRY(cudaStreamSynchronize(o->stream));
        for (int i = 0; i < KFS_STAGE_COUNT; i++) {
            float t = 0;
            if (cudaEventElapsedTime(&t, o->ev[i], o->ev[i + 1]) == cudaSuccess) o->stage_ms[i] = t;
            else cudaGetLastError();
        }
    }
    for (int i = 0; i < KFS_STAGE_COUNT; i++) ms[i] = o->stage_ms[i];
    return KFS_OK;
}

int kfs_orb_last_launches(const kfs_orb *o) { return o ? o->launches : 0; }

} // extern "C"

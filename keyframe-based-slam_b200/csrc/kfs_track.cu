// kfs_track.cu — the callers either side of the two drop-in calls:
//     VisualOdometry::ExtractFeatures   /root/reference/src/lib/VisualOdometry.cpp:21-33
//     VisualOdometry::MatchFeatures     /root/reference/src/lib/VisualOdometry.cpp:35-62
//     KeyFrame (descriptors, keypoints, depth of the last keyframe)   /root/reference/include/lib/KeyFrame.hpp:35-39
// built on the ORB and matcher handles of this library: colour -> gray on the device, a device-resident keyframe cache,
// and ONE kernel after the kNN that does the ratio test, the depth look-up, the pin-hole back-projection and the ordered
// compaction into the 3-D / 2-D correspondence lists solvePnPRansac consumes (:72).
#include "kfs_common.cuh"
#include <new>
#include <string.h>
#include <vector>

// ---- ExtractFeatures :25-28 ----------------------------------------------------------------------------------
// cvtColor(rgb, rgb, RGB2BGR) then cvtColor(rgb, gray, BGR2GRAY): OpenCV's 8-bit path is the 15-bit fixed point
//   gray = (B*3735 + G*19235 + R*9798 + 16384) >> 15      with R = byte 0 of the image the reference receives
// [probe, SURVEY Appendix A "gray probe": bit-exact vs cv2 4.13].  One thread = 4 output pixels = 12 source bytes
// (three aligned 32-bit loads when the row allows it), one 32-bit store; pad bytes of the row replicate the last pixel.
__device__ __forceinline__ uint32_t kfs_gray15(uint32_t r, uint32_t g, uint32_t b) { return (r * 9798u + g * 19235u + b * 3735u + 16384u) >> 15; }

__global__ void __launch_bounds__(256)
k_rgb2gray(const uint8_t *__restrict__ src, long long sstep, long long sfs, uint8_t *__restrict__ dst, int dpitch, long long dfs, int w, int h, int aligned)
{
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    if (x4 >= dpitch || y >= h) return;
    src += (long long)blockIdx.z * sfs; // blockIdx.z = frame of the batch
    dst += (long long)blockIdx.z * dfs;
    const uint8_t *row = src + (long long)y * sstep;
    uint32_t out = 0;
    if (aligned && x4 + 4 <= w) {
        const uint32_t *p = reinterpret_cast<const uint32_t *>(row + 3 * x4);
        const uint32_t a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2); // R0 G0 B0 R1 | G1 B1 R2 G2 | B2 R3 G3 B3
        out = kfs_gray15(a & 0xff, (a >> 8) & 0xff, (a >> 16) & 0xff) |
              kfs_gray15(a >> 24, b & 0xff, (b >> 8) & 0xff) << 8 |
              kfs_gray15((b >> 16) & 0xff, b >> 24, c & 0xff) << 16 |
              kfs_gray15((c >> 8) & 0xff, (c >> 16) & 0xff, c >> 24) << 24;
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int x = min(x4 + j, w - 1);
            out |= kfs_gray15(__ldg(row + 3 * x), __ldg(row + 3 * x + 1), __ldg(row + 3 * x + 2)) << (8 * j);
        }
    }
    *reinterpret_cast<uint32_t *>(dst + (long long)y * dpitch + x4) = out;
}

// nframes colour frames (sfs bytes apart) -> gray frames (dfs bytes apart); src rows 4-byte aligned and sstep % 4 == 0 take the
// three-word path.  Shared with the ORB handle's host-batch pipeline (kfs_orb.cu) and the replay executor.
kfs_status kfs_launch_rgb2gray(const uint8_t *src, long long sstep, long long sfs, uint8_t *dst, int dpitch, long long dfs, int w, int h, int nframes,
                               cudaStream_t st)
{
    const int aligned = ((uintptr_t)src % 4 == 0) && (sstep % 4 == 0) && (sfs % 4 == 0);
    k_rgb2gray<<<dim3(kfs_div_up(dpitch, 128), kfs_div_up(h, 8), nframes), dim3(32, 8), 0, st>>>(src, sstep, sfs, dst, dpitch, dfs, w, h, aligned);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_rgb2gray launch: %s", cudaGetErrorString(e));
    return KFS_OK;
}

// ---- MatchFeatures :45-59 --------------------------------------------------------------------------------------
// One CTA walks the keyframe's kNN rows in query order (the reference's loop order), 1024 rows per round:
//   keep  <=> both neighbours exist and d0 < ratio * d1                                     (:46, fp32)
//   z = (float)(depth(cvRound(pt.y), cvRound(pt.x)) * 0.001)                                (:52, int -> double -> float)
//   y = (float)((pt.y - K(1,2)) * z / K(1,1)),  x likewise with K(0,2), K(0,0)              (:53-54, double -> float)
//   points3D.push_back(Point3d(x, y, z)); points2D.push_back(kpCurr.pt)                     (:56-57, float -> double)
// with IEEE double multiply / divide / subtract (no contraction), so every output equals the CPU's bit for bit.
// A block scan over the keep flags gives each survivor its slot: the lists come out exactly as the reference's
// push_back sequence.
struct TrackArgs {
    const kfs_dmatch *knn;        // [nq][2]
    const kfs_keypoint *kf_kp;    // keyframe keypoints (query side)
    const kfs_keypoint *cur_kp;   // current keypoints (train side)
    const int *d_ncur;            // device count of the current frame (nullptr: n_cur is exact)
    const uint16_t *depth;        // keyframe depth, rows depth_pitch ELEMENTS apart
    long long depth_pitch;
    int depth_w, depth_h;
    double fx, fy, cx, cy;
    float ratio;
    int nq, n_cur, capacity;
    kfs_dmatch *matches;          // [capacity]
    double *p3d, *p2d;            // [capacity][3], [capacity][2]
    int *n_out;                   // total survivors
};

__global__ void __launch_bounds__(1024)
k_ratio_backproject(const TrackArgs A)
{
    __shared__ int s_warp[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    int ncur = A.n_cur;
    if (A.d_ncur) ncur = min(ncur, *A.d_ncur);
    for (int q0 = 0; q0 < A.nq; q0 += 1024) {
        const int q = q0 + tid;
        bool keep = false;
        kfs_dmatch m0 = {0, 0, 0, 0.f};
        if (q < A.nq) {
            const uint4 r0 = *reinterpret_cast<const uint4 *>(A.knn + 2 * (size_t)q), r1 = *reinterpret_cast<const uint4 *>(A.knn + 2 * (size_t)q + 1);
            m0 = *reinterpret_cast<const kfs_dmatch *>(&r0);
            const kfs_dmatch m1 = *reinterpret_cast<const kfs_dmatch *>(&r1);
            // rows with fewer than 2 neighbours (fewer than 2 current keypoints) can never pass; the reference would index
            // knnMatches[i][1] out of bounds there (:46)
            keep = m0.trainIdx >= 0 && m1.trainIdx >= 0 && m0.trainIdx < ncur && m0.distance < __fmul_rn(A.ratio, m1.distance);
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int wsum = s_warp[lane]; // 32 warps
        int incl = wsum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        const int wbase = __shfl_sync(0xffffffffu, incl - wsum, warp);
        const int slot = s_base + wbase + __popc(bal & ((1u << lane) - 1u));
        if (keep && slot < A.capacity) {
            const kfs_keypoint kp = A.kf_kp[m0.queryIdx];
            const kfs_keypoint kc = A.cur_kp[m0.trainIdx];
            const int ix = min(max(__float2int_rn(kp.x), 0), A.depth_w - 1), iy = min(max(__float2int_rn(kp.y), 0), A.depth_h - 1);
            const int d = (int)A.depth[(long long)iy * A.depth_pitch + ix];
            const float z = __double2float_rn(__dmul_rn((double)d, 0.001));
            const float y = __double2float_rn(__ddiv_rn(__dmul_rn(__dsub_rn((double)kp.y, A.cy), (double)z), A.fy));
            const float x = __double2float_rn(__ddiv_rn(__dmul_rn(__dsub_rn((double)kp.x, A.cx), (double)z), A.fx));
            A.matches[slot] = m0;
            A.p3d[3 * (size_t)slot + 0] = (double)x;
            A.p3d[3 * (size_t)slot + 1] = (double)y;
            A.p3d[3 * (size_t)slot + 2] = (double)z;
            A.p2d[2 * (size_t)slot + 0] = (double)kc.x;
            A.p2d[2 * (size_t)slot + 1] = (double)kc.y;
        }
        __syncthreads();
        if (tid == 0) s_base += total;
        __syncthreads();
    }
    if (tid == 0) *A.n_out = s_base;
}

// ---- handle ------------------------------------------------------------------------------------------------------
struct KfSlot { // one cached keyframe
    uint8_t *desc = nullptr;
    kfs_keypoint *kp = nullptr;
    uint16_t *depth = nullptr;
    size_t depth_elems = 0;
    int n = -1, w = 0, h = 0;
};

struct kfs_frontend {
    int device = 0;
    kfs_orb *orb = nullptr;
    kfs_bf *bf = nullptr;
    cudaStream_t stream = nullptr; // the ORB handle's stream; the matcher is moved onto it
    int cap = 0;                   // keypoint rows per frame (kfs_orb_capacity)
    // colour / gray / mask staging of ExtractFeatures
    uint8_t *d_img = nullptr, *d_gray = nullptr, *d_mask = nullptr;
    size_t img_bytes = 0, gray_bytes = 0, mask_bytes = 0;
    // current frame: device views of the last extraction, or uploads of the caller's rows
    const kfs_keypoint *d_cur_kp = nullptr;
    const uint8_t *d_cur_desc = nullptr;
    const int *d_cur_count = nullptr;
    int cur_n = -1, cur_w = 0, cur_h = 0;
    uint8_t *d_up_desc = nullptr;
    kfs_keypoint *d_up_kp = nullptr;
    // keyframe cache (KeyFrame.hpp:35-39): a ring of the last `window` keyframes (Map::GetNKeyFrames, Map.cpp:72-78);
    // ring[head] is the newest one (Map::GetLastKeyFrame, Map.cpp:68-70)
    std::vector<KfSlot> ring;
    int window = 1, head = -1, count = 0;
    int window_wanted = 0;         // ring size to rebuild after a failed kfs_frontend_set_capacity left the handle unsized
    // window-matching outputs: [window][cap][2] kNN rows, [window][cap] survivors, [window] counts
    kfs_dmatch *d_wknn = nullptr, *d_wgood = nullptr;
    int *d_wn = nullptr;
    int *h_wn = nullptr; // pinned
    // match outputs
    kfs_dmatch *d_knn = nullptr, *d_good = nullptr;
    double *d_p3d = nullptr, *d_p2d = nullptr;
    int *d_nout = nullptr;
    int *h_nout = nullptr; // pinned
    // kfs_frontend_track: keypoints / descriptors go home on a side stream while the matcher already runs on `stream`
    cudaStream_t s_copy = nullptr;
    cudaEvent_t ev_extracted = nullptr, ev_copied = nullptr;
    int launches = 0;
};

static kfs_status fe_reserve(uint8_t **p, size_t *have, size_t need)
{
    if (need <= *have) return KFS_OK;
    if (*p) { cudaFree(*p); *p = nullptr; *have = 0; }
    const size_t cap = kfs_align_up(need + need / 4, 1 << 16);
    KFS_CUDA_TRY(cudaMalloc(p, cap));
    *have = cap;
    return KFS_OK;
}

static void fe_free_ring(kfs_frontend *fe)
{
    for (KfSlot &k : fe->ring) {
        if (k.desc) cudaFree(k.desc);
        if (k.kp) cudaFree(k.kp);
        if (k.depth) cudaFree(k.depth);
    }
    fe->ring.clear();
    if (fe->d_wknn) cudaFree(fe->d_wknn);
    if (fe->d_wgood) cudaFree(fe->d_wgood);
    if (fe->d_wn) cudaFree(fe->d_wn);
    fe->d_wknn = fe->d_wgood = nullptr;
    fe->d_wn = nullptr;
    fe->head = -1;
    fe->count = 0;
}

// (Re)builds the ring for `window` keyframes; every cached keyframe is dropped.
static kfs_status fe_alloc_ring(kfs_frontend *fe, int window)
{
    fe_free_ring(fe);
    fe->window = window;
    fe->ring.resize((size_t)window);
    const size_t n = (size_t)fe->cap;
    cudaError_t e = cudaSuccess;
    for (KfSlot &k : fe->ring) {
        if (e == cudaSuccess) e = cudaMalloc(&k.desc, n * 32);
        if (e == cudaSuccess) e = cudaMalloc(&k.kp, n * sizeof(kfs_keypoint));
    }
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_wknn, (size_t)window * n * 2 * sizeof(kfs_dmatch));
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_wgood, (size_t)window * n * sizeof(kfs_dmatch));
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_wn, (size_t)window * sizeof(int));
    if (e != cudaSuccess) {
        fe_free_ring(fe);
        fe->window = 0;
        KFS_FAIL(KFS_ERR_CUDA, "keyframe ring of %d: %s", window, cudaGetErrorString(e));
    }
    return KFS_OK;
}

// Buffers with one row per keypoint: sized from the extractor's capacity (kfs_orb_capacity), rebuilt by kfs_frontend_set_capacity.
static kfs_status fe_alloc_rows(kfs_frontend *fe)
{
    void **bufs[] = {(void **)&fe->d_up_desc, (void **)&fe->d_up_kp, (void **)&fe->d_knn, (void **)&fe->d_good, (void **)&fe->d_p3d, (void **)&fe->d_p2d};
    for (void **b : bufs) if (*b) { cudaFree(*b); *b = nullptr; }
    fe->cap = kfs_orb_capacity(fe->orb);
    const size_t n = (size_t)fe->cap;
    cudaError_t e = cudaMalloc(&fe->d_up_desc, n * 32);
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_up_kp, n * sizeof(kfs_keypoint));
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_knn, n * 2 * sizeof(kfs_dmatch));
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_good, n * sizeof(kfs_dmatch));
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_p3d, n * 3 * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_p2d, n * 2 * sizeof(double));
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "kfs_frontend: row buffers for capacity %d: %s", fe->cap, cudaGetErrorString(e));
    return KFS_OK;
}

// Slot the next keyframe goes into (committed by fe_commit_slot once its uploads are enqueued).
static KfSlot *fe_next_slot(kfs_frontend *fe) { return fe->window > 0 ? &fe->ring[(size_t)((fe->head + 1) % fe->window)] : nullptr; }
static void fe_commit_slot(kfs_frontend *fe)
{
    fe->head = (fe->head + 1) % fe->window;
    if (fe->count < fe->window) fe->count++;
}
static KfSlot *fe_newest(kfs_frontend *fe) { return fe->count > 0 ? &fe->ring[(size_t)fe->head] : nullptr; }
// i-th keyframe of Map::GetNKeyFrames' order (0 = oldest held)
static KfSlot *fe_window_slot(kfs_frontend *fe, int i) { return &fe->ring[(size_t)((fe->head - fe->count + 1 + i + 2 * fe->window) % fe->window)]; }

// A kfs_frontend_set_capacity (or set_window) that ran out of device memory leaves the handle UNSIZED: no row buffers, no keyframe
// ring, capacity 0.  Every computing entry point refuses such a handle instead of running with null device buffers; a later
// successful kfs_frontend_set_capacity makes it usable again.
#define KFS_FE_REQUIRE_SIZED(fe)                                                                                                    \
    if ((fe)->cap <= 0 || (fe)->window <= 0)                                                                                        \
    KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_frontend handle is unsized (an earlier kfs_frontend_set_capacity / set_window failed): call "   \
                                  "kfs_frontend_set_capacity again")
static void fe_unsize(kfs_frontend *fe)
{
    void **bufs[] = {(void **)&fe->d_up_desc, (void **)&fe->d_up_kp, (void **)&fe->d_knn, (void **)&fe->d_good, (void **)&fe->d_p3d, (void **)&fe->d_p2d};
    for (void **b : bufs) if (*b) { cudaFree(*b); *b = nullptr; }
    fe_free_ring(fe);
    fe->window = 0;
    fe->cap = 0;
    fe->cur_n = -1; fe->d_cur_desc = nullptr; fe->d_cur_kp = nullptr; fe->d_cur_count = nullptr;
    cudaGetLastError();
}

extern "C" {

void kfs_frontend_destroy(kfs_frontend *fe)
{
    if (!fe) return;
    KfsDeviceGuard guard(fe->device);
    if (fe->stream) cudaStreamSynchronize(fe->stream);
    if (fe->bf) kfs_bf_destroy(fe->bf);
    if (fe->orb) kfs_orb_destroy(fe->orb);
    void *bufs[] = {fe->d_img, fe->d_gray, fe->d_mask, fe->d_up_desc, fe->d_up_kp, fe->d_knn, fe->d_good, fe->d_p3d, fe->d_p2d, fe->d_nout};
    for (void *b : bufs) if (b) cudaFree(b);
    fe_free_ring(fe);
    if (fe->h_nout) cudaFreeHost(fe->h_nout);
    if (fe->s_copy) { cudaStreamSynchronize(fe->s_copy); cudaStreamDestroy(fe->s_copy); }
    if (fe->ev_extracted) cudaEventDestroy(fe->ev_extracted);
    if (fe->ev_copied) cudaEventDestroy(fe->ev_copied);
    if (fe->h_wn) cudaFreeHost(fe->h_wn);
    delete fe;
}

kfs_status kfs_frontend_create(const kfs_orb_params *params, int device, kfs_frontend **out)
{
    if (!out) KFS_FAIL(KFS_ERR_INVALID_ARG, "out is NULL");
    *out = nullptr;
    kfs_frontend *fe = new (std::nothrow) kfs_frontend();
    if (!fe) KFS_FAIL(KFS_ERR_INVALID_ARG, "out of host memory");
    fe->device = device;
    kfs_status s = kfs_orb_create(params, device, &fe->orb);
    if (!s) s = kfs_bf_create(KFS_NORM_HAMMING, 0, device, &fe->bf);
    if (s) { kfs_frontend_destroy(fe); return s; }
    KfsDeviceGuard guard(device);
    if (!guard.ok) { kfs_frontend_destroy(fe); KFS_FAIL(KFS_ERR_CUDA, "cudaSetDevice(%d) failed", device); }
    fe->stream = (cudaStream_t)kfs_orb_get_stream(fe->orb);
    kfs_bf_set_stream(fe->bf, (void *)fe->stream);
    cudaError_t e = cudaMallocHost(&fe->h_nout, 2 * sizeof(int));
    if (e == cudaSuccess) e = cudaMallocHost(&fe->h_wn, KFS_MAX_WINDOW * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_nout, sizeof(int));
    if (e != cudaSuccess) {
        kfs_frontend_destroy(fe);
        KFS_FAIL(KFS_ERR_CUDA, "kfs_frontend_create: %s", cudaGetErrorString(e));
    }
    s = fe_alloc_rows(fe);
    if (s) { kfs_frontend_destroy(fe); return s; }
    s = fe_alloc_ring(fe, 1);
    if (s) { kfs_frontend_destroy(fe); return s; }
    *out = fe;
    return KFS_OK;
}

int kfs_frontend_capacity(const kfs_frontend *fe) { return fe ? fe->cap : 0; }

// Rows per frame the handle can hold (kfs_orb_set_capacity of the extractor + every per-row buffer + the keyframe ring).  The
// cached keyframes and the current frame are dropped, like kfs_frontend_set_window: call it before tracking starts, or after
// KFS_ERR_CAPACITY told how many rows a tie-rich frame needs.
kfs_status kfs_frontend_set_capacity(kfs_frontend *fe, int capacity)
{
    if (!fe || capacity < 1) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_frontend_set_capacity: bad arguments");
    KFS_DEVICE_GUARD(fe->device);
    KFS_CUDA_TRY(cudaStreamSynchronize(fe->stream));
    const int window = fe->window > 0 ? fe->window : (fe->window_wanted > 0 ? fe->window_wanted : 1);
    kfs_status s = kfs_orb_set_capacity(fe->orb, capacity);
    fe->cur_n = -1; fe->d_cur_desc = nullptr; fe->d_cur_kp = nullptr; fe->d_cur_count = nullptr;
    if (!s) s = fe_alloc_rows(fe);
    if (!s) s = fe_alloc_ring(fe, window);
    if (s) { fe->window_wanted = window; fe_unsize(fe); return s; } // the error text of the failing step stays in kfs_last_error()
    return KFS_OK;
}
kfs_status kfs_frontend_set_blur_fma(kfs_frontend *fe, int fma)
{
    if (!fe) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle is NULL");
    return kfs_orb_set_blur_fma(fe->orb, fma);
}
int kfs_frontend_last_launches(const kfs_frontend *fe) { return fe ? fe->launches : 0; }

// Validates the frame, uploads it (+ mask), converts colour and runs the ORB chain; nothing synchronises.  On return the
// device views of the current frame are set; cur_n is still unknown (-2 = "on the device, count pending").
static kfs_status fe_enqueue_extract(kfs_frontend *fe, const uint8_t *image, size_t step, int w, int h, int channels,
                                     const uint8_t *mask, size_t mask_step, const char *who)
{
    if (channels != 1 && channels != 3) KFS_FAIL(KFS_ERR_UNSUPPORTED, "%s: %d channels (the reference handles 1 or 3, VisualOdometry.cpp:27)", who, channels);
    if (w <= 0 || h <= 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "%s: empty image", who);
    if (!image || step < (size_t)w * channels) KFS_FAIL(KFS_ERR_INVALID_ARG, "image is NULL or step < w*channels");
    if (mask && mask_step < (size_t)w) KFS_FAIL(KFS_ERR_INVALID_ARG, "mask_step < w");
    cudaStream_t st = fe->stream;
    const size_t gpitch = kfs_align_up((size_t)w, 128);
    kfs_status s = fe_reserve(&fe->d_gray, &fe->gray_bytes, gpitch * h);
    if (s) return s;
    fe->launches = 0;
    if (channels == 3) {
        const size_t ipitch = kfs_align_up((size_t)w * 3, 16);
        s = fe_reserve(&fe->d_img, &fe->img_bytes, ipitch * h);
        if (s) return s;
        KFS_CUDA_TRY(cudaMemcpy2DAsync(fe->d_img, ipitch, image, step, (size_t)w * 3, h, cudaMemcpyHostToDevice, st));
        s = kfs_launch_rgb2gray(fe->d_img, (long long)ipitch, 0, fe->d_gray, (int)gpitch, 0, w, h, 1, st);
        if (s) return s;
        fe->launches++;
    } else {
        KFS_CUDA_TRY(cudaMemcpy2DAsync(fe->d_gray, gpitch, image, step, (size_t)w, h, cudaMemcpyHostToDevice, st));
    }
    const uint8_t *d_mask = nullptr;
    if (mask) {
        s = fe_reserve(&fe->d_mask, &fe->mask_bytes, gpitch * h);
        if (s) return s;
        KFS_CUDA_TRY(cudaMemcpy2DAsync(fe->d_mask, gpitch, mask, mask_step, (size_t)w, h, cudaMemcpyHostToDevice, st));
        d_mask = fe->d_mask;
    }
    s = kfs_orb_extract_device(fe->orb, fe->d_gray, gpitch, gpitch * h, w, h, 1, d_mask, gpitch, gpitch * h);
    if (s) return s;
    fe->launches += kfs_orb_last_launches(fe->orb);
    int dcap = 0, nb = 0;
    s = kfs_orb_device_results(fe->orb, &fe->d_cur_kp, &fe->d_cur_desc, &fe->d_cur_count, &dcap, &nb);
    if (s) return s;
    fe->cur_w = w; fe->cur_h = h;
    return KFS_OK;
}

kfs_status kfs_frontend_extract_features(kfs_frontend *fe, const uint8_t *image, size_t step, int w, int h, int channels,
                                         const uint8_t *mask, size_t mask_step,
                                         kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out)
{
    if (!fe || !n_out) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle or n_out is NULL");
    KFS_FE_REQUIRE_SIZED(fe);
    *n_out = 0;
    if (channels != 1 && channels != 3) KFS_FAIL(KFS_ERR_UNSUPPORTED, "ExtractFeatures: %d channels (the reference handles 1 or 3, VisualOdometry.cpp:27)", channels);
    if (w < 0 || h < 0 || capacity < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "negative size");
    if (w == 0 || h == 0) { fe->cur_n = 0; return KFS_OK; }
    if (capacity > 0 && (!kp_out || !desc_out)) KFS_FAIL(KFS_ERR_INVALID_ARG, "output buffers are NULL");
    KFS_DEVICE_GUARD(fe->device);
    kfs_status s = fe_enqueue_extract(fe, image, step, w, h, channels, mask, mask_step, "ExtractFeatures");
    if (s) return s;
    s = kfs_orb_fetch_results(fe->orb, kp_out, desc_out, capacity, n_out);
    fe->cur_n = *n_out < fe->cap ? *n_out : fe->cap;
    return s;
}

static kfs_status fe_upload_depth(kfs_frontend *fe, KfSlot *k, const uint16_t *depth, size_t depth_step, int w, int h)
{
    if (!depth || w <= 0 || h <= 0 || depth_step < (size_t)w * 2) KFS_FAIL(KFS_ERR_INVALID_ARG, "depth is NULL, empty or depth_step < 2*w");
    const size_t elems = (size_t)w * h;
    if (elems > k->depth_elems) {
        if (k->depth) { KFS_CUDA_TRY(cudaStreamSynchronize(fe->stream)); cudaFree(k->depth); k->depth = nullptr; k->depth_elems = 0; }
        KFS_CUDA_TRY(cudaMalloc(&k->depth, elems * 2));
        k->depth_elems = elems;
    }
    KFS_CUDA_TRY(cudaMemcpy2DAsync(k->depth, (size_t)w * 2, depth, depth_step, (size_t)w * 2, h, cudaMemcpyHostToDevice, fe->stream));
    k->w = w; k->h = h;
    return KFS_OK;
}

kfs_status kfs_frontend_promote_keyframe(kfs_frontend *fe, const uint16_t *depth, size_t depth_step)
{
    if (!fe) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle is NULL");
    KFS_FE_REQUIRE_SIZED(fe);
    if (fe->cur_n < 0 || !fe->d_cur_desc) KFS_FAIL(KFS_ERR_INVALID_ARG, "promote_keyframe: no frame has been extracted on this handle yet");
    KFS_DEVICE_GUARD(fe->device);
    KfSlot *k = fe_next_slot(fe);
    kfs_status s = fe_upload_depth(fe, k, depth, depth_step, fe->cur_w, fe->cur_h);
    if (s) return s;
    if (fe->cur_n > 0) {
        KFS_CUDA_TRY(cudaMemcpyAsync(k->desc, fe->d_cur_desc, (size_t)fe->cur_n * 32, cudaMemcpyDeviceToDevice, fe->stream));
        KFS_CUDA_TRY(cudaMemcpyAsync(k->kp, fe->d_cur_kp, (size_t)fe->cur_n * sizeof(kfs_keypoint), cudaMemcpyDeviceToDevice, fe->stream));
    }
    // the caller may reuse its depth buffer for the next frame right away (a pinned buffer makes the upload truly asynchronous)
    KFS_CUDA_TRY(cudaStreamSynchronize(fe->stream));
    k->n = fe->cur_n;
    fe_commit_slot(fe);
    return KFS_OK;
}

kfs_status kfs_frontend_set_keyframe(kfs_frontend *fe, const uint8_t *desc, const kfs_keypoint *kp, int n,
                                     const uint16_t *depth, size_t depth_step, int w, int h)
{
    if (!fe) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle is NULL");
    KFS_FE_REQUIRE_SIZED(fe);
    if (n < 0 || n > fe->cap) KFS_FAIL(KFS_ERR_CAPACITY, "set_keyframe: %d rows, handle capacity %d", n, fe->cap);
    if (n > 0 && (!desc || !kp)) KFS_FAIL(KFS_ERR_INVALID_ARG, "desc or kp is NULL");
    KFS_DEVICE_GUARD(fe->device);
    KfSlot *k = fe_next_slot(fe);
    kfs_status s = fe_upload_depth(fe, k, depth, depth_step, w, h);
    if (s) return s;
    if (n > 0) {
        KFS_CUDA_TRY(cudaMemcpyAsync(k->desc, desc, (size_t)n * 32, cudaMemcpyHostToDevice, fe->stream));
        KFS_CUDA_TRY(cudaMemcpyAsync(k->kp, kp, (size_t)n * sizeof(kfs_keypoint), cudaMemcpyHostToDevice, fe->stream));
    }
    KFS_CUDA_TRY(cudaStreamSynchronize(fe->stream)); // the caller may reuse its (pageable) buffers right away
    k->n = n;
    fe_commit_slot(fe);
    return KFS_OK;
}

// kNN against keyframe `kf` + ratio test / back-projection + the device-to-host copies of the count and the lists; nothing
// synchronises.  count_on_device: `nt` is only a capacity, the live row count is *d_cnt (a frame whose extraction is still in
// flight).
static kfs_status fe_enqueue_match(kfs_frontend *fe, KfSlot *kf, const kfs_keypoint *d_kp, const uint8_t *d_desc, const int *d_cnt, int nt,
                                   bool count_on_device, const double *intrinsics, float ratio, int capacity,
                                   kfs_dmatch *matches_out, double *points3d_out, double *points2d_out)
{
    // the reference indexes knnMatches[i][1] unguarded (:46): with fewer than 2 current keypoints it is undefined there;
    // here such rows simply never pass
    kfs_bf_pair pair;
    pair.query = kf->desc; pair.nq = kf->n; pair.d_nq = nullptr;
    pair.train = d_desc; pair.nt = nt; pair.d_nt = count_on_device ? d_cnt : nullptr;
    pair.out = fe->d_knn; pair.counts = nullptr;
    cudaStream_t st = fe->stream;
    kfs_status s = kfs_bf_knn_match_pairs_device(fe->bf, &pair, 1, 2);
    if (s) return s;
    fe->launches += kfs_bf_last_launches(fe->bf) > 0 ? 1 : 0;
    TrackArgs A;
    A.knn = fe->d_knn; A.kf_kp = kf->kp; A.cur_kp = d_kp; A.d_ncur = d_cnt;
    A.depth = kf->depth; A.depth_pitch = kf->w; A.depth_w = kf->w; A.depth_h = kf->h;
    A.fx = intrinsics[0]; A.fy = intrinsics[1]; A.cx = intrinsics[2]; A.cy = intrinsics[3];
    A.ratio = ratio; A.nq = kf->n; A.n_cur = nt;
    A.capacity = capacity < fe->cap ? capacity : fe->cap;
    A.matches = fe->d_good; A.p3d = fe->d_p3d; A.p2d = fe->d_p2d; A.n_out = fe->d_nout;
    k_ratio_backproject<<<1, 1024, 0, st>>>(A);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_ratio_backproject launch: %s", cudaGetErrorString(e));
    fe->launches++;
    // one round trip: the survivor count and the lists' upper bound (at most one survivor per keyframe row) travel together
    KFS_CUDA_TRY(cudaMemcpyAsync(fe->h_nout, fe->d_nout, sizeof(int), cudaMemcpyDeviceToHost, st));
    const int m = kf->n < A.capacity ? kf->n : A.capacity;
    if (m > 0) {
        KFS_CUDA_TRY(cudaMemcpyAsync(matches_out, fe->d_good, (size_t)m * sizeof(kfs_dmatch), cudaMemcpyDeviceToHost, st));
        KFS_CUDA_TRY(cudaMemcpyAsync(points3d_out, fe->d_p3d, (size_t)m * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
        KFS_CUDA_TRY(cudaMemcpyAsync(points2d_out, fe->d_p2d, (size_t)m * 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    }
    return KFS_OK;
}

kfs_status kfs_frontend_match_features(kfs_frontend *fe, const uint8_t *cur_desc, const kfs_keypoint *cur_kp, int n_cur,
                                       const double *intrinsics, float ratio,
                                       kfs_dmatch *matches_out, double *points3d_out, double *points2d_out,
                                       int capacity, int *n_out)
{
    if (!fe || !n_out || !intrinsics) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle, intrinsics or n_out is NULL");
    KFS_FE_REQUIRE_SIZED(fe);
    *n_out = 0;
    KfSlot *kf = fe_newest(fe);
    if (!kf) KFS_FAIL(KFS_ERR_INVALID_ARG, "match_features: no keyframe cached (promote_keyframe / set_keyframe first)");
    if (capacity < 0 || (capacity > 0 && (!matches_out || !points3d_out || !points2d_out))) KFS_FAIL(KFS_ERR_INVALID_ARG, "output buffers");
    if (intrinsics[0] == 0.0 || intrinsics[1] == 0.0) KFS_FAIL(KFS_ERR_INVALID_ARG, "fx or fy is zero");
    KFS_DEVICE_GUARD(fe->device);
    cudaStream_t st = fe->stream;
    const kfs_keypoint *d_kp;
    const uint8_t *d_desc;
    const int *d_cnt = nullptr;
    int nt;
    if (cur_desc || cur_kp) {
        if (!cur_desc || !cur_kp || n_cur < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "cur_desc / cur_kp must both be given with n_cur >= 0");
        if (n_cur > fe->cap) KFS_FAIL(KFS_ERR_CAPACITY, "match_features: %d current rows, handle capacity %d", n_cur, fe->cap);
        if (n_cur > 0) {
            KFS_CUDA_TRY(cudaMemcpyAsync(fe->d_up_desc, cur_desc, (size_t)n_cur * 32, cudaMemcpyHostToDevice, st));
            KFS_CUDA_TRY(cudaMemcpyAsync(fe->d_up_kp, cur_kp, (size_t)n_cur * sizeof(kfs_keypoint), cudaMemcpyHostToDevice, st));
        }
        d_kp = fe->d_up_kp; d_desc = fe->d_up_desc; nt = n_cur;
    } else {
        if (fe->cur_n < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "match_features: no current frame on the device (extract_features first)");
        d_kp = fe->d_cur_kp; d_desc = fe->d_cur_desc; d_cnt = fe->d_cur_count; nt = fe->cur_n;
    }
    fe->launches = 0;
    if (kf->n == 0) return KFS_OK; // knnMatch of an empty query: no rows
    kfs_status s = fe_enqueue_match(fe, kf, d_kp, d_desc, d_cnt, nt, false, intrinsics, ratio, capacity, matches_out, points3d_out, points2d_out);
    if (s) return s;
    KFS_CUDA_TRY(cudaStreamSynchronize(st));
    const int n = *fe->h_nout;
    *n_out = n;
    if (n > capacity) KFS_FAIL(KFS_ERR_CAPACITY, "match_features: %d survivors, output capacity %d", n, capacity);
    return KFS_OK;
}

// VisualOdometry::Track's front half (:138 ExtractFeatures, :144 MatchFeatures) in ONE call with ONE host synchronisation:
// upload, colour conversion, the ORB chain, kNN against the cached keyframe, ratio test + back-projection and every
// device-to-host copy are enqueued back to back.  Without a cached keyframe only the extraction runs (*n_matches_out = -1).
kfs_status kfs_frontend_track(kfs_frontend *fe, const uint8_t *image, size_t step, int w, int h, int channels,
                              const uint8_t *mask, size_t mask_step, const double *intrinsics, float ratio,
                              kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out,
                              kfs_dmatch *matches_out, double *points3d_out, double *points2d_out, int match_capacity, int *n_matches_out)
{
    if (!fe || !n_out || !n_matches_out || !intrinsics) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle, intrinsics, n_out or n_matches_out is NULL");
    KFS_FE_REQUIRE_SIZED(fe);
    *n_out = 0; *n_matches_out = -1;
    if (capacity < 0 || match_capacity < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "negative capacity");
    if (capacity > 0 && (!kp_out || !desc_out)) KFS_FAIL(KFS_ERR_INVALID_ARG, "output buffers are NULL");
    if (match_capacity > 0 && (!matches_out || !points3d_out || !points2d_out)) KFS_FAIL(KFS_ERR_INVALID_ARG, "match output buffers are NULL");
    if (intrinsics[0] == 0.0 || intrinsics[1] == 0.0) KFS_FAIL(KFS_ERR_INVALID_ARG, "fx or fy is zero");
    KFS_DEVICE_GUARD(fe->device);
    cudaStream_t st = fe->stream;
    kfs_status s = fe_enqueue_extract(fe, image, step, w, h, channels, mask, mask_step, "Track");
    if (s) return s;
    const int cp = capacity < fe->cap ? capacity : fe->cap;
    if (!fe->s_copy) {
        KFS_CUDA_TRY(cudaStreamCreateWithFlags(&fe->s_copy, cudaStreamNonBlocking));
        KFS_CUDA_TRY(cudaEventCreateWithFlags(&fe->ev_extracted, cudaEventDisableTiming));
        KFS_CUDA_TRY(cudaEventCreateWithFlags(&fe->ev_copied, cudaEventDisableTiming));
    }
    // the frame's own results leave on the side stream; the matcher does not queue behind their copies
    KFS_CUDA_TRY(cudaEventRecord(fe->ev_extracted, st));
    KFS_CUDA_TRY(cudaStreamWaitEvent(fe->s_copy, fe->ev_extracted, 0));
    KFS_CUDA_TRY(cudaMemcpyAsync(fe->h_nout + 1, fe->d_cur_count, sizeof(int), cudaMemcpyDeviceToHost, fe->s_copy));
    if (cp > 0) {
        KFS_CUDA_TRY(cudaMemcpyAsync(kp_out, fe->d_cur_kp, (size_t)cp * sizeof(kfs_keypoint), cudaMemcpyDeviceToHost, fe->s_copy));
        KFS_CUDA_TRY(cudaMemcpyAsync(desc_out, fe->d_cur_desc, (size_t)cp * 32, cudaMemcpyDeviceToHost, fe->s_copy));
    }
    KFS_CUDA_TRY(cudaEventRecord(fe->ev_copied, fe->s_copy));
    KfSlot *kf = fe_newest(fe);
    const bool matched = kf && kf->n > 0;
    if (matched) {
        s = fe_enqueue_match(fe, kf, fe->d_cur_kp, fe->d_cur_desc, fe->d_cur_count, fe->cap, true, intrinsics, ratio, match_capacity,
                             matches_out, points3d_out, points2d_out);
        if (s) { cudaStreamSynchronize(fe->s_copy); return s; } // nothing may land in the caller's buffers after the return
    }
    KFS_CUDA_TRY(cudaStreamWaitEvent(st, fe->ev_copied, 0)); // join: still ONE host synchronisation per frame
    KFS_CUDA_TRY(cudaStreamSynchronize(st));
    const int n = fe->h_nout[1];
    *n_out = n;
    fe->cur_n = n < fe->cap ? n : fe->cap;
    if (kf) *n_matches_out = matched ? fe->h_nout[0] : 0;
    if (n > cp) KFS_FAIL(KFS_ERR_CAPACITY, "Track: %d keypoints, output capacity %d (handle capacity %d)", n, capacity, fe->cap);
    if (matched && fe->h_nout[0] > match_capacity) KFS_FAIL(KFS_ERR_CAPACITY, "Track: %d survivors, output capacity %d", fe->h_nout[0], match_capacity);
    return KFS_OK;
}

// ---- keyframe window: Map::GetNKeyFrames (Map.cpp:72-78) -----------------------------------------------------------
kfs_status kfs_frontend_set_window(kfs_frontend *fe, int window)
{
    if (!fe) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle is NULL");
    if (window < 1 || window > KFS_MAX_WINDOW) KFS_FAIL(KFS_ERR_INVALID_ARG, "set_window: %d not in [1, %d]", window, KFS_MAX_WINDOW);
    KFS_DEVICE_GUARD(fe->device);
    KFS_CUDA_TRY(cudaStreamSynchronize(fe->stream));
    if (fe->cap <= 0) { fe->window_wanted = window; KFS_FAIL(KFS_ERR_INVALID_ARG, "set_window: the handle is unsized (a kfs_frontend_set_capacity failed); size it first"); }
    return fe_alloc_ring(fe, window);
}
int kfs_frontend_window(const kfs_frontend *fe) { return fe ? fe->window : 0; }
int kfs_frontend_window_count(const kfs_frontend *fe) { return fe ? fe->count : 0; }

kfs_status kfs_frontend_match_window(kfs_frontend *fe, const uint8_t *cur_desc, int n_cur, float ratio,
                                     int *kf_rows_out, kfs_dmatch *knn_out, int knn_rows,
                                     kfs_dmatch *good_out, int *n_good_out, int capacity, int *n_kf_out)
{
    if (!fe || !n_kf_out) KFS_FAIL(KFS_ERR_INVALID_ARG, "handle or n_kf_out is NULL");
    KFS_FE_REQUIRE_SIZED(fe);
    *n_kf_out = 0;
    if (capacity < 0 || knn_rows < 0 || (capacity > 0 && !good_out) || !n_good_out) KFS_FAIL(KFS_ERR_INVALID_ARG, "match_window: output buffers");
    KFS_DEVICE_GUARD(fe->device);
    cudaStream_t st = fe->stream;
    const uint8_t *d_desc;
    const int *d_cnt = nullptr;
    int nt;
    if (cur_desc) {
        if (n_cur < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "match_window: n_cur < 0");
        if (n_cur > fe->cap) KFS_FAIL(KFS_ERR_CAPACITY, "match_window: %d current rows, handle capacity %d", n_cur, fe->cap);
        if (n_cur > 0) KFS_CUDA_TRY(cudaMemcpyAsync(fe->d_up_desc, cur_desc, (size_t)n_cur * 32, cudaMemcpyHostToDevice, st));
        d_desc = fe->d_up_desc; nt = n_cur;
    } else {
        if (fe->cur_n < 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "match_window: no current frame on the device (extract_features first)");
        d_desc = fe->d_cur_desc; d_cnt = fe->d_cur_count; nt = fe->cur_n;
    }
    fe->launches = 0;
    const int nkf = fe->count;
    *n_kf_out = nkf;
    if (nkf == 0) return KFS_OK;
    const size_t cap = (size_t)fe->cap;
    kfs_bf_pair pairs[KFS_MAX_WINDOW];
    int max_rows = 0;
    for (int i = 0; i < nkf; i++) {
        const KfSlot *k = fe_window_slot(fe, i);
        pairs[i].query = k->desc; pairs[i].nq = k->n; pairs[i].d_nq = nullptr;
        pairs[i].train = d_desc; pairs[i].nt = nt; pairs[i].d_nt = d_cnt;
        pairs[i].out = fe->d_wknn + (size_t)i * cap * 2; pairs[i].counts = nullptr;
        if (kf_rows_out) kf_rows_out[i] = k->n;
        n_good_out[i] = 0;
        max_rows = k->n > max_rows ? k->n : max_rows;
    }
    if (max_rows == 0) return KFS_OK; // every keyframe is empty: knnMatch of an empty query has no rows
    kfs_status s = kfs_bf_knn_match_pairs_device(fe->bf, pairs, nkf, 2);
    if (s) return s;
    fe->launches += kfs_bf_last_launches(fe->bf);
    s = kfs_bf_launch_ratio_table(fe->bf, kfs_bf_device_table(fe->bf), nkf, ratio, fe->d_wgood, (long long)cap, fe->d_wn, st);
    if (s) return s;
    fe->launches++;
    KFS_CUDA_TRY(cudaMemcpyAsync(fe->h_wn, fe->d_wn, (size_t)nkf * sizeof(int), cudaMemcpyDeviceToHost, st));
    bool truncated = false;
    if (knn_out && knn_rows > 0) {
        const int rows = max_rows < knn_rows ? max_rows : knn_rows;
        truncated |= max_rows > knn_rows;
        KFS_CUDA_TRY(cudaMemcpy2DAsync(knn_out, (size_t)knn_rows * 2 * sizeof(kfs_dmatch), fe->d_wknn, cap * 2 * sizeof(kfs_dmatch),
                                       (size_t)rows * 2 * sizeof(kfs_dmatch), (size_t)nkf, cudaMemcpyDeviceToHost, st));
    }
    KFS_CUDA_TRY(cudaStreamSynchronize(st));
    int max_good = 0;
    for (int i = 0; i < nkf; i++) {
        n_good_out[i] = fe->h_wn[i];
        max_good = fe->h_wn[i] > max_good ? fe->h_wn[i] : max_good;
    }
    truncated |= max_good > capacity;
    const int m = max_good < capacity ? max_good : capacity;
    if (m > 0) {
        KFS_CUDA_TRY(cudaMemcpy2DAsync(good_out, (size_t)capacity * sizeof(kfs_dmatch), fe->d_wgood, cap * sizeof(kfs_dmatch),
                                       (size_t)m * sizeof(kfs_dmatch), (size_t)nkf, cudaMemcpyDeviceToHost, st));
        KFS_CUDA_TRY(cudaStreamSynchronize(st));
    }
    if (truncated) KFS_FAIL(KFS_ERR_CAPACITY, "match_window: outputs truncated (max keyframe rows %d vs knn_rows %d, max survivors %d vs capacity %d)",
                            max_rows, knn_rows, max_good, capacity);
    return KFS_OK;
}

} // extern "C"

// kfs_replay.cu — replay executor: the tracker's per-frame loop for batches of frames, kept on the device.
//
// Mirrors /root/reference/src/lib/VisualOdometry.cpp:129-172 (VisualOdometry::Track): ExtractFeatures (:138,
// detectAndCompute :30) -> MatchFeatures against the last keyframe (:144, knnMatch :39, query = keyframe
// descriptors, train = current frame) -> ratio test (:46).  The reference does this one frame at a time through
// host memory; here a batch of frames goes through
//     H2D (chunked) | ORB kernel chain | Hamming kNN on the DEVICE-RESIDENT descriptors | ratio compaction | D2H
// with three arenas: batch i+2 is uploaded while batch i+1 runs its kernels and batch i is downloaded (with two, an arena's
// upload -> kernels -> download cycle is twice the batch period as soon as upload and kernels take equally long).
// Keyframe descriptors never travel back to the device: a keyframe is either a frame of the same batch or an
// entry of the device-resident keyframe cache (KeyFrame::matDescriptors, KeyFrame.hpp:39).
// No host threads: everything is stream-ordered; kfs_replay_submit only enqueues, kfs_replay_wait blocks.
#include "kfs_common.cuh"
#include <string.h>
#include <limits.h>
#include <vector>
#include <new>

constexpr int RP_ARENAS_DEFAULT = 3; // batches in flight: one uploading, one in the kernels, one downloading
constexpr int RP_MAX_ARENAS = KFS_REPLAY_MAX_ARENAS;

struct ReplaySlot {
    bool in_flight = false;
    int batch = 0, ticket = -1;
    kfs_replay_io io;
    cudaEvent_t ev_knn = nullptr, ev_done = nullptr;
    kfs_dmatch *d_knn = nullptr, *d_good = nullptr;
    int *d_small = nullptr;   // [2][max_batch]: n_good | n_query
    int *h_small = nullptr;   // pinned mirror
    BfPairDev *d_table = nullptr;     // device: built by k_replay_table
    std::vector<int> last_plan;       // plan keyframes (+ capacity) d_table was built from
    const uint8_t *last_desc = nullptr;
    std::vector<uint8_t> skip;  // per frame: keypoints / descriptors not wanted on the host
    int rows = 0;               // kNN / survivor rows per frame the eager copies carried
};

struct kfs_replay {
    kfs_replay_params p;
    int device = 0;
    int arenas = RP_ARENAS_DEFAULT;
    kfs_orb *orb[RP_MAX_ARENAS] = {};
    kfs_bf *bf = nullptr;
    cudaStream_t compute = nullptr;
    ReplaySlot slot[RP_MAX_ARENAS];
    uint8_t *d_kf_desc = nullptr; // [slots][cap][32]
    int *d_kf_count = nullptr;    // [slots]
    int cap = 0;
    int seq = 0;
    int launches = 0;
};

// The pair table of a submit, built ON the device from the plan.  The plan travels as kernel arguments (RP_PLAN_CHUNK frames per
// launch), never as a host-to-device copy: a small cudaMemcpyAsync on the compute stream is ordered behind this batch's kernels and
// sits in the copy engine's queue in front of the NEXT batch's frame upload, which then cannot overlap those kernels (cfg3: 3.1 ms per
// submit instead of 2.2).  query = keyframe descriptors (same batch or cache slot), train = the frame's own (VisualOdometry.cpp:39).
constexpr int RP_PLAN_CHUNK = 512;
struct RpPlanArg { int keyframe[RP_PLAN_CHUNK]; };
__global__ void k_replay_table(const __grid_constant__ RpPlanArg plan, int i0, int n, BfPairDev *__restrict__ table, const uint8_t *d_desc,
                               const int *d_cnt, int cap, const uint8_t *d_kf_desc, const int *d_kf_count, kfs_dmatch *d_knn)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int i = i0 + j, k = plan.keyframe[j];
    BfPairDev P;
    P.train = d_desc + (size_t)i * cap * 32; P.nt = cap; P.d_nt = d_cnt + i;
    P.out = d_knn + (size_t)i * cap * 2; P.counts = nullptr;
    if (k == KFS_REPLAY_NO_MATCH) { P.query = d_desc; P.nq = 0; P.d_nq = nullptr; P.nt = 0; }
    else if (k >= 0) { P.query = d_desc + (size_t)k * cap * 32; P.nq = cap; P.d_nq = d_cnt + k; }
    else { const int sl = -(k + 1); P.query = d_kf_desc + (size_t)sl * cap * 32; P.nq = cap; P.d_nq = d_kf_count + sl; }
    table[i] = P;
}

// A frame that becomes a keyframe for later submits: its descriptors and count into a cache slot (a kernel for the same reason: a
// device-to-device cudaMemcpyAsync may be served by a copy engine and queue in front of an upload or a download)
__global__ void k_replay_store(const uint4 *__restrict__ src, const int *__restrict__ src_cnt, uint4 *__restrict__ dst, int *__restrict__ dst_cnt, int n16)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) dst[i] = src[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) *dst_cnt = *src_cnt;
}

__global__ void k_replay_nquery(const BfPairDev *__restrict__ pairs, int npairs, int *__restrict__ n_query)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npairs) return;
    const BfPairDev P = pairs[i];
    int nq = P.nq;
    if (P.d_nq) nq = min(nq, *P.d_nq);
    n_query[i] = nq;
}

static kfs_status replay_finish(kfs_replay *r, int b)
{
    ReplaySlot &S = r->slot[b];
    if (!S.in_flight) return KFS_OK;
    KFS_CUDA_TRY(cudaEventSynchronize(S.ev_done));
    S.in_flight = false;
    kfs_status rc = kfs_orb_finish_host_batch(r->orb[b], S.batch, S.io.n_kp);
    const int mb = r->p.max_batch, cap = r->cap;
    const int cp = S.io.capacity < cap ? S.io.capacity : cap;
    cudaStream_t s_out;
    kfs_orb_internal_streams(r->orb[b], nullptr, &s_out);
    bool late = false;
    for (int i = 0; i < S.batch; i++) {
        const int ng = S.h_small[i], nq = S.h_small[mb + i];
        if (S.io.n_good) S.io.n_good[i] = ng;
        if (S.io.n_query) S.io.n_query[i] = nq;
        // a keyframe that kept more rows than the eager copies carried (ties): the rest of its kNN rows / survivors now
        const int lq = nq < cp ? nq : cp, lg = ng < cp ? ng : cp;
        if (S.io.knn && lq > S.rows) {
            KFS_CUDA_TRY(cudaMemcpyAsync(S.io.knn + ((size_t)i * S.io.capacity + S.rows) * 2, S.d_knn + ((size_t)i * cap + S.rows) * 2,
                                         (size_t)(lq - S.rows) * 2 * sizeof(kfs_dmatch), cudaMemcpyDeviceToHost, s_out));
            late = true;
        }
        if (S.io.good && r->p.ratio > 0.f && lg > S.rows) {
            KFS_CUDA_TRY(cudaMemcpyAsync(S.io.good + (size_t)i * S.io.capacity + S.rows, S.d_good + (size_t)i * cap + S.rows,
                                         (size_t)(lg - S.rows) * sizeof(kfs_dmatch), cudaMemcpyDeviceToHost, s_out));
            late = true;
        }
    }
    if (late) KFS_CUDA_TRY(cudaStreamSynchronize(s_out));
    return rc;
}

extern "C" {

void kfs_replay_default_params(kfs_replay_params *p)
{
    if (!p) return;
    kfs_orb_default_params(&p->orb);
    p->max_batch = 64;
    p->keyframe_slots = 8;
    p->ratio = 0.6f; /* VisualOdometry.cpp:46 */
    p->chunk_frames = 0;
    p->arenas = 0;
}

void kfs_replay_destroy(kfs_replay *r)
{
    if (!r) return;
    KfsDeviceGuard guard(r->device);
    if (r->compute) cudaStreamSynchronize(r->compute);
    for (int b = 0; b < RP_MAX_ARENAS; b++) {
        ReplaySlot &S = r->slot[b];
        if (S.ev_done) { if (S.in_flight) cudaEventSynchronize(S.ev_done); cudaEventDestroy(S.ev_done); }
        if (S.ev_knn) cudaEventDestroy(S.ev_knn);
        if (S.d_knn) cudaFree(S.d_knn);
        if (S.d_good) cudaFree(S.d_good);
        if (S.d_small) cudaFree(S.d_small);
        if (S.h_small) cudaFreeHost(S.h_small);
        if (S.d_table) cudaFree(S.d_table);
        if (r->orb[b]) kfs_orb_destroy(r->orb[b]);
    }
    if (r->bf) kfs_bf_destroy(r->bf);
    if (r->d_kf_desc) cudaFree(r->d_kf_desc);
    if (r->d_kf_count) cudaFree(r->d_kf_count);
    if (r->compute) cudaStreamDestroy(r->compute);
    delete r;
}

kfs_status kfs_replay_create(const kfs_replay_params *params, int device, kfs_replay **out)
{
    if (!params || !out) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_replay_create: NULL argument");
    *out = nullptr;
    if (params->max_batch < 1 || params->max_batch > 65535 || params->keyframe_slots < 1 || params->chunk_frames < 0)
        KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_replay_create: 1 <= max_batch <= 65535 and keyframe_slots >= 1 required");
    if (params->arenas != 0 && (params->arenas < 2 || params->arenas > RP_MAX_ARENAS))
        KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_replay_create: arenas must be 0 (default) or 2..%d", RP_MAX_ARENAS);
    kfs_replay *r = new (std::nothrow) kfs_replay();
    if (!r) KFS_FAIL(KFS_ERR_INVALID_ARG, "out of host memory");
    r->p = *params;
    r->device = device;
    r->arenas = params->arenas ? params->arenas : RP_ARENAS_DEFAULT;
    kfs_status s = KFS_OK;
    for (int b = 0; b < r->arenas && !s; b++) s = kfs_orb_create(&params->orb, device, &r->orb[b]);
    if (!s) s = kfs_bf_create(KFS_NORM_HAMMING, 0, device, &r->bf);
    if (s) { kfs_replay_destroy(r); return s; }
    KfsDeviceGuard guard(device);
    if (!guard.ok) { kfs_replay_destroy(r); KFS_FAIL(KFS_ERR_CUDA, "cudaSetDevice(%d) failed", device); }
    r->cap = kfs_orb_capacity(r->orb[0]);
    const size_t mb = (size_t)params->max_batch, cap = (size_t)r->cap;
#define RP_TRY(expr)                                                                                       \
    do {                                                                                                   \
        cudaError_t e__ = (expr);                                                                          \
        if (e__ != cudaSuccess) {                                                                          \
            kfs_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__));          \
            kfs_replay_destroy(r);                                                                         \
            return KFS_ERR_CUDA;                                                                           \
        }                                                                                                  \
    } while (0)
    RP_TRY(cudaStreamCreateWithFlags(&r->compute, cudaStreamNonBlocking));
    for (int b = 0; b < r->arenas; b++) {
        ReplaySlot &S = r->slot[b];
        kfs_orb_set_stream(r->orb[b], (void *)r->compute);
        kfs_orb_internal_start_fence(r->orb[b], false); // arena reuse is ordered by replay_finish instead
        kfs_orb_set_pipeline_chunk(r->orb[b], params->chunk_frames > 0 ? params->chunk_frames : params->max_batch);
        RP_TRY(cudaEventCreateWithFlags(&S.ev_knn, cudaEventDisableTiming));
        RP_TRY(cudaEventCreateWithFlags(&S.ev_done, cudaEventDisableTiming));
        RP_TRY(cudaMalloc((void **)&S.d_knn, mb * cap * 2 * sizeof(kfs_dmatch)));
        RP_TRY(cudaMalloc((void **)&S.d_good, mb * cap * sizeof(kfs_dmatch)));
        RP_TRY(cudaMalloc((void **)&S.d_small, 2 * mb * sizeof(int)));
        RP_TRY(cudaMallocHost((void **)&S.h_small, 2 * mb * sizeof(int)));
        RP_TRY(cudaMalloc((void **)&S.d_table, mb * sizeof(BfPairDev)));
    }
    RP_TRY(cudaMalloc((void **)&r->d_kf_desc, (size_t)params->keyframe_slots * cap * 32));
    RP_TRY(cudaMalloc((void **)&r->d_kf_count, (size_t)params->keyframe_slots * sizeof(int)));
    RP_TRY(cudaMemset(r->d_kf_count, 0, (size_t)params->keyframe_slots * sizeof(int)));
#undef RP_TRY
    *out = r;
    return KFS_OK;
}

int kfs_replay_capacity(const kfs_replay *r) { return r ? r->cap : 0; }
int kfs_replay_arenas(const kfs_replay *r) { return r ? r->arenas : 0; }
int kfs_replay_last_launches(const kfs_replay *r) { return r ? r->launches : 0; }

kfs_status kfs_replay_submit(kfs_replay *r, const kfs_replay_io *io, const kfs_replay_frame *plan, int batch, int *ticket)
{
    if (!r || !io || !plan || !ticket) KFS_FAIL(KFS_ERR_INVALID_ARG, "kfs_replay_submit: NULL argument");
    if (batch < 1 || batch > r->p.max_batch) KFS_FAIL(KFS_ERR_INVALID_ARG, "batch %d outside 1..max_batch=%d", batch, r->p.max_batch);
    if (!io->n_kp) KFS_FAIL(KFS_ERR_INVALID_ARG, "io->n_kp is NULL");
    if (io->w <= 0 || io->h <= 0) KFS_FAIL(KFS_ERR_INVALID_ARG, "empty frames");
    for (int i = 0; i < batch; i++) {
        const int k = plan[i].keyframe;
        if (k != KFS_REPLAY_NO_MATCH && (k >= batch || (k < 0 && -(k + 1) >= r->p.keyframe_slots)))
            KFS_FAIL(KFS_ERR_INVALID_ARG, "plan[%d].keyframe = %d: not a frame of this batch nor a cache slot", i, k);
        if (plan[i].store_slot < -1 || plan[i].store_slot >= r->p.keyframe_slots)
            KFS_FAIL(KFS_ERR_INVALID_ARG, "plan[%d].store_slot = %d out of range", i, plan[i].store_slot);
        if (plan[i].flags & ~KFS_REPLAY_SKIP_ROWS) KFS_FAIL(KFS_ERR_INVALID_ARG, "plan[%d].flags = %d: unknown bits", i, plan[i].flags);
    }
    if (io->channels != 0 && io->channels != 1 && io->channels != 3)
        KFS_FAIL(KFS_ERR_UNSUPPORTED, "io->channels = %d (the reference handles 1 or 3, VisualOdometry.cpp:27)", io->channels);
    KFS_DEVICE_GUARD(r->device);
    const int b = r->seq % r->arenas;
    ReplaySlot &S = r->slot[b];
    kfs_status s = replay_finish(r, b); // the arena is reused: its previous batch must be complete (results already delivered)
    if (s && s != KFS_ERR_CAPACITY) return s;
    kfs_orb *o = r->orb[b];
    // 1. upload + extraction (+ D2H of keypoints / descriptors / counts), nothing synchronises
    bool any_skip = false;
    S.skip.assign((size_t)batch, 0);
    for (int i = 0; i < batch; i++) if (plan[i].flags & KFS_REPLAY_SKIP_ROWS) { S.skip[(size_t)i] = 1; any_skip = true; }
    KfsHostBatch hb = {};
    hb.image = io->gray; hb.step = io->step; hb.frame_stride = io->frame_stride; hb.w = io->w; hb.h = io->h;
    hb.channels = io->channels == 3 ? 3 : 1; hb.batch = batch;
    hb.mask = io->mask; hb.mask_step = io->mask_step; hb.mask_frame_stride = io->mask_frame_stride;
    hb.kp_out = io->kp; hb.desc_out = io->desc; hb.capacity = io->capacity;
    hb.out_rows = r->p.orb.nfeatures;
    hb.skip_rows = any_skip ? S.skip.data() : nullptr;
    s = kfs_orb_enqueue_host_batch(o, hb);
    if (s) return s;
    r->launches = kfs_orb_last_launches(o);
    const kfs_keypoint *d_kp; const uint8_t *d_desc; const int *d_cnt; int cap = 0, nb = 0;
    s = kfs_orb_device_results(o, &d_kp, &d_desc, &d_cnt, &cap, &nb);
    if (s) return s;
    cudaStream_t s_out;
    kfs_orb_internal_streams(o, nullptr, &s_out);
    const bool want_knn = io->knn != nullptr || io->good != nullptr;
    if (want_knn) {
        // 2. pair table: query = keyframe descriptors (same batch or cache), train = the frame's own; rebuilt (by a kernel, from
        //    the plan in its arguments) only when the plan or the arena's buffers changed
        std::vector<int> &key = S.last_plan;
        bool same = (int)key.size() == batch + 1 && key[(size_t)batch] == cap && S.last_desc == d_desc;
        for (int i = 0; same && i < batch; i++) same = key[(size_t)i] == plan[i].keyframe;
        if (!same) {
            for (int i0 = 0; i0 < batch; i0 += RP_PLAN_CHUNK) {
                RpPlanArg pa;
                const int n = batch - i0 < RP_PLAN_CHUNK ? batch - i0 : RP_PLAN_CHUNK;
                for (int j = 0; j < n; j++) pa.keyframe[j] = plan[i0 + j].keyframe;
                for (int j = n; j < RP_PLAN_CHUNK; j++) pa.keyframe[j] = KFS_REPLAY_NO_MATCH;
                k_replay_table<<<kfs_div_up(n, 128), 128, 0, r->compute>>>(pa, i0, n, S.d_table, d_desc, d_cnt, cap, r->d_kf_desc, r->d_kf_count, S.d_knn);
                if (cudaGetLastError() != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_replay_table launch failed");
                r->launches++;
            }
            key.resize((size_t)batch + 1);
            for (int i = 0; i < batch; i++) key[(size_t)i] = plan[i].keyframe;
            key[(size_t)batch] = cap;
            S.last_desc = d_desc;
        }
        // 3. matching (+ ratio test) on the device-resident descriptors
        s = kfs_bf_launch_table(r->bf, S.d_table, batch, cap, 2, r->compute);
        if (s) return s;
        r->launches++;
        const int mb = r->p.max_batch;
        if (io->good && r->p.ratio > 0.f) {
            s = kfs_bf_launch_ratio_table(r->bf, S.d_table, batch, r->p.ratio, S.d_good, (long long)cap, S.d_small, r->compute);
            if (s) return s;
            r->launches++;
        } else {
            KFS_CUDA_TRY(cudaMemsetAsync(S.d_small, 0, sizeof(int) * (size_t)mb, r->compute));
        }
        k_replay_nquery<<<kfs_div_up(batch, 128), 128, 0, r->compute>>>(S.d_table, batch, S.d_small + mb);
        if (cudaGetLastError() != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_replay_nquery launch failed");
        r->launches++;
    }
    // 4. frames that become keyframes for later batches: descriptors + count into the cache (after this batch's
    //    matching, so a cache reference in this batch still sees the previous keyframe)
    for (int i = 0; i < batch; i++) {
        const int sl = plan[i].store_slot;
        if (sl < 0) continue;
        k_replay_store<<<kfs_div_up(cap * 2, 256), 256, 0, r->compute>>>(reinterpret_cast<const uint4 *>(d_desc + (size_t)i * cap * 32), d_cnt + i,
                                                                         reinterpret_cast<uint4 *>(r->d_kf_desc + (size_t)sl * cap * 32), r->d_kf_count + sl, cap * 2);
        if (cudaGetLastError() != cudaSuccess) KFS_FAIL(KFS_ERR_CUDA, "k_replay_store launch failed");
        r->launches++;
    }
    KFS_CUDA_TRY(cudaEventRecord(S.ev_knn, r->compute));
    KFS_CUDA_TRY(cudaStreamWaitEvent(s_out, S.ev_knn, 0));
    S.rows = 0;
    if (want_knn) {
        const int cp = io->capacity < cap ? io->capacity : cap;
        const int rows = r->p.orb.nfeatures < cp ? r->p.orb.nfeatures : cp; // what a keyframe without ties can fill; the rest: replay_finish
        S.rows = rows;
        if (io->knn && rows > 0)
            KFS_CUDA_TRY(cudaMemcpy2DAsync(io->knn, (size_t)io->capacity * 2 * sizeof(kfs_dmatch), S.d_knn, (size_t)cap * 2 * sizeof(kfs_dmatch),
                                           (size_t)rows * 2 * sizeof(kfs_dmatch), (size_t)batch, cudaMemcpyDeviceToHost, s_out));
        if (io->good && rows > 0 && r->p.ratio > 0.f)
            KFS_CUDA_TRY(cudaMemcpy2DAsync(io->good, (size_t)io->capacity * sizeof(kfs_dmatch), S.d_good, (size_t)cap * sizeof(kfs_dmatch),
                                           (size_t)rows * sizeof(kfs_dmatch), (size_t)batch, cudaMemcpyDeviceToHost, s_out));
        KFS_CUDA_TRY(cudaMemcpyAsync(S.h_small, S.d_small, 2 * sizeof(int) * (size_t)r->p.max_batch, cudaMemcpyDeviceToHost, s_out));
    } else {
        memset(S.h_small, 0, 2 * sizeof(int) * (size_t)r->p.max_batch);
    }
    KFS_CUDA_TRY(cudaEventRecord(S.ev_done, s_out));
    S.in_flight = true;
    S.batch = batch;
    S.io = *io;
    S.ticket = r->seq;
    *ticket = r->seq;
    r->seq++;
    return KFS_OK;
}

kfs_status kfs_replay_wait(kfs_replay *r, int ticket)
{
    if (!r) KFS_FAIL(KFS_ERR_INVALID_ARG, "replay handle is NULL");
    KFS_DEVICE_GUARD(r->device);
    for (int b = 0; b < r->arenas; b++)
        if (r->slot[b].in_flight && r->slot[b].ticket == ticket) return replay_finish(r, b);
    if (ticket >= 0 && ticket < r->seq) return KFS_OK; // already delivered (its arena was reused, or waited before)
    KFS_FAIL(KFS_ERR_INVALID_ARG, "unknown ticket %d", ticket);
}

} // extern "C"

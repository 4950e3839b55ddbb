/*
 * kfslam_frontend.h — C ABI of the B200-native SLAM front-end (ORB detect+describe, Hamming kNN).
 *
 * Drop-in boundary for the two OpenCV calls the reference tracker makes per frame
 * (angelo-nutu/Keyframe-Based-SLAM):
 *   - cv::ORB::create(...)                       /root/reference/src/lib/VisualOdometry.cpp:10-16
 *   - ptrExtractor->detectAndCompute(gray, mask, kp, desc)                        ... :30
 *   - cv::BFMatcher::create(cv::NORM_HAMMING)                                     ... :18
 *   - ptrMatcher->knnMatch(keyframe->matDescriptors, dpCurrImg, knnMatches, 2)    ... :39
 *   - the ratio test  d0 < 0.6f * d1                                              ... :46
 * Plain pointers and sizes only; no C++/torch types cross this boundary.  The C++ adapter that gives
 * the tracker the cv::-shaped methods back is include/kfslam_frontend_cv.hpp; INTEGRATION.md shows the
 * three-line change in VisualOdometry.{hpp,cpp}.
 *
 * Threading contract (mirrors the reference: VisualOdometry.cpp:138-144, single caller thread): a handle
 * is NOT thread-safe; distinct handles are independent (one per host thread / per GPU).  Host-buffer
 * entry points block until results are in host memory.  *_device entry points are asynchronous on the
 * handle's stream.
 *
 * There is no CPU fallback: every entry point that computes needs a CUDA device (sm_100a build) and
 * returns KFS_ERR_NO_DEVICE / KFS_ERR_CUDA otherwise.
 *
 * Environment switches (read when a handle is created or first sized; results are identical for every
 * setting, they select mechanisms for A/B measurements — DESIGN.md section 4): KFS_ORB_GRAPHS,
 * KFS_ORB_TMA, KFS_ORB_FUSED_PYRAMID, KFS_ORB_FAST_HALF, KFS_ORB_PDL, KFS_ORB_DESCRIBE2, KFS_ORB_RESIZE_TMA (0 = off),
 * KFS_BF_SPLIT (1..16),
 * KFS_BF_CLUSTER (0 / 1 / 2).
 */
#ifndef KFSLAM_FRONTEND_H
#define KFSLAM_FRONTEND_H

#include <stddef.h>
#include <stdint.h>
#include <limits.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define KFS_API
#else
#define KFS_API __attribute__((visibility("default")))
#endif

typedef enum kfs_status {
    KFS_OK = 0,
    KFS_ERR_INVALID_ARG = 1, /* null pointer, negative size, ... (OpenCV would throw cv::Exception)      */
    KFS_ERR_UNSUPPORTED = 2, /* a cv::ORB / cv::BFMatcher option outside the reference's usage           */
    KFS_ERR_CAPACITY = 3,    /* output capacity too small; n_out holds the required count (data truncated) */
    KFS_ERR_CUDA = 4,        /* CUDA runtime error; see kfs_last_error()                                  */
    KFS_ERR_NO_DEVICE = 5    /* no CUDA device visible: there is no CPU path                              */
} kfs_status;

/* cv::ORB::ScoreType */
enum { KFS_ORB_HARRIS_SCORE = 0, KFS_ORB_FAST_SCORE = 1 };
/* cv::NormTypes value used by the reference (cv::NORM_HAMMING == 6) */
enum { KFS_NORM_HAMMING = 6 };

/* Arguments of cv::ORB::create, same order and meaning (VisualOdometry.cpp:10-16 passes
 * 3000, 1.2f, 8, 31, 0, 2, HARRIS_SCORE, 31, 20).  Supported: any nfeatures >= 1, scale_factor > 1,
 * 1 <= nlevels <= 16, fast_threshold 1..254; edge_threshold = 31, first_level = 0, wta_k = 2,
 * score_type = HARRIS, patch_size = 31 are the only accepted values (others -> KFS_ERR_UNSUPPORTED). */
typedef struct kfs_orb_params {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;
    int32_t edge_threshold;
    int32_t first_level;
    int32_t wta_k;
    int32_t score_type;
    int32_t patch_size;
    int32_t fast_threshold;
} kfs_orb_params;

/* Layout-identical to cv::KeyPoint (28 bytes). */
typedef struct kfs_keypoint {
    float x, y;      /* pt, level-0 pixel coordinates                 */
    float size;      /* patch_size * scale(octave)                    */
    float angle;     /* degrees, [0, 360)                             */
    float response;  /* Harris response                               */
    int32_t octave;  /* pyramid level                                 */
    int32_t class_id; /* -1                                           */
} kfs_keypoint;

/* Layout-identical to cv::DMatch (16 bytes). */
typedef struct kfs_dmatch {
    int32_t queryIdx, trainIdx, imgIdx;
    float distance;
} kfs_dmatch;

typedef struct kfs_orb kfs_orb; /* replaces cv::Ptr<cv::ORB>       (VisualOdometry.hpp:28) */
typedef struct kfs_bf kfs_bf;   /* replaces cv::Ptr<cv::BFMatcher> (VisualOdometry.hpp:29) */

/* ---- library ------------------------------------------------------------------------------------- */
KFS_API int kfs_version(void);                 /* major*10000 + minor*100 + patch */
KFS_API const char *kfs_last_error(void);      /* thread-local, never NULL         */
KFS_API int kfs_device_count(void);            /* 0 if no CUDA device              */

/* ---- ORB: replaces cv::ORB::create + detectAndCompute (VisualOdometry.cpp:10-16, :30) -------------- */
KFS_API void kfs_orb_default_params(kfs_orb_params *p); /* OpenCV defaults: 500,1.2,8,31,0,2,HARRIS,31,20 */
KFS_API kfs_status kfs_orb_create(const kfs_orb_params *params, int device, kfs_orb **out);
KFS_API void kfs_orb_destroy(kfs_orb *orb);

/* Upper bound on keypoints per frame the handle stores (nfeatures + tie slack); a caller capacity of
 * this many entries never truncates unless ties exceed the slack.  kfs_orb_set_capacity changes it
 * (must be called before the first extraction or it re-allocates). */
KFS_API int kfs_orb_capacity(const kfs_orb *orb);
KFS_API kfs_status kfs_orb_set_capacity(kfs_orb *orb, int capacity);

/* One frame, host buffers in and out; blocks.  gray: h rows of w bytes, `step` bytes apart.
 * mask: NULL (== cv::noArray()) or h x w bytes, non-zero = keep.  Keypoints come out grouped by
 * ascending octave and, inside an octave, sorted by (y, x) — deterministic run to run.
 * kp_out[capacity], desc_out[capacity*32].  w == 0 or h == 0 -> KFS_OK with *n_out = 0. */
KFS_API kfs_status kfs_orb_detect_and_compute(kfs_orb *orb, const uint8_t *gray, size_t step, int w, int h,
                                              const uint8_t *mask, size_t mask_step,
                                              kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out);

/* `batch` frames of one shape, host buffers; frame f starts at gray + f*frame_stride (mask likewise,
 * mask may be NULL).  kp_out[batch][capacity], desc_out[batch][capacity][32], n_out[batch].  Blocks. */
KFS_API kfs_status kfs_orb_detect_and_compute_batch(kfs_orb *orb, const uint8_t *gray, size_t step, size_t frame_stride,
                                                    int w, int h, int batch,
                                                    const uint8_t *mask, size_t mask_step, size_t mask_frame_stride,
                                                    kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out);

/* Device-resident variant: d_gray / d_mask are device pointers on the handle's device; asynchronous on
 * the handle's stream; results stay on the device (kfs_orb_device_results) until fetched. */
KFS_API kfs_status kfs_orb_extract_device(kfs_orb *orb, const uint8_t *d_gray, size_t step, size_t frame_stride,
                                          int w, int h, int batch,
                                          const uint8_t *d_mask, size_t mask_step, size_t mask_frame_stride);
/* Device views of the last extraction: d_kp[batch][capacity], d_desc[batch][capacity][32], d_counts[batch]
 * (d_counts[f] may exceed capacity when ties overflowed; only min(count, capacity) entries are valid). */
KFS_API kfs_status kfs_orb_device_results(kfs_orb *orb, const kfs_keypoint **d_kp, const uint8_t **d_desc,
                                          const int **d_counts, int *capacity, int *batch);
/* Waits for the stream and copies the last extraction's results to host buffers (layout as in _batch). */
KFS_API kfs_status kfs_orb_fetch_results(kfs_orb *orb, kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out);
KFS_API kfs_status kfs_orb_sync(kfs_orb *orb);

/* Stream plumbing: the handle creates its own non-blocking stream; a caller that already has one
 * (e.g. torch.cuda.current_stream().cuda_stream) can install it.  Passing NULL restores the own stream. */
KFS_API kfs_status kfs_orb_set_stream(kfs_orb *orb, void *cuda_stream);
KFS_API void *kfs_orb_get_stream(kfs_orb *orb);
/* kfs_orb_detect_and_compute_batch cuts the batch into chunks of frames that flow through three streams (H2D of
 * chunk i+1 | kernels of chunk i | D2H of chunk i-1).  frames = chunk size; 0 = automatic (batch/4, whole batch when
 * batch < 8).  Use pinned host buffers (cudaHostAlloc / cudaHostRegister) for the copies to overlap. */
KFS_API kfs_status kfs_orb_set_pipeline_chunk(kfs_orb *orb, int frames);

/* Stage taps (parity tests and profiling; not needed by the tracker). */
KFS_API kfs_status kfs_orb_level_info(kfs_orb *orb, int level, int *w, int *h, int *quota, float *scale);
KFS_API kfs_status kfs_orb_download_level(kfs_orb *orb, int frame, int level, uint8_t *out, size_t out_step);
/* FAST corners after NMS + border + mask filter of (frame, level): packed score<<24 | y<<12 | x, unordered. */
/* a1.9: ORB blurs every level with OpenCV's float sepFilter2D path, whose last bit follows the CPU
 * dispatch: fused multiply-adds with the default AVX2/AVX-512 build (fma = 1, the default here), separate multiply and add with
 * cv::setUseOptimized(false) or on a host without FMA (fma = 0; 6 of 921600 pixels of a 720p frame differ by 1).  Pick the one
 * the CPU path you compare against uses. */
KFS_API kfs_status kfs_orb_set_blur_fma(kfs_orb *orb, int fma);
/* Parity / debugging: the blurred copy of a level (rows and columns closer than 12 px to the border are never sampled by a
 * descriptor and are not REFLECT_101-exact). */
KFS_API kfs_status kfs_orb_download_blurred_level(kfs_orb *orb, int frame, int level, uint8_t *out, size_t out_step);
KFS_API kfs_status kfs_orb_download_fast_corners(kfs_orb *orb, int frame, int level, uint32_t *out, int capacity, int *n_out);
/* Per-stage device time of the last extraction when profiling is on (cudaEvent pairs on the handle's
 * stream). Stage ids: KFS_STAGE_*.  ms[KFS_STAGE_COUNT]. */
enum { KFS_STAGE_INGEST = 0, KFS_STAGE_PYRAMID = 1, KFS_STAGE_FAST = 2, KFS_STAGE_SELECT_HARRIS = 3,
       KFS_STAGE_SELECT_SORT = 4, KFS_STAGE_BLUR = 5, KFS_STAGE_DESCRIBE = 6, KFS_STAGE_COUNT = 7 };
KFS_API kfs_status kfs_orb_set_profiling(kfs_orb *orb, int on);
KFS_API kfs_status kfs_orb_stage_times(kfs_orb *orb, float *ms);
/* Number of kernel launches the last extraction issued (bench.py's gpu_launches). */
KFS_API int kfs_orb_last_launches(const kfs_orb *orb);

/* ---- BFMatcher: replaces cv::BFMatcher::create + knnMatch (VisualOdometry.cpp:18, :39) -------------- */
KFS_API kfs_status kfs_bf_create(int norm_type, int cross_check, int device, kfs_bf **out);
KFS_API void kfs_bf_destroy(kfs_bf *bf);

/* knnMatch(query, train, matches, k): query nq x 32 bytes, train nt x 32 bytes, contiguous rows, host.
 * out[nq*k]: row q holds counts[q] = min(k, nt) entries sorted by ascending distance, ties by ascending
 * trainIdx (OpenCV's insertion order); unused tail entries of a row are {q, -1, -1, FLT_MAX}.
 * counts may be NULL.  k must be 1 or 2; nt < 2^18 (OpenCV's own limit).  Blocks. */
KFS_API kfs_status kfs_bf_knn_match(kfs_bf *bf, const uint8_t *query, int nq, const uint8_t *train, int nt, int k,
                                    kfs_dmatch *out, int *counts);

/* One independent (query, train) problem.  For device pairs every pointer is a device pointer.
 * d_nq / d_nt, when non-NULL, point to device ints holding the live row counts (clamped to nq / nt, which
 * then act as capacities) — lets descriptors produced by kfs_orb_extract_device be matched with no host
 * round trip. */
typedef struct kfs_bf_pair {
    const uint8_t *query; int32_t nq; const int32_t *d_nq;
    const uint8_t *train; int32_t nt; const int32_t *d_nt;
    kfs_dmatch *out;      /* nq*k entries */
    int32_t *counts;      /* nq entries or NULL */
} kfs_bf_pair;

/* npairs independent problems in one launch (window matching, offline replay).  Host buffers; blocks.
 * d_nq/d_nt must be NULL here. */
KFS_API kfs_status kfs_bf_knn_match_pairs(kfs_bf *bf, const kfs_bf_pair *pairs, int npairs, int k);
/* Same, device buffers, asynchronous on the handle's stream. */
KFS_API kfs_status kfs_bf_knn_match_pairs_device(kfs_bf *bf, const kfs_bf_pair *pairs, int npairs, int k);

/* Fused ratio test (VisualOdometry.cpp:46): compacts rows with 2 entries and d0 < ratio*d1, in query
 * order, into good[0..*n_good).  Host buffers; blocks.  The tracker may keep doing this itself. */
KFS_API kfs_status kfs_bf_knn_match_ratio(kfs_bf *bf, const uint8_t *query, int nq, const uint8_t *train, int nt,
                                          float ratio, kfs_dmatch *good, int *n_good);

/* Descriptor rows that kfs_orb_detect_and_compute[_batch] of `orb` just wrote to host memory are still on the device.  A matcher
 * attached to that extractor matches such rows where they are instead of uploading them again: a query / train pointer of
 * kfs_bf_knn_match[_pairs] is recognised when it is row 0 of a frame of that extractor's LAST host-buffer call's desc_out and
 * its row count EQUALS that frame's keypoint count (the tracker's knnMatch(keyframe, dpCurrImg): the train side,
 * VisualOdometry.cpp:39) and its first and last row still hold the delivered bytes (checked: a buffer reallocated at the same
 * address or filtered in place is uploaded).  Rows in between are the caller's promise; any other pointer is uploaded
 * as before.  orb == NULL detaches; the extractor must outlive the attachment and live on the same device.
 * kfs_bf_last_aliased: descriptor matrices of the last host-buffer call that were matched in place. */
KFS_API kfs_status kfs_bf_attach_orb(kfs_bf *bf, kfs_orb *orb);
KFS_API int kfs_bf_last_aliased(const kfs_bf *bf);
KFS_API kfs_status kfs_bf_set_stream(kfs_bf *bf, void *cuda_stream);
KFS_API void *kfs_bf_get_stream(kfs_bf *bf);
KFS_API kfs_status kfs_bf_sync(kfs_bf *bf);
KFS_API int kfs_bf_last_launches(const kfs_bf *bf);

/* Matching for callers that must not allocate.  (No counterpart in the reference: cv::BFMatcher allocates freely; this entry
 * point exists for the hosts described below.)
 *
 * Every other matcher call goes through a kfs_bf handle, and a handle grows device buffers with cudaMalloc whenever a call is
 * larger than the last one (pair table, staging, the scratch of split launches).  Some hosts cannot live with that:
 *   - a tracker thread with a per-frame deadline that shares the GPU with a mapping / bundle-adjustment thread: cudaMalloc and
 *     cudaFree serialise against all work on the device, so one growing buffer stalls both threads in the middle of a frame;
 *   - a process whose HBM was handed, whole, to another allocator at start-up (a framework's caching pool, a fixed robotics
 *     memory budget): the driver has nothing left to give, cudaMalloc fails, and the descriptors the caller wants matched
 *     already sit in memory from that pool;
 *   - code that is handed a stream and device pointers by its own caller and may not create CUDA objects of its own.
 * For them this call takes no handle and creates nothing — no cudaMalloc / cudaFree / cudaMallocHost, no stream, no event.  The
 * caller passes one (query, train) problem in device memory and the result rows in device memory; the scratch of the split
 * launch (partial top-2 lists, arrival counters, the kernel's pair table) is a fixed array compiled into the library,
 * about 2 MiB, which every GPU that loads the library holds until the process exits.
 *
 * Limits and rules:
 *   - nq <= KFS_BF_STATIC_MAX_QUERIES (the array is sized for that many query rows; more -> KFS_ERR_INVALID_ARG with a
 *     message naming the limit — split the query rows over several calls, each is independent); nt < 2^18; k = 1 or 2;
 *     all pointers 16-byte aligned device pointers; d_counts may be NULL.
 *   - Results are those of kfs_bf_knn_match_pairs_device for the same rows, bit for bit (same kernel, same tie rule).
 *   - Asynchronous on `cuda_stream` (NULL = the default stream): synchronise that stream before reading d_out.
 *   - There is ONE array per GPU, shared by every caller in the process, and it keeps what the previous call left in it (the
 *     call clears what has to start at zero itself).  Two calls whose kernels can overlap on the GPU — different streams
 *     without an ordering between them, or two host threads — race on it and produce garbage.  Keep all calls for one
 *     device on one stream, or order them with events.
 *   - The first call for a device looks up the array's device address (and the device's SM count) and remembers them; make
 *     that first call during start-up, outside any stream capture.  Later calls only launch two kernels. */
#define KFS_BF_STATIC_MAX_QUERIES 16384
KFS_API kfs_status kfs_bf_knn_match_device_static(int device, void *cuda_stream, const uint8_t *d_query, int nq,
                                                  const uint8_t *d_train, int nt, int k, kfs_dmatch *d_out, int *d_counts);

/* ---- Replay executor: the tracker's per-frame loop on batches of frames ------------------------------
 * Replaces the body of VisualOdometry::Track (VisualOdometry.cpp:129-172) up to the PnP call: ExtractFeatures
 * (:138 -> detectAndCompute :30), MatchFeatures against the last keyframe (:144 -> knnMatch :39, query = keyframe
 * descriptors, train = current frame) and the ratio test (:46).  Frames come from host memory; descriptors stay on
 * the device between extraction and matching, keyframes of earlier batches live in a device-resident cache
 * (KeyFrame::matDescriptors, KeyFrame.hpp:39), and three arenas (kfs_replay_arenas) rotate over the batches: kfs_replay_submit only
 * enqueues (H2D, kernels, D2H are stream-ordered), kfs_replay_wait blocks until that batch's results are in the
 * host buffers.  At most kfs_replay_arenas() = 3 batches are in flight (upload | kernels | download); submitting one more first
 * completes the oldest.
 * Use pinned host buffers (cudaHostAlloc / cudaHostRegister) for the copies to overlap with the kernels. */
typedef struct kfs_replay kfs_replay;
typedef struct kfs_replay_params {
    kfs_orb_params orb;      /* extractor parameters (VisualOdometry.cpp:10-16) */
    int32_t max_batch;       /* frames per submit */
    int32_t keyframe_slots;  /* entries of the device-resident keyframe descriptor cache (>= 1) */
    float ratio;             /* ratio-test threshold (reference: 0.6f, :46); <= 0 disables the fused `good` output */
    int32_t chunk_frames;    /* frames per H2D/kernel/D2H chunk inside one batch; 0 = the whole batch (batches already
                              * overlap each other; smaller chunks only cut the latency of a lone batch) */
    int32_t arenas;          /* batches that can be in flight: 2..KFS_REPLAY_MAX_ARENAS, 0 = the default (3) */
} kfs_replay_params;
#define KFS_REPLAY_MAX_ARENAS 6
#define KFS_REPLAY_NO_MATCH INT32_MIN
#define KFS_REPLAY_SKIP_ROWS 1   /* kfs_replay_frame.flags: this frame's keypoints / descriptors stay on the device */
typedef struct kfs_replay_frame {
    int32_t keyframe;   /* >= 0: index of the frame of THIS batch that is this frame's keyframe; < 0: -(1 + slot) of the
                         * keyframe cache (content stored by an EARLIER submit); KFS_REPLAY_NO_MATCH: extract only */
    int32_t store_slot; /* >= 0: after this batch's matching, keep this frame's descriptors in that cache slot; -1: no */
    int32_t flags;      /* 0, or KFS_REPLAY_SKIP_ROWS: io.kp / io.desc rows of this frame are not written (n_kp still is) — the
                         * tracker keeps keypoints and descriptors only of frames that become keyframes (VisualOdometry.cpp:155-165) */
} kfs_replay_frame;
typedef struct kfs_replay_io { /* host buffers; must stay valid until kfs_replay_wait returns */
    const uint8_t *gray; size_t step, frame_stride; int32_t w, h;   /* batch frames: 8-bit gray, or R,G,B bytes when channels == 3 */
    kfs_keypoint *kp; uint8_t *desc; int32_t capacity; int32_t *n_kp; /* [batch][capacity] (x32 B for desc), [batch] */
    kfs_dmatch *knn;    /* [batch][capacity][2]: row q = the keyframe's descriptor q; NULL: not wanted */
    int32_t *n_query;   /* [batch]: valid rows of knn = the keyframe's keypoint count; may be NULL */
    kfs_dmatch *good;   /* [batch][capacity]: ratio-test survivors in query order; NULL: not wanted */
    int32_t *n_good;    /* [batch]; may be NULL */
    const uint8_t *mask; size_t mask_step, mask_frame_stride; /* NULL (cv::noArray()) or one h x w mask per frame, non-zero = keep
                         * (the reference passes one on every call: VisualOdometry.cpp:30, main.cpp:31-34) */
    int32_t channels;   /* 0 or 1: gray; 3: R,G,B as the reference receives it (:25-28), step = bytes per row */
} kfs_replay_io;
/* Only rows that can be live travel device -> host: min(capacity, nfeatures) rows per frame right behind the kernels (a frame
 * keeps more than nfeatures keypoints only after ties on a selection threshold; those extra rows — and the kNN rows of such a
 * keyframe — are fetched by kfs_replay_wait when a count says so).  Rows past the returned counts are scratch. */
KFS_API void kfs_replay_default_params(kfs_replay_params *p); /* ORB defaults, max_batch 64, 8 slots, ratio 0.6, 3 arenas */
KFS_API kfs_status kfs_replay_create(const kfs_replay_params *params, int device, kfs_replay **out);
KFS_API void kfs_replay_destroy(kfs_replay *r);
KFS_API int kfs_replay_capacity(const kfs_replay *r);           /* per-frame capacity the outputs should have */
KFS_API int kfs_replay_arenas(const kfs_replay *r);             /* batches that can be in flight (submitted, not yet waited for) */
KFS_API kfs_status kfs_replay_submit(kfs_replay *r, const kfs_replay_io *io, const kfs_replay_frame *plan, int batch, int *ticket);
KFS_API kfs_status kfs_replay_wait(kfs_replay *r, int ticket);
KFS_API int kfs_replay_last_launches(const kfs_replay *r);      /* kernels launched by the last submit */

/* ---- tracker front-end: VisualOdometry::ExtractFeatures + MatchFeatures as the reference calls them --------
 * (the callers either side of the two drop-in calls.)
 *   ExtractFeatures  /root/reference/src/lib/VisualOdometry.cpp:21-33   cvtColor(RGB2BGR), cvtColor(BGR2GRAY) when the
 *                    image has 3 channels, then detectAndCompute(gray, mask)
 *   MatchFeatures    /root/reference/src/lib/VisualOdometry.cpp:35-62   knnMatch(keyframe, current, 2), ratio test
 *                    d0 < 0.6f*d1 (:46), depth look-up and pin-hole back-projection of the KEYFRAME keypoint (:52-54),
 *                    2-D point of the current keypoint (:57)
 *   KeyFrame         /root/reference/include/lib/KeyFrame.hpp:35-39     descriptors, keypoints and the uint16 depth
 *                    image of the last keyframe: cached on the device, uploaded once per keyframe
 * One handle owns an ORB extractor, a matcher, the colour/gray staging and the keyframe cache; one stream. */
typedef struct kfs_frontend kfs_frontend;
KFS_API kfs_status kfs_frontend_create(const kfs_orb_params *params, int device, kfs_frontend **out);
KFS_API void kfs_frontend_destroy(kfs_frontend *fe);
KFS_API int kfs_frontend_capacity(const kfs_frontend *fe);   /* rows kp_out / desc_out / the match outputs should have */
/* Changes that capacity (default nfeatures + nfeatures/4 + 64: ties on the selection thresholds may keep more than nfeatures
 * keypoints).  Drops the cached keyframes and the current frame.  A frame that overflows makes
 * extract_features / track return KFS_ERR_CAPACITY with *n_out = the rows it needs (cv::ORB would have returned them all). */
KFS_API kfs_status kfs_frontend_set_capacity(kfs_frontend *fe, int capacity);
/* kfs_orb_set_blur_fma for the tracker handle's extractor: 1 = OpenCV's default AVX2 / AVX-512 blur (fused
 * multiply-adds, the default), 0 = the multiply-then-add build.  A tracker built against an OpenCV without FMA dispatch sets 0 to
 * keep bit-identical descriptors.  Cached keyframes stay valid (they were described with the variant of their time). */
KFS_API kfs_status kfs_frontend_set_blur_fma(kfs_frontend *fe, int fma);

/* ExtractFeatures(rgb, mask): image = h rows of w pixels x `channels` bytes (1 = already gray, 3 = R,G,B byte order as the
 * reference receives it), rows `step` bytes apart, host memory.  Gray = (R*9798 + G*19235 + B*3735 + 16384) >> 15, which
 * is what OpenCV's RGB2BGR + BGR2GRAY pair computes, evaluated on the device.  Outputs as kfs_orb_detect_and_compute;
 * the descriptors and keypoints also stay on the device as "the current frame" for the calls below. */
KFS_API kfs_status kfs_frontend_extract_features(kfs_frontend *fe, const uint8_t *image, size_t step, int w, int h, int channels,
                                                 const uint8_t *mask, size_t mask_step,
                                                 kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out);
/* The current frame becomes the keyframe (device-to-device copy of its descriptors and keypoints) and `depth`
 * (h x w uint16 millimetres, rows depth_step BYTES apart, host) is uploaded next to them.  Blocks until the depth image has
 * left the caller's buffer. */
KFS_API kfs_status kfs_frontend_promote_keyframe(kfs_frontend *fe, const uint16_t *depth, size_t depth_step);
/* Same, from host copies of a keyframe (n rows of descriptors and keypoints, a w x h depth image). */
KFS_API kfs_status kfs_frontend_set_keyframe(kfs_frontend *fe, const uint8_t *desc, const kfs_keypoint *kp, int n,
                                             const uint16_t *depth, size_t depth_step, int w, int h);
/* MatchFeatures: query = the cached keyframe, train = the current frame (cur_desc / cur_kp host rows, or both NULL to use
 * the device-resident result of the last kfs_frontend_extract_features).  intrinsics = {fx, fy, cx, cy} (K(0,0), K(1,1),
 * K(0,2), K(1,2)).  Survivors of d0 < ratio*d1 come out in query order: matches_out[i] (the best cv::DMatch),
 * points3d_out[3*i..] = (x, y, z) and points2d_out[2*i..] = current keypoint pt, all as the reference's doubles
 * (computed through the same float intermediates).  *n_out = survivors (may exceed capacity: then only `capacity` rows
 * were written and the status is KFS_ERR_CAPACITY).  Rows past *n_out (up to min(capacity, keyframe rows)) are scratch:
 * count and lists come back in one device-to-host round trip. */
KFS_API kfs_status kfs_frontend_match_features(kfs_frontend *fe, const uint8_t *cur_desc, const kfs_keypoint *cur_kp, int n_cur,
                                               const double *intrinsics, float ratio,
                                               kfs_dmatch *matches_out, double *points3d_out, double *points2d_out,
                                               int capacity, int *n_out);
/* The front half of VisualOdometry::Track (/root/reference/src/lib/VisualOdometry.cpp:129-172) for one frame in ONE call with
 * ONE host synchronisation: ExtractFeatures (:138) and MatchFeatures against the cached keyframe (:144) are enqueued back to
 * back (upload, colour conversion, ORB chain, kNN, ratio test + back-projection, every result copy).  Arguments as in
 * kfs_frontend_extract_features and kfs_frontend_match_features (the current frame is the one just extracted).
 * *n_matches_out = -1 when no keyframe is cached yet (only the extraction ran: the tracker then promotes this frame,
 * VisualOdometry.cpp:105-123).  Rows past the returned counts are scratch. */
KFS_API kfs_status kfs_frontend_track(kfs_frontend *fe, const uint8_t *image, size_t step, int w, int h, int channels,
                                      const uint8_t *mask, size_t mask_step, const double *intrinsics, float ratio,
                                      kfs_keypoint *kp_out, uint8_t *desc_out, int capacity, int *n_out,
                                      kfs_dmatch *matches_out, double *points3d_out, double *points2d_out, int match_capacity,
                                      int *n_matches_out);

/* ---- keyframe window --------------------------------------------------------
 *   Map::GetNKeyFrames(win)   /root/reference/src/lib/Map.cpp:72-78    the last `win` keyframes, oldest first, or all of
 *                             them while fewer exist
 *   LocalBundleAdjustment     /root/reference/include/lib/Optimizers.hpp:70 (window 5), src/lib/Optimizers.cpp:59
 *   Map::CreateMapPoints      /root/reference/src/lib/Map.cpp:16-65    consumes (keyframe, keyframe) DMatch lists
 * The handle keeps the last `window` keyframes (descriptors, keypoints, depth) in a device ring; promote_keyframe /
 * set_keyframe push into it, match_features keeps using the newest one (Map::GetLastKeyFrame, Map.cpp:68-70).
 * kfs_frontend_set_window drops every cached keyframe.  1 <= window <= KFS_MAX_WINDOW; a new handle has window 1. */
#define KFS_MAX_WINDOW 64
KFS_API kfs_status kfs_frontend_set_window(kfs_frontend *fe, int window);
KFS_API int kfs_frontend_window(const kfs_frontend *fe);         /* ring size                                         */
KFS_API int kfs_frontend_window_count(const kfs_frontend *fe);   /* keyframes held = min(keyframes pushed, window)   */
/* Current frame against every keyframe of the window in ONE kNN launch (one (keyframe, frame) pair per keyframe, as
 * configs 4 and 5 of BASELINE.json shape it) plus one ratio-test launch.  For keyframe i of GetNKeyFrames' order
 * (i = 0 oldest): knnMatch(keyframe_i descriptors (query), current descriptors (train), 2) exactly as
 * VisualOdometry.cpp:39, then d0 < ratio*d1 (:46) with ordered compaction.
 *   cur_desc / n_cur   host rows of the current frame, or NULL to use the device-resident result of the last
 *                      kfs_frontend_extract_features
 *   kf_rows_out[i]     query rows of keyframe i                                         (KFS_MAX_WINDOW ints or NULL)
 *   knn_out            [n_kf][knn_rows][2] raw kNN rows (unused entries {q,-1,-1,FLT_MAX})   (NULL: not wanted)
 *   good_out           [n_kf][capacity] survivors in query order; n_good_out[i] = survivors of keyframe i
 * Returns KFS_ERR_CAPACITY (outputs truncated) when a keyframe has more rows than knn_rows or more survivors than
 * capacity.  *n_kf_out = keyframes matched (0 with an empty ring: no error, like an empty vecKeyFrames). */
KFS_API kfs_status kfs_frontend_match_window(kfs_frontend *fe, const uint8_t *cur_desc, int n_cur, float ratio,
                                             int *kf_rows_out, kfs_dmatch *knn_out, int knn_rows,
                                             kfs_dmatch *good_out, int *n_good_out, int capacity, int *n_kf_out);
KFS_API int kfs_frontend_last_launches(const kfs_frontend *fe);

/* ---- microbenchmarks used by bench.py for the integer-pipe roofline -------------------- */
/* Dependency-free POPC.32 (what=0) or LOP3.32 (what=1) issue-rate probe on `device`; returns 1e9 ops/s. */
KFS_API kfs_status kfs_pipe_peak(int device, int what, double *gops);

#ifdef __cplusplus
}
#endif
#endif /* KFSLAM_FRONTEND_H */

/* track_oracle.c — CPU restatement of the two tracker methods that wrap the hot path.  TEST INFRASTRUCTURE ONLY: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may build, link or call this file.
 *
 *   kfo_rgb2gray        /root/reference/src/lib/VisualOdometry.cpp:25-28   cvtColor(rgb, rgb, RGB2BGR);
 *                       cvtColor(rgb, gray, BGR2GRAY) — OpenCV's 8-bit path; pinned against cv2 4.13 by
 *                       tests/golden/track_gray.npz (tools/make_golden_track.py), bit for bit.
 *   kfo_match_features  /root/reference/src/lib/VisualOdometry.cpp:45-59   the loop over knnMatches: ratio test (:46),
 *                       depth look-up (:52), pin-hole back-projection (:53-54), the two push_backs (:56-57).  The
 *                       reference's own source for these lines cannot be compiled here (it needs OpenCV headers), so
 *                       this part is pinned by construction only: the same C float / double expressions, in the same
 *                       order, compiled with -ffp-contract=off ("parity unpinned" for lines 52-57; the kNN rows that
 *                       feed it are pinned against cv2 by knn_oracle.c's goldens).
 *   kfo_match_window    /root/reference/src/lib/Map.cpp:72-78 (Map::GetNKeyFrames: the last `win` keyframes, oldest first,
 *                       or all of them while fewer exist) composed with VisualOdometry.cpp:39 (knnMatch(keyframe, current,
 *                       2)) and :46 (ratio test) per keyframe — the (frame, keyframe) pairs of BASELINE.json configs 4/5.
 *                       Pinned against cv2 4.13 by tests/golden/window_cfg4.npz (tools/make_golden_window.py).
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

typedef struct { float x, y, size, angle, response; int32_t octave, class_id; } kfo_keypoint; /* == cv::KeyPoint */
typedef struct { int32_t queryIdx, trainIdx, imgIdx; float distance; } kfo_dmatch;            /* == cv::DMatch   */

/* :25-28.  rgb: h rows of w x 3 bytes (R, G, B as the reference names them).  After RGB2BGR the R byte sits in channel 2,
 * which BGR2GRAY weighs with 9798 (0.299), G with 19235 (0.587), B with 3735 (0.114), 15-bit fixed point, +16384, >> 15. */
void kfo_rgb2gray(const uint8_t *rgb, size_t step, int w, int h, uint8_t *gray, size_t gstep)
{
    for (int y = 0; y < h; y++) {
        const uint8_t *s = rgb + (size_t)y * step;
        uint8_t *d = gray + (size_t)y * gstep;
        for (int x = 0; x < w; x++) {
            const unsigned r = s[3 * x], g = s[3 * x + 1], b = s[3 * x + 2];
            d[x] = (uint8_t)((r * 9798u + g * 19235u + b * 3735u + 16384u) >> 15);
        }
    }
}

static int kfo_cvround(float v) { return (int)lrintf(v); } /* cvRound: round half to even (SSE cvtss2si) */

/* :45-59.  knn: nq rows of 2 (row q's entries; trainIdx < 0 marks a missing neighbour, in which case the reference's
 * unguarded knnMatches[i][1] would be out of bounds — such rows are skipped here).  K = {fx, fy, cx, cy}.
 * Returns the number of survivors; writes at most `capacity` of them. */
int kfo_match_features(const kfo_dmatch *knn, int nq, const kfo_keypoint *kf_kp, const kfo_keypoint *cur_kp,
                       const uint16_t *depth, size_t depth_step_bytes, const double *K, float ratio,
                       kfo_dmatch *matches, double *p3d, double *p2d, int capacity)
{
    const double fx = K[0], fy = K[1], cx = K[2], cy = K[3];
    int n = 0;
    for (int i = 0; i < nq; i++) {
        const kfo_dmatch m0 = knn[2 * i], m1 = knn[2 * i + 1];
        if (m0.trainIdx < 0 || m1.trainIdx < 0) continue;
        if (m0.distance < ratio * m1.distance) {                                           /* :46 */
            const kfo_keypoint kpPrev = kf_kp[m0.queryIdx], kpCurr = cur_kp[m0.trainIdx];  /* :49-50 */
            const uint16_t *row = (const uint16_t *)((const uint8_t *)depth + (size_t)kfo_cvround(kpPrev.y) * depth_step_bytes);
            float z = row[kfo_cvround(kpPrev.x)] * 0.001;                                  /* :52  int * double -> float */
            float y = (kpPrev.y - cy) * z / fy;                                            /* :53  double arithmetic -> float */
            float x = (kpPrev.x - cx) * z / fx;                                            /* :54 */
            if (n < capacity) {
                matches[n] = m0;                                                           /* :48 */
                p3d[3 * n] = x; p3d[3 * n + 1] = y; p3d[3 * n + 2] = z;                    /* :56 Point3d(x, y, z) */
                p2d[2 * n] = kpCurr.x; p2d[2 * n + 1] = kpCurr.y;                          /* :57 kpCurr.pt */
            }
            n++;
        }
    }
    return n;
}

/* knn_oracle.c */
int kfo_knn_match(const uint8_t *query, int nq, const uint8_t *train, int nt, int desc_bytes, int k, kfo_dmatch *out, int *counts);
int kfo_ratio_test(const kfo_dmatch *knn, const int *counts, int nq, int k, float ratio, kfo_dmatch *good);

/* Map.cpp:72-78 + VisualOdometry.cpp:39,46.  kf_desc[j] / kf_rows[j]: descriptors of ALL n_kf keyframes of the map in
 * creation order (Map::vecKeyFrames); the window is its last `win` entries (all of them if n_kf < win).  For window entry i
 * (0 = oldest): knn_out + i*row_stride*2 receives the k=2 rows, good_out + i*row_stride the survivors, n_good[i] their count;
 * counts_scratch holds row_stride ints.  Returns the number of keyframes in the window, or -1 on bad kNN arguments. */
int kfo_match_window(const uint8_t *const *kf_desc, const int *kf_rows, int n_kf, int win, const uint8_t *cur_desc, int n_cur, float ratio,
                     kfo_dmatch *knn_out, kfo_dmatch *good_out, int *n_good, int *counts_scratch, int row_stride)
{
    const int first = n_kf >= win ? n_kf - win : 0;                                        /* Map.cpp:73-77 */
    int w = 0;
    for (int j = first; j < n_kf; j++, w++) {
        kfo_dmatch *knn = knn_out + (size_t)w * row_stride * 2;
        if (kfo_knn_match(kf_desc[j], kf_rows[j], cur_desc, n_cur, 32, 2, knn, counts_scratch)) return -1;   /* :39 */
        n_good[w] = kfo_ratio_test(knn, counts_scratch, kf_rows[j], 2, ratio, good_out + (size_t)w * row_stride);   /* :46 */
    }
    return w;
}

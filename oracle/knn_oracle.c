/*
 * oracle/knn_oracle.c — TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * Plain-C CPU restatement of the reference's matching call
 *     this->ptrMatcher->knnMatch(keyframe->matDescriptors, dpCurrImg, knnMatches, 2);
 *         /root/reference/src/lib/VisualOdometry.cpp:39   (BFMatcher(NORM_HAMMING) built at :18)
 * and of the ratio test at :46.  The arithmetic lives in OpenCV (unpinned dependency, not vendored):
 * BFMatcher::knnMatchImpl -> batchDistance(NORM_HAMMING, K) as pinned in SURVEY.md §8(a) row a2:
 * Hamming distance = popcount of the 256-bit XOR; the K smallest are kept by insertion with strict
 * '<' while the train index ascends (ties -> lower trainIdx first); rows hold min(K, Nt) entries.
 * PINNED against opencv-python-headless 4.13.0.92 through tests/golden (tools/make_golden.py).
 */
#include <stdint.h>
#include <string.h>

typedef struct { int32_t queryIdx, trainIdx, imgIdx; float distance; } kfo_dmatch; /* == cv::DMatch */

static inline int kfo_hamming(const uint8_t *a, const uint8_t *b, int nbytes)
{
    int d = 0, i = 0;
    for (; i + 8 <= nbytes; i += 8) {
        uint64_t x, y;
        memcpy(&x, a + i, 8);
        memcpy(&y, b + i, 8);
        d += __builtin_popcountll(x ^ y);
    }
    for (; i < nbytes; i++) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
    return d;
}

/* out: nq*k entries (row q at out + q*k), counts[q] = min(k, nt). Returns 0, or -1 on bad args. */
int kfo_knn_match(const uint8_t *query, int nq, const uint8_t *train, int nt, int nbytes, int k,
                  kfo_dmatch *out, int *counts)
{
    if (k < 1 || k > 8 || nq < 0 || nt < 0 || nt >= (1 << 18)) return -1;
    for (int q = 0; q < nq; q++) {
        int bd[8], bi[8], n = 0;
        const uint8_t *qd = query + (size_t)q * nbytes;
        for (int t = 0; t < nt; t++) {
            int d = kfo_hamming(qd, train + (size_t)t * nbytes, nbytes);
            int pos = n;
            while (pos > 0 && d < bd[pos - 1]) pos--;      /* strict '<': earlier index wins ties */
            if (pos >= k) continue;
            int last = n < k ? n : k - 1;
            for (int j = last; j > pos; j--) { bd[j] = bd[j - 1]; bi[j] = bi[j - 1]; }
            bd[pos] = d; bi[pos] = t;
            if (n < k) n++;
        }
        for (int j = 0; j < n; j++) {
            kfo_dmatch *m = out + (size_t)q * k + j;
            m->queryIdx = q; m->trainIdx = bi[j]; m->imgIdx = 0; m->distance = (float)bd[j];
        }
        if (counts) counts[q] = n;
    }
    return 0;
}

/* Ratio test of VisualOdometry.cpp:46: keep row q iff it has 2 entries and d0 < ratio*d1 (fp32). */
int kfo_ratio_test(const kfo_dmatch *knn, const int *counts, int nq, int k, float ratio, kfo_dmatch *good)
{
    int n = 0;
    for (int q = 0; q < nq; q++) {
        if (counts[q] < 2) continue;
        const kfo_dmatch *r = knn + (size_t)q * k;
        float rhs = ratio * r[1].distance;
        if (r[0].distance < rhs) good[n++] = r[0];
    }
    return n;
}

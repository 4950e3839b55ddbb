// kfslam_frontend_cv.hpp — header-only C++ adapter: gives the reference tracker back the exact call surface
// it uses today, on top of the C ABI (kfslam_frontend.h).
//
//   reference (include/lib/VisualOdometry.hpp:28-29, src/lib/VisualOdometry.cpp:10-18,30,39)      drop-in
//   ------------------------------------------------------------------------------------------   ---------------------------
//   cv::Ptr<cv::ORB>       ptrExtractor = cv::ORB::create(3000,1.2f,8,31,0,2,HARRIS_SCORE,31,20)  kfs::ORB::create(... same ...)
//   ptrExtractor->detectAndCompute(gray, mask, kpCurrImg, dpCurrImg)                              same call, same types
//   cv::Ptr<cv::BFMatcher> ptrMatcher   = cv::BFMatcher::create(cv::NORM_HAMMING)                 kfs::BFMatcher::create(cv::NORM_HAMMING)
//   ptrMatcher->knnMatch(keyframe->matDescriptors, dpCurrImg, knnMatches, 2)                      same call, same types
//
// Error behaviour mirrors OpenCV's: bad arguments / unsupported options throw cv::Exception (the reference never
// catches, VisualOdometry.cpp:30,39).  Output semantics mirror OpenCV's: `keypoints` is resized, `descriptors`
// is (re)created as n x 32 CV_8U or released when n == 0; knnMatch rows hold min(k, Nt) DMatch entries.
// Keypoint order: ascending octave, then (y, x) inside an octave (OpenCV leaves the in-level order unspecified).
#pragma once
#if __has_include(<opencv2/core.hpp>) && !defined(KFS_FORCE_CV_MOCK) && !defined(KFS_CV_MOCK)
#include <opencv2/core.hpp>
#define KFS_HAVE_OPENCV 1
#elif defined(KFS_CV_MOCK) || defined(KFS_FORCE_CV_MOCK)
#include "kfslam_cv_mock.hpp" // test scaffolding (tests/cpp/, add -I tests/cpp): stand-ins for the OpenCV types when its headers are absent
#define KFS_HAVE_OPENCV 0
#else
#error "kfslam_frontend_cv.hpp needs the OpenCV headers (<opencv2/core.hpp> not found). This repository's own tests build against stand-in types: add -I tests/cpp -DKFS_CV_MOCK"
#endif
#include "kfslam_frontend.h"
#include <cstring>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace kfs {

static_assert(sizeof(cv::KeyPoint) == sizeof(kfs_keypoint), "cv::KeyPoint must be layout-identical to kfs_keypoint");
static_assert(sizeof(cv::DMatch) == sizeof(kfs_dmatch), "cv::DMatch must be layout-identical to kfs_dmatch");

namespace detail {
[[noreturn]] inline void raise(kfs_status st, const char *where)
{
    const std::string msg = std::string(where) + ": " + kfs_last_error();
#if KFS_HAVE_OPENCV
    throw cv::Exception(st == KFS_ERR_UNSUPPORTED ? cv::Error::StsNotImplemented : cv::Error::StsBadArg, msg, where, __FILE__, __LINE__);
#else
    throw cv::Exception((int)st, msg);
#endif
}
#if KFS_HAVE_OPENCV
inline cv::Mat as_mat(cv::InputArray a) { return a.getMat(); }
inline void make_desc(cv::OutputArray d, int n) { if (n) d.create(n, 32, CV_8U); else d.release(); }
inline cv::Mat out_mat(cv::OutputArray d) { return d.getMat(); }
// keep the first n rows of a matrix created with more, WITHOUT moving them (the matcher recognises the rows by their address)
inline void shrink_desc(cv::OutputArray d, int n) { if (n) { cv::Mat m = d.getMat(); d.getMatRef() = m.rowRange(0, n); } else d.release(); }
// getMatRef() is defined for a plain cv::Mat output only; cv::UMat, std::vector<uchar>, cv::Mat_<uchar> ... (all of which cv::ORB
// accepts) get their rows through a staging matrix and copyTo instead, and are then simply never matched in place
inline bool desc_in_place(cv::OutputArray d) { return d.kind() == cv::_InputArray::MAT; }
inline void copy_desc(const cv::Mat &stage, int n, cv::OutputArray d) { if (n) stage.rowRange(0, n).copyTo(d); else d.release(); }
#else
inline const cv::Mat &as_mat(cv::InputArray a) { return a; }
inline void make_desc(cv::OutputArray d, int n) { if (n) d.create(n, 32, cv::CV_8U); else d.release(); }
inline cv::Mat &out_mat(cv::OutputArray d) { return d; }
inline void shrink_desc(cv::OutputArray d, int n) { if (n) d.rows = n; else d.release(); }
inline bool desc_in_place(cv::OutputArray) { return true; }
inline void copy_desc(const cv::Mat &, int, cv::OutputArray) {}
#endif
} // namespace detail

// Stand-in for cv::ORB (Feature2D).  Not thread-safe, like the reference's single-threaded use.
class ORB {
public:
    enum ScoreType { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    static cv::Ptr<ORB> create(int nfeatures = 500, float scaleFactor = 1.2f, int nlevels = 8, int edgeThreshold = 31,
                               int firstLevel = 0, int WTA_K = 2, int scoreType = HARRIS_SCORE, int patchSize = 31,
                               int fastThreshold = 20, int device = 0)
    {
        kfs_orb_params p{nfeatures, scaleFactor, nlevels, edgeThreshold, firstLevel, WTA_K, scoreType, patchSize, fastThreshold};
        kfs_orb *h = nullptr;
        const kfs_status st = kfs_orb_create(&p, device, &h);
        if (st != KFS_OK) detail::raise(st, "kfs::ORB::create");
        return cv::Ptr<ORB>(new ORB(h));
    }
    ~ORB() { kfs_orb_destroy(h_); }
    ORB(const ORB &) = delete;
    ORB &operator=(const ORB &) = delete;

    // follow the other OpenCV dispatch of the per-level blur (cv::setUseOptimized(false) / a host without FMA)
    void setBlurFMA(bool fma)
    {
        const kfs_status st = kfs_orb_set_blur_fma(h_, fma ? 1 : 0);
        if (st != KFS_OK) detail::raise(st, "kfs::ORB::setBlurFMA");
    }
    // cv::Feature2D::detectAndCompute(image, mask, keypoints, descriptors, useProvidedKeypoints=false)
    void detectAndCompute(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint> &keypoints,
                          cv::OutputArray descriptors, bool useProvidedKeypoints = false)
    {
        if (useProvidedKeypoints) { kfs_status st = KFS_ERR_UNSUPPORTED; (void)st; throw_unsupported("useProvidedKeypoints=true is not built"); }
        const auto img = detail::as_mat(image);
        const auto msk = detail::as_mat(mask);
        if (img.empty()) { keypoints.clear(); detail::make_desc(descriptors, 0); return; }
        if (img.channels() != 1) throw_unsupported("image must be 8-bit single channel (the tracker converts to gray first, VisualOdometry.cpp:28)");
        if (!msk.empty() && (msk.rows != img.rows || msk.cols != img.cols)) throw_unsupported("mask size differs from image size");
        int cap = kfs_orb_capacity(h_);
        for (;;) {
            // the descriptors are written straight into the output matrix (created with the handle's capacity, cut to n rows
            // in place afterwards): no staging copy, and an attached kfs::BFMatcher finds the rows it is handed next
            // (knnMatch(keyframe, dpCurrImg), VisualOdometry.cpp:39) still on the device by their host address
            keypoints.resize((size_t)cap);
            const bool in_place = detail::desc_in_place(descriptors);
            cv::Mat stage;
            if (in_place) detail::make_desc(descriptors, cap);
            else detail::make_desc(stage, cap);
            auto d = in_place ? detail::out_mat(descriptors) : stage;
            int n = 0;
            const kfs_status st = kfs_orb_detect_and_compute(h_, img.ptr(0), (size_t)img.step, img.cols, img.rows,
                                                             msk.empty() ? nullptr : msk.ptr(0), msk.empty() ? 0 : (size_t)msk.step,
                                                             reinterpret_cast<kfs_keypoint *>(keypoints.data()), d.ptr(0), cap, &n);
            if (st == KFS_ERR_CAPACITY) { // ties pushed the count over the handle's slack: grow and redo
                if (kfs_orb_set_capacity(h_, n + 64) != KFS_OK) {
                    keypoints.clear();
                    detail::make_desc(descriptors, 0);
                    detail::raise(KFS_ERR_CUDA, "kfs::ORB::detectAndCompute");
                }
                cap = kfs_orb_capacity(h_);
                continue;
            }
            if (st != KFS_OK) { // leave the outputs empty, not `cap` rows of whatever the buffers held
                keypoints.clear();
                detail::make_desc(descriptors, 0);
                detail::raise(st, "kfs::ORB::detectAndCompute");
            }
            keypoints.resize((size_t)n);
            if (in_place) detail::shrink_desc(descriptors, n);
            else detail::copy_desc(stage, n, descriptors);
            return;
        }
    }
    kfs_orb *handle() const { return h_; }

private:
    explicit ORB(kfs_orb *h) : h_(h) {}
    [[noreturn]] static void throw_unsupported(const char *what)
    {
#if KFS_HAVE_OPENCV
        throw cv::Exception(cv::Error::StsNotImplemented, what, "kfs::ORB", __FILE__, __LINE__);
#else
        throw cv::Exception((int)KFS_ERR_UNSUPPORTED, what);
#endif
    }
    kfs_orb *h_;
};

// Stand-in for cv::BFMatcher (DescriptorMatcher).
class BFMatcher {
public:
    static cv::Ptr<BFMatcher> create(int normType = cv::NORM_L2, bool crossCheck = false, int device = 0)
    {
        kfs_bf *h = nullptr;
        const kfs_status st = kfs_bf_create(normType, crossCheck ? 1 : 0, device, &h);
        if (st != KFS_OK) detail::raise(st, "kfs::BFMatcher::create");
        return cv::Ptr<BFMatcher>(new BFMatcher(h));
    }
    ~BFMatcher() { kfs_bf_destroy(h_); }
    BFMatcher(const BFMatcher &) = delete;
    BFMatcher &operator=(const BFMatcher &) = delete;

    // cv::DescriptorMatcher::knnMatch(query, train, matches, k, mask=noArray(), compactResult=false)
    void knnMatch(cv::InputArray queryDescriptors, cv::InputArray trainDescriptors,
                  std::vector<std::vector<cv::DMatch>> &matches, int k,
                  cv::InputArray mask = cv::noArray(), bool compactResult = false)
    {
        (void)compactResult;
        const auto q = detail::as_mat(queryDescriptors);
        const auto t = detail::as_mat(trainDescriptors);
        const auto m = detail::as_mat(mask);
        if (!m.empty()) { kfs_status st = KFS_ERR_UNSUPPORTED; (void)st; throw cv::Exception(
#if KFS_HAVE_OPENCV
            cv::Error::StsNotImplemented, "match masks are not built", "kfs::BFMatcher::knnMatch", __FILE__, __LINE__
#else
            (int)KFS_ERR_UNSUPPORTED, "match masks are not built"
#endif
        ); }
        matches.clear();
        const int nq = q.empty() ? 0 : q.rows, nt = t.empty() ? 0 : t.rows;
        if (nq == 0) return;                               // OpenCV: empty query -> empty result
        if (q.cols != 32 || (nt && t.cols != 32)) detail::raise((kfs_set_err("descriptors must be n x 32 CV_8U"), KFS_ERR_INVALID_ARG), "kfs::BFMatcher::knnMatch");
        std::vector<uint8_t> qb, tb;
        const uint8_t *qp = pack(q, qb), *tp = nt ? pack(t, tb) : nullptr;
        flat_.resize((size_t)nq * k);
        counts_.resize((size_t)nq);
        const kfs_status st = kfs_bf_knn_match(h_, qp, nq, tp, nt, k, reinterpret_cast<kfs_dmatch *>(flat_.data()), counts_.data());
        if (st != KFS_OK) detail::raise(st, "kfs::BFMatcher::knnMatch");
        matches.resize((size_t)nq);
        for (int i = 0; i < nq; i++) matches[(size_t)i].assign(flat_.begin() + (size_t)i * k, flat_.begin() + (size_t)i * k + counts_[(size_t)i]);
    }
    kfs_bf *handle() const { return h_; }
    // Descriptor rows `orb` has just returned are matched where they still are on the device instead of being uploaded again
    // (kfs_bf_attach_orb); the extractor must outlive this matcher or be detached with attach(nullptr).
    void attach(const cv::Ptr<ORB> &orb)
    {
        const kfs_status st = kfs_bf_attach_orb(h_, orb ? orb->handle() : nullptr);
        if (st != KFS_OK) detail::raise(st, "kfs::BFMatcher::attach");
    }

private:
    explicit BFMatcher(kfs_bf *h) : h_(h) {}
    static int kfs_set_err(const char *) { return 0; }
    template <typename M> static const uint8_t *pack(const M &m, std::vector<uint8_t> &buf)
    {
        if (m.isContinuous()) return m.ptr(0);
        buf.resize((size_t)m.rows * 32);
        for (int r = 0; r < m.rows; r++) std::memcpy(buf.data() + (size_t)r * 32, m.ptr(r), 32);
        return buf.data();
    }
    kfs_bf *h_;
    std::vector<cv::DMatch> flat_;
    std::vector<int> counts_;
};

// The tracker's two methods around the drop-in calls, with the reference's own signatures:
//   std::pair<std::vector<cv::KeyPoint>, cv::Mat> VisualOdometry::ExtractFeatures(cv::Mat rgb, cv::Mat mask)        (VisualOdometry.cpp:21-33)
//   std::pair<std::vector<cv::Point3d>, std::vector<cv::Point2d>>
//       VisualOdometry::MatchFeatures(cv::Mat dpCurrImg, std::vector<cv::KeyPoint> kpCurrImg, std::vector<cv::DMatch>& matches)   (:35-62)
// The keyframe the reference fetches with map->GetLastKeyFrame() (:36; descriptors, keypoints, depth — KeyFrame.hpp:35-39)
// lives in the handle's device cache: SetKeyFrame / PromoteKeyFrame replace the KeyFrame construction at :105-123.
class TrackerFrontEnd {
public:
    static cv::Ptr<TrackerFrontEnd> create(int nfeatures = 3000, float scaleFactor = 1.2f, int nlevels = 8, int fastThreshold = 20, int device = 0)
    {
        kfs_orb_params p;
        kfs_orb_default_params(&p);
        p.nfeatures = nfeatures; p.scale_factor = scaleFactor; p.nlevels = nlevels; p.fast_threshold = fastThreshold;
        kfs_frontend *h = nullptr;
        const kfs_status st = kfs_frontend_create(&p, device, &h);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::create");
        return cv::Ptr<TrackerFrontEnd>(new TrackerFrontEnd(h));
    }
    ~TrackerFrontEnd() { kfs_frontend_destroy(h_); }
    TrackerFrontEnd(const TrackerFrontEnd &) = delete;
    TrackerFrontEnd &operator=(const TrackerFrontEnd &) = delete;

    // rows per frame (default nfeatures + nfeatures/4 + 64); drops the cached keyframes: call before tracking starts
    void SetCapacity(int rows)
    {
        const kfs_status st = kfs_frontend_set_capacity(h_, rows);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::SetCapacity");
    }
    int Capacity() const { return kfs_frontend_capacity(h_); }
    // K(0,0), K(1,1), K(0,2), K(1,2) of the reference's camera matrix (VisualOdometry.cpp:53-54)
    void setIntrinsics(double fx, double fy, double cx, double cy) { k_[0] = fx; k_[1] = fy; k_[2] = cx; k_[3] = cy; }
    void setRatio(float r) { ratio_ = r; } // 0.6f in the reference (:46)

    std::pair<std::vector<cv::KeyPoint>, cv::Mat> ExtractFeatures(const cv::Mat &rgb, const cv::Mat &mask)
    {
        std::vector<cv::KeyPoint> kp;
        cv::Mat desc;
        const int cap = kfs_frontend_capacity(h_);
        kp.resize((size_t)cap);
        std::vector<uint8_t> d((size_t)cap * 32);
        int n = 0;
        const kfs_status st = kfs_frontend_extract_features(h_, rgb.data, (size_t)rgb.step, rgb.cols, rgb.rows, rgb.channels(),
                                                            mask.empty() ? nullptr : mask.data, mask.empty() ? 0 : (size_t)mask.step,
                                                            reinterpret_cast<kfs_keypoint *>(kp.data()), d.data(), cap, &n);
        // a tie-rich frame that overflows the handle is an error here exactly as in Track(): never a silently shorter keypoint set
        // (kfs::ORB::detectAndCompute grows and retries; this handle also owns the keyframe ring, so growing is SetCapacity's job)
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::ExtractFeatures");
        kp.resize((size_t)n);
        detail::make_desc(desc, n);
        if (n) {
            for (int i = 0; i < n; i++) std::memcpy(desc.ptr(i), d.data() + (size_t)i * 32, 32);
        }
        return {kp, desc};
    }
    // the frame just extracted becomes the keyframe; depth = CV_16UC1 millimetres (KeyFrame::matDepth)
    void PromoteKeyFrame(const cv::Mat &depth)
    {
        const kfs_status st = kfs_frontend_promote_keyframe(h_, reinterpret_cast<const uint16_t *>(depth.data), (size_t)depth.step);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::PromoteKeyFrame");
    }
    void SetKeyFrame(const cv::Mat &descriptors, const std::vector<cv::KeyPoint> &keypoints, const cv::Mat &depth)
    {
        std::vector<uint8_t> d((size_t)descriptors.rows * 32);
        for (int i = 0; i < descriptors.rows; i++) std::memcpy(d.data() + (size_t)i * 32, descriptors.ptr(i), 32);
        const kfs_status st = kfs_frontend_set_keyframe(h_, d.data(), reinterpret_cast<const kfs_keypoint *>(keypoints.data()), (int)keypoints.size(),
                                                        reinterpret_cast<const uint16_t *>(depth.data), (size_t)depth.step, depth.cols, depth.rows);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::SetKeyFrame");
    }
    std::pair<std::vector<cv::Point3d>, std::vector<cv::Point2d>> MatchFeatures(const cv::Mat &dpCurrImg, const std::vector<cv::KeyPoint> &kpCurrImg,
                                                                               std::vector<cv::DMatch> &matches)
    {
        const int cap = kfs_frontend_capacity(h_);
        std::vector<cv::DMatch> m((size_t)cap);
        std::vector<double> p3((size_t)cap * 3), p2((size_t)cap * 2);
        std::vector<uint8_t> d((size_t)dpCurrImg.rows * 32);
        for (int i = 0; i < dpCurrImg.rows; i++) std::memcpy(d.data() + (size_t)i * 32, dpCurrImg.ptr(i), 32);
        int n = 0;
        const bool resident = dpCurrImg.empty() && kpCurrImg.empty(); // match the frame ExtractFeatures left on the device
        const kfs_status st = kfs_frontend_match_features(h_, resident ? nullptr : d.data(),
                                                          resident ? nullptr : reinterpret_cast<const kfs_keypoint *>(kpCurrImg.data()),
                                                          (int)kpCurrImg.size(), k_, ratio_, reinterpret_cast<kfs_dmatch *>(m.data()), p3.data(), p2.data(), cap, &n);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::MatchFeatures");
        std::vector<cv::Point3d> points3D;
        std::vector<cv::Point2d> points2D;
        for (int i = 0; i < n; i++) {
            matches.push_back(m[(size_t)i]);                                                                  // :48
            points3D.push_back(cv::Point3d(p3[3 * (size_t)i], p3[3 * (size_t)i + 1], p3[3 * (size_t)i + 2])); // :56
            points2D.push_back(cv::Point2d(p2[2 * (size_t)i], p2[2 * (size_t)i + 1]));                        // :57
        }
        return {points3D, points2D};
    }

    // ExtractFeatures + MatchFeatures of one frame with one host synchronisation (VisualOdometry.cpp:138,144).  Returns false
    // while no keyframe is cached (only keypoints / descriptors are filled: promote this frame).
    bool Track(const cv::Mat &rgb, const cv::Mat &mask, std::vector<cv::KeyPoint> &keypoints, cv::Mat &descriptors,
               std::vector<cv::Point3d> &points3D, std::vector<cv::Point2d> &points2D, std::vector<cv::DMatch> &matches)
    {
        const int cap = kfs_frontend_capacity(h_);
        keypoints.resize((size_t)cap);
        std::vector<uint8_t> d((size_t)cap * 32);
        std::vector<cv::DMatch> m((size_t)cap);
        std::vector<double> p3((size_t)cap * 3), p2((size_t)cap * 2);
        int n = 0, nm = -1;
        const kfs_status st = kfs_frontend_track(h_, rgb.data, (size_t)rgb.step, rgb.cols, rgb.rows, rgb.channels(), mask.empty() ? nullptr : mask.data,
                                                 mask.empty() ? 0 : (size_t)mask.step, k_, ratio_, reinterpret_cast<kfs_keypoint *>(keypoints.data()), d.data(), cap,
                                                 &n, reinterpret_cast<kfs_dmatch *>(m.data()), p3.data(), p2.data(), cap, &nm);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::Track");
        keypoints.resize((size_t)n);
        detail::make_desc(descriptors, n);
        for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), d.data() + (size_t)i * 32, 32);
        points3D.clear(); points2D.clear(); matches.clear();
        for (int i = 0; i < nm; i++) {
            matches.push_back(m[(size_t)i]);
            points3D.push_back(cv::Point3d(p3[3 * (size_t)i], p3[3 * (size_t)i + 1], p3[3 * (size_t)i + 2]));
            points2D.push_back(cv::Point2d(p2[2 * (size_t)i], p2[2 * (size_t)i + 1]));
        }
        return nm >= 0;
    }

    // ---- keyframe window: Map::GetNKeyFrames (src/lib/Map.cpp:72-78; window 5 at include/lib/Optimizers.hpp:70) ----
    void SetWindow(int window)
    {
        const kfs_status st = kfs_frontend_set_window(h_, window);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::SetWindow");
    }
    int WindowCount() const { return kfs_frontend_window_count(h_); }
    // dpCurrImg (empty: the frame ExtractFeatures left on the device) against every cached keyframe, oldest first:
    // result[i] = ratio-test survivors of knnMatch(keyframe_i, current, 2), the DMatch lists Map::CreateMapPoints consumes
    std::vector<std::vector<cv::DMatch>> MatchWindow(const cv::Mat &dpCurrImg)
    {
        const int cap = kfs_frontend_capacity(h_), cnt = kfs_frontend_window_count(h_);
        std::vector<cv::DMatch> good((size_t)cap * (size_t)(cnt > 0 ? cnt : 1));
        std::vector<uint8_t> d((size_t)dpCurrImg.rows * 32);
        for (int i = 0; i < dpCurrImg.rows; i++) std::memcpy(d.data() + (size_t)i * 32, dpCurrImg.ptr(i), 32);
        int n_good[KFS_MAX_WINDOW], nkf = 0;
        const kfs_status st = kfs_frontend_match_window(h_, dpCurrImg.empty() ? nullptr : d.data(), dpCurrImg.rows, ratio_, nullptr, nullptr, 0,
                                                        reinterpret_cast<kfs_dmatch *>(good.data()), n_good, cap, &nkf);
        if (st != KFS_OK) detail::raise(st, "kfs::TrackerFrontEnd::MatchWindow");
        std::vector<std::vector<cv::DMatch>> out((size_t)nkf);
        for (int i = 0; i < nkf; i++) out[(size_t)i].assign(good.begin() + (size_t)i * cap, good.begin() + (size_t)i * cap + n_good[i]);
        return out;
    }

private:
    explicit TrackerFrontEnd(kfs_frontend *h) : h_(h) {}
    kfs_frontend *h_;
    double k_[4] = {1, 1, 0, 0};
    float ratio_ = 0.6f;
};

} // namespace kfs

// tests/cpp/adapter_test.cpp — drives the C++ drop-in exactly the way the reference tracker drives OpenCV
// (VisualOdometry.cpp:10-18 construction, :30 detectAndCompute, :39 knnMatch, :46 ratio test) and dumps the
// results for the Python test to compare with the golden vectors.  Usage:
//   adapter_test <w> <h> <nfeatures> <frameA.raw> <frameB.raw> <out.bin>     (needs a GPU)
//   adapter_test --abi-only                                                  (no GPU: construction must fail loudly)
#include "kfslam_frontend_cv.hpp"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>

static std::vector<uint8_t> slurp(const char *path)
{
    std::ifstream f(path, std::ios::binary);
    return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

int main(int argc, char **argv)
{
    if (argc == 2 && !std::strcmp(argv[1], "--abi-only")) {
        std::printf("kfs_version=%d devices=%d\n", kfs_version(), kfs_device_count());
        if (kfs_device_count() == 0) {
            try { kfs::ORB::create(3000, 1.2f, 8, 31, 0, 2, kfs::ORB::HARRIS_SCORE, 31, 20); }
            catch (const cv::Exception &e) { std::printf("no device -> cv::Exception: %s\n", e.what()); return 0; }
            return 2; // must not silently fall back to a CPU path
        }
        return 0;
    }
    if (argc == 10 && !std::strcmp(argv[1], "--track")) {
        // adapter_test --track <w> <h> <nfeatures> <rgbA.raw> <rgbB.raw> <depthA.raw> <out.bin> <fx,fy,cx,cy>
        // the reference's Track() front half: ExtractFeatures(keyframe), KeyFrame creation, ExtractFeatures(current),
        // MatchFeatures (VisualOdometry.cpp:138, :105-123, :144) through kfs::TrackerFrontEnd
        const int w = std::atoi(argv[2]), h = std::atoi(argv[3]), nf = std::atoi(argv[4]);
        auto a = slurp(argv[5]), b = slurp(argv[6]), dep = slurp(argv[7]);
        if ((int)a.size() != w * h * 3 || (int)b.size() != w * h * 3 || (int)dep.size() != w * h * 2) { std::fprintf(stderr, "bad sizes\n"); return 65; }
        double k[4];
        if (std::sscanf(argv[9], "%lf,%lf,%lf,%lf", &k[0], &k[1], &k[2], &k[3]) != 4) return 66;
        cv::Ptr<kfs::TrackerFrontEnd> fe = kfs::TrackerFrontEnd::create(nf);
        fe->setIntrinsics(k[0], k[1], k[2], k[3]);
        fe->SetWindow(5);                                                    // Optimizers.hpp:70
        cv::Mat rgbA(h, w, cv::CV_8UC3, a.data()), rgbB(h, w, cv::CV_8UC3, b.data()), depth(h, w, cv::CV_16UC1, dep.data()), mask;
        auto fa = fe->ExtractFeatures(rgbA, mask);
        fe->PromoteKeyFrame(depth);
        auto fb = fe->ExtractFeatures(rgbB, mask);
        std::vector<cv::DMatch> matches;
        auto pts = fe->MatchFeatures(fb.second, fb.first, matches);          // host rows, like the reference passes them
        std::vector<cv::DMatch> matches2;
        auto pts2 = fe->MatchFeatures(cv::Mat(), {}, matches2);              // device-resident current frame
        // keyframe window (Map::GetNKeyFrames): with one keyframe cached the window's only list is MatchFeatures' list
        auto win = fe->MatchWindow(cv::Mat());
        auto win2 = fe->MatchWindow(fb.second);
        const bool win_ok = fe->WindowCount() == 1 && win.size() == 1 && win2.size() == 1 && win[0].size() == matches.size() &&
                            win2[0].size() == matches.size() && !std::memcmp(win[0].data(), matches.data(), matches.size() * sizeof(cv::DMatch)) &&
                            !std::memcmp(win2[0].data(), matches.data(), matches.size() * sizeof(cv::DMatch));
        const bool same = win_ok && matches.size() == matches2.size() && pts.first.size() == pts2.first.size() &&
                          !std::memcmp(pts.first.data(), pts2.first.data(), pts.first.size() * sizeof(cv::Point3d)) &&
                          !std::memcmp(matches.data(), matches2.data(), matches.size() * sizeof(cv::DMatch));
        std::ofstream o(argv[8], std::ios::binary);
        const int hdr[4] = {(int)fa.first.size(), (int)fb.first.size(), (int)matches.size(), same ? 1 : 0};
        o.write((const char *)hdr, sizeof(hdr));
        o.write((const char *)fb.first.data(), fb.first.size() * sizeof(cv::KeyPoint));
        o.write((const char *)matches.data(), matches.size() * sizeof(cv::DMatch));
        o.write((const char *)pts.first.data(), pts.first.size() * sizeof(cv::Point3d));
        o.write((const char *)pts.second.data(), pts.second.size() * sizeof(cv::Point2d));
        return 0;
    }
    if (argc != 7) { std::fprintf(stderr, "usage\n"); return 64; }
    const int w = std::atoi(argv[1]), h = std::atoi(argv[2]), nf = std::atoi(argv[3]);
    auto a = slurp(argv[4]), b = slurp(argv[5]);
    if ((int)a.size() != w * h || (int)b.size() != w * h) { std::fprintf(stderr, "bad frame size\n"); return 65; }

    // --- the reference's constructor body (VisualOdometry.cpp:10-18), types swapped for the drop-in
    cv::Ptr<kfs::ORB> ptrExtractor = kfs::ORB::create(nf, 1.2f, 8, 31, 0, 2, kfs::ORB::HARRIS_SCORE, 31, 20);
    cv::Ptr<kfs::BFMatcher> ptrMatcher = kfs::BFMatcher::create(cv::NORM_HAMMING);

    // --- ExtractFeatures (VisualOdometry.cpp:21-33) on the keyframe and on the current frame
    cv::Mat grayA(h, w, cv::CV_8UC1, a.data()), grayB(h, w, cv::CV_8UC1, b.data()), mask;
    std::vector<cv::KeyPoint> kpA, kpB;
    cv::Mat dpA, dpB;
    ptrExtractor->detectAndCompute(grayA, mask, kpA, dpA);
    ptrExtractor->detectAndCompute(grayB, mask, kpB, dpB);

    // --- MatchFeatures (VisualOdometry.cpp:35-62): query = keyframe, train = current; ratio 0.6
    std::vector<std::vector<cv::DMatch>> knnMatches;
    ptrMatcher->knnMatch(dpA, dpB, knnMatches, 2);
    // the same call with the matcher attached to the extractor: dpCurrImg (the frame just extracted) is matched where it
    // still is on the device, the keyframe's rows (an older extraction) are uploaded as before; identical rows come back
    ptrMatcher->attach(ptrExtractor);
    std::vector<std::vector<cv::DMatch>> knnAttached;
    ptrMatcher->knnMatch(dpA, dpB, knnAttached, 2);
    const int aliased = kfs_bf_last_aliased(ptrMatcher->handle());
    bool attached_same = knnAttached.size() == knnMatches.size();
    for (size_t i = 0; attached_same && i < knnMatches.size(); i++)
        attached_same = knnAttached[i].size() == knnMatches[i].size() &&
                        !std::memcmp(knnAttached[i].data(), knnMatches[i].data(), knnMatches[i].size() * sizeof(cv::DMatch));
    ptrMatcher->attach(nullptr);
    std::vector<cv::DMatch> matches;
    for (size_t i = 0; i < knnMatches.size(); i++)
        if (knnMatches[i].size() >= 2 && knnMatches[i][0].distance < 0.6f * knnMatches[i][1].distance) matches.push_back(knnMatches[i][0]);

    // --- error behaviour: unsupported options throw like OpenCV
    int threw = 0;
    try { kfs::ORB::create(500, 1.2f, 8, 31, 0, 4); } catch (const cv::Exception &) { threw++; }
    try { kfs::BFMatcher::create(cv::NORM_L2); } catch (const cv::Exception &) { threw++; }
    cv::Mat empty; std::vector<cv::KeyPoint> kpE; cv::Mat dpE;
    ptrExtractor->detectAndCompute(empty, mask, kpE, dpE);
    if (!kpE.empty() || !dpE.empty()) threw = -100;

    std::ofstream out(argv[6], std::ios::binary);
    auto put32 = [&](int v) { out.write((const char *)&v, 4); };
    put32((int)kpA.size()); put32((int)kpB.size()); put32((int)knnMatches.size()); put32((int)matches.size());
    put32(threw + 16 * aliased + (attached_same ? 256 : 0)); // bits 0-3: exceptions seen, 4-7: matrices matched in place, 8: attached == plain
    out.write((const char *)kpA.data(), (std::streamsize)(kpA.size() * sizeof(cv::KeyPoint)));
    for (int r = 0; r < dpA.rows; r++) out.write((const char *)dpA.ptr(r), 32);
    out.write((const char *)kpB.data(), (std::streamsize)(kpB.size() * sizeof(cv::KeyPoint)));
    for (int r = 0; r < dpB.rows; r++) out.write((const char *)dpB.ptr(r), 32);
    for (auto &row : knnMatches) { put32((int)row.size()); out.write((const char *)row.data(), (std::streamsize)(row.size() * sizeof(cv::DMatch))); }
    out.write((const char *)matches.data(), (std::streamsize)(matches.size() * sizeof(cv::DMatch)));
    std::printf("kpA=%zu kpB=%zu knn=%zu good=%zu threw=%d\n", kpA.size(), kpB.size(), knnMatches.size(), matches.size(), threw);
    return 0;
}

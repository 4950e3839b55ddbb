"""GPU parity (-m gpu) of the tracker wrappers (kfs_frontend_*: VisualOdometry::ExtractFeatures / MatchFeatures,
/root/reference/src/lib/VisualOdometry.cpp:21-62) through the C ABI, against cv2 goldens and the oracle.  Everything is
integer or IEEE double/float arithmetic evaluated in the reference's order: the bar is bit-exact."""
import numpy as np
import pytest

from _track_util import make_depth, make_rgb
from _util import golden

pytestmark = pytest.mark.gpu
K = np.array([[525.3, 0, 319.7], [0, 524.9, 239.2], [0, 0, 1]])


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


def test_extract_features_colour_matches_cv2_golden(kfs):
    g = golden("track_cfg1")
    fe = kfs.TrackerFrontEnd(nfeatures=1000)
    kp, desc = fe.extractFeatures(make_rgb(480, 640, 21))
    assert np.array_equal(kp, g["kf_kp"]) and np.array_equal(desc, g["kf_desc"])
    assert fe.last_launches >= 8   # colour conversion + the ORB chain (7 launches with the one-launch pyramid) ran on the device


def test_track_two_frames_device_resident(kfs, oracle):
    """extract(kf) -> promoteKeyframe(depth) -> extract(cur) -> matchFeatures(): cv2 golden kNN rows + oracle lines 45-59."""
    g = golden("track_cfg1")
    depth = make_depth(480, 640, 5)
    fe = kfs.TrackerFrontEnd(nfeatures=1000)
    fe.extractFeatures(make_rgb(480, 640, 21))
    fe.promoteKeyframe(depth)
    kp, desc = fe.extractFeatures(make_rgb(480, 640, 21, shift=3))
    assert np.array_equal(kp, g["cur_kp"]) and np.array_equal(desc, g["cur_desc"])
    for ratio in (0.6, 0.75):
        p3, p2, m = fe.matchFeatures(K, ratio=ratio)
        r3, r2, rm = oracle.match_features(g["knn"], g["kf_kp"], g["cur_kp"], depth, K, ratio)
        assert len(rm) > 50
        assert np.array_equal(m, rm) and np.array_equal(p3, r3) and np.array_equal(p2, r2)
    assert fe.last_launches == 2  # kNN + fused ratio / back-projection


def test_match_features_host_rows_and_edges(kfs, oracle):
    g = golden("track_cfg1")
    depth = make_depth(480, 640, 6)
    fe = kfs.TrackerFrontEnd(nfeatures=1000)
    with pytest.raises(kfs.KfsError):
        fe.matchFeatures(K, g["cur_desc"], g["cur_kp"])          # no keyframe yet
    fe.setKeyframe(g["kf_desc"], g["kf_kp"], depth)
    p3, p2, m = fe.matchFeatures(K, g["cur_desc"], g["cur_kp"])
    r3, r2, rm = oracle.match_features(g["knn"], g["kf_kp"], g["cur_kp"], depth, K, 0.6)
    assert np.array_equal(m, rm) and np.array_equal(p3, r3) and np.array_equal(p2, r2)
    # 0, 1 and 2 current keypoints: fewer than 2 neighbours can never pass the ratio test
    for n in (0, 1, 2):
        knn, cnt = oracle.knn_match(g["kf_desc"], g["cur_desc"][:n], 2)
        rr = oracle.match_features(knn, g["kf_kp"], g["cur_kp"][:n], depth, K, 0.95)
        got = fe.matchFeatures(K, g["cur_desc"][:n], g["cur_kp"][:n], ratio=0.95)
        assert all(np.array_equal(a, b) for a, b in zip(got, rr))
    # empty keyframe
    fe.setKeyframe(g["kf_desc"][:0], g["kf_kp"][:0], depth)
    assert len(fe.matchFeatures(K, g["cur_desc"], g["cur_kp"])[2]) == 0


def test_extract_features_gray_mask_and_odd_sizes(kfs, oracle):
    fe = kfs.TrackerFrontEnd(nfeatures=700)
    P = oracle.orb_params(nfeatures=700)
    rgb = make_rgb(257, 333, 14)                      # odd width: unaligned rows, pad columns
    mask = np.zeros((257, 333), np.uint8)
    mask[40:200, 50:300] = 7
    for m in (None, mask):
        kp, desc = fe.extractFeatures(rgb, m)
        rk, rd = oracle.detect_and_compute(oracle.rgb2gray(rgb), m, P)
        assert len(rk) > 100 and np.array_equal(kp, rk) and np.array_equal(desc, rd)
    gray = oracle.rgb2gray(rgb)
    kp, desc = fe.extractFeatures(gray)               # 1 channel: used as is (VisualOdometry.cpp:29)
    rk, rd = oracle.detect_and_compute(gray, None, P)
    assert np.array_equal(kp, rk) and np.array_equal(desc, rd)
    with pytest.raises(kfs.KfsError):
        fe.extractFeatures(np.zeros((64, 64, 4), np.uint8))


def test_track_720p_nf3000_reference_configuration(kfs, oracle):
    """The reference's own extractor configuration (nfeatures = 3000, VisualOdometry.cpp:11) on a 1280x720 colour pair."""
    depth = make_depth(720, 1280, 9)
    K7 = np.array([[910.1, 0, 640.4], [0, 909.6, 360.7], [0, 0, 1]])
    a, b = make_rgb(720, 1280, 31), make_rgb(720, 1280, 31, shift=4)
    fe = kfs.TrackerFrontEnd(nfeatures=3000)
    P = oracle.orb_params(nfeatures=3000)
    ka, da = fe.extractFeatures(a)
    fe.promoteKeyframe(depth)
    kb, db = fe.extractFeatures(b)
    rka, rda = oracle.detect_and_compute(oracle.rgb2gray(a), None, P)
    rkb, rdb = oracle.detect_and_compute(oracle.rgb2gray(b), None, P)
    assert np.array_equal(ka, rka) and np.array_equal(da, rda) and np.array_equal(kb, rkb) and np.array_equal(db, rdb)
    knn, _ = oracle.knn_match(rda, rdb, 2)
    got = fe.matchFeatures(K7)
    ref = oracle.match_features(knn, rka, rkb, depth, K7, 0.6)
    assert len(ref[2]) > 300 and all(np.array_equal(x, y) for x, y in zip(got, ref))


def test_track_one_call_per_frame(kfs, oracle):
    """kfs_frontend_track = ExtractFeatures + MatchFeatures of a frame in one call / one synchronisation: a 6-frame sequence
    with a keyframe promotion in the middle, against the two-call path's oracle (detect_and_compute, knn_match, match_features)."""
    fe = kfs.TrackerFrontEnd(nfeatures=800)
    P = oracle.orb_params(nfeatures=800)
    depth = [make_depth(480, 640, 70 + i) for i in range(2)]
    frames = [make_rgb(480, 640, 61, shift=s) for s in (0, 2, 4, 7, 9, 12)]
    ref = [oracle.detect_and_compute(oracle.rgb2gray(f), None, P) for f in frames]
    kf = None
    for i, f in enumerate(frames):
        kp, desc, p3, p2, m = fe.track(f, K)
        assert np.array_equal(kp, ref[i][0]) and np.array_equal(desc, ref[i][1])
        if kf is None:
            assert m is None                       # no keyframe yet: extraction only
        else:
            knn, _ = oracle.knn_match(ref[kf][1], ref[i][1], 2)
            r3, r2, rm = oracle.match_features(knn, ref[kf][0], ref[i][0], depth[0 if kf == 0 else 1], K, 0.6)
            assert len(rm) > 30 and np.array_equal(m, rm) and np.array_equal(p3, r3) and np.array_equal(p2, r2)
        if i in (0, 3):                            # this frame becomes the keyframe (VisualOdometry.cpp:105-123)
            fe.promoteKeyframe(depth[0 if i == 0 else 1])
            kf = i
    # masked + gray input, and the two-call path agrees with the fused call on the same frame
    gray = oracle.rgb2gray(frames[5])
    mask = np.zeros((480, 640), np.uint8)
    mask[50:400, 60:600] = 255
    kp, desc, p3, p2, m = fe.track(gray, K, mask=mask, ratio=0.75)
    k2, d2 = fe.extractFeatures(gray, mask)
    q3, q2, qm = fe.matchFeatures(K, ratio=0.75)
    assert np.array_equal(kp, k2) and np.array_equal(desc, d2) and np.array_equal(m, qm) and np.array_equal(p3, q3) and np.array_equal(p2, q2)
    rk, rd = oracle.detect_and_compute(gray, mask, P)
    assert np.array_equal(kp, rk) and np.array_equal(desc, rd)

"""CPU: the tracker-wrapper oracle (oracle/track_oracle.c) against cv2 goldens and against an independent numpy
restatement of VisualOdometry.cpp:45-59."""
import hashlib

import numpy as np

from _track_util import make_depth, make_rgb
from _util import golden


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_rgb2gray_matches_cv2_golden(oracle):
    g = golden("track_cfg1")
    h, w, seed = g["meta"]["small"]
    assert np.array_equal(oracle.rgb2gray(make_rgb(h, w, seed)), g["small_gray"])
    for key, digest in g["meta"]["sha"].items():
        hw, seed = key.split("_")
        h, w = map(int, hw.split("x"))
        assert sha(oracle.rgb2gray(make_rgb(h, w, int(seed)))) == digest, key


def test_extract_features_oracle_matches_cv2_golden(oracle):
    """ExtractFeatures = gray conversion + detectAndCompute; keypoints/descriptors of the colour pair equal cv2's."""
    g = golden("track_cfg1")
    P = oracle.orb_params(nfeatures=1000)
    for shift, kk, dk in ((0, "kf_kp", "kf_desc"), (3, "cur_kp", "cur_desc")):
        kp, desc = oracle.detect_and_compute(oracle.rgb2gray(make_rgb(480, 640, 21, shift=shift)), None, P)
        assert np.array_equal(kp, g[kk]) and np.array_equal(desc, g[dk])
    m, c = oracle.knn_match(g["kf_desc"], g["cur_desc"], 2)
    assert np.array_equal(m, g["knn"])


def numpy_match_features(knn, kf_kp, cur_kp, depth, K, ratio):
    """Independent restatement of :45-59 with explicit numpy scalar types (float = np.float32, double = np.float64)."""
    p3, p2, ms = [], [], []
    fx, fy, cx, cy = (np.float64(K[0, 0]), np.float64(K[1, 1]), np.float64(K[0, 2]), np.float64(K[1, 2]))
    for i in range(len(knn)):
        m0, m1 = knn[i, 0], knn[i, 1]
        if m0["trainIdx"] < 0 or m1["trainIdx"] < 0:
            continue
        if np.float32(m0["distance"]) < np.float32(ratio) * np.float32(m1["distance"]):
            kp, kc = kf_kp[m0["queryIdx"]], cur_kp[m0["trainIdx"]]
            d = int(depth[int(np.rint(kp["y"])), int(np.rint(kp["x"]))])
            z = np.float32(np.float64(d) * np.float64(0.001))
            y = np.float32((np.float64(kp["y"]) - cy) * np.float64(z) / fy)
            x = np.float32((np.float64(kp["x"]) - cx) * np.float64(z) / fx)
            p3.append((np.float64(x), np.float64(y), np.float64(z)))
            p2.append((np.float64(kc["x"]), np.float64(kc["y"])))
            ms.append(m0)
    return np.array(p3, np.float64).reshape(-1, 3), np.array(p2, np.float64).reshape(-1, 2), np.array(ms, dtype=knn.dtype)


def test_match_features_oracle_vs_numpy_restatement(oracle):
    g = golden("track_cfg1")
    depth = make_depth(480, 640, 5)
    K = np.array([[525.3, 0, 319.7], [0, 524.9, 239.2], [0, 0, 1]])
    for ratio in (0.6, 0.75, 0.9):
        p3, p2, m = oracle.match_features(g["knn"], g["kf_kp"], g["cur_kp"], depth, K, ratio)
        r3, r2, rm = numpy_match_features(g["knn"], g["kf_kp"], g["cur_kp"], depth, K, ratio)
        assert len(m) == len(rm) and (ratio > 0.6 or len(m) > 50)
        assert np.array_equal(p3, r3) and np.array_equal(p2, r2) and np.array_equal(m, rm)
    # rows with a missing neighbour never pass
    knn = g["knn"].copy()
    knn["trainIdx"][:, 1] = -1
    assert len(oracle.match_features(knn, g["kf_kp"], g["cur_kp"], depth, K, 0.9)[2]) == 0

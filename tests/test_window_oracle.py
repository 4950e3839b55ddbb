"""CPU: the keyframe-window oracle (oracle/track_oracle.c: kfo_match_window = Map::GetNKeyFrames, Map.cpp:72-78, composed
with knnMatch :39 and the ratio test :46) against cv2 goldens (tests/golden/window_cfg4.npz, tools/make_golden_window.py)."""
import numpy as np

from _util import golden

NKF = 7


def _kfs(g):
    return [g["kf%d_desc" % i] for i in range(NKF)]


def test_window_oracle_matches_cv2_golden(oracle):
    g = golden("window_cfg4")
    for win in (1, 3, 5, 7, 10):                       # 5 = the reference's BA window (Optimizers.hpp:70); 10 = config 4
        for tag, ratio in (("060", 0.6), ("075", 0.75)):
            knn, good = oracle.match_window(_kfs(g), g["cur_desc"], win, ratio)
            first = max(0, NKF - win)                  # GetNKeyFrames: the last `win`, oldest first; all of them if fewer exist
            assert len(knn) == len(good) == NKF - first
            for i in range(len(knn)):
                assert np.array_equal(knn[i], g["kf%d_knn" % (first + i)])
                assert np.array_equal(good[i], g["kf%d_good%s" % (first + i, tag)])


def test_window_oracle_empty_map_and_empty_frames(oracle):
    g = golden("window_cfg4")
    knn, good = oracle.match_window([], g["cur_desc"], 5)
    assert knn == [] and good == []
    knn, good = oracle.match_window([g["kf0_desc"][:0], g["kf1_desc"]], g["cur_desc"], 5)
    assert len(knn[0]) == 0 and len(good[0]) == 0 and np.array_equal(good[1], g["kf1_good060"])
    for n in (0, 1):                                   # fewer than 2 current keypoints: no row can pass (:46 needs two neighbours)
        knn, good = oracle.match_window(_kfs(g)[:2], g["cur_desc"][:n], 5, 0.95)
        assert all(len(x) == 0 for x in good)

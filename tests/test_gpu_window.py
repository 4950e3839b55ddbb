"""GPU parity (-m gpu) of the keyframe-window matcher (kfs_frontend_set_window / kfs_frontend_match_window, SURVEY §8f row 3:
Map::GetNKeyFrames, /root/reference/src/lib/Map.cpp:72-78, x knnMatch VisualOdometry.cpp:39 x ratio test :46) through the
C ABI, against cv2 goldens (tests/golden/window_cfg4.npz) and the oracle.  Integer work: the bar is bit-exact."""
import numpy as np
import pytest

from _track_util import make_depth, make_rgb
from _util import golden

pytestmark = pytest.mark.gpu
NKF = 7
K = np.array([[525.3, 0, 319.7], [0, 524.9, 239.2], [0, 0, 1]])


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


def _push_all(kfs, fe, g, depth):
    for i in range(NKF):
        d = g["kf%d_desc" % i]
        fe.setKeyframe(d, np.zeros(len(d), kfs.KP_DTYPE), depth)


def test_window_matches_cv2_golden(kfs):
    g = golden("window_cfg4")
    depth = make_depth(480, 640, 3)
    fe = kfs.TrackerFrontEnd(nfeatures=300)
    assert fe.window == 1 and fe.windowCount == 0
    assert fe.matchWindow(g["cur_desc"]) == []            # empty map: no keyframes, no error
    for win in (1, 3, 5, 7, 10):
        fe.setWindow(win)
        assert fe.windowCount == 0
        _push_all(kfs, fe, g, depth)
        first = max(0, NKF - win)
        assert fe.windowCount == NKF - first
        for tag, ratio in (("060", 0.6), ("075", 0.75)):
            good, knn = fe.matchWindow(g["cur_desc"], ratio, want_knn=True)
            assert len(good) == NKF - first
            for i in range(len(good)):
                assert np.array_equal(knn[i], g["kf%d_knn" % (first + i)])
                assert np.array_equal(good[i], g["kf%d_good%s" % (first + i, tag)])
        assert fe.last_launches == 2                      # one kNN launch for the whole window + one ratio launch


def test_window_fills_up_like_get_n_keyframes(kfs):
    """While fewer than `win` keyframes exist GetNKeyFrames returns all of them (Map.cpp:77)."""
    g = golden("window_cfg4")
    depth = make_depth(480, 640, 3)
    fe = kfs.TrackerFrontEnd(nfeatures=300)
    fe.setWindow(5)
    for i in range(NKF):
        d = g["kf%d_desc" % i]
        fe.setKeyframe(d, np.zeros(len(d), kfs.KP_DTYPE), depth)
        good = fe.matchWindow(g["cur_desc"])
        first = max(0, i + 1 - 5)
        assert len(good) == i + 1 - first
        for j, m in enumerate(good):
            assert np.array_equal(m, g["kf%d_good060" % (first + j)])


def test_window_device_resident_track(kfs, oracle):
    """extract -> promoteKeyframe x4 (window 3) -> extract(current) -> matchWindow() on the device-resident descriptors;
    matchFeatures keeps using the newest keyframe (Map::GetLastKeyFrame)."""
    fe = kfs.TrackerFrontEnd(nfeatures=700)
    fe.setWindow(3)
    P = oracle.orb_params(nfeatures=700)
    descs, kps, depths = [], [], []
    for s in range(4):
        kp, desc = fe.extractFeatures(make_rgb(480, 640, 41, shift=s))
        rk, rd = oracle.detect_and_compute(oracle.rgb2gray(make_rgb(480, 640, 41, shift=s)), None, P)
        assert np.array_equal(kp, rk) and np.array_equal(desc, rd)
        depths.append(make_depth(480, 640, 50 + s))
        fe.promoteKeyframe(depths[-1])
        descs.append(desc); kps.append(kp)
    ckp, cdesc = fe.extractFeatures(make_rgb(480, 640, 41, shift=6))
    good, knn = fe.matchWindow(None, 0.6, want_knn=True)
    rknn, rgood = oracle.match_window(descs, cdesc, 3, 0.6)
    assert len(good) == 3 and sum(len(x) for x in rgood) > 200
    for i in range(3):
        assert np.array_equal(knn[i], rknn[i]) and np.array_equal(good[i], rgood[i])
    got = fe.matchFeatures(K)
    k2, _ = oracle.knn_match(descs[-1], cdesc, 2)
    ref = oracle.match_features(k2, kps[-1], ckp, depths[-1], K, 0.6)
    assert len(ref[2]) > 50 and all(np.array_equal(a, b) for a, b in zip(got, ref))


def test_window_config4_shape_and_edges(kfs, oracle):
    """BASELINE.json config 4: one frame (2000 descriptors) against 10 keyframes of 2000 = 2000 x 20000 compares, tie-rich
    random descriptors; ragged and empty keyframes; capacity status; bad window sizes."""
    rng = np.random.default_rng(7)
    base = rng.integers(0, 256, (2000, 32), dtype=np.uint8)
    depth = make_depth(64, 64, 1)

    def noisy(n):
        d = base[rng.permutation(2000)[:n]].copy()
        flip = rng.random(d.shape) < 0.08
        d[flip] ^= rng.integers(1, 256, int(flip.sum()), dtype=np.uint8)
        return d

    fe = kfs.TrackerFrontEnd(nfeatures=2000)
    fe.setWindow(10)
    kfd = [noisy(n) for n in (2000, 2000, 1999, 0, 1, 2000, 130, 2000, 2000, 2000, 2000, 777)]   # 12 pushed, 10 kept
    for d in kfd:
        fe.setKeyframe(d, np.zeros(len(d), kfs.KP_DTYPE), depth)
    cur = noisy(2000)
    for ratio in (0.6, 0.75):
        good, knn = fe.matchWindow(cur, ratio, want_knn=True)
        rknn, rgood = oracle.match_window(kfd, cur, 10, ratio)
        assert len(good) == 10 and sum(len(x) for x in rgood) > 5000
        for i in range(10):
            assert np.array_equal(knn[i], rknn[i]) and np.array_equal(good[i], rgood[i])
    for n in (0, 1):                                       # fewer than 2 current rows: nothing can pass
        assert all(len(x) == 0 for x in fe.matchWindow(cur[:n], 0.95))
    # output capacity smaller than the survivors / rows -> KFS_ERR_CAPACITY
    import ctypes as C
    L = kfs.lib()
    small = np.zeros((10, 8), kfs.DM_DTYPE)
    ng = (C.c_int * 64)()
    nk = C.c_int()
    st = L.kfs_frontend_match_window(fe._h, cur.ctypes.data, len(cur), 0.6, None, None, 0, small.ctypes.data, ng, 8, C.byref(nk))
    assert st == 3 and nk.value == 10 and ng[0] > 8   # counts are exact, rows truncated
    for bad in (0, -1, 65):
        with pytest.raises(kfs.KfsError):
            fe.setWindow(bad)


def test_window_config5_full_size(kfs, oracle):
    """BASELINE.json config 5 at full size on one GPU: a 3840x2160 frame, nfeatures = 10000, k = 2 + ratio 0.75 against a
    32-keyframe window = 32 pairs of up to 10000 x 10000 (OpenCV forbids one 320000-row train matrix).  The current frame is
    extracted on the device and matched where it lies; every pair is checked against the oracle, bit for bit."""
    synth = __import__("importlib").import_module("keyframe-based-slam_b200.synth")
    frame = synth.make_stream(1, 2160, 3840, seed=2, pad=32)[0]
    fe = kfs.TrackerFrontEnd(nfeatures=10000)
    fe.setWindow(32)
    kp, desc = fe.extractFeatures(frame)
    assert len(kp) >= 9900
    rng = np.random.default_rng(11)
    depth = make_depth(32, 32, 2)
    kfd = []
    for i in range(34):                                    # 34 pushed, the last 32 stay
        n = len(desc) if i % 5 else int(rng.integers(1, len(desc)))
        d = desc[rng.permutation(len(desc))[:n]].copy()
        flip = rng.random(d.shape) < 0.05
        d[flip] ^= rng.integers(1, 256, int(flip.sum()), dtype=np.uint8)
        kfd.append(d)
        fe.setKeyframe(d, np.zeros(n, kfs.KP_DTYPE), depth)
    assert fe.windowCount == 32
    good, knn = fe.matchWindow(None, 0.75, want_knn=True)  # device-resident current frame
    assert fe.last_launches == 2
    rknn, rgood = oracle.match_window(kfd, desc, 32, 0.75)
    assert len(good) == 32 and sum(len(x) for x in rgood) > 100000
    for i in range(32):
        assert np.array_equal(knn[i], rknn[i]), "kNN rows of window keyframe %d differ" % i
        assert np.array_equal(good[i], rgood[i]), "ratio survivors of window keyframe %d differ" % i

"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI (include/kfslam_frontend.h) via the
cv2-shaped host mirror, against (1) golden vectors produced by the real reference implementation (cv2 4.13,
tools/make_golden.py) and (2) the oracle (oracle/*.c) on seeded inputs.

Bars (BASELINE.json north_star): pyramid pixels and kNN idx/dist bit-exact; keypoint recall >= 99 % within
0.5 px, orientation within 1e-3 rad, identical descriptors on matched keypoints.  What is asserted here is
stricter wherever the implementation is in fact exact: identical keypoint sets, bit-identical fields,
angles and descriptors."""
import ctypes as C
import os

import numpy as np
import pytest

from _util import ORB_GOLDENS, ROOT, compare_keypoints, golden, golden_input, sha, synth

pytestmark = pytest.mark.gpu

ANGLE_TOL_RAD = 1e-3   # north_star tolerance
RECALL_MIN = 0.99      # north_star tolerance (keypoints compared on integer level coordinates => 0 px error)


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


# ---------------------------------------------------------------- ORB vs cv2 goldens ---------------
@pytest.mark.parametrize("name", ORB_GOLDENS)
def test_orb_matches_cv2_golden(kfs, oracle, name):
    g = golden(name)
    meta = g["meta"]
    gray, mask = golden_input(name)
    orb = kfs.ORB_create(meta["nfeatures"], 1.2, 8, 31, 0, 2, kfs.ORB_HARRIS_SCORE, 31, 20)
    kp, desc = orb.detectAndCompute(gray, mask)
    if desc is None:
        desc = np.zeros((0, 32), np.uint8)
    # a1.1 pyramid pixels: bit-exact on every level
    for l in range(8):
        w, h, quota, scale = orb.level_info(l)
        assert (w, h) == tuple(oracle.level_sizes(meta["w"], meta["h"])[l])
        assert sha(orb.download_level(0, l)) == meta["pyr_sha"][l], "pyramid level %d differs" % l
    # a1.3/a1.4 FAST corners after NMS + border (+mask): identical set and scores
    opyr = oracle.pyramid(gray)
    omask = None
    if mask is not None:
        omask = [np.where(mask != 0, 255, 0).astype(np.uint8)]
        for l in range(1, 8):
            m = oracle.resize_linear_exact(omask[-1], opyr[l].shape[1], opyr[l].shape[0])
            omask.append(np.where(m == 255, 255, 0).astype(np.uint8))
    for l in range(8):
        w, h, _, _ = orb.level_info(l)
        got = orb.download_fast_corners(0, l)
        if w <= 62 or h <= 62:
            assert len(got) == 0
            continue
        ref = oracle.fast9_detect(opyr[l], 20)
        keep = (ref[:, 0] >= 31) & (ref[:, 0] < w - 31) & (ref[:, 1] >= 31) & (ref[:, 1] < h - 31)
        if omask is not None:
            keep &= omask[l][ref[:, 1], ref[:, 0]] != 0
        ref = ref[keep]
        ref = ref[np.lexsort((ref[:, 0], ref[:, 1]))]
        assert np.array_equal(got, ref), "FAST corner set differs on level %d" % l
    # keypoints / descriptors
    r = compare_keypoints(g["kp"], g["desc"], kp, desc)
    assert r["recall"] >= RECALL_MIN and r["n"] == r["n_ref"], r
    assert r["max_angle_err_rad"] <= ANGLE_TOL_RAD, r
    assert r["desc_equal"] == r["n_common"], r
    # stricter: the implementation is exact and its output order is (octave, y, x) like the golden's
    assert np.array_equal(kp, g["kp"]), r
    assert np.array_equal(desc, g["desc"])


def test_orb_empty_and_tiny_inputs(kfs):
    orb = kfs.ORB_create(500)
    n = (C.c_int * 1)(7)
    st = kfs.lib().kfs_orb_detect_and_compute(orb._h, None, 0, 0, 0, None, 0, None, None, 0, n)
    assert st == 0 and n[0] == 0          # empty image: OpenCV returns early with no keypoints
    kp, desc = orb.detectAndCompute(np.full((40, 50), 128, np.uint8))
    assert len(kp) == 0 and desc is None
    kp, desc = orb.detectAndCompute(np.full((480, 640), 77, np.uint8))  # flat image: no corners
    assert len(kp) == 0 and desc is None


def test_orb_batch_equals_single_and_is_deterministic(kfs):
    frames = synth.make_stream(5, 240, 320, seed=9)
    orb = kfs.ORB_create(400)
    kps, descs = orb.detectAndComputeBatch(frames)
    kps2, descs2 = orb.detectAndComputeBatch(frames)
    single = kfs.ORB_create(400)
    for f in range(5):
        assert np.array_equal(kps[f], kps2[f]) and np.array_equal(descs[f], descs2[f])
        k1, d1 = single.detectAndCompute(frames[f])
        assert np.array_equal(kps[f], k1) and np.array_equal(descs[f], d1)
    assert len(kps[0]) > 100


def test_orb_batch_with_masks_vs_oracle(kfs, oracle):
    frames = synth.make_stream(3, 200, 333, seed=4)
    masks = np.zeros_like(frames)
    masks[0, 40:160, 40:300] = 255
    masks[1] = 1
    masks[2, :, :150] = 9
    orb = kfs.ORB_create(300)
    kps, descs = orb.detectAndComputeBatch(frames, masks)
    for f in range(3):
        rk, rd = oracle.detect_and_compute(frames[f], masks[f], oracle.orb_params(nfeatures=300))
        assert np.array_equal(kps[f], rk) and np.array_equal(descs[f], rd)


def test_orb_capacity_status(kfs):
    gray, _ = golden_input("cfg1_480p_a")
    orb = kfs.ORB_create(1000)
    cap = 100
    kp = np.zeros(cap, kfs.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8); n = (C.c_int * 1)()
    st = kfs.lib().kfs_orb_detect_and_compute(orb._h, gray.ctypes.data, 640, 640, 480, None, 0, kp.ctypes.data, desc.ctypes.data, cap, n)
    assert st == 3 and n[0] == 1000       # KFS_ERR_CAPACITY, n_out = required count
    assert np.array_equal(kp, golden("cfg1_480p_a")["kp"][:cap])   # truncated prefix is still correct


def test_orb_strided_input(kfs, oracle):
    gray, _ = golden_input("odd_333x257")
    big = np.zeros((257, 400), np.uint8); big[:, :333] = gray
    orb = kfs.ORB_create(700)
    cap = orb.capacity
    kp = np.zeros(cap, kfs.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8); n = (C.c_int * 1)()
    st = kfs.lib().kfs_orb_detect_and_compute(orb._h, big.ctypes.data, 400, 333, 257, None, 0, kp.ctypes.data, desc.ctypes.data, cap, n)
    assert st == 0
    g = golden("odd_333x257")
    assert np.array_equal(kp[:n[0]], g["kp"]) and np.array_equal(desc[:n[0]], g["desc"])


def test_orb_rejects_unsupported(kfs):
    with pytest.raises(kfs.KfsError) as e:
        kfs.ORB_create(500, 1.2, 8, 31, 0, 4)     # WTA_K = 4
    assert e.value.status == 2
    with pytest.raises(kfs.KfsError) as e:
        kfs.ORB_create(500, 1.2, 8, 31, 0, 2, 1)  # FAST_SCORE
    assert e.value.status == 2
    with pytest.raises(kfs.KfsError) as e:
        kfs.BFMatcher_create(4)                   # NORM_L2
    assert e.value.status == 2


# ---------------------------------------------------------------- kNN -----------------------------
def _check_knn(kfs, oracle, q, t, k=2):
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    m, c = bf.knnMatch(q, t, k)
    rm, rc = oracle.knn_match(q, t, k)
    assert np.array_equal(c, rc)
    kk = min(k, len(t))
    for f in ("queryIdx", "trainIdx", "imgIdx", "distance"):
        assert np.array_equal(m[f][:, :kk], rm[f][:, :kk]), f
    return m, c


def test_knn_matches_cv2_golden(kfs):
    z = golden("knn_cases")
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    for case in ("rand", "ties", "nt1", "nt2", "nq1"):
        q, t, idx, dist = z[case + "_q"], z[case + "_t"], z[case + "_idx"], z[case + "_dist"]
        m, c = bf.knnMatch(q, t, 2)
        kk = min(2, len(t))
        assert (c == kk).all()
        assert np.array_equal(m["trainIdx"][:, :kk], idx[:, :kk]), case
        assert np.array_equal(m["distance"][:, :kk], dist[:, :kk]), case
        assert (m["imgIdx"][:, :kk] == 0).all() and (m["queryIdx"][:, 0] == np.arange(len(q))).all()
        if kk < 2:
            assert (m["trainIdx"][:, kk:] == -1).all()


def test_knn_cfg1_pair_golden_and_ratio(kfs):
    a, b, z = golden("cfg1_480p_a"), golden("cfg1_480p_b"), golden("cfg1_knn")
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    m, c = bf.knnMatch(a["desc"], b["desc"], 2)   # query = keyframe, train = current (VisualOdometry.cpp:39)
    assert np.array_equal(m["trainIdx"], z["idx"]) and np.array_equal(m["distance"], z["dist"])
    good = bf.knnMatchRatio(a["desc"], b["desc"], 0.6)
    ref = [(i, int(z["idx"][i, 0])) for i in range(len(z["idx"])) if z["dist"][i, 0] < np.float32(0.6) * z["dist"][i, 1]]
    assert [(int(x["queryIdx"]), int(x["trainIdx"])) for x in good] == ref


@pytest.mark.parametrize("nq,nt", [(1, 1), (5, 0), (0, 5), (33, 511), (257, 513), (1000, 1000), (2000, 2000), (3000, 3000), (300, 5000)])
def test_knn_vs_oracle_shapes(kfs, oracle, nq, nt):
    q = synth.random_descriptors(nq, seed=nq + 1)
    t = synth.random_descriptors(nt, seed=nt + 7)
    _check_knn(kfs, oracle, q, t, 2)
    if nt:
        _check_knn(kfs, oracle, q, t, 1)


def test_knn_tie_rich_real_descriptors(kfs, oracle):
    g = golden("ties_periodic")
    m, c = _check_knn(kfs, oracle, g["desc"], g["desc"][::-1].copy(), 2)
    assert (m["distance"][:, 0] == 0).all()


def test_knn_window_pairs_and_limits(kfs, oracle):
    q = synth.random_descriptors(700, seed=1)
    trains = [synth.random_descriptors(n, seed=10 + n) for n in (640, 1, 0, 1300)]
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    res = bf.knnMatchPairs([(q, t) for t in trains], 2)
    for (m, c), t in zip(res, trains):
        rm, rc = oracle.knn_match(q, t, 2)
        kk = min(2, len(t))
        assert np.array_equal(c, rc)
        assert np.array_equal(m["trainIdx"][:, :kk], rm["trainIdx"][:, :kk]) and np.array_equal(m["distance"][:, :kk], rm["distance"][:, :kk])
    with pytest.raises(kfs.KfsError):
        bf.knnMatch(q, q, 3)
    big = np.zeros((1 << 18, 32), np.uint8)
    with pytest.raises(kfs.KfsError):
        bf.knnMatch(q[:4], big, 2)            # OpenCV asserts train rows < 2^18 too


def test_knn_full_size_properties(kfs):
    """BASELINE config 5 shape per pair (10000 x 10000): size-independent properties instead of the oracle."""
    n = 10000
    t = synth.random_descriptors(n, seed=99)
    perm = np.random.default_rng(5).permutation(n)
    q = t[perm].copy()
    q[:, 0] ^= 1                                 # every query is 1 bit away from exactly one train row
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    m, c = bf.knnMatch(q, t, 2)
    assert (c == 2).all()
    assert np.array_equal(m["trainIdx"][:, 0], perm) and (m["distance"][:, 0] == 1).all()
    assert (m["distance"][:, 1] >= m["distance"][:, 0]).all() and (m["trainIdx"][:, 1] != m["trainIdx"][:, 0]).all()
    # second-best distances are true Hamming distances of the reported rows
    rows = np.arange(0, n, 97)
    x = np.bitwise_xor(q[rows], t[m["trainIdx"][rows, 1]])
    assert np.array_equal(np.unpackbits(x, axis=1).sum(1).astype(np.float32), m["distance"][rows, 1])


# ---------------------------------------------------------------- end to end ----------------------
def test_end_to_end_720p_stream_vs_oracle(kfs, oracle):
    """BASELINE config 2 shape: 1280x720, nf=2000, extract + kNN vs keyframe, device-produced descriptors."""
    frames = synth.make_stream(3, 720, 1280, seed=21)
    orb = kfs.ORB_create(2000)
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    kps, descs = orb.detectAndComputeBatch(frames)
    P = oracle.orb_params(nfeatures=2000)
    for f in range(3):
        rk, rd = oracle.detect_and_compute(frames[f], None, P)
        assert np.array_equal(kps[f], rk) and np.array_equal(descs[f], rd)
    m, c = bf.knnMatch(descs[0], descs[2], 2)
    rm, rc = oracle.knn_match(descs[0], descs[2], 2)
    assert np.array_equal(m, rm) and np.array_equal(c, rc)
    assert len(oracle.ratio_test(rm, rc, 0.6)) > 200


def test_4k_frame_vs_oracle(kfs, oracle):
    """BASELINE config 5 frame shape: 3840x2160, nf=10000."""
    frame = synth.make_stream(1, 2160, 3840, seed=2, pad=32)[0]
    orb = kfs.ORB_create(10000)
    kp, desc = orb.detectAndCompute(frame)
    rk, rd = oracle.detect_and_compute(frame, None, oracle.orb_params(nfeatures=10000))
    r = compare_keypoints(rk, rd, kp, desc)
    assert r["recall"] >= RECALL_MIN and r["max_angle_err_rad"] <= ANGLE_TOL_RAD and r["desc_equal"] == r["n_common"], r
    assert np.array_equal(kp, rk) and np.array_equal(desc, rd)
    assert len(kp) >= 9900


def test_orb_graph_replay_same_handle_many_calls(kfs, oracle):
    """The kernel chain is replayed as a CUDA graph from the third call of a launch configuration on (kfs_orb_run): the
    results must keep following the DATA (new frames, new masks), survive a mask appearing / disappearing, a batch and a size
    change in between (arena re-allocation drops the graphs), and stay bit-exact against the oracle throughout."""
    orb = kfs.ORB_create(400)
    P = oracle.orb_params(nfeatures=400)
    frames = synth.make_stream(9, 240, 320, seed=77)
    mask = np.zeros((240, 320), np.uint8)
    mask[30:200, 40:280] = 255
    assert orb.last_launches == 0

    def check(img, m=None):
        kp, desc = orb.detectAndCompute(img, m)
        rk, rd = oracle.detect_and_compute(img, m, P)
        assert len(rk) > 50 and np.array_equal(kp, rk) and np.array_equal(desc, rd)
    for f in range(5):                       # eager, capture, replay, replay, replay
        check(frames[f])
    n_plain = orb.last_launches
    assert n_plain >= 7                      # a replayed graph still reports the kernels it launched (7 with the one-launch pyramid)
    for f in range(4):                       # the masked chain is another graph
        check(frames[f], mask)
    assert orb.last_launches > n_plain
    check(frames[5])                         # back to the unmasked graph
    assert orb.last_launches == n_plain
    kps, descs = orb.detectAndComputeBatch(frames[:3])          # batch change: chunked pipeline, different frame counts
    for f in range(3):
        rk, rd = oracle.detect_and_compute(frames[f], None, P)
        assert np.array_equal(kps[f], rk) and np.array_equal(descs[f], rd)
    big = synth.make_stream(4, 300, 400, seed=78)               # size change: arena re-allocated, graphs dropped
    for f in range(4):
        check(big[f])
    for f in range(6, 9):                    # and the first shape again
        check(frames[f])

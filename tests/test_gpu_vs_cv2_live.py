"""GPU vs the reference implementation run LIVE on the same box (-m gpu; skipped when cv2 is not importable): cv2's
ORB::detectAndCompute / BFMatcher::knnMatch — the OpenCV code behind VisualOdometry.cpp:30,39 — called directly on inputs at
the sizes where the committed goldens stop (they cover <= 720p and <= 1000 x 1000 rows; above that the GPU was checked against
the oracle only).  Bit-exact on every field, like the golden tests.

The per-level blur of OpenCV follows the CPU's FMA dispatch; tests/test_blur_dispatch.py's probe tells which
variant this host's cv2 uses and the handle is set accordingly."""
import numpy as np
import pytest

from _track_util import make_refset_mask, make_rgb
from _util import synth

cv2 = pytest.importorskip("cv2")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


@pytest.fixture(scope="module")
def host_fma(oracle):
    """True when this host's cv2 blurs with fused multiply-adds (the default AVX2 / AVX-512 dispatch), False for multiply + add."""
    img = synth.make_stream(1, 720, 1280, seed=100)[0]
    k = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
    live = cv2.sepFilter2D(img, -1, k, k, borderType=cv2.BORDER_REFLECT_101)
    if np.array_equal(live, oracle.blur7(img, use_fma=True)):
        return True
    assert np.array_equal(live, oracle.blur7(img, use_fma=False)), "the local cv2 matches neither blur variant"
    return False


def cv_orb(gray, mask, nf):
    orb = cv2.ORB_create(nf, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
    kps, desc = orb.detectAndCompute(gray, mask)
    import kfslam_b200 as k
    arr = np.array([(p.pt[0], p.pt[1], p.size, p.angle, p.response, p.octave, p.class_id) for p in kps], dtype=k.KP_DTYPE)
    order = np.lexsort((arr["x"], arr["y"], arr["octave"]))
    return arr[order], desc[order]


def cv_knn(q, t):
    import kfslam_b200 as k
    rows = cv2.BFMatcher.create(cv2.NORM_HAMMING).knnMatch(q, t, 2)
    return np.array([[(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in r] for r in rows], dtype=k.DM_DTYPE)


def test_live_720p_masked_nf3000_rgb(kfs, host_fma):
    """The reference's operating point on a frame no golden holds (another seed)."""
    rgb, mask = make_rgb(720, 1280, 52, shift=2), make_refset_mask(720, 1280, 53)
    gray = cv2.cvtColor(cv2.cvtColor(rgb, cv2.COLOR_RGB2BGR), cv2.COLOR_BGR2GRAY)
    rk, rd = cv_orb(gray, mask, 3000)
    fe = kfs.TrackerFrontEnd(nfeatures=3000)
    fe.set_blur_fma(host_fma)          # the handle follows this host's cv2 (round 2 skipped the leg on a non-FMA host)
    kp, desc = fe.extractFeatures(rgb, mask)
    assert len(kp) == len(rk) and np.array_equal(kp, rk) and np.array_equal(desc, rd)


def test_live_4k_nf10000(kfs, host_fma):
    """Config 5's frame: 3840x2160, nfeatures = 10000."""
    f = synth.make_stream(1, 2160, 3840, seed=91)[0]
    rk, rd = cv_orb(f, None, 10000)
    orb = kfs.ORB_create(10000)
    orb.set_blur_fma(host_fma)
    kp, desc = orb.detectAndCompute(f)
    assert len(kp) == len(rk) > 9000 and np.array_equal(kp, rk) and np.array_equal(desc, rd)


@pytest.mark.parametrize("nq,nt,seed", [(2000, 2000, 1), (3000, 3000, 2), (10000, 10000, 3), (2000, 20000, 4)])
def test_live_knn_large(kfs, nq, nt, seed):
    """kNN shapes of configs 2, 4 and 5 on tie-rich random descriptors (low-entropy bytes make equal distances common)."""
    rng = np.random.default_rng(seed)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q[:, 8:] &= 0x11; t[:, 8:] &= 0x11          # few live bits -> many exact distance ties: "lower trainIdx wins" is exercised
    t[nt // 2:nt // 2 + nq // 4] = q[:nq // 4]  # exact duplicates (distance 0)
    ref = cv_knn(q, t)
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    m, c = bf.knnMatch(q, t, 2)
    assert (c == 2).all() and np.array_equal(m, ref)


def test_live_window_cfg4_shape(kfs, host_fma):
    """Current frame vs a window of keyframes (Map::GetNKeyFrames, Map.cpp:72-78): survivors of every pair equal cv2's."""
    frames = synth.make_stream(7, 720, 1280, seed=61)
    fe = kfs.TrackerFrontEnd(nfeatures=2000)
    fe.setWindow(5)
    depth = np.full((720, 1280), 1000, np.uint16)
    descs = []
    fe.set_blur_fma(host_fma)
    for f in frames[:6]:
        _, d = fe.extractFeatures(f)
        fe.promoteKeyframe(depth)
        descs.append(cv_orb(f, None, 2000)[1])
        assert np.array_equal(d, descs[-1])
    _, cur = fe.extractFeatures(frames[6])
    good, knn = fe.matchWindow(ratio=0.75, want_knn=True)
    assert len(good) == 5
    for i, kf in enumerate(descs[1:6]):           # the ring holds the last 5 of the 6 keyframes, oldest first
        ref = cv_knn(kf, cur)
        assert np.array_equal(knn[i], ref)
        keep = ref["distance"][:, 0] < np.float32(0.75) * ref["distance"][:, 1]
        assert np.array_equal(good[i], ref[keep, 0])

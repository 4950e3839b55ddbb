"""CPU: the oracle re-checked LIVE against the reference implementation wherever cv2 is importable (it is in the build container
and on the GPU box): fresh seeds, so this is not a replay of the committed goldens.  ORB::detectAndCompute and
BFMatcher::knnMatch are the OpenCV calls behind VisualOdometry.cpp:30 and :39."""
import numpy as np
import pytest

from _track_util import make_refset_mask
from _util import synth

cv2 = pytest.importorskip("cv2")

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def cv_orb(gray, mask, nf):
    kps, desc = cv2.ORB_create(nf, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20).detectAndCompute(gray, mask)
    arr = np.array([(p.pt[0], p.pt[1], p.size, p.angle, p.response, p.octave, p.class_id) for p in kps], dtype=KP)
    order = np.lexsort((arr["x"], arr["y"], arr["octave"]))
    return arr[order], desc[order]


def host_fma(oracle):
    img = synth.make_stream(1, 240, 320, seed=100)[0]
    k = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
    live = cv2.sepFilter2D(img, -1, k, k, borderType=cv2.BORDER_REFLECT_101)
    return bool(np.array_equal(live, oracle.blur7(img, use_fma=True)))


@pytest.mark.parametrize("h,w,nf,seed,masked", [(720, 1280, 3000, 201, True), (480, 640, 1000, 202, False), (600, 801, 1500, 203, True)])
def test_oracle_orb_equals_live_cv2(oracle, h, w, nf, seed, masked):
    f = synth.make_stream(1, h, w, seed=seed)[0]
    mask = make_refset_mask(h, w, seed) if masked else None
    rk, rd = cv_orb(f, mask, nf)
    kp, desc = oracle.detect_and_compute(f, mask, oracle.orb_params(nfeatures=nf), blur_fma=host_fma(oracle))
    assert np.array_equal(kp, rk) and np.array_equal(desc, rd)


def test_oracle_knn_equals_live_cv2(oracle):
    rng = np.random.default_rng(7)
    q = rng.integers(0, 256, (1500, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (2100, 32), dtype=np.uint8)
    q[:, 6:] &= 0x21; t[:, 6:] &= 0x21
    rows = cv2.BFMatcher.create(cv2.NORM_HAMMING).knnMatch(q, t, 2)
    m, c = oracle.knn_match(q, t, 2)
    assert (c == 2).all()
    assert np.array_equal(m["trainIdx"], np.array([[d.trainIdx for d in r] for r in rows], np.int32))
    assert np.array_equal(m["distance"], np.array([[d.distance for d in r] for r in rows], np.float32))

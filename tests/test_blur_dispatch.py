"""the per-level 7x7 blur follows OpenCV's CPU dispatch in its last bit (FMA kernels by default, multiply + add with
cv2.setUseOptimized(False) or on a host without FMA).  Both variants are built; cv2 digests of both
(tests/golden/blur_dispatch.npz, tools/make_golden_blur.py) pin the oracle on CPU, and the GPU kernel (kfs_orb_set_blur_fma) is
compared with the oracle for both.  The test prints which variant the cv2 on this machine matches."""
import hashlib

import numpy as np
import pytest

from _util import golden, synth


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_oracle_blur_matches_cv2_under_both_dispatches(oracle):
    g = golden("blur_dispatch")
    img = synth.make_stream(1, 720, 1280, seed=100)[0]
    fma, mad = oracle.blur7(img, use_fma=True), oracle.blur7(img, use_fma=False)
    assert _sha(fma) == g["meta"]["sha_fma"] and _sha(mad) == g["meta"]["sha_muladd"]
    d = g["diff_yx"]
    assert len(d) == g["meta"]["n_diff"] > 0 and int((fma != mad).sum()) == len(d)
    assert np.array_equal(fma[d[:, 0], d[:, 1]], g["diff_fma"]) and np.array_equal(mad[d[:, 0], d[:, 1]], g["diff_muladd"])


@pytest.mark.gpu
def test_gpu_blur_both_dispatches(oracle, capsys):
    import kfslam_b200 as kfs
    g = golden("blur_dispatch")
    img = synth.make_stream(1, 720, 1280, seed=100)[0]
    orb = kfs.ORB_create(2000)
    P = oracle.orb_params(nfeatures=2000)
    got = {}
    for name, fma in (("fma", True), ("muladd", False), ("fma", True)):
        orb.set_blur_fma(fma)
        for rep in range(3):                               # eager, capture, graph replay: the mode must survive all three
            kp, desc = orb.detectAndCompute(img)
        rk, rd = oracle.detect_and_compute(img, None, P, blur_fma=fma)
        assert np.array_equal(kp, rk) and np.array_equal(desc, rd), name
        pyr = oracle.pyramid(img)
        for lvl in (0, 3, 7):
            b = orb.download_blurred_level(0, lvl)
            ref = oracle.blur7(pyr[lvl], use_fma=fma)
            assert np.array_equal(b[12:-12, 12:-12], ref[12:-12, 12:-12]), (name, lvl)
            if lvl == 0:
                got[name] = b
    d = g["diff_yx"]
    inside = (d[:, 0] >= 12) & (d[:, 0] < 720 - 12) & (d[:, 1] >= 12) & (d[:, 1] < 1280 - 12)
    d = d[inside]
    assert len(d) > 0                                      # the two dispatches really differ inside the sampled area
    assert np.array_equal(got["fma"][d[:, 0], d[:, 1]], g["diff_fma"][inside])          # cv2's own pixels at those places
    assert np.array_equal(got["muladd"][d[:, 0], d[:, 1]], g["diff_muladd"][inside])
    assert int((got["fma"][12:-12, 12:-12] != got["muladd"][12:-12, 12:-12]).sum()) == len(d)
    try:                                                   # which variant does the cv2 on THIS machine produce?
        import cv2
        k = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
        live = cv2.sepFilter2D(img, -1, k, k, borderType=cv2.BORDER_REFLECT_101)
        which = [n for n in ("fma", "muladd") if np.array_equal(live[12:-12, 12:-12], got[n][12:-12, 12:-12])]
        with capsys.disabled():
            print("\n[blur dispatch] cv2 %s on this host matches the %s variant" % (cv2.__version__, " and ".join(which) or "NEITHER"))
        assert which, "the local cv2 matches neither blur variant"
    except ImportError:
        pass


@pytest.mark.gpu
def test_gpu_tracker_handle_follows_blur_dispatch(oracle):
    """kfs_frontend_set_blur_fma: the tracker handle's extractor can be switched like a plain kfs_orb, so a tracker built against an
    OpenCV without FMA dispatch keeps bit-identical descriptors."""
    import kfslam_b200 as kfs
    img = synth.make_stream(1, 720, 1280, seed=100)[0]
    P = oracle.orb_params(nfeatures=2000)
    fe = kfs.TrackerFrontEnd(nfeatures=2000)
    for fma in (True, False, True):   # (the 6 pixels that differ between the variants on this frame flip no descriptor bit: the blurred
        fe.set_blur_fma(fma)          # levels themselves are compared in test_gpu_blur_both_dispatches through the kfs_orb handle)
        kp, desc = fe.extractFeatures(img)
        rk, rd = oracle.detect_and_compute(img, None, P, blur_fma=fma)
        assert np.array_equal(kp, rk) and np.array_equal(desc, rd), fma
    fe.close()

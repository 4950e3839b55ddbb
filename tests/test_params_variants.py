"""ORB parameter sets other than the reference's: scaleFactor (incl. exactly 2.0 and > 2 taps), nlevels (1 ... 12),
fastThreshold and a tiny nfeatures are run-time parameters of cv::ORB::create (VisualOdometry.cpp:10-16).  cv2 digests
(tests/golden/params_variants.npz, tools/make_golden_params.py) pin the oracle on CPU; the GPU path is compared with
the oracle bit for bit."""
import hashlib

import numpy as np
import pytest

from _util import golden, synth


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _sorted(kp, desc):
    o = np.lexsort((kp["x"], kp["y"], kp["octave"]))
    return kp[o], desc[o]


def _sets():
    return golden("params_variants")["meta"]["sets"]


def test_oracle_matches_cv2_digests_for_parameter_variants(oracle):
    img = synth.make_pair(480, 640, seed=5)[0]
    for s in _sets():
        nf, sf, nl, thr = s["params"]
        kp, desc = _sorted(*oracle.detect_and_compute(img, None, oracle.orb_params(nfeatures=nf, scale_factor=sf, nlevels=nl, fast_threshold=thr)))
        assert len(kp) == s["n"] and _sha(kp) == s["kp_sha"] and _sha(desc) == s["desc_sha"], s["params"]


@pytest.mark.gpu
def test_gpu_matches_oracle_for_parameter_variants(oracle):
    import kfslam_b200 as kfs
    img = synth.make_pair(480, 640, seed=5)[0]
    for s in _sets():
        nf, sf, nl, thr = s["params"]
        orb = kfs.ORB_create(nf, sf, nl, 31, 0, 2, kfs.ORB_HARRIS_SCORE, 31, thr)
        for rep in range(3):                       # third call = CUDA-graph replay
            kp, desc = _sorted(*orb.detectAndCompute(img))
            assert len(kp) == s["n"] and _sha(kp) == s["kp_sha"] and _sha(desc) == s["desc_sha"], (s["params"], rep)
        orb.close()

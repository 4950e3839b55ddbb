"""CPU: the oracle at the reference's own operating point — cv::ORB::create(3000, ...) (VisualOdometry.cpp:10-16), R,G,B input
converted at :25-28, a mask on every call (:30) — against the cv2 golden tests/golden/refset_720p.npz
(tools/make_golden_refsetting.py).  Pins the oracle for masks at 1280x720, where the other mask goldens stop at 640x480."""
import hashlib

import numpy as np

from _track_util import make_refset_mask, make_rgb
from _util import golden


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def refset_inputs():
    g = golden("refset_720p")
    h, w = g["meta"]["shape"]
    seed = g["meta"]["seed"]
    out = {}
    for name, shift in (("kf", 0), ("cur", 4)):
        rgb, mask = make_rgb(h, w, seed, shift=shift), make_refset_mask(h, w, seed + shift)
        assert sha(rgb) == g["meta"]["sha"][name + "_rgb"] and sha(mask) == g["meta"]["sha"][name + "_mask"], "input generator drifted"
        out[name] = (rgb, mask)
    return g, out


def test_oracle_matches_cv2_at_the_reference_setting(oracle):
    g, inp = refset_inputs()
    P = oracle.orb_params(nfeatures=g["meta"]["nfeatures"])
    for name in ("kf", "cur"):
        rgb, mask = inp[name]
        kp, desc = oracle.detect_and_compute(oracle.rgb2gray(rgb), mask, P)
        assert np.array_equal(kp, g[name + "_kp"]) and np.array_equal(desc, g[name + "_desc"]), name
    m, c = oracle.knn_match(g["kf_desc"], g["cur_desc"], 2)
    assert np.array_equal(m, g["knn"]) and (c == 2).all()

#!/usr/bin/env python3
"""Golden vectors for the tracker wrappers, produced HERE by the real reference implementation
(cv2 4.13 = the OpenCV code behind cv::cvtColor / detectAndCompute / knnMatch at VisualOdometry.cpp:25-30,39) and committed
as tests/golden/track_*.npz.  The inputs are regenerated from seeds by tests/_track_util.py (make_rgb / make_depth)."""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _track_util import make_rgb  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gray_cv2(rgb):
    bgr = cv2.cvtColor(rgb, cv2.COLOR_RGB2BGR)       # VisualOdometry.cpp:25
    return cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)     # :28


out = {}
small = make_rgb(61, 97, seed=11)
out["small_gray"] = gray_cv2(small)
meta = {"cv2": cv2.__version__, "small": [61, 97, 11], "sha": {}}
for (h, w, seed) in [(480, 640, 12), (720, 1280, 13), (333, 257, 14)]:
    meta["sha"]["%dx%d_%d" % (h, w, seed)] = sha(gray_cv2(make_rgb(h, w, seed)))
# full ExtractFeatures on a 480p colour frame pair + knnMatch (the kNN rows MatchFeatures starts from)
orb = cv2.ORB_create(1000, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
bf = cv2.BFMatcher.create(cv2.NORM_HAMMING)
res = []
for shift in (0, 3):  # keyframe, then the same scene 3 px to the right (so that the ratio test has survivors)
    kp, desc = orb.detectAndCompute(gray_cv2(make_rgb(480, 640, 21, shift=shift)), None)
    arr = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave, k.class_id) for k in kp],
                   dtype=[("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
    order = np.lexsort((arr["x"], arr["y"], arr["octave"]))
    res.append((arr[order], desc[order]))
out["kf_kp"], out["kf_desc"] = res[0]
out["cur_kp"], out["cur_desc"] = res[1]
knn = bf.knnMatch(res[0][1], res[1][1], 2)
out["knn"] = np.array([[(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in row] for row in knn],
                      dtype=[("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])
out["meta"] = np.array(json.dumps(meta))
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "track_cfg1.npz"), **out)
print("written", {k: getattr(v, "shape", None) for k, v in out.items()})

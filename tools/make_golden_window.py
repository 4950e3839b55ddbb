#!/usr/bin/env python3
"""Golden vectors for the keyframe-window matcher (BASELINE.json configs 4/5 in miniature), produced HERE
by the real reference implementation (cv2 4.13: detectAndCompute + BFMatcher.knnMatch, the calls at
VisualOdometry.cpp:30,39) and committed as tests/golden/window_cfg4.npz.  Seven 480p keyframes (the same scene shifted by
0..6 px) and one current frame; for every keyframe the k=2 rows of knnMatch(keyframe, current) and the survivors of
d0 < ratio*d1 (:46) for ratio 0.6 (reference) and 0.75 (config 5).  Inputs are regenerated from seeds by
tests/_track_util.py (make_rgb)."""
import json
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _track_util import make_rgb  # noqa: E402

KP = [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]
DM = [("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")]
NF, SEED, NKF, CUR_SHIFT = 300, 33, 7, 8


def gray_cv2(rgb):
    return cv2.cvtColor(cv2.cvtColor(rgb, cv2.COLOR_RGB2BGR), cv2.COLOR_BGR2GRAY)   # VisualOdometry.cpp:25-28


def extract(orb, shift):
    kp, desc = orb.detectAndCompute(gray_cv2(make_rgb(480, 640, SEED, shift=shift)), None)   # :30
    arr = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave, k.class_id) for k in kp], dtype=KP)
    order = np.lexsort((arr["x"], arr["y"], arr["octave"]))
    return arr[order], desc[order]


orb = cv2.ORB_create(NF, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
bf = cv2.BFMatcher.create(cv2.NORM_HAMMING)
out = {"meta": np.array(json.dumps({"cv2": cv2.__version__, "nfeatures": NF, "seed": SEED, "keyframe_shifts": list(range(NKF)),
                                    "cur_shift": CUR_SHIFT, "size": [480, 640]}))}
cur_kp, cur_desc = extract(orb, CUR_SHIFT)
out["cur_desc"] = cur_desc
for i in range(NKF):
    kp, desc = extract(orb, i)
    out["kf%d_desc" % i] = desc
    knn = bf.knnMatch(desc, cur_desc, 2)                                                      # :39
    rows = np.array([[(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in row] for row in knn], dtype=DM)
    out["kf%d_knn" % i] = rows
    for tag, ratio in (("060", 0.6), ("075", 0.75)):
        good = [row[0] for row in knn if row[0].distance < np.float32(ratio) * np.float32(row[1].distance)]   # :46
        out["kf%d_good%s" % (i, tag)] = np.array([(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in good], dtype=DM)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "window_cfg4.npz"), **out)
print("written", {k: getattr(v, "shape", None) for k, v in out.items()})

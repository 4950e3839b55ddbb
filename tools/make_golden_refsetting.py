#!/usr/bin/env python3
"""Golden vectors for the reference's own operating point: cv::ORB::create(3000, ...) (VisualOdometry.cpp:10-16), an R,G,B frame
converted at :25-28, a mask on EVERY call (:30, fed from src/executable/main.cpp:31-34), knnMatch(keyframe, current, 2) (:39).
Produced HERE by cv2 4.13 (the OpenCV code the reference links) and committed as tests/golden/refset_720p.npz; the inputs are
regenerated from seeds by tests/_track_util.py (make_rgb / make_refset_mask)."""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _track_util import make_refset_mask, make_rgb  # noqa: E402

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
DM = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gray_cv2(rgb):
    return cv2.cvtColor(cv2.cvtColor(rgb, cv2.COLOR_RGB2BGR), cv2.COLOR_BGR2GRAY)


H, W, NF = 720, 1280, 3000
orb = cv2.ORB_create(NF, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
bf = cv2.BFMatcher.create(cv2.NORM_HAMMING)
out, meta = {}, {"cv2": cv2.__version__, "shape": [H, W], "nfeatures": NF, "seed": 31, "sha": {}}
res = []
for name, shift in (("kf", 0), ("cur", 4)):
    rgb, mask = make_rgb(H, W, 31, shift=shift), make_refset_mask(H, W, 31 + shift)
    meta["sha"][name + "_rgb"], meta["sha"][name + "_mask"] = sha(rgb), sha(mask)
    kp, desc = orb.detectAndCompute(gray_cv2(rgb), mask)
    arr = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave, k.class_id) for k in kp], dtype=KP)
    order = np.lexsort((arr["x"], arr["y"], arr["octave"]))
    res.append((arr[order], desc[order]))
    out[name + "_kp"], out[name + "_desc"] = res[-1]
knn = bf.knnMatch(res[0][1], res[1][1], 2)
out["knn"] = np.array([[(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in row] for row in knn], dtype=DM)
out["meta"] = np.array(json.dumps(meta))
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "refset_720p.npz"), **out)
print("written", {k: getattr(v, "shape", None) for k, v in out.items()},
      "survivors(0.6):", int((out["knn"]["distance"][:, 0] < np.float32(0.6) * out["knn"]["distance"][:, 1]).sum()))

#!/usr/bin/env python3
"""Golden digests for ORB parameter sets other than the reference's (scaleFactor, nlevels, fastThreshold, nfeatures are
run-time parameters of cv::ORB::create, VisualOdometry.cpp:10-16), produced HERE by cv2 4.13 and committed as
tests/golden/params_variants.npz: per parameter set the keypoint count and the SHA-256 of the keypoints and descriptors
sorted by (octave, y, x), on the seeded 640x480 frame synth.make_pair(480, 640, seed=5)[0]."""
import hashlib
import importlib
import json
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("keyframe-based-slam_b200.synth")
KP = [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]
# (nfeatures, scaleFactor, nlevels, fastThreshold)
SETS = [(800, 1.5, 5, 20), (600, 2.0, 4, 20), (900, 1.1, 12, 20), (700, 1.2, 8, 7), (700, 1.2, 8, 50), (50, 1.2, 8, 20),
        (1000, 1.3, 1, 20), (500, 1.2, 3, 20), (1200, 1.41, 6, 12)]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


img = synth.make_pair(480, 640, seed=5)[0]
out = []
for nf, sf, nl, thr in SETS:
    kp, desc = cv2.ORB_create(nf, sf, nl, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, thr).detectAndCompute(img, None)
    arr = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave, k.class_id) for k in kp], dtype=KP)
    o = np.lexsort((arr["x"], arr["y"], arr["octave"]))
    out.append({"params": [nf, sf, nl, thr], "n": len(kp), "kp_sha": sha(arr[o]), "desc_sha": sha(desc[o])})
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "params_variants.npz"),
                    meta=np.array(json.dumps({"cv2": cv2.__version__, "frame": "synth.make_pair(480, 640, seed=5)[0]", "sets": out})))
print(json.dumps(out, indent=1))

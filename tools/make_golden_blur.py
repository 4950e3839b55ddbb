#!/usr/bin/env python3
"""Golden digests of the per-level 7x7 blur under BOTH OpenCV dispatches, produced HERE by cv2 4.13:
cv2.sepFilter2D(level, -1, g, g, borderType=BORDER_REFLECT_101) with cv2.setUseOptimized(True) (AVX2/AVX-512 FMA kernels,
the default) and with cv2.setUseOptimized(False) (plain C++: multiply, then add).  Committed as tests/golden/blur_dispatch.npz;
input = synth.make_stream(1, 720, 1280, seed=100)[0] (the bench texture)."""
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


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


img = synth.make_stream(1, 720, 1280, seed=100)[0]
g = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
res = {}
for name, opt in (("fma", True), ("muladd", False)):
    cv2.setUseOptimized(opt)
    res[name] = cv2.sepFilter2D(img, -1, g, g, borderType=cv2.BORDER_REFLECT_101)
cv2.setUseOptimized(True)
diff = np.argwhere(res["fma"] != res["muladd"])
meta = {"cv2": cv2.__version__, "frame": "synth.make_stream(1, 720, 1280, seed=100)[0]", "sha_fma": sha(res["fma"]), "sha_muladd": sha(res["muladd"]),
        "n_diff": int(len(diff)), "cpu": cv2.getCPUFeaturesLine() if hasattr(cv2, "getCPUFeaturesLine") else ""}
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "blur_dispatch.npz"), meta=np.array(json.dumps(meta)), diff_yx=diff.astype(np.int32),
                    diff_fma=res["fma"][diff[:, 0], diff[:, 1]], diff_muladd=res["muladd"][diff[:, 0], diff[:, 1]])
print(meta)

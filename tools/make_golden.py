#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the REAL reference implementation of the hot path: OpenCV
(`cv2`, opencv-python-headless 4.13.0.92), i.e. exactly the code behind
/root/reference/src/lib/VisualOdometry.cpp:30 (ORB::detectAndCompute) and :39 (BFMatcher::knnMatch).

Run in the build container (cv2 importable).  The GPU box never needs cv2 or /root/reference: tests
read the committed .npz files.  Inputs are regenerated from seeds by keyframe-based-slam_b200/synth.py;
each golden stores the SHA-256 of its input frame so a drifting generator is caught, not silently used.
"""
import hashlib, importlib, json, os, sys
import numpy as np
import cv2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("keyframe-based-slam_b200.synth")
OUT = os.path.join(ROOT, "tests", "golden")
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
               ("octave", "<i4"), ("class_id", "<i4")])


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def scales(n=8):
    return [np.float32(np.float64(np.float32(1.2)) ** l) for l in range(n)]


def level_sizes(w, h, n=8):
    out = []
    for l in range(n):
        inv = np.float32(1.0) / scales(n)[l]
        out.append((int(np.rint(np.float64(np.float32(w) * inv))), int(np.rint(np.float64(np.float32(h) * inv)))))
    return out


def cv_orb(gray, mask, nf):
    orb = cv2.ORB_create(nf, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
    kps, desc = orb.detectAndCompute(gray, mask)
    arr = np.zeros(len(kps), KP)
    for i, k in enumerate(kps):
        arr[i] = (k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave, k.class_id)
    desc = desc if desc is not None else np.zeros((0, 32), np.uint8)
    sc = scales()
    key = np.array([(int(k["octave"]) << 40) | (int(round(float(k["y"]) / sc[k["octave"]])) << 20) |
                    int(round(float(k["x"]) / sc[k["octave"]])) for k in arr], dtype=np.int64)
    order = np.argsort(key, kind="stable")
    return arr[order], desc[order]


def cv_knn(q, t, k=2):
    m = cv2.BFMatcher.create(cv2.NORM_HAMMING).knnMatch(q, t, k) if len(q) and len(t) else [[] for _ in range(len(q))]
    idx = np.full((len(q), k), -1, np.int32); dist = np.full((len(q), k), -1, np.float32)
    for i, row in enumerate(m):
        for j, d in enumerate(row):
            assert d.queryIdx == i and d.imgIdx == 0
            idx[i, j] = d.trainIdx; dist[i, j] = d.distance
    return idx, dist


def orb_case(name, gray, mask, nf, store_pyramid_pixels=False):
    kp, desc = cv_orb(gray, mask, nf)
    h, w = gray.shape
    pyr_sha, fast_counts, fast_sha = [], [], []
    prev = gray
    pyr_small = {}
    for l, (lw, lh) in enumerate(level_sizes(w, h)):
        if l:
            prev = cv2.resize(prev, (lw, lh), interpolation=cv2.INTER_LINEAR_EXACT)
        pyr_sha.append(sha(prev))
        if store_pyramid_pixels and l in (1, 7):
            pyr_small["pyr%d" % l] = prev
        if lw >= 7 and lh >= 7:
            f = cv2.FastFeatureDetector_create(20, True).detect(prev)
            c = np.array(sorted((int(k.pt[1]), int(k.pt[0]), int(k.response)) for k in f), np.int32).reshape(-1, 3)
        else:
            c = np.zeros((0, 3), np.int32)
        fast_counts.append(len(c)); fast_sha.append(sha(c))
    meta = dict(name=name, w=w, h=h, nfeatures=nf, frame_sha=sha(gray), mask_sha=sha(mask) if mask is not None else "",
                pyr_sha=pyr_sha, fast_counts=fast_counts, fast_sha=fast_sha, n=len(kp), cv2=cv2.__version__)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), kp=kp, desc=desc, meta=json.dumps(meta), **pyr_small)
    print(name, "n =", len(kp), "fast", fast_counts)
    return kp, desc


def main():
    os.makedirs(OUT, exist_ok=True)
    # config 1: 640x480 pair, nf=1000, kNN k=2 (query = keyframe = frame A, train = current = frame B)
    a, b = synth.make_pair(480, 640, seed=3)
    kpa, da = orb_case("cfg1_480p_a", a, None, 1000, store_pyramid_pixels=True)
    kpb, db = orb_case("cfg1_480p_b", b, None, 1000)
    idx, dist = cv_knn(da, db)
    np.savez_compressed(os.path.join(OUT, "cfg1_knn.npz"), idx=idx, dist=dist)
    # masks (reference always passes one: VisualOdometry.cpp:30)
    m = np.zeros((480, 640), np.uint8); m[100:400, 50:600] = 255; m[200:250, 200:300] = 0
    orb_case("cfg1_480p_masked", a, m, 1000)
    m2 = (np.random.default_rng(0).random((480, 640)) > 0.3).astype(np.uint8) * 7
    orb_case("cfg1_480p_mask_nonbinary", a, m2, 1000)
    # ties, tiny / ragged sizes
    orb_case("ties_periodic", synth.periodic_frame(), None, 1000)
    orb_case("small_160x120", np.ascontiguousarray(a[:120, :160]), None, 500)
    orb_case("small_200x90", np.ascontiguousarray(a[:90, :200]), None, 500)
    orb_case("small_64x64", np.ascontiguousarray(a[:64, :64]), None, 500)
    orb_case("odd_333x257", np.ascontiguousarray(a[:257, :333]), None, 700)
    # config 2 shape: 1280x720, nf=2000 and the reference's own nf=3000
    f = synth.make_stream(2, 720, 1280, seed=1)
    orb_case("cfg2_720p_nf2000", f[0], None, 2000)
    orb_case("cfg2_720p_nf3000", f[1], None, 3000)
    # kNN known answers: random, tie-rich, ragged
    rng = np.random.default_rng(11)
    cases = {}
    q = rng.integers(0, 256, (300, 32), dtype=np.uint8); t = rng.integers(0, 256, (517, 32), dtype=np.uint8)
    cases["rand"] = (q, t)
    tq = rng.integers(0, 2, (200, 32), dtype=np.uint8) * 255; tt = np.repeat(tq[:40], 5, 0)[rng.permutation(200)]
    cases["ties"] = (tq, tt)
    cases["nt1"] = (q[:7], t[:1]); cases["nt2"] = (q[:7], t[:2]); cases["nq1"] = (q[:1], t)
    out = {}
    for k, (qq, tt_) in cases.items():
        i, d = cv_knn(qq, tt_)
        out[k + "_q"] = qq; out[k + "_t"] = tt_; out[k + "_idx"] = i; out[k + "_dist"] = d
    np.savez_compressed(os.path.join(OUT, "knn_cases.npz"), **out)
    print("golden written to", OUT)


if __name__ == "__main__":
    main()

"""ctypes front for oracle/libkforacle.so (orb_oracle.c + knn_oracle.c).  TEST INFRASTRUCTURE ONLY.

Restates /root/reference/src/lib/VisualOdometry.cpp:30 (cv::ORB::detectAndCompute) and :39
(cv::BFMatcher::knnMatch) + :46 (ratio test); see the headers of the two C files for the pinning story.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
DM_DTYPE = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])
assert KP_DTYPE.itemsize == 28 and DM_DTYPE.itemsize == 16


class OrbParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int),
                ("edge_threshold", C.c_int), ("first_level", C.c_int), ("wta_k", C.c_int),
                ("score_type", C.c_int), ("patch_size", C.c_int), ("fast_threshold", C.c_int)]


def build(force=False):
    so = os.path.join(_HERE, "libkforacle.so")
    srcs = [os.path.join(_HERE, f) for f in ("orb_oracle.c", "knn_oracle.c", "track_oracle.c", "Makefile")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        u8p = C.POINTER(C.c_uint8)
        L.kfo_level_scale.restype = C.c_float
        L.kfo_level_scale.argtypes = [C.c_float, C.c_int, C.c_int]
        L.kfo_level_size.argtypes = [C.c_int, C.c_int, C.c_float, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.kfo_level_quotas.argtypes = [C.c_int, C.c_float, C.c_int, C.POINTER(C.c_int)]
        L.kfo_resize_linear_exact.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_int, C.c_int, C.c_size_t]
        L.kfo_fast9_detect.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, C.c_int,
                                       C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
        L.kfo_fast9_scoremap.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, u8p, C.c_size_t]
        L.kfo_harris.restype = C.c_float
        L.kfo_harris.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int]
        L.kfo_ic_angle.restype = C.c_float
        L.kfo_ic_angle.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int]
        L.kfo_ic_moments.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.kfo_fast_atan2.restype = C.c_float
        L.kfo_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.kfo_blur7.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_size_t, C.c_int]
        L.kfo_brief.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_float, u8p]
        L.kfo_orb_detect_and_compute_ex.argtypes = [C.POINTER(OrbParams), u8p, C.c_size_t, C.c_int, C.c_int,
                                                    u8p, C.c_size_t, C.c_void_p, u8p, C.c_int,
                                                    C.POINTER(C.c_int), C.POINTER(u8p), C.c_int]
        L.kfo_knn_match.argtypes = [u8p, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int)]
        L.kfo_ratio_test.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_float, C.c_void_p]
        L.kfo_rgb2gray.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, u8p, C.c_size_t]
        L.kfo_match_features.restype = C.c_int
        L.kfo_match_features.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_double), C.c_float,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.kfo_match_window.restype = C.c_int
        L.kfo_match_window.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int, C.c_int, u8p, C.c_int, C.c_float,
                                       C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
        _LIB = L
    return _LIB


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def orb_params(nfeatures=500, scale_factor=1.2, nlevels=8, edge_threshold=31, first_level=0, wta_k=2,
               score_type=0, patch_size=31, fast_threshold=20):
    return OrbParams(nfeatures, scale_factor, nlevels, edge_threshold, first_level, wta_k, score_type,
                     patch_size, fast_threshold)


def level_sizes(w, h, scale_factor=1.2, nlevels=8):
    out = []
    for l in range(nlevels):
        a, b = C.c_int(), C.c_int()
        lib().kfo_level_size(w, h, scale_factor, l, C.byref(a), C.byref(b))
        out.append((a.value, b.value))
    return out


def level_scales(scale_factor=1.2, nlevels=8):
    return [lib().kfo_level_scale(scale_factor, l, 0) for l in range(nlevels)]


def level_quotas(nfeatures, scale_factor=1.2, nlevels=8):
    q = (C.c_int * nlevels)()
    lib().kfo_level_quotas(nfeatures, scale_factor, nlevels, q)
    return list(q)


def resize_linear_exact(src, dw, dh):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().kfo_resize_linear_exact(_u8(src), src.shape[1], src.shape[0], src.strides[0], _u8(dst), dw, dh, dst.strides[0])
    return dst


def pyramid(gray, scale_factor=1.2, nlevels=8):
    lv = [np.ascontiguousarray(gray, dtype=np.uint8)]
    for (w, h) in level_sizes(gray.shape[1], gray.shape[0], scale_factor, nlevels)[1:]:
        lv.append(resize_linear_exact(lv[-1], w, h))
    return lv


def fast9_detect(img, thr=20):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    cap = img.size // 4 + 16
    xs = (C.c_int * cap)(); ys = (C.c_int * cap)(); ss = (C.c_int * cap)()
    n = lib().kfo_fast9_detect(_u8(img), img.shape[1], img.shape[0], img.strides[0], thr, xs, ys, ss, cap)
    return np.stack([np.frombuffer(xs, np.int32)[:n], np.frombuffer(ys, np.int32)[:n], np.frombuffer(ss, np.int32)[:n]], 1).copy()


def harris(img, x, y):
    return lib().kfo_harris(_u8(img), img.strides[0], int(x), int(y))


def ic_angle(img, x, y):
    return lib().kfo_ic_angle(_u8(img), img.strides[0], int(x), int(y))


def blur7(img, use_fma=True):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    dst = np.empty_like(img)
    lib().kfo_blur7(_u8(img), img.shape[1], img.shape[0], img.strides[0], _u8(dst), dst.strides[0], int(use_fma))
    return dst


def brief(blurred, cx, cy, angle_deg):
    d = np.zeros(32, np.uint8)
    lib().kfo_brief(_u8(blurred), blurred.strides[0], int(cx), int(cy), float(angle_deg), _u8(d))
    return d


def detect_and_compute(gray, mask=None, params=None, capacity=None, want_pyramid=False, blur_fma=True):
    """-> (keypoints[KP_DTYPE] sorted by (octave, y, x), descriptors uint8[n,32]) (+ pyramid list)."""
    P = params or orb_params()
    gray = np.ascontiguousarray(gray, dtype=np.uint8)
    h, w = gray.shape
    cap = capacity or (2 * P.nfeatures + 4096)
    kp = np.zeros(cap, KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    n = C.c_int()
    if mask is not None:
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        mp, ms = _u8(mask), mask.strides[0]
    else:
        mp, ms = None, 0
    pyr, pyr_ptrs = None, None
    if want_pyramid:
        pyr = [np.zeros((lh, lw), np.uint8) for (lw, lh) in level_sizes(w, h, P.scale_factor, P.nlevels)]
        pyr_ptrs = (C.POINTER(C.c_uint8) * P.nlevels)(*[_u8(p) for p in pyr])
    rc = lib().kfo_orb_detect_and_compute_ex(C.byref(P), _u8(gray), gray.strides[0], w, h, mp, ms,
                                             kp.ctypes.data, _u8(desc), cap, C.byref(n), pyr_ptrs, int(blur_fma))
    if rc < 0:
        raise ValueError("unsupported ORB parameters")
    if rc == 1:
        return detect_and_compute(gray, mask, P, n.value + 16, want_pyramid, blur_fma)
    out = (kp[:n.value].copy(), desc[:n.value].copy())
    return out + (pyr,) if want_pyramid else out


def knn_match(query, train, k=2):
    """-> (matches[DM_DTYPE] shape (nq,k), counts int32[nq])"""
    query = np.ascontiguousarray(query, dtype=np.uint8).reshape(-1, 32) if query.size else np.zeros((0, 32), np.uint8)
    train = np.ascontiguousarray(train, dtype=np.uint8).reshape(-1, 32) if train.size else np.zeros((0, 32), np.uint8)
    nq, nt = query.shape[0], train.shape[0]
    out = np.zeros((nq, k), DM_DTYPE)
    counts = np.zeros(nq, np.int32)
    rc = lib().kfo_knn_match(_u8(query), nq, _u8(train), nt, 32, k, out.ctypes.data, counts.ctypes.data_as(C.POINTER(C.c_int)))
    if rc:
        raise ValueError("bad kNN arguments (k in 1..8, Nt < 2^18)")
    return out, counts


def ratio_test(matches, counts, ratio=0.6):
    nq, k = matches.shape
    good = np.zeros(nq, DM_DTYPE)
    n = lib().kfo_ratio_test(matches.ctypes.data, counts.ctypes.data_as(C.POINTER(C.c_int)), nq, k, ratio, good.ctypes.data)
    return good[:n].copy()


def rgb2gray(rgb):
    """VisualOdometry.cpp:25-28 (RGB2BGR + BGR2GRAY): uint8 [H,W,3] (R,G,B) -> uint8 [H,W]."""
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    h, w, _ = rgb.shape
    out = np.zeros((h, w), np.uint8)
    lib().kfo_rgb2gray(_u8(rgb), rgb.strides[0], w, h, _u8(out), out.strides[0])
    return out


def match_features(knn, kf_kp, cur_kp, depth, K, ratio=0.6):
    """VisualOdometry.cpp:45-59 on kNN rows (nq,2): -> (points3D [n,3] f64, points2D [n,2] f64, matches DM_DTYPE [n])."""
    knn = np.ascontiguousarray(knn, dtype=DM_DTYPE).reshape(-1, 2)
    kf_kp = np.ascontiguousarray(kf_kp, dtype=KP_DTYPE)
    cur_kp = np.ascontiguousarray(cur_kp, dtype=KP_DTYPE)
    depth = np.ascontiguousarray(depth, dtype=np.uint16)
    K = np.asarray(K, dtype=np.float64)
    intr = (C.c_double * 4)(K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    nq = knn.shape[0]
    m = np.zeros(nq, DM_DTYPE)
    p3 = np.zeros((nq, 3), np.float64)
    p2 = np.zeros((nq, 2), np.float64)
    n = lib().kfo_match_features(knn.ctypes.data, nq, kf_kp.ctypes.data, cur_kp.ctypes.data, depth.ctypes.data, depth.strides[0], intr, ratio,
                                 m.ctypes.data, p3.ctypes.data, p2.ctypes.data, nq)
    return p3[:n].copy(), p2[:n].copy(), m[:n].copy()


def match_window(kf_descs, cur_desc, win, ratio=0.6):
    """Map::GetNKeyFrames(win) (Map.cpp:72-78) x knnMatch (:39) x ratio test (:46): kf_descs = descriptors of every keyframe
    of the map in creation order -> (list of k=2 rows per window keyframe, list of survivor arrays), oldest first."""
    kfs = [np.ascontiguousarray(d, dtype=np.uint8).reshape(-1, 32) for d in kf_descs]
    cur = np.ascontiguousarray(cur_desc, dtype=np.uint8).reshape(-1, 32)
    n = len(kfs)
    stride = max([len(d) for d in kfs] + [1])
    ptrs = (C.c_void_p * max(n, 1))(*[d.ctypes.data for d in kfs])
    rows = (C.c_int * max(n, 1))(*[len(d) for d in kfs])
    nw = min(n, win)
    knn = np.zeros((max(nw, 1), stride, 2), DM_DTYPE)
    good = np.zeros((max(nw, 1), stride), DM_DTYPE)
    ngood = (C.c_int * max(nw, 1))()
    scratch = (C.c_int * stride)()
    got = lib().kfo_match_window(ptrs, rows, n, win, _u8(cur), len(cur), ratio, knn.ctypes.data, good.ctypes.data, ngood, scratch, stride)
    if got < 0:
        raise ValueError("bad kNN arguments")
    first = n - got
    return [knn[i, :len(kfs[first + i])].copy() for i in range(got)], [good[i, :ngood[i]].copy() for i in range(got)]

"""Shared helpers for the parity tests: golden loading, input regeneration, keypoint-set comparison."""
import hashlib
import importlib
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
synth = importlib.import_module("keyframe-based-slam_b200.synth")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    out = {k: z[k] for k in z.files if k != "meta"}
    if "meta" in z.files:
        out["meta"] = json.loads(str(z["meta"]))
    return out


_cache = {}


def golden_input(name):
    """Regenerate the (gray, mask) a golden was made from; verified against the stored SHA-256."""
    if name in _cache:
        return _cache[name]
    a, b = synth.make_pair(480, 640, seed=3)
    if name == "cfg1_480p_a":
        g, m = a, None
    elif name == "cfg1_480p_b":
        g, m = b, None
    elif name == "cfg1_480p_masked":
        m = np.zeros((480, 640), np.uint8); m[100:400, 50:600] = 255; m[200:250, 200:300] = 0
        g = a
    elif name == "cfg1_480p_mask_nonbinary":
        g, m = a, (np.random.default_rng(0).random((480, 640)) > 0.3).astype(np.uint8) * 7
    elif name == "ties_periodic":
        g, m = synth.periodic_frame(), None
    elif name == "small_160x120":
        g, m = np.ascontiguousarray(a[:120, :160]), None
    elif name == "small_200x90":
        g, m = np.ascontiguousarray(a[:90, :200]), None
    elif name == "small_64x64":
        g, m = np.ascontiguousarray(a[:64, :64]), None
    elif name == "odd_333x257":
        g, m = np.ascontiguousarray(a[:257, :333]), None
    elif name == "cfg2_720p_nf2000":
        g, m = synth.make_stream(2, 720, 1280, seed=1)[0], None
    elif name == "cfg2_720p_nf3000":
        g, m = synth.make_stream(2, 720, 1280, seed=1)[1], None
    else:
        raise KeyError(name)
    meta = golden(name)["meta"]
    assert sha(g) == meta["frame_sha"], "synthetic generator drifted for %s" % name
    if m is not None:
        assert sha(m) == meta["mask_sha"]
    _cache[name] = (g, m)
    return g, m


ORB_GOLDENS = ["cfg1_480p_a", "cfg1_480p_b", "cfg1_480p_masked", "cfg1_480p_mask_nonbinary", "ties_periodic",
               "small_160x120", "small_200x90", "small_64x64", "odd_333x257", "cfg2_720p_nf2000", "cfg2_720p_nf3000"]


def level_scales(n=8, sf=1.2):
    return [np.float32(np.float64(np.float32(sf)) ** l) for l in range(n)]


def kp_keys(kp, sf=1.2, nlevels=8):
    """(octave, y_l, x_l) integer key per keypoint (compare keypoints as sets on this key)."""
    sc = np.array(level_scales(nlevels, sf), np.float32)
    s = sc[kp["octave"]]
    return list(zip(kp["octave"].tolist(), np.rint(kp["y"] / s).astype(np.int64).tolist(),
                    np.rint(kp["x"] / s).astype(np.int64).tolist()))


def compare_keypoints(ref_kp, ref_desc, kp, desc):
    """-> dict(recall, n_ref, n, exact_fields, max_angle_err_rad, desc_equal, n_common)."""
    ka, kb = kp_keys(ref_kp), kp_keys(kp)
    ia = {k: i for i, k in enumerate(ka)}
    common = [(ia[k], j) for j, k in enumerate(kb) if k in ia]
    ai = np.array([c[0] for c in common], np.int64); bj = np.array([c[1] for c in common], np.int64)
    out = dict(n_ref=len(ka), n=len(kb), n_common=len(common), recall=(len(common) / len(ka)) if len(ka) else 1.0)
    if len(common):
        A, B = ref_kp[ai], kp[bj]
        out["exact_fields"] = int(sum(int(np.array_equal(A[f], B[f])) for f in ("x", "y", "size", "octave", "class_id", "response")))
        d = np.abs(A["angle"].astype(np.float64) - B["angle"].astype(np.float64))
        d = np.minimum(d, 360.0 - d)
        out["max_angle_err_rad"] = float(np.deg2rad(d.max()))
        out["angle_exact"] = int((A["angle"] == B["angle"]).sum())
        out["desc_equal"] = int((ref_desc[ai] == desc[bj]).all(axis=1).sum())
    else:
        out.update(exact_fields=6, max_angle_err_rad=0.0, angle_exact=0, desc_equal=0)
    return out

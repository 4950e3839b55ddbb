"""GPU test of the C++ drop-in (include/kfslam_frontend_cv.hpp): tests/cpp/adapter_test.cpp replays the reference's
own call sequence (VisualOdometry.cpp:10-18, :30, :39, :46) through the adapter; results must equal the cv2 goldens."""
import os
import struct
import subprocess

import numpy as np
import pytest

from _util import ROOT, golden, golden_input

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
DM = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])


def _build():
    exe = os.path.join(ROOT, "tests", "cpp", "adapter_test")
    src = os.path.join(ROOT, "tests", "cpp", "adapter_test.cpp")
    lib = os.path.join(ROOT, "keyframe-based-slam_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "cpp"), "-DKFS_CV_MOCK", src, "-L", lib, "-lkfslam_frontend",
                           "-Wl,-rpath," + lib, "-o", exe])
    return exe


def test_adapter_compiles_and_fails_loudly_without_gpu():
    exe = _build()
    out = subprocess.run([exe, "--abi-only"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr


@pytest.mark.gpu
def test_adapter_reference_call_sequence(tmp_path):
    exe = _build()
    a, _ = golden_input("cfg1_480p_a")
    b, _ = golden_input("cfg1_480p_b")
    fa, fb, fo = tmp_path / "a.raw", tmp_path / "b.raw", tmp_path / "out.bin"
    a.tofile(fa); b.tofile(fb)
    r = subprocess.run([exe, "640", "480", "1000", str(fa), str(fb), str(fo)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    buf = open(fo, "rb").read()
    na, nb, nknn, ngood, threw = struct.unpack_from("<5i", buf, 0)
    off = 20
    kpa = np.frombuffer(buf, KP, na, off); off += na * 28
    da = np.frombuffer(buf, np.uint8, na * 32, off).reshape(na, 32); off += na * 32
    kpb = np.frombuffer(buf, KP, nb, off); off += nb * 28
    db = np.frombuffer(buf, np.uint8, nb * 32, off).reshape(nb, 32); off += nb * 32
    ga, gb, gk = golden("cfg1_480p_a"), golden("cfg1_480p_b"), golden("cfg1_knn")
    assert np.array_equal(kpa, ga["kp"]) and np.array_equal(da, ga["desc"])
    assert np.array_equal(kpb, gb["kp"]) and np.array_equal(db, gb["desc"])
    # 2 exceptions for unsupported options; with the matcher attached to the extractor exactly ONE matrix (dpCurrImg, the
    # frame just extracted) was matched in place on the device, and the rows equal the plain call's
    assert nknn == na and (threw & 15) == 2 and (threw >> 4) & 15 == 1 and threw & 256
    idx = np.zeros((nknn, 2), np.int32); dist = np.zeros((nknn, 2), np.float32)
    for i in range(nknn):
        (cnt,) = struct.unpack_from("<i", buf, off); off += 4
        assert cnt == 2
        row = np.frombuffer(buf, DM, 2, off); off += 32
        assert (row["queryIdx"] == i).all() and (row["imgIdx"] == 0).all()
        idx[i] = row["trainIdx"]; dist[i] = row["distance"]
    assert np.array_equal(idx, gk["idx"]) and np.array_equal(dist, gk["dist"])
    good = np.frombuffer(buf, DM, ngood, off)
    ref = [(i, int(gk["idx"][i, 0])) for i in range(nknn) if gk["dist"][i, 0] < np.float32(0.6) * gk["dist"][i, 1]]
    assert [(int(g["queryIdx"]), int(g["trainIdx"])) for g in good] == ref


@pytest.mark.gpu
def test_adapter_tracker_front_half(tmp_path, oracle):
    """kfs::TrackerFrontEnd (ExtractFeatures / PromoteKeyFrame / MatchFeatures with the reference's signatures) vs the oracle."""
    from _track_util import make_depth, make_rgb
    exe = _build()
    a, b, depth = make_rgb(480, 640, 21), make_rgb(480, 640, 21, shift=3), make_depth(480, 640, 5)
    fa, fb, fd, fo = tmp_path / "a.rgb", tmp_path / "b.rgb", tmp_path / "d.u16", tmp_path / "out.bin"
    a.tofile(fa); b.tofile(fb); depth.tofile(fd)
    r = subprocess.run([exe, "--track", "640", "480", "1000", str(fa), str(fb), str(fd), str(fo), "525.3,524.9,319.7,239.2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    buf = open(fo, "rb").read()
    na, nb, nm, same = struct.unpack_from("<4i", buf, 0)
    off = 16
    kpb = np.frombuffer(buf, KP, nb, off); off += nb * 28
    m = np.frombuffer(buf, DM, nm, off); off += nm * 16
    p3 = np.frombuffer(buf, np.float64, nm * 3, off).reshape(nm, 3); off += nm * 24
    p2 = np.frombuffer(buf, np.float64, nm * 2, off).reshape(nm, 2)
    g = golden("track_cfg1")
    K = np.array([[525.3, 0, 319.7], [0, 524.9, 239.2], [0, 0, 1]])
    r3, r2, rm = oracle.match_features(g["knn"], g["kf_kp"], g["cur_kp"], depth, K, 0.6)
    assert na == len(g["kf_kp"]) and np.array_equal(kpb, g["cur_kp"]) and same == 1
    assert nm == len(rm) > 50 and np.array_equal(m, rm) and np.array_equal(p3, r3) and np.array_equal(p2, r2)

"""CPU suite: the device math header (csrc/kfs_math.cuh) compiled with g++ and checked against the oracle.
Guards the bit-level contracts (FAST packed score + quick reject, Harris, fastAtan2, blur order, rBRIEF
rounding) without a GPU; the kernels themselves are checked by the -m gpu tests."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from _util import ROOT, golden, golden_input

SO = os.path.join(ROOT, "tests", "_cpu_shim.so")


@pytest.fixture(scope="module")
def shim():
    src = os.path.join(ROOT, "tests", "cpu_shim.cpp")
    hdr = os.path.join(ROOT, "keyframe-based-slam_b200", "csrc", "kfs_math.cuh")
    if not os.path.exists(SO) or max(os.path.getmtime(src), os.path.getmtime(hdr)) > os.path.getmtime(SO):
        subprocess.check_call(["g++", "-O2", "-fPIC", "-shared", "-std=c++17", "-ffp-contract=off", "-march=x86-64-v3",
                               "-Wno-unknown-pragmas", "-o", SO, src])
    L = C.CDLL(SO)
    u8 = C.POINTER(C.c_uint8)
    L.shim_scoremap.restype = C.c_long
    L.shim_scoremap.argtypes = [u8, C.c_int, C.c_int, C.c_long, C.c_int, u8, C.c_long, C.c_int]
    L.shim_harris.restype = C.c_float
    L.shim_harris.argtypes = [u8, C.c_long, C.c_int, C.c_int]
    L.shim_atan2.restype = C.c_float
    L.shim_atan2.argtypes = [C.c_float, C.c_float]
    L.shim_ic_angle.restype = C.c_float
    L.shim_ic_angle.argtypes = [u8, C.c_long, C.c_int, C.c_int]
    L.shim_blur_patch.argtypes = [u8, C.c_long, C.c_int, C.c_int, u8]
    L.shim_brief.argtypes = [u8, C.c_float, u8]
    L.shim_float_key.restype = C.c_uint32
    L.shim_float_key.argtypes = [C.c_float]
    return L


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


@pytest.mark.parametrize("name", ["cfg1_480p_a", "ties_periodic"])
def test_fast_score_and_quick_reject(shim, oracle, name):
    gray, _ = golden_input(name)
    rng = np.random.default_rng(5)
    noisy = rng.integers(0, 256, (97, 131), dtype=np.uint8)  # saturating corners: v +- t crosses 0 / 255
    for img in (gray, oracle.pyramid(gray)[3], noisy):
        img = np.ascontiguousarray(img)
        h, w = img.shape
        ref = np.zeros_like(img)
        oracle.lib().kfo_fast9_scoremap(_u8(img), w, h, img.strides[0], 20, _u8(ref), ref.strides[0])
        for use_quick in (0, 1):
            out = np.zeros_like(img)
            rej = shim.shim_scoremap(_u8(img), w, h, img.strides[0], 20, _u8(out), out.strides[0], use_quick)
            assert rej >= 0
            assert np.array_equal(out, ref)
        if name == "cfg1_480p_a" and img is gray:
            assert rej > 0.8 * img.size  # on natural texture the quick test rejects the bulk


def test_fast_threshold_extremes(shim, oracle):
    img = np.random.default_rng(1).integers(0, 256, (64, 80), dtype=np.uint8)
    for thr in (1, 100, 254):
        ref = np.zeros_like(img); out = np.zeros_like(img)
        oracle.lib().kfo_fast9_scoremap(_u8(img), 80, 64, 80, thr, _u8(ref), 80)
        assert shim.shim_scoremap(_u8(img), 80, 64, 80, thr, _u8(out), 80, 1) >= 0
        assert np.array_equal(out, ref)


def test_harris_angle_blur_brief(shim, oracle):
    gray, _ = golden_input("cfg1_480p_a")
    g = golden("cfg1_480p_a")
    pyr = oracle.pyramid(gray)
    blurred = [oracle.blur7(p) for p in pyr]
    sc = oracle.level_scales()
    rng = np.random.default_rng(0)
    idx = rng.choice(len(g["kp"]), 300, replace=False)
    for i in idx:
        k = g["kp"][i]
        l = int(k["octave"])
        x, y = int(round(float(k["x"]) / sc[l])), int(round(float(k["y"]) / sc[l]))
        img = pyr[l]
        assert shim.shim_harris(_u8(img), img.strides[0], x, y) == k["response"]
        assert shim.shim_ic_angle(_u8(img), img.strides[0], x, y) == k["angle"]
        patch = np.zeros((37, 37), np.uint8)
        shim.shim_blur_patch(_u8(img), img.strides[0], x, y, _u8(patch))
        assert np.array_equal(patch, blurred[l][y - 18:y + 19, x - 18:x + 19])
        d = np.zeros(32, np.uint8)
        shim.shim_brief(_u8(patch), float(k["angle"]), _u8(d))
        assert np.array_equal(d, g["desc"][i])


def test_atan2_matches_oracle(shim, oracle):
    rng = np.random.default_rng(3)
    ys = rng.integers(-200000, 200000, 20000).astype(np.float32)
    xs = rng.integers(-200000, 200000, 20000).astype(np.float32)
    xs[:50] = 0; ys[50:100] = 0; xs[100] = ys[100] = 0
    for y, x in zip(ys.tolist(), xs.tolist()):
        assert shim.shim_atan2(y, x) == oracle.lib().kfo_fast_atan2(y, x)


def test_float_key_is_order_preserving(shim):
    v = np.array([-np.inf, -1e30, -1.5, -1e-30, -0.0, 0.0, 1e-38, 1e-7, 3.25, 1e30, np.inf], np.float32)
    keys = [shim.shim_float_key(float(x)) for x in v]
    assert all(a <= b for a, b in zip(keys, keys[1:]))
    assert keys[3] < keys[5] and keys[5] < keys[6]

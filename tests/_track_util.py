"""Seeded inputs of the tracker-wrapper tests: colour frames (R,G,B byte order) and uint16 depth images."""
import importlib

import numpy as np

synth = importlib.import_module("keyframe-based-slam_b200.synth")


def make_rgb(h, w, seed, shift=0):
    """Three differently textured channels so that a swapped or mis-weighted channel changes the gray image."""
    a, b = synth.make_pair(h, w, seed=seed)
    rng = np.random.default_rng(seed + 1000)
    r = np.roll(a, shift, axis=1)
    g = np.roll(b, shift, axis=1)
    bl = (np.roll(a, shift + 5, axis=0).astype(np.int32) // 2 + rng.integers(0, 128, (h, w))).astype(np.uint8)
    return np.ascontiguousarray(np.stack([r, g, bl], axis=2))


def make_depth(h, w, seed):
    """Smooth ramp + noise in millimetres, with zero holes like a real depth camera."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    d = 800 + 3 * xx + 2 * yy + rng.integers(0, 50, (h, w))
    d[rng.random((h, w)) < 0.02] = 0
    return d.astype(np.uint16)


def make_refset_mask(h, w, seed):
    """A mask like the one the reference's camera feed delivers with every frame (main.cpp:31-34): a valid region with holes
    and a ragged border; values 0 / 255."""
    rng = np.random.default_rng(seed)
    m = np.full((h, w), 255, np.uint8)
    m[:, :40] = 0
    m[:25, :] = 0
    for _ in range(30):
        x0, y0 = int(rng.integers(0, w - 80)), int(rng.integers(0, h - 60))
        m[y0:y0 + int(rng.integers(8, 60)), x0:x0 + int(rng.integers(8, 80))] = 0
    yy, xx = np.mgrid[0:h, 0:w]
    m[(yy - h * 0.6) ** 2 + (xx - w * 0.7) ** 2 < (h * 0.12) ** 2] = 0
    return m

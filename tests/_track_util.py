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
